// Hook into the reference's PackedFunc registry using ONLY libdgl.so's C ABI
// (include/dgl/runtime/c_runtime_api.h:255-402 of the reference): the symbols are looked up
// with dlsym on the handle the caller passes, so libdgl_b200.so has no link-time dependency on
// libdgl.so and no dependency on its C++ std::function layout.
//
// Registered globals:
//   b200._CAPI_DGLB200Version()                      -> int
//   b200._CAPI_DGLB200LaunchCount()                  -> int
//   b200._CAPI_DGLB200CsrBuild(n_rows, n_cols, nnz, row, col, eid, order, indptr, indices, eids,
//                              workspace, workspace_bytes, stream)      (device pointers as ints)
//   b200._CAPI_DGLB200SpMMCsr(n_rows, x_rows, nnz, indptr, indices, eids, msg_op, reduce_op,
//                             dtype, x, feat, w, w_width, y, arg, stream)
//   b200._CAPI_DGLB200DegreeBucketing(msg_ids, vids, recv_ids)  -> PackedFunc(i) -> NDArray
//   runtime.degree_bucketing._CAPI_DGLDegreeBucketing            (override, same signature as
//        src/scheduler/scheduler_apis.cc:16-23: three CPU int64 NDArrays in, indexed getter out)
#include <dlfcn.h>

#include <algorithm>
#include <cstring>
#include <string>
#include <vector>

#include "common.cuh"

#include "packed_common.cuh"

namespace dglb200 {
namespace packed {

LibDgl g_dgl;

int fail(const char* where) {
  // forward our thread-local message into libdgl's (DGLAPISetLastError, c_runtime_api.h:157)
  std::string msg = std::string(where) + ": " + dglb200_last_error();
  if (g_dgl.APISetLastError) g_dgl.APISetLastError(msg.c_str());
  return -1;
}

int64_t as_int(const RefValue& v, int code) {
  return (code == kRefHandle || code == kRefArrayHandle) ? (int64_t)(intptr_t)v.v_handle
         : code == kRefFloat ? (int64_t)v.v_float64 : v.v_int64;
}
void* as_ptr(const RefValue& v, int code) {
  if (code == kRefNull) return nullptr;
  return (code == kRefInt) ? (void*)(intptr_t)v.v_int64 : v.v_handle;
}
int ret_int(void* ret, int64_t value) {
  RefValue v;
  v.v_int64 = value;
  int code = kRefInt;
  return g_dgl.CFuncSetReturn(ret, &v, &code, 1);
}

int register_one(const char* name, PackedCFunc fn, int override_existing) {
  void* h = nullptr;
  if (g_dgl.FuncCreateFromCFunc(fn, nullptr, nullptr, &h) != 0) {
    set_error("DGLFuncCreateFromCFunc failed for %s", name);
    return -1;
  }
  if (g_dgl.FuncRegisterGlobal(name, h, override_existing) != 0) {
    set_error("DGLFuncRegisterGlobal failed for %s", name);
    g_dgl.FuncFree(h);
    return -1;
  }
  g_dgl.FuncFree(h);  // the registry holds its own copy (src/runtime/registry.cc:129-135)
  return 0;
}

}  // namespace packed

using namespace packed;
namespace {

int pf_version(RefValue*, int*, int, void* ret, void*) { return ret_int(ret, dglb200_abi_version()); }
int pf_launch_count(RefValue*, int*, int, void* ret, void*) { return ret_int(ret, dglb200_launch_count()); }

int pf_csr_build(RefValue* a, int* c, int n, void* ret, void*) {
  if (n != 13) { set_error("expects 13 arguments, got %d", n); return fail("b200._CAPI_DGLB200CsrBuild"); }
  int rc = dglb200_csr_build(as_int(a[0], c[0]), as_int(a[1], c[1]), as_int(a[2], c[2]),
                             (const int64_t*)as_ptr(a[3], c[3]), (const int64_t*)as_ptr(a[4], c[4]),
                             (const int64_t*)as_ptr(a[5], c[5]), (int)as_int(a[6], c[6]),
                             (int64_t*)as_ptr(a[7], c[7]), (int32_t*)as_ptr(a[8], c[8]),
                             (int32_t*)as_ptr(a[9], c[9]), as_ptr(a[10], c[10]),
                             (size_t)as_int(a[11], c[11]), as_ptr(a[12], c[12]));
  if (rc != 0) return fail("b200._CAPI_DGLB200CsrBuild");
  return ret_int(ret, 0);
}

int pf_spmm(RefValue* a, int* c, int n, void* ret, void*) {
  if (n != 16) { set_error("expects 16 arguments, got %d", n); return fail("b200._CAPI_DGLB200SpMMCsr"); }
  int rc = dglb200_spmm_csr(as_int(a[0], c[0]), as_int(a[1], c[1]), as_int(a[2], c[2]),
                            (const int64_t*)as_ptr(a[3], c[3]), (const int32_t*)as_ptr(a[4], c[4]),
                            (const int32_t*)as_ptr(a[5], c[5]), (int)as_int(a[6], c[6]),
                            (int)as_int(a[7], c[7]), (int)as_int(a[8], c[8]), as_ptr(a[9], c[9]),
                            as_int(a[10], c[10]), as_ptr(a[11], c[11]), as_int(a[12], c[12]),
                            as_ptr(a[13], c[13]), (int32_t*)as_ptr(a[14], c[14]), nullptr,
                            as_ptr(a[15], c[15]));
  if (rc != 0) return fail("b200._CAPI_DGLB200SpMMCsr");
  return ret_int(ret, 0);
}

// ---- degree bucketing with the reference's calling convention --------------------------------

struct BucketResult {
  RefDLTensor* arr[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
};
void bucket_result_free(void* p) {
  BucketResult* r = static_cast<BucketResult*>(p);
  for (auto* t : r->arr)
    if (t) g_dgl.ArrayFree(t);
  delete r;
}
// the indexed getter the reference returns (ConvertNDArrayVectorToPackedFunc, c_api_common.cc:25-35)
int pf_bucket_get(RefValue* a, int* c, int n, void* ret, void* resource) {
  BucketResult* r = static_cast<BucketResult*>(resource);
  const int64_t which = n >= 1 ? as_int(a[0], c[0]) : -1;
  if (which < 0 || which >= 5) { set_error("invalid choice %lld", (long long)which); return fail("degree_bucketing getter"); }
  RefValue v;
  v.v_handle = r->arr[which];
  int code = kRefNDArrayContainer;  // RetValue takes its own reference (packed_func.h:748-751)
  return g_dgl.CFuncSetReturn(ret, &v, &code, 1);
}

bool valid_id_array(const RefDLTensor* t) {  // IsValidIdArray, src/c_api_common.h:35-38
  return t && t->device_type == 1 && t->ndim == 1 && t->code == 0 && t->bits == 64;
}

struct DevBuf {
  void* p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
  int alloc(size_t n) { return cudaMalloc(&p, n ? n : 1) == cudaSuccess ? 0 : -1; }
};

int pf_degree_bucketing(RefValue* a, int* c, int n, void* ret, void*) {
  const char* who = "runtime.degree_bucketing._CAPI_DGLDegreeBucketing[b200]";
  if (n != 3) { set_error("expects 3 arguments, got %d", n); return fail(who); }
  const RefDLTensor* t[3];
  for (int i = 0; i < 3; ++i) {
    t[i] = static_cast<const RefDLTensor*>(a[i].v_handle);
    if (!(c[i] == kRefArrayHandle || c[i] == kRefNDArrayContainer) || !valid_id_array(t[i])) {
      set_error("argument %d is not a CPU int64 1-D array", i);
      return fail(who);
    }
  }
  const int64_t M = t[0]->shape[0], R = t[2]->shape[0];
  if (t[1]->shape[0] != M) { set_error("msg_ids and vids differ in length"); return fail(who); }
  const int64_t* h_mid = (const int64_t*)((char*)t[0]->data + t[0]->byte_offset);
  const int64_t* h_vid = (const int64_t*)((char*)t[1]->data + t[1]->byte_offset);
  const int64_t* h_rcv = (const int64_t*)((char*)t[2]->data + t[2]->byte_offset);
  int64_t bound = 1;
  for (int64_t i = 0; i < M; ++i) bound = std::max(bound, h_vid[i] + 1);
  for (int64_t i = 0; i < R; ++i) bound = std::max(bound, h_rcv[i] + 1);
  const int64_t cap = dglb200_bucket_capacity(M);
  size_t ws_bytes = 0;
  if (dglb200_degree_bucketing_workspace_bytes(M, R, &ws_bytes) != 0) return fail(who);
  DevBuf d_mid, d_vid, d_rcv, d_degs, d_nids, d_nsec, d_mids, d_msec, d_ws;
  if (d_mid.alloc(M * 8) || d_vid.alloc(M * 8) || d_rcv.alloc(R * 8) || d_degs.alloc(cap * 8) ||
      d_nids.alloc((M + R) * 8) || d_nsec.alloc(cap * 8) || d_mids.alloc(M * 8) ||
      d_msec.alloc(cap * 8) || d_ws.alloc(ws_bytes)) {
    set_error("cudaMalloc failed: %s", cudaGetErrorString(cudaGetLastError()));
    return fail(who);
  }
  cudaMemcpy(d_mid.p, h_mid, M * 8, cudaMemcpyHostToDevice);
  cudaMemcpy(d_vid.p, h_vid, M * 8, cudaMemcpyHostToDevice);
  cudaMemcpy(d_rcv.p, h_rcv, R * 8, cudaMemcpyHostToDevice);
  int64_t counts[3] = {0, 0, 0};
  if (dglb200_degree_bucketing(M, R, bound, (const int64_t*)d_mid.p, (const int64_t*)d_vid.p,
                               (const int64_t*)d_rcv.p, (int64_t*)d_degs.p, (int64_t*)d_nids.p,
                               (int64_t*)d_nsec.p, (int64_t*)d_mids.p, (int64_t*)d_msec.p, counts,
                               d_ws.p, ws_bytes, nullptr) != 0)
    return fail(who);
  BucketResult* res = new BucketResult();
  const int64_t sizes[5] = {counts[0], counts[1], counts[0], M, counts[2]};
  void* src[5] = {d_degs.p, d_nids.p, d_nsec.p, d_mids.p, d_msec.p};
  for (int i = 0; i < 5; ++i) {
    // int64 CPU arrays, like IdArray::Empty(..., vids->dtype, vids->ctx) in scheduler.cc:49-53
    if (g_dgl.ArrayAlloc(&sizes[i], 1, 0, 64, 1, 1, 0, &res->arr[i]) != 0) { bucket_result_free(res); return -1; }
    if (sizes[i] > 0 &&
        cudaMemcpy(res->arr[i]->data, src[i], sizes[i] * 8, cudaMemcpyDeviceToHost) != cudaSuccess) {
      set_error("cudaMemcpy D2H failed: %s", cudaGetErrorString(cudaGetLastError()));
      bucket_result_free(res);
      return fail(who);
    }
  }
  void* getter = nullptr;
  if (g_dgl.FuncCreateFromCFunc(pf_bucket_get, res, bucket_result_free, &getter) != 0) return -1;
  RefValue v;
  v.v_handle = getter;
  int code = kRefFuncHandle;  // RetValue copies the PackedFunc (packed_func.h:740-743)
  const int rc = g_dgl.CFuncSetReturn(ret, &v, &code, 1);
  g_dgl.FuncFree(getter);
  return rc;
}

}  // namespace
}  // namespace dglb200

using namespace dglb200;

extern "C" int dglb200_register_packed_funcs(void* libdgl_handle, int override_reference) {
  clear_error();
  DGLB200_REQUIRE(libdgl_handle != nullptr, "register_packed_funcs: libdgl handle is NULL");
#define DGLB200_SYM(field, name)                                                       \
  *reinterpret_cast<void**>(&g_dgl.field) = dlsym(libdgl_handle, name);               \
  DGLB200_REQUIRE(g_dgl.field != nullptr, "register_packed_funcs: libdgl.so lacks %s", name)
  DGLB200_SYM(FuncCreateFromCFunc, "DGLFuncCreateFromCFunc");
  DGLB200_SYM(FuncRegisterGlobal, "DGLFuncRegisterGlobal");
  DGLB200_SYM(FuncFree, "DGLFuncFree");
  DGLB200_SYM(CFuncSetReturn, "DGLCFuncSetReturn");
  DGLB200_SYM(APISetLastError, "DGLAPISetLastError");
  DGLB200_SYM(ArrayAlloc, "DGLArrayAlloc");
  DGLB200_SYM(ArrayFree, "DGLArrayFree");
  DGLB200_SYM(FuncGetGlobal, "DGLFuncGetGlobal");
  DGLB200_SYM(FuncCall, "DGLFuncCall");
#undef DGLB200_SYM
  int n = 0;
  if (register_one("b200._CAPI_DGLB200Version", pf_version, 1) != 0) return -1; ++n;
  if (register_one("b200._CAPI_DGLB200LaunchCount", pf_launch_count, 1) != 0) return -1; ++n;
  if (register_one("b200._CAPI_DGLB200CsrBuild", pf_csr_build, 1) != 0) return -1; ++n;
  if (register_one("b200._CAPI_DGLB200SpMMCsr", pf_spmm, 1) != 0) return -1; ++n;
  if (register_one("b200._CAPI_DGLB200DegreeBucketing", pf_degree_bucketing, 1) != 0) return -1; ++n;
  if (override_reference) {
    if (register_one("runtime.degree_bucketing._CAPI_DGLDegreeBucketing", pf_degree_bucketing, 1) != 0) return -1;
    ++n;
    // device-resident graph index behind the reference's own names (packed_graph.cu)
    const int g = register_graph_overrides();
    if (g < 0) return -1;
    n += g;
  }
  return n;
}
