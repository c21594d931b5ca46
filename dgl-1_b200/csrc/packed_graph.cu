// Device-resident graph index behind the reference's OWN PackedFunc names:
//   graph_index._CAPI_DGLGraphGetAdj      (src/graph/graph_apis.cc:474-482; return shape :39-49)
//   graph_index._CAPI_DGLGraphInEdges_2   (:284)      graph_index._CAPI_DGLGraphOutEdges_2  (:300)
//   graph_index._CAPI_DGLGraphEdges       (:308)
//   graph_index._CAPI_DGLGraphInDegrees   (:324)      graph_index._CAPI_DGLGraphOutDegrees  (:340)
//   graph_index._CAPI_DGLGraphFindEdges   (:268)      graph_index._CAPI_DGLGraphEdgeIds     (:259)
//   graph_index._CAPI_DGLGraphInEdges_1 / OutEdges_1 (:276,292)   InDegree / OutDegree (:316,332)
// Same arity, same argument types (graph handle, CPU int64 id arrays, strings), same return order
// (indexed PackedFunc getters over CPU int64 NDArrays), so python/dgl/graph_index.py needs no change.
// The answers come from CSRs that live on the GPU: the first query on a graph handle fetches its edge
// list ONCE through the reference's original _CAPI_DGLGraphEdges, builds the in-/out-edge CSRs with
// kernel (4) and keeps them keyed by the handle; the mutators (AddVertices / AddEdge / AddEdges / Clear /
// Free) are wrapped so that they drop the cached index before calling the original.
// READ-ONLY handles (ImmutableGraph: other slot orders - immutable_graph.cc:206-237, 458-493) are passed
// through to the reference's originals by these six; what runs on the GPU for them is the neighbour sampler
// (packed_sampling.cuh: _CAPI_DGLGraphUniformSampling*).
#include <mutex>
#include <unordered_map>
#include <vector>

#include "packed_common.cuh"

namespace dglb200 {
namespace packed {
namespace {

struct DevMem {
  void* p = nullptr;
  DevMem() = default;
  DevMem(const DevMem&) = delete;
  DevMem& operator=(const DevMem&) = delete;
  ~DevMem() { reset(); }
  void reset() { if (p) cudaFree(p); p = nullptr; }
  int alloc(size_t bytes) {
    reset();
    return cudaMalloc(&p, bytes ? bytes : 8) == cudaSuccess ? 0 : -1;
  }
  template <typename T> T* as() const { return static_cast<T*>(p); }
};

struct DevCsr {
  DevMem ptr, idx, eid;   // int64[n+1], int32[e], int32[e]
};

struct DevIndex {
  int64_t n = -1, e = -1;
  DevMem src, dst;        // int64[e], edge-id order
  DevCsr in, out;         // rows = dst / rows = src, slots in ascending edge id (Graph::GetAdj order)
  DevCsr by_src_dst;      // rows = src, slots sorted by dst then edge id: Edges("srcdst"); built on demand
  bool has_sorted = false;
};

std::mutex g_mu;
std::unordered_map<void*, DevIndex*> g_index;
struct Originals {
  void *edges = nullptr, *num_vertices = nullptr, *num_edges = nullptr;
} g_orig;

int get_global(const char* name, void** out) {
  *out = nullptr;
  if (g_dgl.FuncGetGlobal(name, out) != 0 || *out == nullptr) {
    set_error("global %s is not registered in libdgl", name);
    return -1;
  }
  return 0;
}

// call a PackedFunc of libdgl with (handle[, one more value]); the caller owns a returned func / array
int call1(void* fn, void* handle, const RefValue* extra, int extra_code, RefValue* ret, int* rcode) {
  RefValue args[2];
  int codes[2];
  args[0].v_handle = handle;
  codes[0] = kRefHandle;
  int n = 1;
  if (extra != nullptr) {
    args[1] = *extra;
    codes[1] = extra_code;
    n = 2;
  }
  return g_dgl.FuncCall(fn, args, codes, n, ret, rcode);
}

int64_t call_count(void* fn, void* handle) {
  RefValue ret;
  int rcode = kRefNull;
  if (call1(fn, handle, nullptr, 0, &ret, &rcode) != 0 || rcode != kRefInt) return -1;
  return ret.v_int64;
}

int build_csr(const int64_t* row, const int64_t* col, int64_t n, int64_t e, int order, DevCsr* out) {
  size_t ws_bytes = 0;
  if (dglb200_csr_build_workspace_bytes(n, e, &ws_bytes) != 0) return -1;
  DevMem ws;
  if (out->ptr.alloc((size_t)(n + 1) * 8) || out->idx.alloc((size_t)e * 4) || out->eid.alloc((size_t)e * 4) ||
      ws.alloc(ws_bytes)) {
    set_error("cudaMalloc failed: %s", cudaGetErrorString(cudaGetLastError()));
    return -1;
  }
  return dglb200_csr_build(n, n, e, row, col, nullptr, order, out->ptr.as<int64_t>(), out->idx.as<int32_t>(),
                           out->eid.as<int32_t>(), ws.p, ws_bytes, nullptr);
}

// The cached index of a graph handle, (re)built when the graph's vertex / edge counts changed.
DevIndex* get_index(void* handle) {
  const int64_t n = call_count(g_orig.num_vertices, handle);
  const int64_t e = call_count(g_orig.num_edges, handle);
  if (n < 0 || e < 0) {
    set_error("cannot query the graph handle (NumVertices / NumEdges failed)");
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(g_mu);
  auto it = g_index.find(handle);
  if (it != g_index.end() && it->second->n == n && it->second->e == e) return it->second;
  if (it != g_index.end()) {
    delete it->second;
    g_index.erase(it);
  }
  // edge list in edge-id order from the reference's own graph object (one-off per graph state)
  RefValue order;
  order.v_str = "";
  RefValue ret;
  int rcode = kRefNull;
  if (call1(g_orig.edges, handle, &order, kRefStr, &ret, &rcode) != 0 || rcode != kRefFuncHandle) {
    set_error("the reference's _CAPI_DGLGraphEdges failed");
    return nullptr;
  }
  void* getter = ret.v_handle;
  RefDLTensor* arr[2] = {nullptr, nullptr};
  bool ok = true;
  for (int i = 0; i < 2 && ok; ++i) {
    RefValue a, r;
    int c = kRefInt, rc = kRefNull;
    a.v_int64 = i;
    ok = g_dgl.FuncCall(getter, &a, &c, 1, &r, &rc) == 0 && rc == kRefNDArrayContainer;
    if (ok) arr[i] = static_cast<RefDLTensor*>(r.v_handle);
  }
  g_dgl.FuncFree(getter);
  DevIndex* ix = new DevIndex();
  ix->n = n;
  ix->e = e;
  if (ok) ok = ix->src.alloc((size_t)e * 8) == 0 && ix->dst.alloc((size_t)e * 8) == 0;
  if (ok && e > 0) {
    ok = cudaMemcpy(ix->src.p, (char*)arr[0]->data + arr[0]->byte_offset, (size_t)e * 8, cudaMemcpyHostToDevice) == cudaSuccess &&
         cudaMemcpy(ix->dst.p, (char*)arr[1]->data + arr[1]->byte_offset, (size_t)e * 8, cudaMemcpyHostToDevice) == cudaSuccess;
  }
  for (auto* t : arr)
    if (t) g_dgl.ArrayFree(t);
  if (ok) ok = build_csr(ix->dst.as<int64_t>(), ix->src.as<int64_t>(), n, e, 0, &ix->in) == 0 &&
               build_csr(ix->src.as<int64_t>(), ix->dst.as<int64_t>(), n, e, 0, &ix->out) == 0;
  if (!ok) {
    if (dglb200_last_error()[0] == '\0') set_error("building the device graph index failed");
    delete ix;
    return nullptr;
  }
  g_index[handle] = ix;
  return ix;
}

void drop_index(void* handle) {
  std::lock_guard<std::mutex> lock(g_mu);
  auto it = g_index.find(handle);
  if (it != g_index.end()) {
    delete it->second;
    g_index.erase(it);
  }
}

// ---- results: CPU int64 NDArrays behind an indexed getter (ConvertEdgeArrayToPackedFunc /
// ConvertAdjToPackedFunc, graph_apis.cc:21-49) -------------------------------------------------------

struct ArrayList {
  std::vector<RefDLTensor*> arr;
};
void array_list_free(void* p) {
  ArrayList* r = static_cast<ArrayList*>(p);
  for (auto* t : r->arr)
    if (t) g_dgl.ArrayFree(t);
  delete r;
}
int pf_list_get(RefValue* a, int* c, int n, void* ret, void* resource) {
  ArrayList* r = static_cast<ArrayList*>(resource);
  const int64_t which = n >= 1 ? as_int(a[0], c[0]) : -1;
  if (which < 0 || which >= (int64_t)r->arr.size()) {
    set_error("invalid choice %lld", (long long)which);
    return fail("graph index getter");
  }
  RefValue v;
  v.v_handle = r->arr[which];
  int code = kRefNDArrayContainer;
  return g_dgl.CFuncSetReturn(ret, &v, &code, 1);
}
RefDLTensor* cpu_ids(int64_t len) {
  RefDLTensor* t = nullptr;
  if (g_dgl.ArrayAlloc(&len, 1, 0, 64, 1, 1, 0, &t) != 0) return nullptr;
  return t;
}
// new CPU int64 array filled from `len` int64 on the device
RefDLTensor* from_device(const int64_t* dev, int64_t len) {
  RefDLTensor* t = cpu_ids(len);
  if (t && len > 0 && cudaMemcpy(t->data, dev, (size_t)len * 8, cudaMemcpyDeviceToHost) != cudaSuccess) {
    set_error("cudaMemcpy D2H failed: %s", cudaGetErrorString(cudaGetLastError()));
    g_dgl.ArrayFree(t);
    return nullptr;
  }
  return t;
}
RefDLTensor* iota(int64_t len) {
  RefDLTensor* t = cpu_ids(len);
  if (t)
    for (int64_t i = 0; i < len; ++i) static_cast<int64_t*>(t->data)[i] = i;
  return t;
}
int return_list(ArrayList* res, void* ret) {
  for (auto* t : res->arr)
    if (t == nullptr) {
      array_list_free(res);
      return -1;
    }
  void* getter = nullptr;
  if (g_dgl.FuncCreateFromCFunc(pf_list_get, res, array_list_free, &getter) != 0) return -1;
  RefValue v;
  v.v_handle = getter;
  int code = kRefFuncHandle;
  const int rc = g_dgl.CFuncSetReturn(ret, &v, &code, 1);
  g_dgl.FuncFree(getter);
  return rc;
}

bool valid_ids(const RefDLTensor* t) {  // IsValidIdArray, src/c_api_common.h:35-38
  return t && t->device_type == 1 && t->ndim == 1 && t->code == 0 && t->bits == 64;
}

}  // namespace
}  // namespace packed
}  // namespace dglb200

#include "packed_sampling.cuh"
#include "packed_graph_funcs.cuh"
