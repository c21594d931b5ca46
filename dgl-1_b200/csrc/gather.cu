// Row gather: out[i, :] = x[idx[i], :].  READ_ROW of the reference's executors
// (python/dgl/runtime/ir/executor.py:365-394 -> F.gather_row, backend/pytorch/tensor.py:108-109)
// and the send-side packing of the compacted halo exchange (dgl_b200/dist.py).  HBM-bound:
// n_out * row_bytes read + written; 128-bit accesses, one group of lanes per row.
#include <dlfcn.h>

#include <cstring>

#include "common.cuh"

namespace dglb200 {

template <int BYTES_PER_LANE>
__global__ void __launch_bounds__(256)
gather_rows_kernel(int64_t n_out, const int64_t* __restrict__ idx, int64_t x_rows,
                   const char* __restrict__ x, int64_t row_bytes, char* __restrict__ out,
                   int lanes_per_row, int32_t* __restrict__ bad) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t row = tid / lanes_per_row;
  const int lane = (int)(tid % lanes_per_row);
  if (row >= n_out) return;
  const int64_t s = idx[row];
  if (s < 0 || s >= x_rows) {
    if (lane == 0) atomicExch(bad, 1);
    return;
  }
  const char* src = x + s * row_bytes;
  char* dst = out + row * row_bytes;
  for (int64_t b = (int64_t)lane * BYTES_PER_LANE; b < row_bytes; b += (int64_t)lanes_per_row * BYTES_PER_LANE) {
    if (BYTES_PER_LANE == 16) {
      *reinterpret_cast<uint4*>(dst + b) = __ldg(reinterpret_cast<const uint4*>(src + b));
    } else if (BYTES_PER_LANE == 4) {
      *reinterpret_cast<uint32_t*>(dst + b) = __ldg(reinterpret_cast<const uint32_t*>(src + b));
    } else {
      dst[b] = src[b];
    }
  }
}

}  // namespace dglb200

using namespace dglb200;

extern "C" int dglb200_gather_rows(int64_t n_out, const int64_t* idx, int64_t x_rows, const void* x,
                                   int64_t row_bytes, void* out, int32_t* bad_flag, void* stream_) {
  clear_error();
  DGLB200_REQUIRE(n_out >= 0 && x_rows >= 0 && row_bytes >= 1, "gather_rows: bad sizes");
  if (n_out == 0) return 0;
  DGLB200_REQUIRE(idx && x && out, "gather_rows: NULL pointer");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  auto al = [](const void* p, int a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; };
  const int unit = (row_bytes % 16 == 0 && al(x, 16) && al(out, 16)) ? 16
                   : (row_bytes % 4 == 0 && al(x, 4) && al(out, 4)) ? 4 : 1;
  int lanes = 1;
  while (lanes < 32 && (int64_t)lanes * unit < row_bytes) lanes <<= 1;
  const int64_t threads = n_out * lanes;
  const int64_t blocks = ceil_div(threads, 256);
  DGLB200_REQUIRE(blocks < (1ll << 31), "gather_rows: grid too large");
  const char* xs = static_cast<const char*>(x);
  char* os = static_cast<char*>(out);
  if (unit == 16)
    gather_rows_kernel<16><<<(unsigned)blocks, 256, 0, stream>>>(n_out, idx, x_rows, xs, row_bytes, os, lanes, bad_flag);
  else if (unit == 4)
    gather_rows_kernel<4><<<(unsigned)blocks, 256, 0, stream>>>(n_out, idx, x_rows, xs, row_bytes, os, lanes, bad_flag);
  else
    gather_rows_kernel<1><<<(unsigned)blocks, 256, 0, stream>>>(n_out, idx, x_rows, xs, row_bytes, os, lanes, bad_flag);
  DGLB200_POST_LAUNCH("gather_rows_kernel");
  return 0;
}

// ---- CUDA IPC: map a peer process's device allocation into this process, on the CURRENT device ----
// (cudaIpcMemLazyEnablePeerAccess enables peer access from the current device to the exporter's
// device).  Used by dgl_b200/dist.py for the NVLink peer-load exchange.
extern "C" int dglb200_ipc_get_handle(const void* dev_ptr, void* handle64_out, int64_t* offset_out) {
  clear_error();
  DGLB200_REQUIRE(dev_ptr && handle64_out && offset_out, "ipc_get_handle: NULL pointer");
  // base of the cudaMalloc allocation that contains dev_ptr (torch sub-allocates its blocks):
  // cuMemGetAddressRange of the driver API, resolved at run time so that the library has no
  // link-time dependency on libcuda
  typedef int (*GetRange)(unsigned long long*, size_t*, unsigned long long);
  static GetRange get_range = nullptr;
  if (get_range == nullptr) {
    void* drv = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
    DGLB200_REQUIRE(drv != nullptr, "ipc_get_handle: cannot load libcuda.so.1");
    get_range = reinterpret_cast<GetRange>(dlsym(drv, "cuMemGetAddressRange_v2"));
    DGLB200_REQUIRE(get_range != nullptr, "ipc_get_handle: libcuda lacks cuMemGetAddressRange_v2");
  }
  unsigned long long base = 0;
  size_t size = 0;
  const int rc = get_range(&base, &size, (unsigned long long)(uintptr_t)dev_ptr);
  DGLB200_REQUIRE(rc == 0, "ipc_get_handle: cuMemGetAddressRange failed (%d)", rc);
  cudaIpcMemHandle_t h;
  DGLB200_CUDA(cudaIpcGetMemHandle(&h, reinterpret_cast<void*>((uintptr_t)base)));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64_out, &h, 64);
  *offset_out = (int64_t)((unsigned long long)(uintptr_t)dev_ptr - base);
  return 0;
}

extern "C" int dglb200_ipc_open(const void* handle64, void** ptr_out) {
  clear_error();
  DGLB200_REQUIRE(handle64 && ptr_out, "ipc_open: NULL pointer");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  DGLB200_CUDA(cudaIpcOpenMemHandle(ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

extern "C" int dglb200_ipc_close(void* ptr) {
  clear_error();
  DGLB200_CUDA(cudaIpcCloseMemHandle(ptr));
  return 0;
}
