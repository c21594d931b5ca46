// Row gather: out[i, :] = x[idx[i], :].  READ_ROW of the reference's executors
// (python/dgl/runtime/ir/executor.py:365-394 -> F.gather_row, backend/pytorch/tensor.py:108-109)
// and the send-side packing of the compacted halo exchange (dgl_b200/dist.py).  HBM-bound:
// n_out * row_bytes read + written; 128-bit accesses, one group of lanes per row.
#include <dlfcn.h>

#include <cstring>
#include <type_traits>

#include "common.cuh"

namespace dglb200 {

// Units of work are (row, 16-byte lane) pairs, walked with a grid-stride loop, four independent
// loads in flight per thread.  `max_ctas` bounds the grid: a pull over NVLink needs bytes in
// flight (~1 MB for 700 GB/s at ~1.5 us), not SMs, so a small persistent grid saturates the link
// and leaves the SMs to the SpMM that runs concurrently on another stream.
template <int UNIT>
__global__ void __launch_bounds__(256)
gather_rows_kernel(int64_t n_out, const int64_t* __restrict__ idx, int64_t x_rows,
                   const char* __restrict__ x, int64_t x_ld_bytes, int64_t row_bytes,
                   char* __restrict__ out, int64_t out_ld_bytes, int32_t* __restrict__ bad) {
  using V = typename std::conditional<UNIT == 16, uint4,
                                      typename std::conditional<UNIT == 4, uint32_t, uint8_t>::type>::type;
  const int64_t lanes = row_bytes / UNIT;
  const int64_t total = n_out * lanes;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  constexpr int K = 4;
  for (int64_t u0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; u0 < total; u0 += stride * K) {
    V v[K];
    int64_t dst_off[K];
    bool ok[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const int64_t u = u0 + (int64_t)k * stride;
      ok[k] = u < total;
      if (ok[k]) {
        const int64_t row = u / lanes;
        const int64_t lane = u - row * lanes;
        const int64_t s = idx ? idx[row] : row;
        if (s < 0 || s >= x_rows) {
          atomicExch(bad, 1);
          ok[k] = false;
        } else {
          v[k] = *reinterpret_cast<const V*>(x + s * x_ld_bytes + lane * UNIT);
          dst_off[k] = row * out_ld_bytes + lane * UNIT;
        }
      }
    }
#pragma unroll
    for (int k = 0; k < K; ++k)
      if (ok[k]) *reinterpret_cast<V*>(out + dst_off[k]) = v[k];
  }
}

// Multi-source form: output row i comes from source `seg` with seg_ptr[seg] <= i < seg_ptr[seg+1],
// row idx[i] of that source.  One launch pulls a halo chunk from ALL owners at once (every NVLink
// peer busy simultaneously, no per-owner launch tails).
constexpr int kMaxSources = 16;
struct MultiSrc {
  const char* base[kMaxSources];
  int64_t rows[kMaxSources];
  int64_t seg_ptr[kMaxSources + 1];
  int n_src;
};

__global__ void __launch_bounds__(256)
gather_rows_multi_kernel(const MultiSrc ms, const int64_t* __restrict__ idx, int64_t row_bytes,
                         char* __restrict__ out, int32_t* __restrict__ bad) {
  const int64_t lanes = row_bytes / 16;
  const int64_t total = ms.seg_ptr[ms.n_src] * lanes;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  constexpr int K = 4;
  for (int64_t u0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; u0 < total; u0 += stride * K) {
    uint4 v[K];
    int64_t dst_off[K];
    bool ok[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const int64_t u = u0 + (int64_t)k * stride;
      ok[k] = u < total;
      if (ok[k]) {
        const int64_t row = u / lanes;
        const int64_t lane = u - row * lanes;
        int seg = 0;
#pragma unroll
        for (int q = 1; q < kMaxSources; ++q)
          if (q < ms.n_src && row >= ms.seg_ptr[q]) seg = q;
        const int64_t s = idx[row];
        if (s < 0 || s >= ms.rows[seg]) {
          atomicExch(bad, 1);
          ok[k] = false;
        } else {
          v[k] = *reinterpret_cast<const uint4*>(ms.base[seg] + s * row_bytes + lane * 16);
          dst_off[k] = row * row_bytes + lane * 16;
        }
      }
    }
#pragma unroll
    for (int k = 0; k < K; ++k)
      if (ok[k]) *reinterpret_cast<uint4*>(out + dst_off[k]) = v[k];
  }
}

// Converting row gather for the reduced-precision WIRE mode of the halo exchange (dist.py,
// DGL_B200_WIRE=bf16): TO_BF16 packs out[i, :] = bf16_rn(x[idx[i], :]) (the owner's push over NVLink
// moves half the bytes), !TO_BF16 widens out[i, :] = float(x[idx[i], :]) (the receiver restores the
// fp32 operand the SpMM reads).  A unit of work is 8 consecutive elements of one row: two 16-byte
// fp32 accesses against one 16-byte bf16 access; rows whose width is not a multiple of 8 finish
// element-wise.  Same grid-stride / max_ctas structure as gather_rows_kernel.
template <bool TO_BF16>
__global__ void __launch_bounds__(256)
gather_rows_convert_kernel(int64_t n_out, const int64_t* __restrict__ idx, int64_t x_rows,
                           const void* __restrict__ x_, int64_t x_ld, int64_t feat,
                           void* __restrict__ out_, int64_t out_ld, int vec_ok,
                           int32_t* __restrict__ bad) {
  using In = typename std::conditional<TO_BF16, float, __nv_bfloat16>::type;
  using Out = typename std::conditional<TO_BF16, __nv_bfloat16, float>::type;
  const In* x = static_cast<const In*>(x_);
  Out* out = static_cast<Out*>(out_);
  const int64_t units = (feat + 7) / 8;
  const int64_t total = n_out * units;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; u < total; u += stride) {
    const int64_t row = u / units;
    const int64_t f0 = (u - row * units) * 8;
    const int64_t s = idx ? idx[row] : row;
    if (s < 0 || s >= x_rows) {
      atomicExch(bad, 1);
      continue;
    }
    const In* src = x + s * x_ld + f0;
    Out* dst = out + row * out_ld + f0;
    if (vec_ok && f0 + 8 <= feat) {
      if constexpr (TO_BF16) {
        const float4 a = *reinterpret_cast<const float4*>(src);
        const float4 b = *reinterpret_cast<const float4*>(src + 4);
        uint4 t;
        t.x = pack_bf16(a.x, a.y); t.y = pack_bf16(a.z, a.w);
        t.z = pack_bf16(b.x, b.y); t.w = pack_bf16(b.z, b.w);
        *reinterpret_cast<uint4*>(dst) = t;
      } else {
        const uint4 t = *reinterpret_cast<const uint4*>(src);
        *reinterpret_cast<float4*>(dst) = make_float4(bf16_lo(t.x), bf16_hi(t.x), bf16_lo(t.y), bf16_hi(t.y));
        *reinterpret_cast<float4*>(dst + 4) = make_float4(bf16_lo(t.z), bf16_hi(t.z), bf16_lo(t.w), bf16_hi(t.w));
      }
    } else {
      const int64_t f1 = f0 + 8 < feat ? f0 + 8 : feat;
      for (int64_t f = f0; f < f1; ++f) {
        if constexpr (TO_BF16) dst[f - f0] = __float2bfloat16_rn(src[f - f0]);
        else dst[f - f0] = __bfloat162float(src[f - f0]);
      }
    }
  }
}

}  // namespace dglb200

using namespace dglb200;

extern "C" int dglb200_gather_rows_convert(int64_t n_out, const int64_t* idx, int64_t x_rows,
                                           int to_bf16, const void* x, int64_t x_ld, int64_t feat,
                                           void* out, int64_t out_ld, int32_t* bad_flag,
                                           int max_ctas, void* stream_) {
  clear_error();
  DGLB200_REQUIRE(n_out >= 0 && x_rows >= 0 && feat >= 1, "gather_rows_convert: bad sizes");
  DGLB200_REQUIRE(x_ld >= feat && out_ld >= feat, "gather_rows_convert: leading dimension < feat");
  if (n_out == 0) return 0;
  DGLB200_REQUIRE(x && out && bad_flag, "gather_rows_convert: NULL pointer");
  // 16-byte accesses on both sides: bases aligned, leading dimensions multiples of 8 elements
  const int vec_ok = (reinterpret_cast<uintptr_t>(x) % 16 == 0) && (reinterpret_cast<uintptr_t>(out) % 16 == 0) &&
                     x_ld % 8 == 0 && out_ld % 8 == 0;
  const int64_t total = n_out * ((feat + 7) / 8);
  int64_t blocks = ceil_div(total, 256);
  if (max_ctas > 0 && blocks > max_ctas) blocks = max_ctas;
  if (blocks > 148 * 64) blocks = 148 * 64;      // grid-stride: a few waves of the 148 SMs are enough
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (to_bf16)
    gather_rows_convert_kernel<true><<<(unsigned)blocks, 256, 0, stream>>>(n_out, idx, x_rows, x, x_ld, feat, out,
                                                                         out_ld, vec_ok, bad_flag);
  else
    gather_rows_convert_kernel<false><<<(unsigned)blocks, 256, 0, stream>>>(n_out, idx, x_rows, x, x_ld, feat, out,
                                                                          out_ld, vec_ok, bad_flag);
  DGLB200_POST_LAUNCH("gather_rows_convert_kernel");
  return 0;
}

extern "C" int dglb200_gather_rows_multi(int n_src, const void* const* src_base_host,
                                         const int64_t* src_rows_host, const int64_t* seg_ptr_host,
                                         const int64_t* idx, int64_t row_bytes, void* out,
                                         int32_t* bad_flag, int max_ctas, void* stream_) {
  clear_error();
  DGLB200_REQUIRE(n_src >= 1 && n_src <= kMaxSources, "gather_rows_multi: n_src %d out of range", n_src);
  DGLB200_REQUIRE(src_base_host && src_rows_host && seg_ptr_host, "gather_rows_multi: NULL table");
  DGLB200_REQUIRE(row_bytes >= 16 && row_bytes % 16 == 0, "gather_rows_multi: rows must be multiples of 16 bytes");
  MultiSrc ms;
  ms.n_src = n_src;
  for (int q = 0; q < kMaxSources; ++q) {
    ms.base[q] = q < n_src ? static_cast<const char*>(src_base_host[q]) : nullptr;
    ms.rows[q] = q < n_src ? src_rows_host[q] : 0;
  }
  for (int q = 0; q <= kMaxSources; ++q) ms.seg_ptr[q] = seg_ptr_host[q <= n_src ? q : n_src];
  const int64_t n_out = ms.seg_ptr[n_src];
  if (n_out == 0) return 0;
  DGLB200_REQUIRE(idx && out && bad_flag, "gather_rows_multi: NULL pointer");
  DGLB200_REQUIRE((reinterpret_cast<uintptr_t>(out) & 15) == 0, "gather_rows_multi: out must be 16-byte aligned");
  const int64_t total = n_out * (row_bytes / 16);
  int64_t blocks = ceil_div(total, 256 * 4);
  if (max_ctas > 0 && blocks > max_ctas) blocks = max_ctas;
  DGLB200_REQUIRE(blocks < (1ll << 31), "gather_rows_multi: grid too large");
  gather_rows_multi_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      ms, idx, row_bytes, static_cast<char*>(out), bad_flag);
  DGLB200_POST_LAUNCH("gather_rows_multi_kernel");
  return 0;
}

extern "C" int dglb200_gather_rows_strided(int64_t n_out, const int64_t* idx, int64_t x_rows,
                                           const void* x, int64_t x_ld_bytes, int64_t row_bytes,
                                           void* out, int64_t out_ld_bytes, int32_t* bad_flag,
                                           int max_ctas, void* stream_) {
  clear_error();
  DGLB200_REQUIRE(n_out >= 0 && x_rows >= 0 && row_bytes >= 1, "gather_rows: bad sizes");
  DGLB200_REQUIRE(x_ld_bytes >= row_bytes && out_ld_bytes >= row_bytes, "gather_rows: leading dimension < row");
  if (n_out == 0) return 0;
  DGLB200_REQUIRE(x && out && bad_flag, "gather_rows: NULL pointer");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  auto al = [](const void* p, int64_t ld, int a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0 && ld % a == 0; };
  const int unit = (row_bytes % 16 == 0 && al(x, x_ld_bytes, 16) && al(out, out_ld_bytes, 16)) ? 16
                   : (row_bytes % 4 == 0 && al(x, x_ld_bytes, 4) && al(out, out_ld_bytes, 4)) ? 4 : 1;
  const int64_t total = n_out * (row_bytes / unit);
  int64_t blocks = ceil_div(total, 256 * 4);
  if (max_ctas > 0 && blocks > max_ctas) blocks = max_ctas;
  DGLB200_REQUIRE(blocks < (1ll << 31), "gather_rows: grid too large");
  const char* xs = static_cast<const char*>(x);
  char* os = static_cast<char*>(out);
  if (unit == 16)
    gather_rows_kernel<16><<<(unsigned)blocks, 256, 0, stream>>>(n_out, idx, x_rows, xs, x_ld_bytes, row_bytes, os, out_ld_bytes, bad_flag);
  else if (unit == 4)
    gather_rows_kernel<4><<<(unsigned)blocks, 256, 0, stream>>>(n_out, idx, x_rows, xs, x_ld_bytes, row_bytes, os, out_ld_bytes, bad_flag);
  else
    gather_rows_kernel<1><<<(unsigned)blocks, 256, 0, stream>>>(n_out, idx, x_rows, xs, x_ld_bytes, row_bytes, os, out_ld_bytes, bad_flag);
  DGLB200_POST_LAUNCH("gather_rows_kernel");
  return 0;
}

extern "C" int dglb200_gather_rows(int64_t n_out, const int64_t* idx, int64_t x_rows, const void* x,
                                   int64_t row_bytes, void* out, int32_t* bad_flag, void* stream_) {
  DGLB200_REQUIRE(idx != nullptr, "gather_rows: idx is NULL");
  return dglb200_gather_rows_strided(n_out, idx, x_rows, x, row_bytes, row_bytes, out, row_bytes, bad_flag, 0,
                                     stream_);
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
