// Row gather: out[i, :] = x[idx[i], :].  READ_ROW of the reference's executors
// (python/dgl/runtime/ir/executor.py:365-394 -> F.gather_row, backend/pytorch/tensor.py:108-109)
// and the send-side packing of the compacted halo exchange (dgl_b200/dist.py).  HBM-bound:
// n_out * row_bytes read + written; 128-bit accesses, one group of lanes per row.
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
