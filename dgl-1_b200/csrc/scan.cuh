// Device-wide exclusive prefix sum (hand-written; no CUB): reduce tiles -> scan the tile sums
// recursively -> scan inside tiles with the tile offset.  Used for row pointers, stream
// compaction and radix-sort digit offsets.  HBM-bound: reads the input twice, writes once.
#pragma once
#include "common.cuh"

namespace dglb200 {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;  // 2048 elements per CTA

template <typename TOut>
__device__ __forceinline__ TOut block_exclusive_scan(TOut v, TOut* total, TOut* smem /*[32]*/) {
  // inclusive warp scan, then scan of warp totals
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  TOut inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    TOut t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) smem[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    TOut w = (lane < (int)(blockDim.x >> 5)) ? smem[lane] : TOut(0);
    TOut winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      TOut t = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += t;
    }
    smem[lane] = winc - w;  // exclusive warp offsets
    if (lane == 31) smem[32] = winc;
  }
  __syncthreads();
  TOut r = smem[warp] + inc - v;
  *total = smem[32];
  __syncthreads();
  return r;
}

template <typename TIn, typename TOut>
__global__ void __launch_bounds__(kScanThreads) scan_tile_sums(const TIn* in, int64_t n, TOut* sums) {
  __shared__ TOut smem[33];
  const int64_t base = (int64_t)blockIdx.x * kScanTile;
  TOut s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    int64_t idx = base + (int64_t)i * kScanThreads + threadIdx.x;
    if (idx < n) s += (TOut)in[idx];
  }
  TOut total;
  (void)block_exclusive_scan<TOut>(s, &total, smem);
  if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

// out[i] = offset(tile) + exclusive prefix inside the tile; if out_total != nullptr the CTA
// owning the last element also writes the grand total to out[n] (row-pointer convention).
template <typename TIn, typename TOut>
__global__ void __launch_bounds__(kScanThreads)
scan_tile_apply(const TIn* in, int64_t n, const TOut* tile_offsets, TOut* out, int write_total) {
  __shared__ TOut smem[33];
  const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
  TOut v[kScanItems];
  TOut s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    int64_t idx = base + i;
    v[i] = idx < n ? (TOut)in[idx] : TOut(0);
    s += v[i];
  }
  TOut total;
  TOut excl = block_exclusive_scan<TOut>(s, &total, smem);
  TOut run = excl + (tile_offsets ? tile_offsets[blockIdx.x] : TOut(0));
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    int64_t idx = base + i;
    if (idx < n) out[idx] = run;
    run += v[i];
    if (write_total && idx == n - 1) out[n] = run;
  }
}

inline size_t scan_workspace_bytes(int64_t n, size_t elem) {
  size_t total = 0;
  int64_t m = ceil_div(n > 0 ? n : 1, kScanTile);
  while (true) {
    total += align_up((size_t)m * elem);
    if (m <= 1) break;
    m = ceil_div(m, kScanTile);
  }
  return total + 256;
}

// Exclusive scan of in[0..n) into out[0..n) (+ out[n] = total when write_total).  `in` and
// `out` may alias only if TIn == TOut.  Returns 0 / -1.
template <typename TIn, typename TOut>
int exclusive_scan(const TIn* in, TOut* out, int64_t n, int write_total, Workspace& ws,
                   cudaStream_t stream) {
  if (n <= 0) {
    if (write_total) DGLB200_CUDA(cudaMemsetAsync(out, 0, sizeof(TOut), stream));
    return 0;
  }
  const int64_t tiles = ceil_div(n, kScanTile);
  TOut* sums = nullptr;
  if (tiles > 1) {
    sums = ws.take<TOut>((size_t)tiles);
    DGLB200_REQUIRE(sums != nullptr, "exclusive_scan: workspace too small");
    scan_tile_sums<TIn, TOut><<<(unsigned)tiles, kScanThreads, 0, stream>>>(in, n, sums);
    DGLB200_POST_LAUNCH("scan_tile_sums");
    if (exclusive_scan<TOut, TOut>(sums, sums, tiles, 0, ws, stream) != 0) return -1;
  }
  scan_tile_apply<TIn, TOut><<<(unsigned)tiles, kScanThreads, 0, stream>>>(in, n, sums, out,
                                                                          write_total);
  DGLB200_POST_LAUNCH("scan_tile_apply");
  return 0;
}

}  // namespace dglb200
