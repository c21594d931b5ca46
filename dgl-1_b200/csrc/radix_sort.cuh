// Hand-written stable LSD radix sort of (uint32 key, uint32 value) pairs, 8 bits per pass.
// Stability is what makes the CSR build reproduce the reference's per-row edge order
// (Graph::GetAdj keeps adjacency-list insertion order = ascending edge id,
// src/graph/graph.cc:540-547) and the bucketing keep per-node message order
// (src/scheduler/scheduler.cc:22-25).
//
// Per pass: (a) per-tile digit histogram -> (b) exclusive scan over [digit][tile] ->
// (c) stable scatter.  Inside a tile the order of elements is (warp, round, lane); ranks among
// equal digits come from __match_any_sync + popc, per-warp running counters live in shared
// memory.  HBM-bound: 2 reads + 1 write of the pairs per pass.
#pragma once
#include "common.cuh"
#include "scan.cuh"

namespace dglb200 {

constexpr int kRadixBits = 8;
constexpr int kRadix = 1 << kRadixBits;
constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortRounds = 16;                          // items per thread
constexpr int kSortTile = kSortThreads * kSortRounds;    // 4096 elements per CTA
constexpr int kSortWarpSpan = 32 * kSortRounds;          // 512 contiguous elements per warp

static __global__ void __launch_bounds__(kSortThreads)
radix_hist(const uint32_t* __restrict__ keys, int64_t n, int shift, int32_t* __restrict__ ghist,
           int64_t tiles) {
  __shared__ int32_t hist[kRadix];
  hist[threadIdx.x] = 0;
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t base = (int64_t)blockIdx.x * kSortTile + warp * kSortWarpSpan + lane;
#pragma unroll
  for (int r = 0; r < kSortRounds; ++r) {
    const int64_t i = base + r * 32;
    if (i < n) atomicAdd(&hist[(keys[i] >> shift) & (kRadix - 1)], 1);
  }
  __syncthreads();
  ghist[(int64_t)threadIdx.x * tiles + blockIdx.x] = hist[threadIdx.x];
}

// Stable scatter of one tile.  The tile's pairs are first put in DIGIT ORDER in shared memory (their
// rank inside the tile), then written out by consecutive threads: elements of one digit are
// consecutive in shared memory AND in the output, so a warp's 32 stores fall into a few long runs
// instead of up to 32 different sectors (round 1 stored straight from the ranks: 0.14 of the HBM
// roofline).
static __global__ void __launch_bounds__(kSortThreads)
radix_scatter(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
              uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int64_t n,
              int shift, const int32_t* __restrict__ goff, int64_t tiles) {
  __shared__ int32_t wcnt[kSortWarps][kRadix];
  __shared__ int32_t gbase[kRadix];          // global position of a pair = gbase[digit] + rank in tile
  __shared__ int32_t scan_smem[33];
  __shared__ uint32_t skey[kSortTile];
  __shared__ uint32_t sval[kSortTile];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < kSortWarps * kRadix; i += kSortThreads) (&wcnt[0][0])[i] = 0;
  __syncthreads();

  const int64_t tile0 = (int64_t)blockIdx.x * kSortTile;
  const int64_t base = tile0 + warp * kSortWarpSpan + lane;
  uint32_t key[kSortRounds];
  const uint32_t lt_mask = (1u << lane) - 1;

  // phase 1: per-warp digit counts
#pragma unroll
  for (int r = 0; r < kSortRounds; ++r) {
    const int64_t i = base + r * 32;
    const bool ok = i < n;
    key[r] = ok ? keys_in[i] : 0u;
    const int d = ok ? (int)((key[r] >> shift) & (kRadix - 1)) : kRadix;
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    if (ok && (peers & lt_mask) == 0) wcnt[warp][d] += __popc(peers);
    __syncwarp();
  }
  __syncthreads();

  // phase 2: counts -> start ranks inside the tile (digit-major, warp-major inside a digit)
  {
    const int d = threadIdx.x;
    int32_t tot = 0;
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w) tot += wcnt[w][d];
    int32_t tile_total;
    int32_t run = block_exclusive_scan<int32_t>(tot, &tile_total, scan_smem);
    gbase[d] = goff[(int64_t)d * tiles + blockIdx.x] - run;
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w) {
      const int32_t c = wcnt[w][d];
      wcnt[w][d] = run;
      run += c;
    }
  }
  __syncthreads();

  // phase 3: stable placement in shared memory
#pragma unroll
  for (int r = 0; r < kSortRounds; ++r) {
    const int64_t i = base + r * 32;
    const bool ok = i < n;
    const int d = ok ? (int)((key[r] >> shift) & (kRadix - 1)) : kRadix;
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    int32_t pos = 0;
    if (ok) pos = wcnt[warp][d] + __popc(peers & lt_mask);
    __syncwarp();
    if (ok && (peers & lt_mask) == 0) wcnt[warp][d] += __popc(peers);
    __syncwarp();
    if (ok) {
      skey[pos] = key[r];
      sval[pos] = vals_in ? vals_in[i] : (uint32_t)i;
    }
  }
  __syncthreads();

  // phase 4: coalesced write-out in rank order
  const int tile_n = (int)min((int64_t)kSortTile, n - tile0);
  for (int j = threadIdx.x; j < tile_n; j += kSortThreads) {
    const uint32_t k = skey[j];
    const int32_t pos = gbase[(k >> shift) & (kRadix - 1)] + j;
    keys_out[pos] = k;
    vals_out[pos] = sval[j];
  }
}

inline int bits_for(int64_t bound) {  // number of bits needed for keys in [0, bound)
  int b = 1;
  while (b < 32 && (1ll << b) < bound) ++b;
  return b;
}

inline size_t radix_sort_workspace_bytes(int64_t n) {
  const int64_t tiles = ceil_div(n > 0 ? n : 1, kSortTile);
  return align_up((size_t)kRadix * tiles * sizeof(int32_t)) +
         scan_workspace_bytes(kRadix * tiles, sizeof(int32_t)) + 256;
}

// Sorts n pairs by the low `bits` bits of the key.  vals_in == nullptr means value = index.
// The sorted pairs end up in (keys_out, vals_out); (keys_tmp, vals_tmp) is scratch of the same
// size.  keys_in is never written.  Returns 0 / -1.
inline int radix_sort_pairs(const uint32_t* keys_in, const uint32_t* vals_in, uint32_t* keys_out,
                            uint32_t* vals_out, uint32_t* keys_tmp, uint32_t* vals_tmp, int64_t n,
                            int bits, Workspace& ws, cudaStream_t stream) {
  if (n <= 0) return 0;
  DGLB200_REQUIRE(n < (1ll << 31), "radix_sort: n too large");
  const int passes = bits <= 0 ? 1 : (bits + kRadixBits - 1) / kRadixBits;
  const int64_t tiles = ceil_div(n, kSortTile);
  const size_t mark = ws.used;
  int32_t* ghist = ws.take<int32_t>((size_t)kRadix * tiles);
  DGLB200_REQUIRE(ghist != nullptr, "radix_sort: workspace too small");
  const size_t mark2 = ws.used;
  const uint32_t* kin = keys_in;
  const uint32_t* vin = vals_in;
  for (int p = 0; p < passes; ++p) {
    // the last pass must land in *_out: passes-1-p even -> out, odd -> tmp
    const bool to_out = ((passes - 1 - p) % 2) == 0;
    uint32_t* kout = to_out ? keys_out : keys_tmp;
    uint32_t* vout = to_out ? vals_out : vals_tmp;
    const int shift = p * kRadixBits;
    radix_hist<<<(unsigned)tiles, kSortThreads, 0, stream>>>(kin, n, shift, ghist, tiles);
    DGLB200_POST_LAUNCH("radix_hist");
    ws.used = mark2;  // the scan scratch is reused every pass
    if (exclusive_scan<int32_t, int32_t>(ghist, ghist, (int64_t)kRadix * tiles, 0, ws, stream) != 0)
      return -1;
    radix_scatter<<<(unsigned)tiles, kSortThreads, 0, stream>>>(kin, vin, kout, vout, n, shift,
                                                                ghist, tiles);
    DGLB200_POST_LAUNCH("radix_scatter");
    kin = kout;
    vin = vout;
  }
  ws.used = mark;
  return 0;
}

}  // namespace dglb200
