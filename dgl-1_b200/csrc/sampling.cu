// Uniform neighbour sampling on the device: for a set of rows of a CSR, min(degree, fanout) distinct
// slots per row, uniformly at random, emitted in ascending slot order - one hop of
// ImmutableGraph::SampleSubgraph / GetUniformSample (src/graph/immutable_graph.cc:701-741, 772-874).
// The reference walks its BFS queue on one core and draws with rand_r seeded by time(); here a hop is one
// launch, one WARP per row, and the draws come from a counter-based generator keyed by (seed, row id,
// draw, attempt), so a call is reproducible and independent of the launch geometry.  Rows with
// degree <= fanout are copied whole (the reference's first branch): that case is deterministic and is what
// the golden vectors of the unmodified reference pin bit for bit.
#include "scan.cuh"

namespace dglb200 {

constexpr int kSampleWarps = 4;          // warps (= rows in flight) per CTA
constexpr int kSampleMaxDraw = 128;      // draws a warp holds in shared memory; more: sequential selection

__device__ __forceinline__ uint64_t sample_mix(uint64_t z) {      // splitmix64 finaliser
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// draw number `i`, attempt `a`, of row `r`: uniform in [0, bound)
__device__ __forceinline__ int64_t sample_draw(uint64_t seed, int64_t r, uint32_t i, uint32_t a, int64_t bound) {
  const uint64_t h = sample_mix(sample_mix(seed ^ ((uint64_t)r * 0xD6E8FEB86659FD93ull)) + (((uint64_t)a << 32) | i));
  return (int64_t)__umul64hi(h, (uint64_t)bound);
}

// len[i] = min(degree of rows[i], fanout); bad ids as in select_lens_kernel.
__global__ void __launch_bounds__(256)
sample_lens_kernel(int64_t n_sel, const int64_t* __restrict__ rows, int64_t n_rows,
                   const int64_t* __restrict__ indptr, int64_t fanout, int64_t* __restrict__ len, int32_t* bad) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_sel) return;
  const int64_t r = rows[i];
  if (r < 0 || r >= n_rows) {
    if (bad != nullptr) atomicExch(bad, 1);
    len[i] = 0;
    return;
  }
  const int64_t d = indptr[r + 1] - indptr[r];
  len[i] = d < fanout ? d : fanout;
}

// One warp per selected row.  deg <= fanout: copy.  Otherwise the warp draws the SMALLER of the kept set
// (fanout slots) and the dropped set (deg - fanout slots) - the reference's two branches, :717-729 - by
// rejection: every draw is an independent uniform slot, a draw equal to an earlier-numbered one is redrawn.
// The procedure only looks at which draws are equal, so it is symmetric under any relabelling of the slots:
// every subset of that size is equally likely.  The picks are rank-sorted in shared memory, then written
// (kept set) or skipped (dropped set) in ascending slot order.
__global__ void __launch_bounds__(kSampleWarps * kWarp)
neighbor_sample_kernel(int64_t n_sel, const int64_t* __restrict__ rows, const int64_t* __restrict__ offsets,
                       const int64_t* __restrict__ indptr, const int32_t* __restrict__ indices,
                       const int32_t* __restrict__ eids, int64_t fanout, uint64_t seed,
                       int64_t* __restrict__ out_row, int64_t* __restrict__ out_col,
                       int64_t* __restrict__ out_eid, int64_t* __restrict__ out_slot) {
  __shared__ int64_t s_pick[kSampleWarps][kSampleMaxDraw];
  __shared__ int64_t s_sort[kSampleWarps][kSampleMaxDraw];
  const int lane = threadIdx.x & (kWarp - 1), wid = threadIdx.x / kWarp;
  int64_t* pick = s_pick[wid];
  int64_t* srt = s_sort[wid];
  const int64_t n_warps = (int64_t)gridDim.x * kSampleWarps;
  for (int64_t i = (int64_t)blockIdx.x * kSampleWarps + wid; i < n_sel; i += n_warps) {
    const int64_t r = rows[i];
    const int64_t start = indptr[r], deg = indptr[r + 1] - start, off = offsets[i];
    auto emit = [&](int64_t p, int64_t s) {                 // output position p <- CSR slot s
      if (out_row) out_row[p] = r;
      if (out_col) out_col[p] = indices[s];
      if (out_eid) out_eid[p] = eids ? (int64_t)eids[s] : s;
      if (out_slot) out_slot[p] = s;
    };
    if (deg <= fanout) {
      for (int64_t j = lane; j < deg; j += kWarp) emit(off + j, start + j);
      continue;
    }
    const bool keep = fanout <= deg - fanout;               // draw the kept set, else the dropped one
    const int64_t m = keep ? fanout : deg - fanout;
    if (m > kSampleMaxDraw) {
      // selection sampling (Knuth 3.4.2 S) by one lane: slot j is kept with probability
      // (still needed) / (still to come); emits in ascending order by construction
      if (lane == 0) {
        int64_t need = fanout, p = off;
        for (int64_t j = 0; j < deg && need > 0; ++j) {
          if (sample_draw(seed, r, (uint32_t)j, (uint32_t)(j >> 32) + 1u, deg - j) < need) {
            emit(p++, start + j);
            --need;
          }
        }
      }
      continue;
    }
    const int mi = (int)m;
    uint32_t attempt[kSampleMaxDraw / kWarp] = {0, 0, 0, 0};
    for (int q = lane; q < mi; q += kWarp) pick[q] = sample_draw(seed, r, (uint32_t)q, 0u, deg);
    for (;;) {
      __syncwarp();
      unsigned redo = 0;                                     // bit t: my t-th draw equals an earlier-numbered draw
      for (int q = lane, t = 0; q < mi; q += kWarp, ++t) {
        const int64_t v = pick[q];
        bool dup = false;
        for (int j = 0; j < q; ++j) dup |= (pick[j] == v);
        redo |= dup ? (1u << t) : 0u;
      }
      if (__ballot_sync(0xffffffffu, redo != 0) == 0u) break;
      __syncwarp();                                          // all compares done before any rewrite
      for (int q = lane, t = 0; q < mi; q += kWarp, ++t)
        if (redo & (1u << t)) pick[q] = sample_draw(seed, r, (uint32_t)q, ++attempt[t], deg);
    }
    // rank sort (all picks are distinct)
    for (int q = lane; q < mi; q += kWarp) {
      const int64_t v = pick[q];
      int rank = 0;
      for (int j = 0; j < mi; ++j) rank += (pick[j] < v) ? 1 : 0;
      srt[rank] = v;
    }
    __syncwarp();
    if (keep) {
      for (int q = lane; q < mi; q += kWarp) emit(off + q, start + srt[q]);
    } else {
      // slot j survives unless dropped; it lands at j - (number of dropped slots before it)
      for (int64_t j = lane; j < deg; j += kWarp) {
        int lo = 0, hi = mi;                                 // first index with srt[idx] >= j
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (srt[mid] < j) lo = mid + 1; else hi = mid;
        }
        if (lo < mi && srt[lo] == j) continue;
        emit(off + j - lo, start + j);
      }
    }
    __syncwarp();                                            // pick / srt are reused by the next row
  }
}

}  // namespace dglb200

using namespace dglb200;

extern "C" int dglb200_neighbor_sample_offsets(int64_t n_sel, const int64_t* rows, int64_t n_rows,
                                               const int64_t* indptr, int64_t fanout, int64_t* offsets,
                                               int64_t* total_host, void* workspace, size_t workspace_bytes,
                                               void* stream_) {
  clear_error();
  DGLB200_REQUIRE(n_sel >= 0 && n_rows >= 0 && total_host != nullptr, "neighbor_sample_offsets: bad args");
  DGLB200_REQUIRE(fanout >= 0, "neighbor_sample_offsets: fanout must be >= 0, got %lld", (long long)fanout);
  *total_host = 0;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (n_sel == 0) {
    if (offsets) DGLB200_CUDA(cudaMemsetAsync(offsets, 0, sizeof(int64_t), stream));
    return 0;
  }
  DGLB200_REQUIRE(rows && indptr && offsets, "neighbor_sample_offsets: NULL pointer");
  Workspace ws(workspace, workspace_bytes);
  int64_t* len = ws.take<int64_t>(n_sel + 1);     // last word: the bad-id flag
  DGLB200_REQUIRE(len != nullptr, "neighbor_sample_offsets: workspace too small");
  int32_t* bad = reinterpret_cast<int32_t*>(len + n_sel);
  DGLB200_CUDA(cudaMemsetAsync(bad, 0, sizeof(int64_t), stream));
  sample_lens_kernel<<<(unsigned)ceil_div(n_sel, 256), 256, 0, stream>>>(n_sel, rows, n_rows, indptr, fanout, len, bad);
  DGLB200_POST_LAUNCH("sample_lens_kernel");
  if (exclusive_scan<int64_t, int64_t>(len, offsets, n_sel, 1, ws, stream) != 0) return -1;
  int32_t bad_h = 0;
  DGLB200_CUDA(cudaMemcpyAsync(total_host, offsets + n_sel, sizeof(int64_t), cudaMemcpyDeviceToHost, stream));
  DGLB200_CUDA(cudaMemcpyAsync(&bad_h, bad, sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
  DGLB200_CUDA(cudaStreamSynchronize(stream));
  DGLB200_REQUIRE(bad_h == 0, "Invalid vertex: a seed id is outside [0, %lld)", (long long)n_rows);
  return 0;
}

extern "C" int dglb200_neighbor_sample_fill(int64_t n_sel, const int64_t* rows, const int64_t* offsets,
                                            const int64_t* indptr, const int32_t* indices, const int32_t* eids,
                                            int64_t fanout, uint64_t seed, int64_t* out_row, int64_t* out_col,
                                            int64_t* out_eid, int64_t* out_slot, void* stream_) {
  clear_error();
  DGLB200_REQUIRE(n_sel >= 0 && fanout >= 0, "neighbor_sample_fill: bad sizes");
  if (n_sel == 0) return 0;
  DGLB200_REQUIRE(rows && offsets && indptr, "neighbor_sample_fill: NULL pointer");
  DGLB200_REQUIRE(out_col == nullptr || indices != nullptr, "neighbor_sample_fill: out_col needs indices");
  int64_t blocks = ceil_div(n_sel, kSampleWarps);
  const int64_t cap = (int64_t)kNumSMsB200 * 16;            // 16 CTAs of 128 threads per SM
  if (blocks > cap) blocks = cap;
  neighbor_sample_kernel<<<(unsigned)blocks, kSampleWarps * kWarp, 0, static_cast<cudaStream_t>(stream_)>>>(
      n_sel, rows, offsets, indptr, indices, eids, fanout, seed, out_row, out_col, out_eid, out_slot);
  DGLB200_POST_LAUNCH("neighbor_sample_kernel");
  return 0;
}
