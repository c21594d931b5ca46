// Uniform neighbour sampling on the device: for a set of rows of a CSR, min(degree, fanout) distinct
// slots per row, uniformly at random, emitted in ascending slot order - one hop of
// ImmutableGraph::SampleSubgraph / GetUniformSample (src/graph/immutable_graph.cc:701-741, 772-874).
// The reference walks its BFS queue on one core and draws with rand_r seeded by time(); here a hop is one
// launch, one WARP per row, and the draws come from a counter-based generator keyed by (seed, row id,
// draw, attempt), so a call is reproducible and independent of the launch geometry.  Rows with
// degree <= fanout are copied whole (the reference's first branch): that case is deterministic and is what
// the golden vectors of the unmodified reference pin bit for bit.
#include <string.h>

#include "scan.cuh"

namespace dglb200 {

constexpr int kSampleWarps = 4;          // warps (= rows in flight) per CTA
constexpr int kSampleMaxDraw = 128;      // draws a warp holds in shared memory; more: sequential selection
// picks per hop the whole-call entry point over-allocates for instead of waiting for the exact number (the host
// emulation of tests/native/cuda_emu pays one host thread per CUDA thread of the capacity-sized grids: lower there)
#ifdef DGLB200_HOST_EMU
constexpr int64_t kSampleBoundLimit = 1 << 11;
#else
constexpr int64_t kSampleBoundLimit = 1 << 21;
#endif

__device__ __forceinline__ uint64_t sample_mix(uint64_t z) {      // splitmix64 finaliser
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// draw number `i`, attempt `a`, of row `r`: uniform in [0, bound).  row_key(seed, r) is hoisted out of the draws.
__device__ __forceinline__ uint64_t sample_row_key(uint64_t seed, int64_t r) {
  return sample_mix(seed ^ ((uint64_t)r * 0xD6E8FEB86659FD93ull));
}
__device__ __forceinline__ int64_t sample_draw(uint64_t row_key, uint32_t i, uint32_t a, int64_t bound) {
  return (int64_t)__umul64hi(sample_mix(row_key + (((uint64_t)a << 32) | i)), (uint64_t)bound);
}

// len[i] = min(degree of rows[i], fanout); bad ids as in select_lens_kernel.
__global__ void __launch_bounds__(256)
sample_lens_kernel(int64_t n_sel, const int32_t* __restrict__ n_sel_dev, const int64_t* __restrict__ rows,
                   int64_t n_rows, const int64_t* __restrict__ indptr, int64_t fanout, int64_t* __restrict__ len,
                   int32_t* bad) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_sel) return;
  if (n_sel_dev != nullptr && i >= *n_sel_dev) {     // beyond the live part of an appended frontier
    len[i] = 0;
    return;
  }
  const int64_t r = rows[i];
  if (r < 0 || r >= n_rows) {
    if (bad != nullptr) atomicExch(bad, 1);
    len[i] = 0;
    return;
  }
  const int64_t d = indptr[r + 1] - indptr[r];
  len[i] = d < fanout ? d : fanout;
}

// Discovery state of one sample_subgraph call (MARK): layer[v] = hop at which v was first reached (-1 = not
// yet), bitmap = the same set as bits (compacted at the end into the ascending vertex list), next / next_count =
// the frontier of the following hop, appended warp-aggregated.  The reference keeps an unordered_set + a queue
// (immutable_graph.cc:783-793, 852-860); a whole layer is discovered by one launch here, and within a layer
// every discoverer writes the same value, so the outcome does not depend on who wins the atomicCAS.
struct SampleMark {
  int32_t* layer;
  uint32_t* bitmap;
  int64_t* next;
  int32_t* next_count;
  int32_t hop;
};

__device__ __forceinline__ void mark_append(const SampleMark& mk, bool fresh, int64_t c, int lane) {
  const unsigned m = __ballot_sync(0xffffffffu, fresh);
  if (m == 0u) return;
  const int leader = __ffs((int)m) - 1;
  int base = 0;
  if (lane == leader) base = atomicAdd(mk.next_count, __popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (fresh) {
    mk.next[base + __popc(m & ((1u << lane) - 1u))] = c;
    atomicOr(&mk.bitmap[c >> 5], 1u << (c & 31));
  }
}

// One warp per selected row.  deg <= fanout: copy.  Otherwise the warp draws the SMALLER of the kept set
// (fanout slots) and the dropped set (deg - fanout slots) - the reference's two branches, :717-729 - by
// rejection: every draw is an independent uniform slot, a draw equal to an earlier-numbered one is redrawn.
// The procedure only looks at which draws are equal, so it is symmetric under any relabelling of the slots:
// every subset of that size is equally likely.  The picks are rank-sorted in shared memory, then written
// (kept set) or skipped (dropped set) in ascending slot order.
// n_sel_dev (nullable): the number of valid rows lives on the device (a frontier appended by the previous hop).
template <bool MARK>
__global__ void __launch_bounds__(kSampleWarps * kWarp)
neighbor_sample_kernel(int64_t n_sel, const int32_t* __restrict__ n_sel_dev, const int64_t* __restrict__ rows,
                       const int64_t* __restrict__ offsets, const int64_t* __restrict__ indptr,
                       const int32_t* __restrict__ indices, const int32_t* __restrict__ eids, int64_t fanout,
                       uint64_t seed, int64_t* __restrict__ out_row, int64_t* __restrict__ out_col,
                       int64_t* __restrict__ out_eid, int64_t* __restrict__ out_slot, SampleMark mk) {
  __shared__ int64_t s_pick[kSampleWarps][kSampleMaxDraw];
  __shared__ int64_t s_sort[kSampleWarps][kSampleMaxDraw];
  const int lane = threadIdx.x & (kWarp - 1), wid = threadIdx.x / kWarp;
  int64_t* pick = s_pick[wid];
  int64_t* srt = s_sort[wid];
  if (n_sel_dev != nullptr) {
    const int64_t live = *n_sel_dev;
    if (live < n_sel) n_sel = live;
  }
  const int64_t n_warps = (int64_t)gridDim.x * kSampleWarps;
  for (int64_t i = (int64_t)blockIdx.x * kSampleWarps + wid; i < n_sel; i += n_warps) {
    const int64_t r = rows[i];
    const int64_t start = indptr[r], deg = indptr[r + 1] - start, off = offsets[i];
    // output position p <- CSR slot s; called by all 32 lanes (act: this lane has a pick)
    auto emit = [&](bool act, int64_t p, int64_t s) {
      int64_t c = 0;
      if (act) {
        c = indices[s];
        if (out_row) out_row[p] = r;
        if (out_col) out_col[p] = c;
        if (out_eid) out_eid[p] = eids ? (int64_t)eids[s] : s;
        if (out_slot) out_slot[p] = s;
      }
      if (MARK) {
        const bool fresh = act && atomicCAS(&mk.layer[c], -1, mk.hop + 1) == -1;
        mark_append(mk, fresh, c, lane);
      }
    };
    if (deg <= fanout) {
      for (int64_t b = 0; b < deg; b += kWarp) emit(b + lane < deg, off + b + lane, start + b + lane);
      continue;
    }
    const bool keep = fanout <= deg - fanout;               // draw the kept set, else the dropped one
    const int64_t m = keep ? fanout : deg - fanout;
    const uint64_t rk = sample_row_key(seed, r);
    if (m <= kWarp && deg < (1ll << 31)) {
      // At most one draw per lane: everything stays in registers.  Same draws and the same redraw rule as the
      // shared-memory path below (a draw equal to an EARLIER-numbered one is redrawn), so the picks are identical.
      const int mi = (int)m;
      const unsigned lt = (1u << lane) - 1u;
      uint32_t attempt = 0;
      int pick = lane < mi ? (int)sample_draw(rk, (uint32_t)lane, 0u, deg) : -1 - lane;   // idle lanes: unique
      for (;;) {
        const bool dup = (__match_any_sync(0xffffffffu, pick) & lt) != 0u;
        if (__ballot_sync(0xffffffffu, dup) == 0u) break;
        if (dup) pick = (int)sample_draw(rk, (uint32_t)lane, ++attempt, deg);
      }
      if (keep) {
        int rank = 0;                                        // picks below mine = my place in ascending slot order
        for (int j = 0; j < mi; ++j) rank += (__shfl_sync(0xffffffffu, pick, j) < pick) ? 1 : 0;
        emit(lane < mi, off + rank, start + (lane < mi ? pick : 0));
      } else {
        for (int64_t b = 0; b < deg; b += kWarp) {
          const int j = (int)b + lane;
          int before = 0;                                    // dropped slots below j
          bool dropped = false;
          for (int t = 0; t < mi; ++t) {
            const int v = __shfl_sync(0xffffffffu, pick, t);
            before += (v < j) ? 1 : 0;
            dropped |= (v == j);
          }
          emit(j < deg && !dropped, off + j - before, start + j);
        }
      }
      continue;
    }
    if (m > kSampleMaxDraw) {
      // selection sampling (Knuth 3.4.2 S) by one lane: slot j is kept with probability
      // (still needed) / (still to come); emits in ascending order by construction
      if (lane == 0) {
        int64_t need = fanout, p = off;
        for (int64_t j = 0; j < deg && need > 0; ++j) {
          if (sample_draw(rk, (uint32_t)j, (uint32_t)(j >> 32) + 1u, deg - j) < need) {
            const int64_t s = start + j, c = indices[s];
            if (out_row) out_row[p] = r;
            if (out_col) out_col[p] = c;
            if (out_eid) out_eid[p] = eids ? (int64_t)eids[s] : s;
            if (out_slot) out_slot[p] = s;
            if (MARK && atomicCAS(&mk.layer[c], -1, mk.hop + 1) == -1) {
              mk.next[atomicAdd(mk.next_count, 1)] = c;
              atomicOr(&mk.bitmap[c >> 5], 1u << (c & 31));
            }
            ++p;
            --need;
          }
        }
      }
      __syncwarp();
      continue;
    }
    const int mi = (int)m;
    uint32_t attempt[kSampleMaxDraw / kWarp] = {0, 0, 0, 0};
    for (int q = lane; q < mi; q += kWarp) pick[q] = sample_draw(rk, (uint32_t)q, 0u, deg);
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
        if (redo & (1u << t)) pick[q] = sample_draw(rk, (uint32_t)q, ++attempt[t], deg);
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
      for (int b = 0; b < mi; b += kWarp) {
        const bool act = b + lane < mi;
        emit(act, off + b + lane, start + (act ? srt[b + lane] : 0));
      }
    } else {
      // slot j survives unless dropped; it lands at j - (number of dropped slots before it)
      for (int64_t b = 0; b < deg; b += kWarp) {
        const int64_t j = b + lane;
        int lo = 0, hi = mi;                                 // first index with srt[idx] >= j
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (srt[mid] < j) lo = mid + 1; else hi = mid;
        }
        const bool act = j < deg && !(lo < mi && srt[lo] == j);
        emit(act, off + j - lo, start + j);
      }
    }
    __syncwarp();                                            // pick / srt are reused by the next row
  }
}

// ---- whole-subgraph pipeline: seeds, compaction of the discovered set, placement of the picks ------------

// Seeds (duplicates allowed, immutable_graph.cc:787-793): first occurrence enters layer 0 and the frontier.
__global__ void __launch_bounds__(256)
sample_seeds_kernel(int64_t n_seeds, const int64_t* __restrict__ seeds, int64_t n_rows, int32_t* layer,
                    uint32_t* bitmap, int64_t* frontier, int32_t* count, int32_t* bad) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_seeds) return;
  const int64_t v = seeds[i];
  if (v < 0 || v >= n_rows) {
    atomicExch(bad, 1);
    return;
  }
  if (atomicCAS(&layer[v], -1, 0) == -1) {
    frontier[atomicAdd(count, 1)] = v;
    atomicOr(&bitmap[v >> 5], 1u << (v & 31));
  }
}

__global__ void __launch_bounds__(256)
bitmap_popc_kernel(int64_t n_words, const uint32_t* __restrict__ bitmap, int32_t* __restrict__ cnt) {
  const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w < n_words) cnt[w] = __popc(bitmap[w]);
}

// Ascending vertex list out of the bitmap: word w holds vertices 32w .. 32w+31, its set bits land at
// word_prefix[w] ...  Also the layer of every vertex (sub_vers sorted by id, :876-895).
__global__ void __launch_bounds__(256)
bitmap_emit_kernel(int64_t n_words, const uint32_t* __restrict__ bitmap, const int32_t* __restrict__ word_prefix,
                   const int32_t* __restrict__ layer, int64_t* __restrict__ nodes, int64_t* __restrict__ layers) {
  const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= n_words) return;
  uint32_t bits = bitmap[w];
  int64_t p = word_prefix[w];
  while (bits) {
    const int b = __ffs((int)bits) - 1;
    bits &= bits - 1u;
    const int64_t v = w * 32 + b;
    nodes[p] = v;
    layers[p] = layer[v];
    ++p;
  }
}

// compacted id of vertex v (CompactSubgraph's id_map, :963-983): set bits before it
__device__ __forceinline__ int64_t sub_id(const uint32_t* bitmap, const int32_t* word_prefix, int64_t v) {
  return (int64_t)word_prefix[v >> 5] + __popc(bitmap[v >> 5] & ((1u << (v & 31)) - 1u));
}

// picks of every sampled row -> edge count of its compacted vertex (rows never sampled keep 0, :919-946)
__global__ void __launch_bounds__(256)
place_count_kernel(int64_t n_sel, const int32_t* __restrict__ n_sel_dev, const int64_t* __restrict__ rows,
                   const int64_t* __restrict__ offsets, const uint32_t* __restrict__ bitmap,
                   const int32_t* __restrict__ word_prefix, int64_t* __restrict__ sub_count) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_sel || (n_sel_dev != nullptr && i >= *n_sel_dev)) return;
  sub_count[sub_id(bitmap, word_prefix, rows[i])] = offsets[i + 1] - offsets[i];
}

// One thread per pick: binary search of its row in the hop's offsets, then copy to the row's final place in
// the subgraph's CSR with the column compacted.  The picks of a row keep their (ascending slot) order.
__global__ void __launch_bounds__(256)
place_picks_kernel(int64_t n_sel, int64_t total, const int64_t* __restrict__ rows,
                   const int64_t* __restrict__ offsets, const int64_t* __restrict__ cols,
                   const int64_t* __restrict__ eids, const uint32_t* __restrict__ bitmap,
                   const int32_t* __restrict__ word_prefix, const int64_t* __restrict__ sub_indptr,
                   int32_t* __restrict__ sub_indices, int64_t* __restrict__ sub_rows,
                   int64_t* __restrict__ induced_edges) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= total) return;
  int64_t lo = 0, hi = n_sel;            // last i with offsets[i] <= p
  while (hi - lo > 1) {
    const int64_t mid = (lo + hi) >> 1;
    if (offsets[mid] <= p) lo = mid; else hi = mid;
  }
  const int64_t row = sub_id(bitmap, word_prefix, rows[lo]);
  const int64_t q = sub_indptr[row] + (p - offsets[lo]);
  sub_indices[q] = (int32_t)sub_id(bitmap, word_prefix, cols[p]);
  sub_rows[q] = row;
  induced_edges[q] = eids[p];
}

// leave the persistent state clean for the next call: only what this call touched
__global__ void __launch_bounds__(256)
sample_reset_kernel(int64_t n_sub, const int64_t* __restrict__ nodes, int32_t* layer, uint32_t* bitmap) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_sub) return;
  const int64_t v = nodes[i];
  layer[v] = -1;
  bitmap[v >> 5] = 0u;
}

}  // namespace dglb200

using namespace dglb200;

extern "C" int dglb200_neighbor_sample_offsets(int64_t n_sel, const int64_t* rows, const int32_t* n_sel_dev,
                                               int64_t n_rows, const int64_t* indptr, int64_t fanout,
                                               int64_t* offsets, int64_t* total_host, int64_t* n_sel_host,
                                               void* workspace, size_t workspace_bytes, void* stream_) {
  clear_error();
  DGLB200_REQUIRE(n_sel >= 0 && n_rows >= 0 && total_host != nullptr, "neighbor_sample_offsets: bad args");
  DGLB200_REQUIRE(fanout >= 0, "neighbor_sample_offsets: fanout must be >= 0, got %lld", (long long)fanout);
  *total_host = 0;
  if (n_sel_host) *n_sel_host = n_sel_dev ? 0 : n_sel;
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
  sample_lens_kernel<<<(unsigned)ceil_div(n_sel, 256), 256, 0, stream>>>(n_sel, n_sel_dev, rows, n_rows, indptr,
                                                                         fanout, len, bad);
  DGLB200_POST_LAUNCH("sample_lens_kernel");
  if (exclusive_scan<int64_t, int64_t>(len, offsets, n_sel, 1, ws, stream) != 0) return -1;
  int32_t bad_h = 0, live_h = 0;
  DGLB200_CUDA(cudaMemcpyAsync(total_host, offsets + n_sel, sizeof(int64_t), cudaMemcpyDeviceToHost, stream));
  DGLB200_CUDA(cudaMemcpyAsync(&bad_h, bad, sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
  if (n_sel_dev) DGLB200_CUDA(cudaMemcpyAsync(&live_h, n_sel_dev, sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
  DGLB200_CUDA(cudaStreamSynchronize(stream));
  DGLB200_REQUIRE(bad_h == 0, "Invalid vertex: a seed id is outside [0, %lld)", (long long)n_rows);
  if (n_sel_host && n_sel_dev) *n_sel_host = live_h < n_sel ? live_h : n_sel;
  return 0;
}

extern "C" int dglb200_neighbor_sample_fill(int64_t n_sel, const int64_t* rows, const int32_t* n_sel_dev,
                                            const int64_t* offsets, const int64_t* indptr, const int32_t* indices,
                                            const int32_t* eids, int64_t fanout, uint64_t seed, int64_t* out_row,
                                            int64_t* out_col, int64_t* out_eid, int64_t* out_slot, int32_t* layer,
                                            uint32_t* bitmap, int32_t hop, int64_t* next_frontier,
                                            int32_t* next_count, void* stream_) {
  clear_error();
  DGLB200_REQUIRE(n_sel >= 0 && fanout >= 0, "neighbor_sample_fill: bad sizes");
  if (n_sel == 0) return 0;
  DGLB200_REQUIRE(rows && offsets && indptr, "neighbor_sample_fill: NULL pointer");
  DGLB200_REQUIRE((out_col == nullptr && layer == nullptr) || indices != nullptr,
                  "neighbor_sample_fill: out_col / discovery need indices");
  DGLB200_REQUIRE(layer == nullptr || (bitmap && next_frontier && next_count),
                  "neighbor_sample_fill: discovery needs layer, bitmap, next_frontier and next_count");
  int64_t blocks = ceil_div(n_sel, kSampleWarps);
  const int64_t cap = (int64_t)kNumSMsB200 * 16;            // 16 CTAs of 128 threads per SM
  if (blocks > cap) blocks = cap;
  SampleMark mk{layer, bitmap, next_frontier, next_count, hop};
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (layer != nullptr) {
    neighbor_sample_kernel<true><<<(unsigned)blocks, kSampleWarps * kWarp, 0, stream>>>(
        n_sel, n_sel_dev, rows, offsets, indptr, indices, eids, fanout, seed, out_row, out_col, out_eid, out_slot, mk);
  } else {
    neighbor_sample_kernel<false><<<(unsigned)blocks, kSampleWarps * kWarp, 0, stream>>>(
        n_sel, n_sel_dev, rows, offsets, indptr, indices, eids, fanout, seed, out_row, out_col, out_eid, out_slot, mk);
  }
  DGLB200_POST_LAUNCH("neighbor_sample_kernel");
  return 0;
}

static inline int64_t bitmap_words(int64_t n_rows) { return (n_rows + 31) / 32; }

// Whole call, host side: see include/dgl_b200.h.  Everything the call produces or needs lives in the caller's
// workspace; a workspace that turns out too small (the number of picks is only known hop by hop) ends the call
// with out->needed_bytes set, the persistent state wiped, and the caller retries with a larger one.
extern "C" int dglb200_sample_subgraph(int64_t n_rows, const int64_t* indptr, const int32_t* indices,
                                       const int32_t* eids, int64_t n_seeds, const int64_t* seeds, int64_t fanout,
                                       int32_t num_hops, uint64_t seed, int32_t* layer, uint32_t* bitmap,
                                       int32_t* word_prefix, void* workspace, size_t workspace_bytes,
                                       dglb200_sampled_subgraph_t* out, void* stream_) {
  clear_error();
  DGLB200_REQUIRE(out != nullptr, "sample_subgraph: out is NULL");
  memset(out, 0, sizeof(*out));
  DGLB200_REQUIRE(n_rows > 0 && n_rows < (1ll << 31) && indptr && indices && layer && bitmap && word_prefix,
                  "sample_subgraph: bad graph / state arguments");
  DGLB200_REQUIRE(n_seeds > 0 && seeds, "sample_subgraph: no seed vertices");
  DGLB200_REQUIRE(fanout >= 0 && num_hops >= 0 && num_hops <= 64, "sample_subgraph: fanout >= 0, 0 <= num_hops <= 64");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int64_t words = bitmap_words(n_rows);
  Workspace ws(workspace, workspace_bytes);
  bool dirty = false;
  // capacity failure: wipe the marks (O(n), rare), report the size that would have carried the call this far
  auto too_small = [&](size_t extra) -> int {
    if (dirty) {
      cudaMemsetAsync(layer, 0xFF, sizeof(int32_t) * (size_t)n_rows, stream);
      cudaMemsetAsync(bitmap, 0, sizeof(uint32_t) * (size_t)words, stream);
      cudaStreamSynchronize(stream);
    }
    out->needed_bytes = (ws.used + extra) * 2 + (64u << 10);
    set_error("sample_subgraph: workspace of %zu bytes is too small (needed_bytes = %zu)", workspace_bytes,
              out->needed_bytes);
    return -1;
  };
#define TAKE(var, T, count)                                   \
  T* var = ws.take<T>((size_t)(count));                       \
  if (var == nullptr) return too_small(0)

  const int n_counters = num_hops + 2;                        // [h]: size of hop h's frontier; last: bad-id flag
  TAKE(counters, int32_t, n_counters);
  TAKE(frontier0, int64_t, n_seeds);
  DGLB200_CUDA(cudaMemsetAsync(counters, 0, sizeof(int32_t) * n_counters, stream));
  dirty = true;
  sample_seeds_kernel<<<(unsigned)ceil_div(n_seeds, 256), 256, 0, stream>>>(
      n_seeds, seeds, n_rows, layer, bitmap, frontier0, counters, counters + n_counters - 1);
  DGLB200_POST_LAUNCH("sample_seeds_kernel");

  // n_sel / total: exact when the hop synchronised, else capacities (live counts: n_sel_dev, offsets[n_sel])
  struct Hop { int64_t n_sel, total; const int32_t* n_sel_dev; const int64_t *rows, *offsets, *cols, *pe; };
  Hop hops[64];
  int n_hops = 0;
  const int64_t* frontier = frontier0;
  int64_t cap = n_seeds, total_picks = 0;
  bool exact = true;                                           // every hop returned its sizes to the host
  const int64_t grid_cap = (int64_t)kNumSMsB200 * 16;
  for (int h = 0; h < num_hops && cap > 0; ++h) {
    TAKE(offsets, int64_t, cap + 1);
    const size_t mark0 = ws.used;
    if (ws.used + align_up((size_t)(cap + 1) * 8) + scan_workspace_bytes(cap, 8) > ws.size)
      return too_small(align_up((size_t)(cap + 1) * 8) + scan_workspace_bytes(cap, 8));
    int64_t* len = ws.take<int64_t>(cap + 1);
    sample_lens_kernel<<<(unsigned)ceil_div(cap, 256), 256, 0, stream>>>(cap, counters + h, frontier, n_rows, indptr,
                                                                         fanout, len, nullptr);
    DGLB200_POST_LAUNCH("sample_lens_kernel");
    if (exclusive_scan<int64_t, int64_t>(len, offsets, cap, 1, ws, stream) != 0) return -1;
    // A hop's picks are at most cap * fanout.  While that bound is small the outputs are sized by it and the
    // host does NOT wait for the exact number: the kernels read the live counts from device memory.
    const bool bounded = fanout <= kSampleBoundLimit && cap <= kSampleBoundLimit / (fanout > 0 ? fanout : 1);
    int64_t total = bounded ? cap * fanout : 0, n_sel = cap;
    const int32_t* n_sel_dev = counters + h;
    if (!bounded) {
      int32_t live = 0, bad = 0;
      DGLB200_CUDA(cudaMemcpyAsync(&total, offsets + cap, sizeof(int64_t), cudaMemcpyDeviceToHost, stream));
      DGLB200_CUDA(cudaMemcpyAsync(&live, counters + h, sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
      DGLB200_CUDA(cudaMemcpyAsync(&bad, counters + n_counters - 1, sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
      DGLB200_CUDA(cudaStreamSynchronize(stream));
      if (bad != 0) {
        (void)too_small(0);
        out->needed_bytes = 0;
        set_error("Invalid vertex: a seed id is outside [0, %lld)", (long long)n_rows);
        return -1;
      }
      ws.used = mark0;                                         // the stream is idle: lens + scan scratch is free again
      n_sel = live < cap ? live : cap;
      n_sel_dev = nullptr;
    } else {
      exact = false;
    }
    hops[n_hops] = Hop{n_sel, total, n_sel_dev, frontier, offsets, nullptr, nullptr};
    if (total > 0) {
      if (ws.used + 3 * align_up((size_t)total * 8) > ws.size) return too_small(3 * align_up((size_t)total * 8));
      TAKE(cols, int64_t, total);
      TAKE(pe, int64_t, total);
      TAKE(next, int64_t, total);
      hops[n_hops].cols = cols;
      hops[n_hops].pe = pe;
      int64_t blocks = ceil_div(n_sel, kSampleWarps);
      if (blocks > grid_cap) blocks = grid_cap;
      SampleMark mk{layer, bitmap, next, counters + h + 1, h};
      neighbor_sample_kernel<true><<<(unsigned)blocks, kSampleWarps * kWarp, 0, stream>>>(
          n_sel, n_sel_dev, frontier, offsets, indptr, indices, eids, fanout, seed, nullptr, cols, pe, nullptr, mk);
      DGLB200_POST_LAUNCH("neighbor_sample_kernel");
      frontier = next;
    }
    ++n_hops;
    total_picks += total;
    cap = total;                                               // the next frontier: at most one vertex per pick
  }
  // ---- compaction of the discovered set, then placement of the picks
  if (ws.used + align_up((size_t)words * 4) + scan_workspace_bytes(words, 4) > ws.size)
    return too_small(align_up((size_t)words * 4) + scan_workspace_bytes(words, 4));
  const size_t mark1 = ws.used;
  int32_t* cnt = ws.take<int32_t>(words);
  bitmap_popc_kernel<<<(unsigned)ceil_div(words, 256), 256, 0, stream>>>(words, bitmap, cnt);
  DGLB200_POST_LAUNCH("bitmap_popc_kernel");
  if (exclusive_scan<int32_t, int32_t>(cnt, word_prefix, words, 1, ws, stream) != 0) return -1;
  int32_t n_sub32 = 0, bad = 0;
  int64_t real_total[64];
  for (int h = 0; h < n_hops; ++h) {                           // bounded hops: their pick counts come back only now
    real_total[h] = hops[h].total;
    if (hops[h].n_sel_dev != nullptr)
      DGLB200_CUDA(cudaMemcpyAsync(&real_total[h], hops[h].offsets + hops[h].n_sel, sizeof(int64_t),
                                   cudaMemcpyDeviceToHost, stream));
  }
  DGLB200_CUDA(cudaMemcpyAsync(&n_sub32, word_prefix + words, sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
  DGLB200_CUDA(cudaMemcpyAsync(&bad, counters + n_counters - 1, sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
  DGLB200_CUDA(cudaStreamSynchronize(stream));
  if (bad != 0) {
    (void)too_small(0);
    out->needed_bytes = 0;
    set_error("Invalid vertex: a seed id is outside [0, %lld)", (long long)n_rows);
    return -1;
  }
  ws.used = mark1;
  const int64_t n_sub = n_sub32;
  (void)exact;
  total_picks = 0;
  for (int h = 0; h < n_hops; ++h) total_picks += real_total[h];
  {
    const size_t e8 = align_up((size_t)(total_picks > 0 ? total_picks : 1) * 8), n8 = align_up((size_t)(n_sub + 1) * 8);
    const size_t rest = 3 * e8 + 4 * n8 + scan_workspace_bytes(n_sub + 1, 8);
    if (ws.used + rest > ws.size) return too_small(rest);
  }
  TAKE(nodes, int64_t, n_sub);
  TAKE(layers, int64_t, n_sub);
  TAKE(sub_indptr, int64_t, n_sub + 1);
  TAKE(sub_indices, int32_t, total_picks > 0 ? total_picks : 1);
  TAKE(sub_rows, int64_t, total_picks > 0 ? total_picks : 1);
  TAKE(induced, int64_t, total_picks > 0 ? total_picks : 1);
  if (ws.used + align_up((size_t)(n_sub + 1) * 8) + scan_workspace_bytes(n_sub + 1, 8) > ws.size)
    return too_small(align_up((size_t)(n_sub + 1) * 8) + scan_workspace_bytes(n_sub + 1, 8));
  int64_t* sub_count = ws.take<int64_t>(n_sub + 1);
  bitmap_emit_kernel<<<(unsigned)ceil_div(words, 256), 256, 0, stream>>>(words, bitmap, word_prefix, layer, nodes, layers);
  DGLB200_POST_LAUNCH("bitmap_emit_kernel");
  DGLB200_CUDA(cudaMemsetAsync(sub_count, 0, sizeof(int64_t) * (size_t)(n_sub + 1), stream));
  for (int h = 0; h < n_hops; ++h) {
    if (hops[h].n_sel == 0) continue;
    place_count_kernel<<<(unsigned)ceil_div(hops[h].n_sel, 256), 256, 0, stream>>>(
        hops[h].n_sel, hops[h].n_sel_dev, hops[h].rows, hops[h].offsets, bitmap, word_prefix, sub_count);
    DGLB200_POST_LAUNCH("place_count_kernel");
  }
  if (exclusive_scan<int64_t, int64_t>(sub_count, sub_indptr, n_sub, 1, ws, stream) != 0) return -1;
  for (int h = 0; h < n_hops; ++h) {
    if (hops[h].n_sel == 0 || real_total[h] == 0) continue;
    place_picks_kernel<<<(unsigned)ceil_div(real_total[h], 256), 256, 0, stream>>>(
        hops[h].n_sel, real_total[h], hops[h].rows, hops[h].offsets, hops[h].cols, hops[h].pe, bitmap, word_prefix,
        sub_indptr, sub_indices, sub_rows, induced);
    DGLB200_POST_LAUNCH("place_picks_kernel");
  }
  if (n_sub > 0) {
    sample_reset_kernel<<<(unsigned)ceil_div(n_sub, 256), 256, 0, stream>>>(n_sub, nodes, layer, bitmap);
    DGLB200_POST_LAUNCH("sample_reset_kernel");
  }
#undef TAKE
  out->n_nodes = n_sub;
  out->n_edges = total_picks;
  out->nodes = nodes;
  out->layers = layers;
  out->sub_indptr = sub_indptr;
  out->sub_indices = sub_indices;
  out->sub_rows = sub_rows;
  out->induced_edges = induced;
  out->needed_bytes = ws.used;
  return 0;
}
