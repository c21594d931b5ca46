// Kernel family (4b): degree bucketing and sorted-unique relabelling on the device.
// Replaces sched::DegreeBucketing (src/scheduler/scheduler.cc:13-93) and the
// F.unique / F.sort_1d / build_relabel_map chain of the subset schedulers
// (python/dgl/runtime/scheduler.py:136,341; python/dgl/utils.py:321-361).
// Everything is sort + scan + compaction on uint32 keys (hand-written radix sort, scan.cuh);
// integer work, bit-exact with the canonical form of the reference's output.
#include <cmath>

#include "common.cuh"
#include "radix_sort.cuh"
#include "scan.cuh"

namespace dglb200 {

__global__ void to_u32_checked(const int64_t* __restrict__ in, int64_t n, int64_t bound,
                               uint32_t* __restrict__ out, int32_t* __restrict__ bad) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int64_t v = in[i];
  if (v < 0 || v >= bound) { atomicExch(bad, 1); out[i] = 0; return; }
  out[i] = (uint32_t)v;
}

__global__ void head_flags(const uint32_t* __restrict__ sorted, int64_t n, int32_t* __restrict__ flag) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) flag[i] = (i == 0 || sorted[i] != sorted[i - 1]) ? 1 : 0;
}

// For every run head j (flag set) with rank g = pos[j]: head_key[g] = sorted[j], head_start[g] = j.
// The last thread also closes head_start[count] = n.
__global__ void emit_heads(const uint32_t* __restrict__ sorted, const int32_t* __restrict__ flag,
                           const int32_t* __restrict__ pos, int64_t n,
                           uint32_t* __restrict__ head_key, int32_t* __restrict__ head_start) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (flag[i]) {
    head_key[pos[i]] = sorted[i];
    head_start[pos[i]] = (int32_t)i;
  }
  if (i == n - 1) head_start[pos[i] + flag[i]] = (int32_t)n;
}

__global__ void run_lengths(const int32_t* __restrict__ start, int64_t n_runs, uint32_t* __restrict__ len) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_runs) len[i] = (uint32_t)(start[i + 1] - start[i]);
}

// unique: uniq[g] (int64) and inverse[perm[j]] = g
__global__ void unique_emit(const uint32_t* __restrict__ sorted, const uint32_t* __restrict__ perm,
                            const int32_t* __restrict__ flag, const int32_t* __restrict__ pos,
                            int64_t n, int64_t* __restrict__ uniq, int64_t* __restrict__ inverse) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const int32_t g = pos[j] + flag[j] - 1;
  if (flag[j]) uniq[g] = (int64_t)sorted[j];
  if (inverse) inverse[perm[j]] = g;
}

// buckets: one entry per distinct degree
__global__ void emit_buckets(const uint32_t* __restrict__ bucket_deg, const int32_t* __restrict__ bucket_start,
                             int64_t n_buckets, int64_t* __restrict__ degs, int64_t* __restrict__ nid_section,
                             int64_t* __restrict__ mid_section) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n_buckets) return;
  const int64_t cnt = bucket_start[b + 1] - bucket_start[b];
  degs[b] = bucket_deg[b];
  nid_section[b] = cnt;
  mid_section[b] = (int64_t)bucket_deg[b] * cnt;
}

__global__ void emit_nids(const uint32_t* __restrict__ node_vid, const uint32_t* __restrict__ sorted_g,
                          int64_t n_nodes, int64_t* __restrict__ nids) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n_nodes) nids[k] = (int64_t)node_vid[sorted_g[k]];
}

// mids[q]: q-th message in bucket order.  Node k (in bucket order) owns [out_off[k], out_off[k+1]).
__global__ void emit_mids(const int64_t* __restrict__ out_off, const uint32_t* __restrict__ sorted_g,
                          const int32_t* __restrict__ node_start, const uint32_t* __restrict__ perm,
                          const int64_t* __restrict__ msg_ids, int64_t n_nodes, int64_t n_msgs,
                          int64_t* __restrict__ mids) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n_msgs) return;
  int64_t lo = 0, hi = n_nodes;  // last k with out_off[k] <= q
  while (hi - lo > 1) {
    const int64_t mid = (lo + hi) >> 1;
    if (out_off[mid] <= q) lo = mid; else hi = mid;
  }
  const int64_t t = q - out_off[lo];
  mids[q] = msg_ids[perm[node_start[sorted_g[lo]] + t]];
}

// zero-degree receivers: unique recv ids that are not a key of node_vid (ascending)
__global__ void flag_missing(const uint32_t* __restrict__ recv_u, int64_t n_recv_u,
                             const uint32_t* __restrict__ node_vid, int64_t n_nodes,
                             int32_t* __restrict__ flag) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_recv_u) return;
  const uint32_t key = recv_u[i];
  int64_t lo = 0, hi = n_nodes;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (node_vid[mid] < key) lo = mid + 1; else hi = mid;
  }
  flag[i] = (lo < n_nodes && node_vid[lo] == key) ? 0 : 1;
}

__global__ void compact_missing(const uint32_t* __restrict__ recv_u, const int32_t* __restrict__ flag,
                                const int32_t* __restrict__ pos, int64_t n, int64_t* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && flag[i]) out[pos[i]] = (int64_t)recv_u[i];
}

struct SortedRuns {
  uint32_t* sorted = nullptr;   // [n] sorted keys
  uint32_t* perm = nullptr;     // [n] original positions
  uint32_t* head_key = nullptr; // [n_runs] distinct keys ascending
  int32_t* head_start = nullptr;// [n_runs + 1]
  int32_t* flag = nullptr;
  int32_t* pos = nullptr;
  int64_t n_runs = 0;
};

// sort keys (stable), find runs of equal keys.  Syncs the stream (returns n_runs).
static int sorted_runs(const uint32_t* keys, int64_t n, int bits, Workspace& ws, cudaStream_t stream,
                       SortedRuns* r) {
  r->sorted = ws.take<uint32_t>(n);
  r->perm = ws.take<uint32_t>(n);
  uint32_t* kt = ws.take<uint32_t>(n);
  uint32_t* vt = ws.take<uint32_t>(n);
  r->flag = ws.take<int32_t>(n + 1);
  r->pos = ws.take<int32_t>(n + 1);
  r->head_key = ws.take<uint32_t>(n);
  r->head_start = ws.take<int32_t>(n + 1);
  DGLB200_REQUIRE(r->sorted && r->perm && kt && vt && r->flag && r->pos && r->head_key && r->head_start,
                  "bucketing: workspace too small");
  if (radix_sort_pairs(keys, nullptr, r->sorted, r->perm, kt, vt, n, bits, ws, stream) != 0) return -1;
  const unsigned blocks = (unsigned)ceil_div(n, 256);
  head_flags<<<blocks, 256, 0, stream>>>(r->sorted, n, r->flag);
  DGLB200_POST_LAUNCH("head_flags");
  if (exclusive_scan<int32_t, int32_t>(r->flag, r->pos, n, 1, ws, stream) != 0) return -1;
  emit_heads<<<blocks, 256, 0, stream>>>(r->sorted, r->flag, r->pos, n, r->head_key, r->head_start);
  DGLB200_POST_LAUNCH("emit_heads");
  int32_t cnt = 0;
  DGLB200_CUDA(cudaMemcpyAsync(&cnt, r->pos + n, sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
  DGLB200_CUDA(cudaStreamSynchronize(stream));
  r->n_runs = cnt;
  return 0;
}

static size_t sorted_runs_bytes(int64_t n) {
  return align_up((size_t)n * 4) * 5 + align_up((size_t)(n + 1) * 4) * 3 +
         radix_sort_workspace_bytes(n) + scan_workspace_bytes(n + 1, sizeof(int64_t)) + 4096;
}

static int convert_ids(const int64_t* ids, int64_t n, int64_t bound, uint32_t* out, int32_t* bad,
                       const char* what, cudaStream_t stream) {
  DGLB200_CUDA(cudaMemsetAsync(bad, 0, sizeof(int32_t), stream));
  to_u32_checked<<<(unsigned)ceil_div(n, 256), 256, 0, stream>>>(ids, n, bound, out, bad);
  DGLB200_POST_LAUNCH("to_u32_checked");
  int32_t h = 0;
  DGLB200_CUDA(cudaMemcpyAsync(&h, bad, sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
  DGLB200_CUDA(cudaStreamSynchronize(stream));
  DGLB200_REQUIRE(h == 0, "%s: id out of range [0, %lld)", what, (long long)bound);
  return 0;
}

}  // namespace dglb200

using namespace dglb200;

extern "C" {

int dglb200_unique_workspace_bytes(int64_t n, size_t* bytes_host) {
  DGLB200_REQUIRE(bytes_host != nullptr && n >= 0, "unique_workspace_bytes: bad args");
  *bytes_host = sorted_runs_bytes(n) + align_up((size_t)n * 4) + 1024;
  return 0;
}

int dglb200_unique_sorted(int64_t n, int64_t id_bound, const int64_t* ids, int64_t* uniq,
                          int64_t* inverse, int64_t* counts_host, void* workspace,
                          size_t workspace_bytes, void* stream_) {
  clear_error();
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  DGLB200_REQUIRE(counts_host != nullptr, "unique_sorted: counts_host is NULL");
  counts_host[0] = 0;
  DGLB200_REQUIRE(n >= 0 && n < (1ll << 31) && id_bound >= 0 && id_bound < (1ll << 31),
                  "unique_sorted: n / id_bound must be < 2^31");
  if (n == 0) return 0;
  DGLB200_REQUIRE(ids && uniq, "unique_sorted: NULL pointer");
  Workspace ws(workspace, workspace_bytes);
  uint32_t* keys = ws.take<uint32_t>(n);
  int32_t* bad = ws.take<int32_t>(4);
  DGLB200_REQUIRE(keys && bad, "unique_sorted: workspace too small");
  if (convert_ids(ids, n, id_bound, keys, bad, "unique_sorted", stream) != 0) return -1;
  SortedRuns r;
  if (sorted_runs(keys, n, bits_for(id_bound), ws, stream, &r) != 0) return -1;
  unique_emit<<<(unsigned)ceil_div(n, 256), 256, 0, stream>>>(r.sorted, r.perm, r.flag, r.pos, n, uniq, inverse);
  DGLB200_POST_LAUNCH("unique_emit");
  counts_host[0] = r.n_runs;
  return 0;
}

int64_t dglb200_bucket_capacity(int64_t n_msgs) {
  // B distinct non-zero degrees need at least B(B+1)/2 messages; +1 zero bucket, +1 slack
  return (int64_t)std::floor(std::sqrt(2.0 * (double)(n_msgs > 0 ? n_msgs : 0))) + 3;
}

int dglb200_degree_bucketing_workspace_bytes(int64_t n_msgs, int64_t n_recv, size_t* bytes_host) {
  DGLB200_REQUIRE(bytes_host != nullptr && n_msgs >= 0 && n_recv >= 0, "degree_bucketing_workspace_bytes: bad args");
  size_t b = 0;
  b += align_up((size_t)n_msgs * 4) + align_up((size_t)n_recv * 4) + 1024;  // converted keys
  b += sorted_runs_bytes(n_msgs) * 2;                                        // by vid, by degree
  b += sorted_runs_bytes(n_recv);
  b += align_up((size_t)(n_msgs + 1) * 8) * 2;                               // out_off + deg64
  b += align_up((size_t)(n_recv + 1) * 4) * 2;
  b += scan_workspace_bytes(n_msgs + n_recv + 2, sizeof(int64_t)) * 2;
  *bytes_host = b + 8192;
  return 0;
}

int dglb200_degree_bucketing(int64_t n_msgs, int64_t n_recv, int64_t id_bound,
                             const int64_t* msg_ids, const int64_t* vids, const int64_t* recv_ids,
                             int64_t* degs, int64_t* nids, int64_t* nid_section, int64_t* mids,
                             int64_t* mid_section, int64_t* counts_host, void* workspace,
                             size_t workspace_bytes, void* stream_) {
  clear_error();
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  DGLB200_REQUIRE(counts_host != nullptr, "degree_bucketing: counts_host is NULL");
  counts_host[0] = counts_host[1] = counts_host[2] = 0;
  DGLB200_REQUIRE(n_msgs >= 0 && n_recv >= 0 && n_msgs < (1ll << 31) && n_recv < (1ll << 31) &&
                  id_bound >= 0 && id_bound < (1ll << 31), "degree_bucketing: sizes must be < 2^31");
  DGLB200_REQUIRE(degs && nids && nid_section && (n_msgs == 0 || (mids && mid_section && msg_ids && vids)) &&
                  (n_recv == 0 || recv_ids), "degree_bucketing: NULL pointer");
  Workspace ws(workspace, workspace_bytes);
  int32_t* bad = ws.take<int32_t>(4);
  DGLB200_REQUIRE(bad != nullptr, "degree_bucketing: workspace too small");
  int64_t n_nodes = 0, n_buckets = 0;
  SortedRuns by_vid;
  if (n_msgs > 0) {
    uint32_t* vkeys = ws.take<uint32_t>(n_msgs);
    DGLB200_REQUIRE(vkeys != nullptr, "degree_bucketing: workspace too small");
    if (convert_ids(vids, n_msgs, id_bound, vkeys, bad, "degree_bucketing(vids)", stream) != 0) return -1;
    // messages grouped by destination, input order kept inside a destination
    if (sorted_runs(vkeys, n_msgs, bits_for(id_bound), ws, stream, &by_vid) != 0) return -1;
    n_nodes = by_vid.n_runs;
    // nodes sorted by degree; ascending vid inside a degree (stable, input already by vid)
    uint32_t* node_deg = ws.take<uint32_t>(n_nodes);
    DGLB200_REQUIRE(node_deg != nullptr, "degree_bucketing: workspace too small");
    const unsigned nb = (unsigned)ceil_div(n_nodes, 256);
    run_lengths<<<nb, 256, 0, stream>>>(by_vid.head_start, n_nodes, node_deg);
    DGLB200_POST_LAUNCH("run_lengths");
    SortedRuns by_deg;
    if (sorted_runs(node_deg, n_nodes, bits_for(n_msgs + 1), ws, stream, &by_deg) != 0) return -1;
    n_buckets = by_deg.n_runs;
    DGLB200_REQUIRE(n_buckets + 1 <= dglb200_bucket_capacity(n_msgs), "degree_bucketing: internal capacity error");
    emit_buckets<<<(unsigned)ceil_div(n_buckets, 256), 256, 0, stream>>>(by_deg.head_key, by_deg.head_start,
                                                                         n_buckets, degs, nid_section, mid_section);
    DGLB200_POST_LAUNCH("emit_buckets");
    emit_nids<<<nb, 256, 0, stream>>>(by_vid.head_key, by_deg.perm, n_nodes, nids);
    DGLB200_POST_LAUNCH("emit_nids");
    int64_t* out_off = ws.take<int64_t>(n_nodes + 1);
    DGLB200_REQUIRE(out_off != nullptr, "degree_bucketing: workspace too small");
    if (exclusive_scan<uint32_t, int64_t>(by_deg.sorted, out_off, n_nodes, 1, ws, stream) != 0) return -1;
    emit_mids<<<(unsigned)ceil_div(n_msgs, 256), 256, 0, stream>>>(out_off, by_deg.perm, by_vid.head_start,
                                                                   by_vid.perm, msg_ids, n_nodes, n_msgs, mids);
    DGLB200_POST_LAUNCH("emit_mids");
  }
  // zero-degree bucket: receivers that got no message
  int64_t n_zero = 0;
  if (n_recv > 0) {
    uint32_t* rkeys = ws.take<uint32_t>(n_recv);
    DGLB200_REQUIRE(rkeys != nullptr, "degree_bucketing: workspace too small");
    if (convert_ids(recv_ids, n_recv, id_bound, rkeys, bad, "degree_bucketing(recv_ids)", stream) != 0) return -1;
    SortedRuns rr;
    if (sorted_runs(rkeys, n_recv, bits_for(id_bound), ws, stream, &rr) != 0) return -1;
    const int64_t ru = rr.n_runs;
    int32_t* zflag = ws.take<int32_t>(ru + 1);
    int32_t* zpos = ws.take<int32_t>(ru + 1);
    DGLB200_REQUIRE(zflag && zpos, "degree_bucketing: workspace too small");
    flag_missing<<<(unsigned)ceil_div(ru, 256), 256, 0, stream>>>(rr.head_key, ru, by_vid.head_key, n_nodes, zflag);
    DGLB200_POST_LAUNCH("flag_missing");
    if (exclusive_scan<int32_t, int32_t>(zflag, zpos, ru, 1, ws, stream) != 0) return -1;
    compact_missing<<<(unsigned)ceil_div(ru, 256), 256, 0, stream>>>(rr.head_key, zflag, zpos, ru, nids + n_nodes);
    DGLB200_POST_LAUNCH("compact_missing");
    int32_t z = 0;
    DGLB200_CUDA(cudaMemcpyAsync(&z, zpos + ru, sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
    DGLB200_CUDA(cudaStreamSynchronize(stream));
    n_zero = z;
    if (n_zero > 0) {
      const int64_t tail[2] = {0, n_zero};
      DGLB200_CUDA(cudaMemcpyAsync(degs + n_buckets, &tail[0], sizeof(int64_t), cudaMemcpyHostToDevice, stream));
      DGLB200_CUDA(cudaMemcpyAsync(nid_section + n_buckets, &tail[1], sizeof(int64_t), cudaMemcpyHostToDevice, stream));
      DGLB200_CUDA(cudaStreamSynchronize(stream));  // `tail` lives on this stack frame
    }
  }
  counts_host[0] = n_buckets + (n_zero > 0 ? 1 : 0);
  counts_host[1] = n_nodes + n_zero;
  counts_host[2] = n_buckets;
  return 0;
}

}  // extern "C"
