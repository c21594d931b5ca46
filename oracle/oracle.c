/*
 * oracle.c - plain-C CPU restatement of the reference's builtin message-passing path.
 *
 * TEST INFRASTRUCTURE ONLY: the checker for tests/, __graft_entry__.smoke() and bench.py's CPU
 * legs.  The product (dgl-1_b200/) never links, loads or calls this file.
 *
 * Same algorithms as oracle/oracle.py (numpy), written as scalar loops so that sizes the numpy
 * version cannot finish in seconds (np.add.at over 10^7 edges) stay checkable.  Parity status:
 * PINNED through tests/test_oracle_golden.py (this file == oracle.py == golden vectors made by the
 * unmodified reference).  Paths below are relative to /root/reference.
 *
 * All ids are int64, features float32, like the reference (src/c_api_common.h:35-38).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* Graph::InDegrees, src/graph/graph.cc:412-424 (deg[v] = |in-edge list of v|). */
void oracle_in_degrees(int64_t n, int64_t e, const int64_t* dst, int64_t* deg) {
  memset(deg, 0, (size_t)n * sizeof(int64_t));
  for (int64_t i = 0; i < e; ++i) deg[dst[i]]++;
}

/*
 * Graph::GetAdj(transpose, "csr"), src/graph/graph.cc:525-549: the per-node adjacency lists
 * concatenated; a list holds its edges in insertion order = ascending edge id
 * (Graph::AddEdge, graph.cc:48-63).  key = dst (in-edge CSR) or src (transposed).
 */
void oracle_get_adj_csr(int64_t n, int64_t e, const int64_t* key, const int64_t* other,
                        int64_t* indptr, int64_t* indices, int64_t* eid) {
  memset(indptr, 0, (size_t)(n + 1) * sizeof(int64_t));
  for (int64_t i = 0; i < e; ++i) indptr[key[i] + 1]++;
  for (int64_t v = 0; v < n; ++v) indptr[v + 1] += indptr[v];
  int64_t* fill = (int64_t*)malloc((size_t)(n > 0 ? n : 1) * sizeof(int64_t));
  memcpy(fill, indptr, (size_t)n * sizeof(int64_t));
  for (int64_t i = 0; i < e; ++i) {
    const int64_t p = fill[key[i]]++;
    indices[p] = other[i];
    eid[p] = i;
  }
  free(fill);
}

/*
 * SPMVExecutor.run / SPMVWithDataExecutor.run -> th.sparse.mm on the uncoalesced COO built in
 * edge-id order (python/dgl/runtime/ir/executor.py:452-473,535-566;
 * python/dgl/backend/pytorch/tensor.py:45-51,143-146) and the builtins' UDF semantics
 * (python/dgl/function/message.py:75-91,128-135,173-180):
 *   out[dst_e, f] += msg(e, f), edges visited in ascending edge id, float32 accumulation.
 *   msg_op 0 copy_src: x[src_e, f]; 1 src_mul_edge: x[src_e, f] * w[e, f / (feat / w_width)];
 *   2 copy_edge: w[e, f] (w_width == feat).  Rows without edges stay 0.
 */
void oracle_spmm_sum(int64_t n_out, int64_t e, const int64_t* src, const int64_t* dst, int msg_op,
                     const float* x, const float* w, int64_t feat, int64_t w_width, float* out) {
  memset(out, 0, (size_t)n_out * (size_t)feat * sizeof(float));
  const int64_t group = (msg_op == 1) ? feat / w_width : 1;
  for (int64_t i = 0; i < e; ++i) {
    float* o = out + dst[i] * feat;
    if (msg_op == 0) {
      const float* xs = x + src[i] * feat;
      for (int64_t f = 0; f < feat; ++f) o[f] += xs[f];
    } else if (msg_op == 1) {
      const float* xs = x + src[i] * feat;
      const float* we = w + i * w_width;
      for (int64_t f = 0; f < feat; ++f) o[f] += xs[f] * we[f / group];
    } else {
      const float* we = w + i * feat;
      for (int64_t f = 0; f < feat; ++f) o[f] += we[f];
    }
  }
}

/*
 * fn.max through degree bucketing: F.max(mailbox, 1) (python/dgl/function/reducer.py:45-46;
 * python/dgl/backend/pytorch/tensor.py:95-97); rows without messages come from the zero
 * initializer (python/dgl/runtime/degree_bucketing.py:72-78; python/dgl/init.py:47-61).
 * arg = edge id of the first maximum in ascending edge order, -1 for empty rows (not exposed by
 * the reference; pinned by this restatement only).
 */
void oracle_spmm_max(int64_t n_out, int64_t e, const int64_t* src, const int64_t* dst, int msg_op,
                     const float* x, const float* w, int64_t feat, int64_t w_width, float* out,
                     int64_t* arg) {
  const int64_t total = n_out * feat;
  for (int64_t i = 0; i < total; ++i) { out[i] = 0.f; arg[i] = -1; }
  const int64_t group = (msg_op == 1) ? feat / w_width : 1;
  for (int64_t i = 0; i < e; ++i) {
    float* o = out + dst[i] * feat;
    int64_t* a = arg + dst[i] * feat;
    for (int64_t f = 0; f < feat; ++f) {
      float m;
      if (msg_op == 0) m = x[src[i] * feat + f];
      else if (msg_op == 1) m = x[src[i] * feat + f] * w[i * w_width + f / group];
      else m = w[i * feat + f];
      /* NaN propagates (torch.max): the first NaN message wins and stays */
      if (a[f] < 0 || m > o[f] || (m != m && o[f] == o[f])) { o[f] = m; a[f] = i; }
    }
  }
}

/* Backward of th.sparse.mm wrt the dense operand (autograd of tensor.py:146):
 * dx[src_e, f] += wgt * dy[dst_e, f], ascending edge id. */
void oracle_spmm_grad_x(int64_t n_in, int64_t e, const int64_t* src, const int64_t* dst,
                        const float* dy, const float* w, int64_t feat, int64_t w_width, float* dx) {
  memset(dx, 0, (size_t)n_in * (size_t)feat * sizeof(float));
  const int64_t group = w ? feat / w_width : 1;
  for (int64_t i = 0; i < e; ++i) {
    float* o = dx + src[i] * feat;
    const float* g = dy + dst[i] * feat;
    if (w) {
      const float* we = w + i * w_width;
      for (int64_t f = 0; f < feat; ++f) o[f] += g[f] * we[f / group];
    } else {
      for (int64_t f = 0; f < feat; ++f) o[f] += g[f];
    }
  }
}

/*
 * sched::DegreeBucketing (src/scheduler/scheduler.cc:13-93) in CANONICAL order (the reference
 * iterates std::unordered_map, so bucket / node order is implementation-defined): non-zero
 * buckets by ascending degree, nodes ascending inside a bucket, per-node message ids in input
 * order (scheduler.cc:22-25), then the zero-degree bucket of recv ids without messages (:33-39,
 * :79-85).  id_bound > every id.  Output arrays need capacities: degs/nid_section/mid_section
 * id_bound+1, nids n_msgs+n_recv, mids n_msgs.  counts = {B, K, B'}.
 */
void oracle_degree_bucketing(int64_t n_msgs, int64_t n_recv, int64_t id_bound,
                             const int64_t* msg_ids, const int64_t* vids, const int64_t* recv_ids,
                             int64_t* degs, int64_t* nids, int64_t* nid_section, int64_t* mids,
                             int64_t* mid_section, int64_t* counts) {
  int64_t* deg = (int64_t*)calloc((size_t)id_bound + 1, sizeof(int64_t));
  int64_t* ptr = (int64_t*)calloc((size_t)id_bound + 2, sizeof(int64_t));
  for (int64_t i = 0; i < n_msgs; ++i) deg[vids[i]]++;
  for (int64_t v = 0; v < id_bound; ++v) ptr[v + 1] = ptr[v] + deg[v];
  /* per-node message lists in input order */
  int64_t* node_msgs = (int64_t*)malloc((size_t)(n_msgs > 0 ? n_msgs : 1) * sizeof(int64_t));
  int64_t* fill = (int64_t*)malloc((size_t)(id_bound + 1) * sizeof(int64_t));
  memcpy(fill, ptr, (size_t)(id_bound + 1) * sizeof(int64_t));
  for (int64_t i = 0; i < n_msgs; ++i) node_msgs[fill[vids[i]]++] = msg_ids[i];
  /* histogram of degrees -> buckets in ascending degree */
  int64_t max_deg = 0;
  for (int64_t v = 0; v < id_bound; ++v) if (deg[v] > max_deg) max_deg = deg[v];
  int64_t* hist = (int64_t*)calloc((size_t)max_deg + 2, sizeof(int64_t));
  for (int64_t v = 0; v < id_bound; ++v) if (deg[v] > 0) hist[deg[v]]++;
  int64_t B = 0, K = 0, M = 0, Bnz = 0;
  for (int64_t d = 1; d <= max_deg; ++d) {
    if (hist[d] == 0) continue;
    degs[B] = d;
    nid_section[B] = hist[d];
    mid_section[Bnz++] = d * hist[d];
    ++B;
    for (int64_t v = 0; v < id_bound; ++v) {   /* small cases only: O(B * id_bound) */
      if (deg[v] != d) continue;
      nids[K++] = v;
      for (int64_t k = ptr[v]; k < ptr[v + 1]; ++k) mids[M++] = node_msgs[k];
    }
  }
  /* zero-degree bucket: unique recv ids with no message, ascending */
  char* seen = (char*)calloc((size_t)id_bound + 1, 1);
  for (int64_t i = 0; i < n_recv; ++i) if (deg[recv_ids[i]] == 0) seen[recv_ids[i]] = 1;
  int64_t nz = 0;
  for (int64_t v = 0; v < id_bound; ++v) if (seen[v]) { nids[K++] = v; ++nz; }
  if (nz > 0) { degs[B] = 0; nid_section[B] = nz; ++B; }
  counts[0] = B; counts[1] = K; counts[2] = Bnz;
  free(seen); free(hist); free(fill); free(node_msgs); free(ptr); free(deg);
}

/* Per-destination numerically stable softmax of (E, H) edge scores; the GAT example's
 * un-normalised variant is in oracle.py (examples/pytorch/gat/train.py:61-97). */
void oracle_edge_softmax(int64_t n, int64_t e, const int64_t* dst, int64_t heads,
                         const float* score, float* alpha) {
  float* mx = (float*)malloc((size_t)n * heads * sizeof(float));
  float* den = (float*)calloc((size_t)n * heads, sizeof(float));
  for (int64_t i = 0; i < n * heads; ++i) mx[i] = -INFINITY;
  for (int64_t i = 0; i < e; ++i)
    for (int64_t h = 0; h < heads; ++h)
      if (score[i * heads + h] > mx[dst[i] * heads + h]) mx[dst[i] * heads + h] = score[i * heads + h];
  for (int64_t i = 0; i < e; ++i)
    for (int64_t h = 0; h < heads; ++h) {
      const float x = expf(score[i * heads + h] - mx[dst[i] * heads + h]);
      alpha[i * heads + h] = x;
      den[dst[i] * heads + h] += x;
    }
  for (int64_t i = 0; i < e; ++i)
    for (int64_t h = 0; h < heads; ++h) alpha[i * heads + h] /= den[dst[i] * heads + h];
  free(mx); free(den);
}
