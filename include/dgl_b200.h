/*
 * dgl_b200.h - C ABI of libdgl_b200.so, the B200 (sm_100a) message-passing engine that
 * sits behind DGL 0.1.x's builtin-function path (update_all / send_and_recv / pull with
 * dgl.function.{copy_src, src_mul_edge, copy_edge} -> {sum, max}).
 *
 * Conventions (they follow the reference's C API conventions, src/c_api_common.h:18-53 and
 * include/dgl/runtime/c_runtime_api.h:157-515 of GaiYu0/dgl-1):
 *   - every entry point returns 0 on success and -1 on failure; the message is then
 *     available from dglb200_last_error() (thread-local, like DGLGetLastError,
 *     src/runtime/c_runtime_api.cc:130-148).  Nothing aborts, nothing throws across the ABI.
 *   - plain pointers and sizes only.  All data pointers are DEVICE pointers on the current
 *     CUDA device unless the parameter name ends in _host.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  Kernels are
 *     launched on it and nothing synchronises unless documented ("syncs").
 *   - node / edge ids that cross the reference's surface are int64 (c_api_common.h:35-38,
 *     IsValidIdArray).  Engine-internal structure arrays use int32 column / edge ids
 *     (all configs have N, E < 2^31) and int64 row pointers.
 *   - outputs are caller-allocated (the Python host hands in torch tensors so torch's caching
 *     allocator and stream semantics apply); the library only uses caller-provided workspace
 *     whose size comes from the matching *_workspace_bytes query.
 *
 * Feature layout: node features x are row-major [n, feat] with the trailing dims flattened
 * exactly like SPMVExecutor.run does (python/dgl/runtime/ir/executor.py:452-473).
 */
#ifndef DGL_B200_H_
#define DGL_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DGLB200_ABI_VERSION 1

/* message builtins: python/dgl/function/message.py:47 (src_mul_edge), :102 (copy_src), :145 (copy_edge) */
enum { DGLB200_MSG_COPY_SRC = 0, DGLB200_MSG_SRC_MUL_EDGE = 1, DGLB200_MSG_COPY_EDGE = 2 };
/* reduce builtins: python/dgl/function/reducer.py:52 (sum), :75 (max) */
enum { DGLB200_RED_SUM = 0, DGLB200_RED_MAX = 1 };
/* storage type of features / edge data; accumulation is always fp32 */
enum { DGLB200_F32 = 0, DGLB200_BF16 = 1 };
/* u-op-v edge ops of the SDDMM */
enum {
  DGLB200_SDDMM_DOT = 0, DGLB200_SDDMM_ADD = 1, DGLB200_SDDMM_MUL = 2,
  DGLB200_SDDMM_COPY_U = 3, /* out[e] = u[src]: READ_ROW(nf, u) of _gen_send, runtime/scheduler.py:572-583 */
  DGLB200_SDDMM_COPY_V = 4  /* out[e] = v[dst]: READ_ROW(nf, v); also d(loss)/d(w) of copy_edge->sum */
};
/* edge-softmax flavours of the fused GAT kernel */
enum {
  DGLB200_SOFTMAX_STABLE = 0,  /* exp(s - rowmax) / sum */
  DGLB200_SOFTMAX_GAT_EXAMPLE = 1 /* clamp(exp(s), -10, 10) / sum, as examples/pytorch/gat/train.py:74-97 */
};

/* ---- library ------------------------------------------------------------------------- */

/* ABI version of the loaded library (== DGLB200_ABI_VERSION). */
int dglb200_abi_version(void);
/* Last error message of the calling thread ("" if none).  Mirrors DGLGetLastError. */
const char* dglb200_last_error(void);
/* Number of kernels this library has launched since load (all threads).  bench.py reports it. */
int64_t dglb200_launch_count(void);
/* Compute capability of the current device as major*10+minor; -1 + error if no CUDA device. */
int dglb200_device_cc(void);

/*
 * Kernel-selection override for benchmarking sweeps (cf. the env-var style knobs of the reference,
 * docs/source/env_var.rst).  Keys: "spmm_ucap" (edges per pipeline stage: 1,2,4,8),
 * "spmm_slab_chunks" (16-byte chunks of a feature row per pass: 16,32,64), "spmm_threads"
 * (CTA size: 128,256,512); value 0 restores the automatic choice.  Only the fp32 vectorised
 * copy_src->sum kernel honours them.  "sddmm_dot_generic" = 1 keeps the SDDMM DOT on its generic slot-group
 * kernel; "gat_head_slab" = heads per pass of the fused GAT kernel (0 = one pass; slabs measured slower,
 * see csrc/sddmm.cu); "spmm_head_pass" = 1 / 2: per-head edge weights as scalar-weight passes of one head / as
 * chunk weights over the usual slabs (0 = automatic by head width).  Results never depend on the knobs.
 */
int dglb200_set_tuning(const char* key, int64_t value);

/* ---- (4) graph structure: in-edge CSR / out-edge CSR build --------------------------- */

/*
 * Replaces Graph::GetAdj(transpose, "csr") (src/graph/graph.cc:525-549) and
 * ImmutableGraph::CSR::FromEdges (src/graph/immutable_graph.cc:206-237).
 *
 * Builds the CSR of a COO edge list keyed by `row` (pass dst for the in-edge CSR that
 * aggregation uses, src for the out-edge CSR that its gradient uses).
 *   order == 0: within a row, edges keep ascending position in the input (== ascending edge id
 *               when eid == NULL): the mutable Graph's adjacency-list order (graph.cc:540-547).
 *   order == 1: within a row, sorted by `col`, ties by input position: the immutable graph's
 *               layout (immutable_graph.cc:211-217; std::sort leaves ties unspecified).
 * row/col/eid are int64 device arrays of length nnz (eid may be NULL = identity).
 * Outputs: indptr int64[n_rows+1], indices int32[nnz] (= col), eids int32[nnz].
 * Errors if any id is out of range or n_rows/n_cols/nnz >= 2^31.  Syncs the stream once
 * (range check).
 */
int dglb200_csr_build_workspace_bytes(int64_t n_rows, int64_t nnz, size_t* bytes_host);
int dglb200_csr_build(int64_t n_rows, int64_t n_cols, int64_t nnz,
                      const int64_t* row, const int64_t* col, const int64_t* eid, int order,
                      int64_t* indptr, int32_t* indices, int32_t* eids,
                      void* workspace, size_t workspace_bytes, void* stream);

/* Widen an int32 engine array to the reference's int64 surface (GetAdj returns int64). */
int dglb200_widen_i32_to_i64(int64_t n, const int32_t* in, int64_t* out, void* stream);
/* In-degrees from a row pointer: deg[i] = indptr[i+1]-indptr[i] (Graph::InDegrees, graph.cc:412-424). */
int dglb200_degrees_from_indptr(int64_t n_rows, const int64_t* indptr, int64_t* deg, void* stream);
/* COO of a CSR: row id of every slot (ImmutableGraph::GetAdj "coo", immutable_graph.cc:559-573). */
int dglb200_csr_row_ids(int64_t n_rows, int64_t nnz, const int64_t* indptr, int64_t* row_of_slot,
                        void* stream);

/*
 * Edge lists of a SET of rows of a cached CSR, on the device: Graph::InEdges(IdArray) / OutEdges(IdArray)
 * (src/graph/graph.cc:294-365; PackedFuncs _CAPI_DGLGraphInEdges_2 / OutEdges_2,
 * src/graph/graph_apis.cc:284-306) and the slot lists pull / push / send_and_recv build their
 * sub-adjacency from (python/dgl/runtime/spmv.py:154-227).  Two steps because the output size is data
 * dependent:
 *   csr_select_offsets: offsets[i] = number of slots of rows[0..i) (offsets has n_sel + 1 entries),
 *     *total_host = offsets[n_sel].  An id outside [0, n_rows) is an error ("Invalid vertex", like the
 *     reference's CHECK, graph.cc:299).  Syncs the stream (returns the total).
 *   csr_select_expand: for output position p in [0, total), owned by selection i (offsets[i] <= p):
 *     which[p] = i, slot[p] = indptr[rows[i]] + (p - offsets[i]), out_row[p] = rows[i],
 *     out_col[p] = indices[slot], out_eid[p] = eids[slot] (slot itself when eids == NULL).
 *     Every output is nullable.  Per row the slots keep the CSR's order (ascending edge id on an
 *     order-0 CSR = the reference's adjacency-list order).  Does not sync.
 */
int dglb200_csr_select_workspace_bytes(int64_t n_sel, size_t* bytes_host);
int dglb200_csr_select_offsets(int64_t n_sel, const int64_t* rows, int64_t n_rows, const int64_t* indptr,
                               int64_t* offsets, int64_t* total_host,
                               void* workspace, size_t workspace_bytes, void* stream);
int dglb200_csr_select_expand(int64_t n_sel, int64_t total, const int64_t* rows, const int64_t* offsets,
                              const int64_t* indptr, const int32_t* indices, const int32_t* eids,
                              int64_t* which, int64_t* slot, int64_t* out_row, int64_t* out_col,
                              int64_t* out_eid, void* stream);
/* deg[i] = indptr[rows[i]+1] - indptr[rows[i]] (Graph::InDegrees(IdArray), graph.cc:412-424); an id
 * outside [0, n_rows) sets *bad_flag (nullable) and yields 0.  Does not sync. */
int dglb200_degrees_of(int64_t n_sel, const int64_t* rows, int64_t n_rows, const int64_t* indptr,
                       int64_t* deg, int32_t* bad_flag, void* stream);

/*
 * One hop of uniform neighbour sampling over a CSR - replaces the loop body of
 * ImmutableGraph::SampleSubgraph with GetUniformSample (src/graph/immutable_graph.cc:701-741, 772-874;
 * reached from NeighborSampler, python/dgl/contrib/sampling/sampler.py:196, through
 * graph_index._CAPI_DGLGraphUniformSampling*, src/graph/graph_apis.cc:434-472).  For every selected row,
 * min(degree, fanout) DISTINCT slots, every subset of that size equally likely, emitted in ascending
 * slot order (the reference sorts its picks, :720/:727).  Rows with degree <= fanout are copied whole.
 * Two steps because the output size is data dependent:
 *   neighbor_sample_offsets: offsets[i] = sum of min(degree, fanout) over rows[0..i); *total_host =
 *     offsets[n_sel].  Workspace: dglb200_csr_select_workspace_bytes(n_sel).  An id outside [0, n_rows) is
 *     an error.  n_sel_dev (nullable): the number of live rows is on the device (a frontier appended by the
 *     previous hop's fill); rows beyond it count 0, and *n_sel_host (nullable) receives min(n_sel, live).
 *     Syncs the stream (returns the totals).
 *   neighbor_sample_fill: for the j-th pick (ascending slot s) of selection i, at p = offsets[i] + j:
 *     out_row[p] = rows[i], out_col[p] = indices[s], out_eid[p] = eids[s] (s when eids == NULL),
 *     out_slot[p] = s.  Every output is nullable.  Does not sync.
 *     Discovery (layer != NULL; the hash set + queue of :783-793, 852-860 for a whole layer at once): a
 *     picked column c with layer[c] == -1 gets layer[c] = hop + 1, its bit in `bitmap`, and is appended to
 *     next_frontier[(*next_count)++] (order unspecified; the next hop's samples do not depend on it).
 * The draws come from a counter-based generator keyed by (seed, row id, draw, attempt): the same seed gives
 * the same sample whatever the launch geometry.  The reference seeds rand_r with time(nullptr) (:778), so
 * its samples are not reproducible and parity is pinned on what IS deterministic (degree <= fanout: whole
 * rows, golden vectors of the unmodified reference) plus the distribution (tests).
 */
int dglb200_neighbor_sample_offsets(int64_t n_sel, const int64_t* rows, const int32_t* n_sel_dev, int64_t n_rows,
                                    const int64_t* indptr, int64_t fanout, int64_t* offsets, int64_t* total_host,
                                    int64_t* n_sel_host, void* workspace, size_t workspace_bytes, void* stream);
int dglb200_neighbor_sample_fill(int64_t n_sel, const int64_t* rows, const int32_t* n_sel_dev,
                                 const int64_t* offsets, const int64_t* indptr, const int32_t* indices,
                                 const int32_t* eids, int64_t fanout, uint64_t seed, int64_t* out_row,
                                 int64_t* out_col, int64_t* out_eid, int64_t* out_slot, int32_t* layer,
                                 uint32_t* bitmap, int32_t hop, int64_t* next_frontier, int32_t* next_count,
                                 void* stream);
/*
 * ImmutableGraph::NeighborUniformSample = SampleSubgraph + CompactSubgraph (:772-1008) in one call: the per-hop
 * pair above for every layer of the BFS, then the compaction.  (indptr, indices, eids): the parent's CSR keyed
 * by the vertex whose neighbours are sampled (the immutable graph's in-CSR for neighbor_type "in").
 * Persistent state per parent graph, owned by the caller: layer int32[n_rows] (all -1), bitmap
 * uint32[ceil(n_rows / 32)] (all 0), word_prefix int32[ceil(n_rows / 32) + 1] (scratch).  A call leaves layer /
 * bitmap clean again by resetting only what it touched, so it costs O(subgraph) - apart from the bitmap's
 * popcount scan, n_rows / 32 words.  One call at a time per state.
 * Results are placed IN the caller's workspace (device pointers in *out_host, valid while the workspace is):
 *   nodes[n_nodes] ascending, layers[n_nodes] (hop of first discovery; seeds = 0),
 *   sub_indptr[n_nodes + 1], sub_indices[n_edges] (compacted neighbour), sub_rows[n_edges] (compacted row
 *   vertex), induced_edges[n_edges] (parent edge id): the subgraph's CSR, rows ascending, the picks of a row in
 *   ascending slot order (:919-946); its own edge ids are 0 .. n_edges-1 (:948-949).
 * The number of picks is only known hop by hop: when the workspace is too small the call fails with
 * out_host->needed_bytes > 0 (a size that would have carried it that far, plus slack), the state wiped; retry
 * with a larger workspace.  On success needed_bytes = the bytes used.  Syncs the stream once at the end (the
 * sizes come back to the host), plus once per hop whose bound frontier x fanout exceeds 2^21 picks (such a hop
 * waits for its exact pick count instead of over-allocating).
 */
typedef struct {
  int64_t n_nodes, n_edges;
  int64_t* nodes;
  int64_t* layers;
  int64_t* sub_indptr;
  int32_t* sub_indices;
  int64_t* sub_rows;
  int64_t* induced_edges;
  size_t needed_bytes;
} dglb200_sampled_subgraph_t;
int dglb200_sample_subgraph(int64_t n_rows, const int64_t* indptr, const int32_t* indices, const int32_t* eids,
                            int64_t n_seeds, const int64_t* seeds, int64_t fanout, int32_t num_hops, uint64_t seed,
                            int32_t* layer, uint32_t* bitmap, int32_t* word_prefix, void* workspace,
                            size_t workspace_bytes, dglb200_sampled_subgraph_t* out_host, void* stream);

/* ---- (1)(2) SpMM: generalised message -> reduce over a CSR -------------------------- */

/*
 * Replaces SPMVExecutor.run / SPMVWithDataExecutor.run (python/dgl/runtime/ir/executor.py:
 * 452-473, 535-566) -> F.spmm -> th.sparse.mm (python/dgl/backend/pytorch/tensor.py:143-146)
 * and, for max / copy_edge / non-scalar weights, the degree-bucketing and e2v fallbacks
 * (python/dgl/runtime/scheduler.py:540-570) with the semantics of the builtins' __call__
 * (function/message.py:75-91,128-135,173-180; function/reducer.py:45-46).
 *
 *   y[r, f] = reduce_{p in [indptr[r], indptr[r+1])} msg(p, f)
 *     COPY_SRC      msg = x[indices[p], f]
 *     SRC_MUL_EDGE  msg = x[indices[p], f] * w[e(p), f / (feat / w_width)]   (w_width | feat)
 *     COPY_EDGE     msg = w[e(p), f]                                          (w_width == feat)
 *   e(p) = eids[p], or p itself when eids == NULL.
 *   SUM: rows are accumulated in fp32 in slot order (ascending edge id on an order-0 CSR);
 *        rows longer than the plan's chunk are summed chunk by chunk in a fixed order, so
 *        results are run-to-run deterministic.  Empty rows give 0.
 *   MAX: y = max over the row; empty rows give 0 and arg = -1 (the bucketing path fills
 *        zero-degree nodes from the zero initializer, runtime/degree_bucketing.py:72-78).
 *        If arg != NULL, arg[r, f] = e(p) of the winning slot, ties to the smallest slot.
 * The transpose product for the gradient (kernel 2) is this same call on the CSR keyed by src.
 *
 * `plan` (nullable) is the long-row split table made by dglb200_spmm_plan for this indptr;
 * without it every row is reduced by a single warp group (correct, slower on hub rows).
 * x_rows bounds-checks nothing at run time; it sizes the L2-residency slab choice: when the
 * gathered operand (x_rows * feat) exceeds about half of the L2 and rows are re-used (mean degree
 * >= 8) the feature dimension is processed in slabs narrow enough for one slab of every source
 * row to stay L2-resident (results do not depend on the slab width).
 */
typedef struct dglb200_spmm_plan_t {
  int32_t chunk;              /* edges per work item for long rows (multiple of 32) */
  int32_t n_long_rows;        /* rows with degree > chunk */
  int64_t n_chunks;           /* total work items over long rows */
  const int32_t* long_rows;   /* [n_long_rows] row ids (device) */
  const int64_t* chunk_ptr;   /* [n_long_rows+1] first work item of each long row (device) */
  void* partial;              /* [n_chunks, feat] fp32 scratch (device), >= partial_bytes */
  int32_t* partial_arg;       /* [n_chunks, feat] int32 scratch for MAX with arg (nullable) */
  size_t partial_bytes;
} dglb200_spmm_plan_t;

/*
 * Scan indptr for rows longer than `chunk` and emit the split table.  long_rows/chunk_ptr must
 * hold `capacity` (+1) entries; counts_host[0] = n_long_rows, counts_host[1] = n_chunks.
 * Rows are listed in ascending id.  Syncs the stream (returns counts to the host).
 */
int dglb200_spmm_plan(int64_t n_rows, const int64_t* indptr, int32_t chunk, int64_t capacity,
                      int32_t* long_rows, int64_t* chunk_ptr, int64_t* counts_host,
                      void* workspace, size_t workspace_bytes, void* stream);
int dglb200_spmm_plan_workspace_bytes(int64_t n_rows, size_t* bytes_host);

int dglb200_spmm_csr(int64_t n_rows, int64_t x_rows, int64_t nnz,
                     const int64_t* indptr, const int32_t* indices, const int32_t* eids,
                     int msg_op, int reduce_op, int dtype,
                     const void* x, int64_t feat,
                     const void* w, int64_t w_width,
                     void* y, int32_t* arg,
                     const dglb200_spmm_plan_t* plan, void* stream);

/*
 * Same, for operands that are column blocks of wider row-major matrices: x rows are x_ld elements
 * apart, y rows y_ld (>= feat).  Used to pipeline a feature-slab exchange with the SpMM over that
 * slab (dgl_b200/dist.py); arg, w and the plan's partial buffers stay dense in feat.
 */
int dglb200_spmm_csr_strided(int64_t n_rows, int64_t x_rows, int64_t nnz,
                             const int64_t* indptr, const int32_t* indices, const int32_t* eids,
                             int msg_op, int reduce_op, int dtype,
                             const void* x, int64_t x_ld, int64_t feat,
                             const void* w, int64_t w_width,
                             void* y, int64_t y_ld, int32_t* arg,
                             const dglb200_spmm_plan_t* plan, void* stream);

/*
 * Extended form for the multi-GPU row partition (dgl_b200/dist.py; no counterpart in the reference):
 *   accumulate != 0  SUM only: y += result.  A rank sums the column block of its OWN source rows first
 *                    (no exchange needed) and adds the remote column blocks as their rows arrive.
 *   x_halo != NULL   fp32 copy_src -> sum only: column ids >= halo_first are rows of a SECOND operand kept
 *                    in bfloat16 (row id - halo_first, dense [*, feat], feat a multiple of 8): the halo
 *                    rows stay in the precision they crossed NVLink in (DGL_B200_WIRE=bf16), accumulation
 *                    is fp32; results within the bf16 bar (1e-2), own rows exact.
 *   row_scale / bias / relu   SUM only: the layer epilogue y = relu(row_scale[r] * sum + bias[f]) applied in
 *                    the store (each part optional; with accumulate it is applied to the accumulated value,
 *                    so only the LAST of several accumulating calls may carry it).
 * ext == NULL is dglb200_spmm_csr_strided.
 */
typedef struct dglb200_spmm_ext_t {
  int32_t accumulate;
  int32_t relu;            /* SUM epilogue: y = max(., 0) last */
  const void* x_halo;
  int64_t halo_first;
  const float* row_scale;  /* SUM epilogue: y[r, :] = row_scale[r] * sum (nullable) ... */
  const float* bias;       /* ... + bias[f] (nullable): the fused GCN layer, examples/pytorch/gcn/gcn_spmv.py:55-60 */
} dglb200_spmm_ext_t;
int dglb200_spmm_csr_ext(int64_t n_rows, int64_t x_rows, int64_t nnz,
                         const int64_t* indptr, const int32_t* indices, const int32_t* eids,
                         int msg_op, int reduce_op, int dtype,
                         const void* x, int64_t x_ld, int64_t feat,
                         const void* w, int64_t w_width,
                         void* y, int64_t y_ld, int32_t* arg,
                         const dglb200_spmm_plan_t* plan, const dglb200_spmm_ext_t* ext, void* stream);

/*
 * Allocation-free variant of dglb200_spmm_csr.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  Same semantics as dglb200_spmm_csr with arg == NULL, but
 * the fp32 scratch of the long-row slices (plan->partial) is NOT taken from the caller: it lives in a
 * file-scope `static __device__` array of DGLB200_STATIC_SCRATCH_BYTES inside the library (every GPU that
 * loads the library holds it until exit).  plan->long_rows / chunk_ptr / chunk / n_long_rows / n_chunks are
 * the caller's tables as made by dglb200_spmm_plan; plan->partial, partial_arg and partial_bytes are ignored.
 * The call performs no cudaMalloc / cudaFree and calls no allocating library routine; it does not sync.
 *   - n_chunks * feat * 4 bytes must be <= DGLB200_STATIC_SCRATCH_BYTES, else -1 with a message that
 *     states both sizes (use dglb200_spmm_csr with your own plan->partial then).
 *   - ALL calls on a GPU share the one array, which keeps the previous call's contents (nothing has to start
 *     at zero: every slice is fully written before it is read).  Calls that OVERLAP on the GPU (different
 *     streams, or graph branches that may run concurrently) race on it; calls on one stream are safe.
 *   - MAX with an argmax output is not offered (it would need a second, int32 scratch).
 */
#define DGLB200_STATIC_SCRATCH_BYTES (4u << 20)
int dglb200_spmm_csr_static_scratch(int64_t n_rows, int64_t x_rows, int64_t nnz,
                                    const int64_t* indptr, const int32_t* indices, const int32_t* eids,
                                    int msg_op, int reduce_op, int dtype,
                                    const void* x, int64_t feat,
                                    const void* w, int64_t w_width,
                                    void* y, const dglb200_spmm_plan_t* plan, void* stream);

/*
 * Narrow edge data (a scalar or one weight per head) in edge-id order -> CSR slot order:
 * out[p, :] = w[eids[p], :], p in [0, nnz), rows of row_bytes bytes (even, <= 4096).
 * The SpMM kernels index edge data with e(p) = eids[p], or p itself when eids == NULL; on a large
 * graph the per-slot w[eids[p]] read is a random gather repeated for every feature slab (measured
 * 38 ms instead of 7 ms at 114.6M edges with scalar weights), so callers that reduce with SUM (no
 * arg output, edge ids not needed) permute once with this entry point and pass eids = NULL + out
 * as w.  This is the READ_ROW(edge_frame, eid) of the reference's SPMV_WITH_DATA schedule
 * (python/dgl/runtime/scheduler.py:540-559, runtime/spmv.py:154-227).  Ids outside [0, n_edges)
 * are never dereferenced (the slot is left unwritten); bad_flag (nullable) is set to 1 for them.
 */
int dglb200_edge_data_to_slots(int64_t nnz, const int32_t* eids, int64_t n_edges, int64_t row_bytes,
                               const void* w, void* out, int32_t* bad_flag, void* stream);

/*
 * Gradient of MAX wrt x (and of SUM, with arg == NULL) in gather form over the CSR keyed by
 * src (rows = source nodes, indices = dst, eids = edge id):
 *   dx[u, f] = sum_{p in row u} [arg == NULL or arg[indices[p], f] == e(p)] * wgt * dy[indices[p], f]
 *   wgt = 1, or w[e(p), f / (feat / w_width)] when w != NULL (src_mul_edge).
 * dy_rows = number of rows of dy (destination nodes); it only steers the L2 slab choice.
 * Deterministic (no atomics).  This is what torch's sparse.mm backward provides in the
 * reference (call site backend/pytorch/tensor.py:146).
 */
int dglb200_spmm_csr_backward_x(int64_t n_rows, int64_t dy_rows, int64_t nnz,
                                const int64_t* indptr, const int32_t* indices, const int32_t* eids,
                                int dtype, const void* dy, int64_t feat,
                                const void* w, int64_t w_width, const int32_t* arg,
                                void* dx, const dglb200_spmm_plan_t* plan, void* stream);

/* ---- (3) SDDMM, per-destination edge softmax, fused GAT aggregation ------------------ */

/*
 * Per-edge u-op-v over the in-edge CSR (rows = dst):  for slot p of row r, s = indices[p]:
 *   DOT: out[e(p), h]      = sum_d u[s, h, d] * v[r, h, d]          out is [nnz, heads]
 *   ADD: out[e(p), h, d]   = u[s, h, d] + v[r, h, d]                out is [nnz, heads*dim]
 *   MUL: out[e(p), h, d]   = u[s, h, d] * v[r, h, d]
 *   COPY_U / COPY_V: out[e(p), h, d] = u[s, h, d]  /  v[r, h, d]   (the unused operand may be NULL)
 * DOT is d(loss)/d(w) of SRC_MUL_EDGE (u = x, v = dy) and the u.v attention score
 * the north-star names; ADD is the reference GAT example's a1[src] + a2[dst]
 * (examples/pytorch/gat/train.py:90-92; replaces READ_ROW x2 + EDGE_UDF of
 * runtime/scheduler.py:234-277,572-583).
 */
int dglb200_sddmm_csr(int64_t n_rows, int64_t nnz,
                      const int64_t* indptr, const int32_t* indices, const int32_t* eids,
                      int op, int dtype, const void* u, const void* v,
                      int64_t heads, int64_t dim, void* out, void* stream);

/*
 * Edge softmax over the in-edges of every destination, per head:
 *   alpha[e, h] = exp(score[e, h] - max_row) / sum_row(...)          (stable flavour)
 * score/alpha are [nnz, heads] indexed by edge id e(p).  fp32 only.
 * Backward: dscore[e,h] = alpha[e,h] * (dalpha[e,h] - sum_{row} alpha*dalpha).
 */
int dglb200_edge_softmax_csr(int64_t n_rows, int64_t nnz, const int64_t* indptr,
                             const int32_t* eids, int64_t heads,
                             const float* score, float* alpha, void* stream);
int dglb200_edge_softmax_backward_csr(int64_t n_rows, int64_t nnz, const int64_t* indptr,
                                      const int32_t* eids, int64_t heads,
                                      const float* alpha, const float* dalpha, float* dscore,
                                      void* stream);

/*
 * One launch for the reference GAT layer's edge path
 * (examples/pytorch/gat/train.py:61-97: apply_edges(edge_attention) ->
 *  update_all([src_mul_edge('ft','a_drop','ft'), copy_edge('a','a')], [sum, sum]) -> ft / z):
 *   s[p,h]   = leaky_relu(a_src[indices[p], h] + a_dst[r, h], negative_slope)
 *   a[p,h]   = GAT_EXAMPLE: min(max(exp(s), -10), 10)      STABLE: exp(s - max_row s)
 *   out[r,h,:] = sum_p a[p,h] * ft[indices[p], h, :] / sum_p a[p,h]
 * Optional outputs (nullable): a_out[nnz, heads] unnormalised a by edge id, z_out[n_rows, heads]
 * the normaliser (what the example keeps in edata['a'] / ndata['z']).  Empty rows: out = 0/0 in
 * the reference (NaN); here out = 0 and z = 0 is documented in DESIGN.md.  fp32 only.
 */
int dglb200_gat_aggregate_csr(int64_t n_rows, int64_t nnz,
                              const int64_t* indptr, const int32_t* indices, const int32_t* eids,
                              const float* a_src, const float* a_dst, const float* ft,
                              int64_t heads, int64_t dim, float negative_slope, int flavour,
                              float* out, float* a_out, float* z_out, void* stream);

/* ---- (4b) degree bucketing ----------------------------------------------------------- */

/*
 * Replaces sched::DegreeBucketing (src/scheduler/scheduler.cc:13-93; PackedFunc
 * "runtime.degree_bucketing._CAPI_DGLDegreeBucketing", src/scheduler/scheduler_apis.cc:16-23).
 * In:  msg_ids[M], vids[M] (destination of every message), recv_ids[R]; ids in [0, id_bound).
 * Out (int64, device), in CANONICAL order - the reference's order is std::unordered_map
 * iteration order, i.e. unspecified:
 *   degs[B]         distinct non-zero in-degrees, ascending; then 0 if any recv id got no message
 *   nids[K]         nodes grouped by bucket, ascending id inside a bucket
 *   nid_section[B]  nodes per bucket
 *   mids[M]         message ids grouped by bucket then node; per node in INPUT order
 *   mid_section[B'] deg * bucket size for the non-zero buckets (B' = B or B-1)
 * Capacities the caller must provide: degs/nid_section/mid_section >= dglb200_bucket_capacity(M)
 * entries, nids >= M + R, mids >= M.  counts_host[0..2] = B, K, B'.  Syncs the stream.
 */
int64_t dglb200_bucket_capacity(int64_t n_msgs);
int dglb200_degree_bucketing_workspace_bytes(int64_t n_msgs, int64_t n_recv, size_t* bytes_host);
int dglb200_degree_bucketing(int64_t n_msgs, int64_t n_recv, int64_t id_bound,
                             const int64_t* msg_ids, const int64_t* vids, const int64_t* recv_ids,
                             int64_t* degs, int64_t* nids, int64_t* nid_section,
                             int64_t* mids, int64_t* mid_section, int64_t* counts_host,
                             void* workspace, size_t workspace_bytes, void* stream);

/* ---- subset message passing (send_and_recv / pull): unique-ascending relabel --------- */

/*
 * Replaces F.unique + F.sort_1d + utils.build_relabel_map + the COO build of
 * build_adj_matrix_uv (python/dgl/runtime/scheduler.py:136,341; python/dgl/utils.py:321-361;
 * python/dgl/runtime/spmv.py:154-227).  Sorted unique values of ids[n] (< id_bound):
 * uniq[counts_host[0]] ascending; inverse[i] = position of ids[i] in uniq (nullable).  Syncs.
 */
int dglb200_unique_workspace_bytes(int64_t n, size_t* bytes_host);
int dglb200_unique_sorted(int64_t n, int64_t id_bound, const int64_t* ids,
                          int64_t* uniq, int64_t* inverse, int64_t* counts_host,
                          void* workspace, size_t workspace_bytes, void* stream);

/* ---- row gather --------------------------------------------------------------------------- */

/*
 * out[i, :] = x[idx[i], :] for rows of row_bytes bytes (any dtype).  READ_ROW / F.gather_row of the
 * reference (python/dgl/runtime/ir/executor.py:365-394; python/dgl/backend/pytorch/tensor.py:108-109)
 * and the send-side packing of the multi-GPU halo exchange.  idx int64[n_out] on the device, values
 * in [0, x_rows); an out-of-range id sets *bad_flag (device int32, caller-zeroed; nullable is NOT
 * allowed) and leaves that output row untouched.  Does not sync.
 */
int dglb200_gather_rows(int64_t n_out, const int64_t* idx, int64_t x_rows, const void* x,
                        int64_t row_bytes, void* out, int32_t* bad_flag, void* stream);

/*
 * General form: x rows x_ld_bytes apart, out rows out_ld_bytes apart, row_bytes copied per row;
 * idx == NULL copies row i to row i (a strided 2-D block copy).  max_ctas > 0 bounds the grid
 * (grid-stride loop): pulls from a peer GPU over NVLink need bytes in flight, not SMs, and must
 * leave the SMs to the SpMM running concurrently on another stream.
 */
int dglb200_gather_rows_strided(int64_t n_out, const int64_t* idx, int64_t x_rows,
                                const void* x, int64_t x_ld_bytes, int64_t row_bytes,
                                void* out, int64_t out_ld_bytes, int32_t* bad_flag,
                                int max_ctas, void* stream);

/*
 * Multi-source form (one launch pulls a halo chunk from every owner at once): output row i, with
 * seg_ptr[q] <= i < seg_ptr[q+1], is row idx[i] of source q (src_base[q], src_rows[q] rows of
 * row_bytes bytes, dense).  The three tables are HOST arrays (n_src <= 16 entries, seg_ptr n_src+1);
 * idx / out / bad_flag are device pointers.  row_bytes must be a multiple of 16.
 */
int dglb200_gather_rows_multi(int n_src, const void* const* src_base_host, const int64_t* src_rows_host,
                              const int64_t* seg_ptr_host, const int64_t* idx, int64_t row_bytes,
                              void* out, int32_t* bad_flag, int max_ctas, void* stream);

/*
 * Converting row gather for the reduced-precision WIRE mode of the halo exchange (no counterpart in
 * the reference, which has no multi-GPU layer; SURVEY.md 8e lever 2 "bf16 on the wire, fp32
 * accumulate").  to_bf16 != 0: x is float, out is bf16, out[i, :] = round_to_nearest_even(x[idx[i], :]);
 * to_bf16 == 0: x is bf16, out is float (exact widening).  x_ld / out_ld are leading dimensions in
 * ELEMENTS; idx == NULL converts row i to row i.  `out` may be a peer GPU's mapped memory (push).
 * Off by default (DGL_B200_WIRE=bf16 selects it); the fp32 wire is the parity mode.
 */
int dglb200_gather_rows_convert(int64_t n_out, const int64_t* idx, int64_t x_rows, int to_bf16,
                                const void* x, int64_t x_ld, int64_t feat,
                                void* out, int64_t out_ld, int32_t* bad_flag,
                                int max_ctas, void* stream);

/* ---- CUDA IPC (multi-GPU peer-load exchange) ------------------------------------------------- */

/*
 * The reference has no multi-GPU layer.  These three calls let one process per
 * GPU read another rank's feature block straight over NVLink: the owner exports the BASE of a
 * cudaMalloc allocation (64-byte cudaIpcMemHandle_t), a peer maps it on ITS current device
 * (peer access is enabled lazily by the driver) and passes the mapped pointer as `x` of
 * dglb200_gather_rows / dglb200_spmm_csr.  dev_ptr may point inside an allocation (torch sub-allocates):
 * the handle is that of the containing cudaMalloc block and *offset_out the byte offset of dev_ptr in it.
 */
int dglb200_ipc_get_handle(const void* dev_ptr, void* handle64_out, int64_t* offset_out);
int dglb200_ipc_open(const void* handle64, void** ptr_out);
int dglb200_ipc_close(void* ptr);

/* ---- push + signal exchange over NVLink peer memory (multi-GPU halo / all-gather) ------------- */

/*
 * The engine's own transport for the 1-D row partition (no counterpart in the reference).
 * ONE launch moves this rank's rows into every destination's exchange buffer (peer memory mapped with
 * dglb200_ipc_open) and then sets one flag word per destination; consumers order their kernels behind
 * dglb200_peer_wait on their own flag words.  No NCCL call and no host round trip on the data path.
 *
 *   destination q in [0, n_dst) receives rows idx[seg_ptr[q] .. seg_ptr[q+1]) of x (idx == NULL: the
 *   segment's row i is x row i, i.e. a block copy) at dst_base[q] + i * dst_ld_bytes.  The three tables
 *   are HOST arrays (n_dst <= 16; seg_ptr has n_dst + 1 entries); a destination with an empty segment may
 *   have dst_base NULL (this is how a rank lists ITSELF so that its own flag is set as well).
 *   convert_to_bf16 != 0: x rows are float32 and are rounded to bfloat16 (nearest even) on the way out
 *   (row_bytes counts the float32 bytes, a multiple of 32; destination rows hold row_bytes / 2).
 *   wait_flags (nullable): n_wait LOCAL flag words that must all be >= wait_value before the first store
 *   (the destinations have finished reading what the previous exchange delivered).
 *   signal_flags (nullable): per destination, the flag word that receives signal_value once every store
 *   of the launch is visible system-wide (fence.sys + arrival counter; `counter` is a local int32 that is
 *   0 on entry and is left 0).  Flags only grow; comparisons are wrap-safe.
 *   A wait that sees no signal for 20 s sets *err_flag (nullable) and lets the kernel proceed: a wrong
 *   result that the caller's checks catch, never a hung GPU.  bad_flag as in dglb200_gather_rows.
 *   rotate: rotation of the destination order (pass the rank).  max_ctas: grid bound (0 = 2 per SM).
 */
int dglb200_peer_push(int n_dst, const int64_t* seg_ptr_host, void* const* dst_base_host,
                      int32_t* const* signal_flags_host, const int64_t* idx, int64_t x_rows,
                      const void* x, int64_t x_ld_bytes, int64_t row_bytes, int64_t dst_ld_bytes,
                      int convert_to_bf16, const int32_t* wait_flags, int n_wait, int32_t wait_value,
                      int32_t signal_value, int32_t* counter, int32_t* err_flag, int32_t* bad_flag,
                      int rotate, int max_ctas, void* stream);
/* Stream-ordered signal: once everything enqueued before it on `stream` has completed, `value` is
 * stored to each of the n (<= 16) flag words listed in the HOST array (peer or local memory). */
int dglb200_peer_signal(int n, int32_t* const* flags_host, int32_t value, void* stream);
/* Stream-ordered wait: kernels enqueued after it on `stream` start only once the n LOCAL flag words
 * flags[0..n) are all >= value (spin with back-off on one thread; time-out as above). */
int dglb200_peer_wait(const int32_t* flags, int n, int32_t value, int32_t* err_flag, void* stream);

/* ---- hook into the reference's PackedFunc registry ----------------------------------- */

/*
 * Registers the entry points above as PackedFunc globals of an already-loaded libdgl.so
 * (names "b200._CAPI_*", plus an override of
 * "runtime.degree_bucketing._CAPI_DGLDegreeBucketing" when override_reference != 0) using only
 * libdgl's C ABI: DGLFuncCreateFromCFunc / DGLFuncRegisterGlobal / DGLCFuncSetReturn /
 * DGLAPISetLastError (include/dgl/runtime/c_runtime_api.h:255-350).  `libdgl_handle` is the
 * dlopen handle of libdgl.so (ctypes: CDLL._handle).  Returns the number of globals registered
 * or -1.
 */
int dglb200_register_packed_funcs(void* libdgl_handle, int override_reference);

#ifdef __cplusplus
}
#endif
#endif  /* DGL_B200_H_ */
