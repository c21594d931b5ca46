// Device-resident graph index queries over the cached CSRs: the edge lists of a set of rows
// (Graph::InEdges(IdArray) / OutEdges(IdArray), src/graph/graph.cc:294-365; the slot lists that
// pull / push / send_and_recv build their sub-adjacency from, python/dgl/runtime/spmv.py:154-227) and
// degrees of a set of nodes (Graph::InDegrees, graph.cc:412-424).  Integer work, bit-exact.
#include "scan.cuh"

namespace dglb200 {

// len[i] = degree of rows[i]; an id outside [0, n_rows) raises *bad (the reference CHECKs
// "Invalid vertex", graph.cc:299) and counts as an empty row.
__global__ void __launch_bounds__(256)
select_lens_kernel(int64_t n_sel, const int64_t* __restrict__ rows, int64_t n_rows,
                   const int64_t* __restrict__ indptr, int64_t* __restrict__ len, int32_t* bad) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_sel) return;
  const int64_t r = rows[i];
  if (r < 0 || r >= n_rows) {
    if (bad != nullptr) atomicExch(bad, 1);
    len[i] = 0;
    return;
  }
  len[i] = indptr[r + 1] - indptr[r];
}

// One thread per output slot: binary search of the owning selection in `offsets`, then the slot of
// the CSR.  Outputs are nullable; `which` = position in the selection, `slot` = CSR slot.
__global__ void __launch_bounds__(256)
select_expand_kernel(int64_t n_sel, int64_t total, const int64_t* __restrict__ rows,
                     const int64_t* __restrict__ offsets, const int64_t* __restrict__ indptr,
                     const int32_t* __restrict__ indices, const int32_t* __restrict__ eids,
                     int64_t* __restrict__ which, int64_t* __restrict__ slot,
                     int64_t* __restrict__ out_row, int64_t* __restrict__ out_col,
                     int64_t* __restrict__ out_eid) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= total) return;
  int64_t lo = 0, hi = n_sel;            // last i with offsets[i] <= p
  while (hi - lo > 1) {
    const int64_t mid = (lo + hi) >> 1;
    if (offsets[mid] <= p) lo = mid; else hi = mid;
  }
  const int64_t r = rows[lo];
  const int64_t s = indptr[r] + (p - offsets[lo]);
  if (which) which[p] = lo;
  if (slot) slot[p] = s;
  if (out_row) out_row[p] = r;
  if (out_col) out_col[p] = indices[s];
  if (out_eid) out_eid[p] = eids ? (int64_t)eids[s] : s;
}

}  // namespace dglb200

using namespace dglb200;

extern "C" int dglb200_csr_select_workspace_bytes(int64_t n_sel, size_t* bytes_host) {
  DGLB200_REQUIRE(bytes_host != nullptr && n_sel >= 0, "csr_select_workspace_bytes: bad args");
  *bytes_host = align_up((size_t)(n_sel + 1) * sizeof(int64_t)) + scan_workspace_bytes(n_sel + 1, sizeof(int64_t)) + 1024;
  return 0;
}

extern "C" int dglb200_csr_select_offsets(int64_t n_sel, const int64_t* rows, int64_t n_rows,
                                          const int64_t* indptr, int64_t* offsets, int64_t* total_host,
                                          void* workspace, size_t workspace_bytes, void* stream_) {
  clear_error();
  DGLB200_REQUIRE(n_sel >= 0 && n_rows >= 0 && total_host != nullptr, "csr_select_offsets: bad args");
  *total_host = 0;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (n_sel == 0) {
    if (offsets) DGLB200_CUDA(cudaMemsetAsync(offsets, 0, sizeof(int64_t), stream));
    return 0;
  }
  DGLB200_REQUIRE(rows && indptr && offsets, "csr_select_offsets: NULL pointer");
  Workspace ws(workspace, workspace_bytes);
  int64_t* len = ws.take<int64_t>(n_sel + 1);     // last word: the bad-id flag
  DGLB200_REQUIRE(len != nullptr, "csr_select_offsets: workspace too small");
  int32_t* bad = reinterpret_cast<int32_t*>(len + n_sel);
  DGLB200_CUDA(cudaMemsetAsync(bad, 0, sizeof(int64_t), stream));
  select_lens_kernel<<<(unsigned)ceil_div(n_sel, 256), 256, 0, stream>>>(n_sel, rows, n_rows, indptr, len, bad);
  DGLB200_POST_LAUNCH("select_lens_kernel");
  if (exclusive_scan<int64_t, int64_t>(len, offsets, n_sel, 1, ws, stream) != 0) return -1;
  int32_t bad_h = 0;
  DGLB200_CUDA(cudaMemcpyAsync(total_host, offsets + n_sel, sizeof(int64_t), cudaMemcpyDeviceToHost, stream));
  DGLB200_CUDA(cudaMemcpyAsync(&bad_h, bad, sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
  DGLB200_CUDA(cudaStreamSynchronize(stream));
  DGLB200_REQUIRE(bad_h == 0, "Invalid vertex: an id is outside [0, %lld)", (long long)n_rows);
  return 0;
}

extern "C" int dglb200_csr_select_expand(int64_t n_sel, int64_t total, const int64_t* rows,
                                         const int64_t* offsets, const int64_t* indptr,
                                         const int32_t* indices, const int32_t* eids, int64_t* which,
                                         int64_t* slot, int64_t* out_row, int64_t* out_col,
                                         int64_t* out_eid, void* stream_) {
  clear_error();
  DGLB200_REQUIRE(n_sel >= 0 && total >= 0, "csr_select_expand: bad sizes");
  if (total == 0) return 0;
  DGLB200_REQUIRE(rows && offsets && indptr, "csr_select_expand: NULL pointer");
  DGLB200_REQUIRE(out_col == nullptr || indices != nullptr, "csr_select_expand: out_col needs indices");
  DGLB200_REQUIRE(total < (1ll << 40), "csr_select_expand: too many slots");
  select_expand_kernel<<<(unsigned)ceil_div(total, 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      n_sel, total, rows, offsets, indptr, indices, eids, which, slot, out_row, out_col, out_eid);
  DGLB200_POST_LAUNCH("select_expand_kernel");
  return 0;
}

extern "C" int dglb200_degrees_of(int64_t n_sel, const int64_t* rows, int64_t n_rows, const int64_t* indptr,
                                  int64_t* deg, int32_t* bad_flag, void* stream_) {
  clear_error();
  DGLB200_REQUIRE(n_sel >= 0 && n_rows >= 0, "degrees_of: bad sizes");
  if (n_sel == 0) return 0;
  DGLB200_REQUIRE(rows && indptr && deg, "degrees_of: NULL pointer");
  select_lens_kernel<<<(unsigned)ceil_div(n_sel, 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      n_sel, rows, n_rows, indptr, deg, bad_flag);
  DGLB200_POST_LAUNCH("select_lens_kernel");
  return 0;
}
