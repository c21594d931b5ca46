// Kernel family (4): device-side CSR build from a COO edge list.
// Replaces Graph::GetAdj(.., "csr") (src/graph/graph.cc:525-549, per-row insertion order) and
// ImmutableGraph::CSR::FromEdges (src/graph/immutable_graph.cc:206-237, rows sorted by the
// other endpoint).  Integer work: results are bit-exact with the reference.
#include "common.cuh"
#include "radix_sort.cuh"
#include "scan.cuh"

namespace dglb200 {

// int64 ids -> uint32 keys, range check, row histogram
__global__ void csr_prepare(const int64_t* __restrict__ row, const int64_t* __restrict__ col,
                            int64_t nnz, int64_t n_rows, int64_t n_cols,
                            uint32_t* __restrict__ rowk, uint32_t* __restrict__ colk,
                            int32_t* __restrict__ deg, int32_t* __restrict__ bad) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nnz) return;
  const int64_t r = row[i], c = col[i];
  if (r < 0 || r >= n_rows || c < 0 || c >= n_cols) {
    atomicExch(bad, 1);
    rowk[i] = 0;
    if (colk) colk[i] = 0;
    return;
  }
  rowk[i] = (uint32_t)r;
  if (colk) colk[i] = (uint32_t)c;
  atomicAdd(&deg[r], 1);
}

__global__ void gather_u32(const uint32_t* __restrict__ src, const uint32_t* __restrict__ perm,
                           int64_t n, uint32_t* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = src[perm[i]];
}

// perm[p] = input position of the edge in slot p
__global__ void csr_finalize(const uint32_t* __restrict__ perm, const int64_t* __restrict__ col,
                             const int64_t* __restrict__ eid, int64_t nnz,
                             int32_t* __restrict__ indices, int32_t* __restrict__ eids) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= nnz) return;
  const uint32_t i = perm[p];
  indices[p] = (int32_t)col[i];
  eids[p] = eid ? (int32_t)eid[i] : (int32_t)i;
}

__global__ void widen_kernel(const int32_t* __restrict__ in, int64_t n, int64_t* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = in[i];
}

__global__ void degrees_kernel(const int64_t* __restrict__ indptr, int64_t n, int64_t* __restrict__ deg) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) deg[i] = indptr[i + 1] - indptr[i];
}

// one warp per row writes the row id into every slot of the row
__global__ void row_ids_kernel(const int64_t* __restrict__ indptr, int64_t n_rows,
                               int64_t* __restrict__ out) {
  const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= n_rows) return;
  const int lane = threadIdx.x & 31;
  const int64_t s = indptr[r], e = indptr[r + 1];
  for (int64_t p = s + lane; p < e; p += 32) out[p] = r;
}

}  // namespace dglb200

using namespace dglb200;

extern "C" {

int dglb200_csr_build_workspace_bytes(int64_t n_rows, int64_t nnz, size_t* bytes_host) {
  DGLB200_REQUIRE(bytes_host != nullptr && n_rows >= 0 && nnz >= 0, "csr_build_workspace_bytes: bad args");
  size_t b = 0;
  b += align_up((size_t)(n_rows + 1) * sizeof(int32_t));          // deg
  b += align_up(sizeof(int32_t) * 4);                             // bad flag
  b += align_up((size_t)nnz * sizeof(uint32_t)) * 6;              // rowk, colk, 2x(keys, vals)
  b += radix_sort_workspace_bytes(nnz);
  b += scan_workspace_bytes(n_rows + 1, sizeof(int64_t));
  *bytes_host = b + 4096;
  return 0;
}

int dglb200_csr_build(int64_t n_rows, int64_t n_cols, int64_t nnz, const int64_t* row,
                      const int64_t* col, const int64_t* eid, int order, int64_t* indptr,
                      int32_t* indices, int32_t* eids, void* workspace, size_t workspace_bytes,
                      void* stream_) {
  clear_error();
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  DGLB200_REQUIRE(n_rows >= 0 && n_cols >= 0 && nnz >= 0, "csr_build: negative size");
  DGLB200_REQUIRE(n_rows < (1ll << 31) && n_cols < (1ll << 31) && nnz < (1ll << 31),
                  "csr_build: n_rows / n_cols / nnz must be < 2^31 (got %lld, %lld, %lld)",
                  (long long)n_rows, (long long)n_cols, (long long)nnz);
  DGLB200_REQUIRE(order == 0 || order == 1, "csr_build: order must be 0 or 1");
  DGLB200_REQUIRE(indptr != nullptr, "csr_build: indptr is NULL");
  if (nnz == 0) {
    DGLB200_CUDA(cudaMemsetAsync(indptr, 0, sizeof(int64_t) * (size_t)(n_rows + 1), stream));
    return 0;
  }
  DGLB200_REQUIRE(row && col && indices && eids, "csr_build: NULL pointer");
  Workspace ws(workspace, workspace_bytes);
  int32_t* deg = ws.take<int32_t>(n_rows + 1);
  int32_t* bad = ws.take<int32_t>(4);
  uint32_t* rowk = ws.take<uint32_t>(nnz);
  uint32_t* colk = order == 1 ? ws.take<uint32_t>(nnz) : nullptr;
  uint32_t* ka = ws.take<uint32_t>(nnz);
  uint32_t* va = ws.take<uint32_t>(nnz);
  uint32_t* kb = ws.take<uint32_t>(nnz);
  uint32_t* vb = ws.take<uint32_t>(nnz);
  DGLB200_REQUIRE(deg && bad && rowk && ka && va && kb && vb && (order == 0 || colk),
                  "csr_build: workspace too small (%zu bytes given)", workspace_bytes);
  DGLB200_CUDA(cudaMemsetAsync(deg, 0, sizeof(int32_t) * (size_t)(n_rows + 1), stream));
  DGLB200_CUDA(cudaMemsetAsync(bad, 0, sizeof(int32_t) * 4, stream));
  const unsigned eb = (unsigned)ceil_div(nnz, 256);
  csr_prepare<<<eb, 256, 0, stream>>>(row, col, nnz, n_rows, n_cols, rowk, colk, deg, bad);
  DGLB200_POST_LAUNCH("csr_prepare");
  int bad_host = 0;
  DGLB200_CUDA(cudaMemcpyAsync(&bad_host, bad, sizeof(int), cudaMemcpyDeviceToHost, stream));
  DGLB200_CUDA(cudaStreamSynchronize(stream));
  DGLB200_REQUIRE(bad_host == 0, "csr_build: vertex id out of range [0, %lld) x [0, %lld)",
                  (long long)n_rows, (long long)n_cols);

  const uint32_t* perm = nullptr;
  if (order == 0) {
    if (radix_sort_pairs(rowk, nullptr, ka, va, kb, vb, nnz, bits_for(n_rows), ws, stream) != 0) return -1;
    perm = va;
  } else {
    // LSD over the composite key (row, col): sort by col first, then stably by row
    if (radix_sort_pairs(colk, nullptr, ka, va, kb, vb, nnz, bits_for(n_cols), ws, stream) != 0) return -1;
    gather_u32<<<eb, 256, 0, stream>>>(rowk, va, nnz, colk);  // colk reused: row key in col order
    DGLB200_POST_LAUNCH("gather_u32");
    if (radix_sort_pairs(colk, va, ka, kb, rowk, vb, nnz, bits_for(n_rows), ws, stream) != 0) return -1;
    perm = kb;  // values of the second sort
  }
  csr_finalize<<<eb, 256, 0, stream>>>(perm, col, eid, nnz, indices, eids);
  DGLB200_POST_LAUNCH("csr_finalize");
  if (exclusive_scan<int32_t, int64_t>(deg, indptr, n_rows, 1, ws, stream) != 0) return -1;
  return 0;
}

int dglb200_widen_i32_to_i64(int64_t n, const int32_t* in, int64_t* out, void* stream) {
  clear_error();
  if (n <= 0) return 0;
  DGLB200_REQUIRE(in && out, "widen_i32_to_i64: NULL pointer");
  widen_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(in, n, out);
  DGLB200_POST_LAUNCH("widen_kernel");
  return 0;
}

int dglb200_degrees_from_indptr(int64_t n_rows, const int64_t* indptr, int64_t* deg, void* stream) {
  clear_error();
  if (n_rows <= 0) return 0;
  DGLB200_REQUIRE(indptr && deg, "degrees_from_indptr: NULL pointer");
  degrees_kernel<<<(unsigned)ceil_div(n_rows, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(indptr, n_rows, deg);
  DGLB200_POST_LAUNCH("degrees_kernel");
  return 0;
}

int dglb200_csr_row_ids(int64_t n_rows, int64_t nnz, const int64_t* indptr, int64_t* row_of_slot,
                        void* stream) {
  clear_error();
  if (n_rows <= 0 || nnz <= 0) return 0;
  DGLB200_REQUIRE(indptr && row_of_slot, "csr_row_ids: NULL pointer");
  row_ids_kernel<<<(unsigned)ceil_div(n_rows * 32, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(indptr, n_rows, row_of_slot);
  DGLB200_POST_LAUNCH("row_ids_kernel");
  return 0;
}

}  // extern "C"
