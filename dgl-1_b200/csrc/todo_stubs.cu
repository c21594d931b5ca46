// Temporary: entry points declared in include/dgl_b200.h whose kernels are not written yet.
// They fail loudly (no fallback).  Each stub disappears when its kernel lands.
#include "common.cuh"
using namespace dglb200;
#define NOT_YET(name) do { set_error(name ": not implemented yet"); return -1; } while (0)
extern "C" {
int dglb200_sddmm_csr(int64_t, int64_t, const int64_t*, const int32_t*, const int32_t*, int, int,
                      const void*, const void*, int64_t, int64_t, void*, void*) { NOT_YET("sddmm_csr"); }
int dglb200_edge_softmax_csr(int64_t, int64_t, const int64_t*, const int32_t*, int64_t, const float*,
                             float*, void*) { NOT_YET("edge_softmax_csr"); }
int dglb200_edge_softmax_backward_csr(int64_t, int64_t, const int64_t*, const int32_t*, int64_t,
                                      const float*, const float*, float*, void*) { NOT_YET("edge_softmax_backward_csr"); }
int dglb200_gat_aggregate_csr(int64_t, int64_t, const int64_t*, const int32_t*, const int32_t*,
                              const float*, const float*, const float*, int64_t, int64_t, float, int,
                              float*, float*, float*, void*) { NOT_YET("gat_aggregate_csr"); }
int64_t dglb200_bucket_capacity(int64_t) { return 0; }
int dglb200_degree_bucketing_workspace_bytes(int64_t, int64_t, size_t*) { NOT_YET("degree_bucketing"); }
int dglb200_degree_bucketing(int64_t, int64_t, int64_t, const int64_t*, const int64_t*, const int64_t*,
                             int64_t*, int64_t*, int64_t*, int64_t*, int64_t*, int64_t*, void*, size_t,
                             void*) { NOT_YET("degree_bucketing"); }
int dglb200_unique_workspace_bytes(int64_t, size_t*) { NOT_YET("unique_sorted"); }
int dglb200_unique_sorted(int64_t, int64_t, const int64_t*, int64_t*, int64_t*, int64_t*, void*, size_t,
                          void*) { NOT_YET("unique_sorted"); }
int dglb200_register_packed_funcs(void*, int) { NOT_YET("register_packed_funcs"); }
}
