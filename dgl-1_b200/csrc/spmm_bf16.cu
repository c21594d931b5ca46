// bf16-storage / fp32-accumulate instantiations of the SpMM kernel (see spmm_kernel.cuh).
#include "spmm_kernel.cuh"
namespace dglb200 {
int launch_spmm_bf16(const SpmmArgs& a, int msg, int red, bool vec, cudaStream_t stream) {
  return vec ? launch_ops<__nv_bfloat16, 8>(a, msg, red, stream)
             : launch_ops<__nv_bfloat16, 1>(a, msg, red, stream);
}
}  // namespace dglb200
