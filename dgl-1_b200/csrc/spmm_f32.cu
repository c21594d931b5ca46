// fp32 instantiations of the SpMM kernel (see spmm_kernel.cuh).
#include "spmm_kernel.cuh"
namespace dglb200 {
int launch_spmm_f32(const SpmmArgs& a, int msg, int red, bool vec, cudaStream_t stream) {
  return vec ? launch_ops<float, 4>(a, msg, red, stream) : launch_ops<float, 1>(a, msg, red, stream);
}
}  // namespace dglb200
