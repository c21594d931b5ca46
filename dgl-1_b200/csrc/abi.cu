// Library-level entry points of libdgl_b200.so: version, thread-local last error, launch
// counter.  Error convention follows the reference (src/runtime/c_runtime_api.cc:130-148).
#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "common.cuh"

namespace dglb200 {

namespace {
thread_local char tls_error[1024] = {0};
std::atomic<int64_t> g_launches{0};
}  // namespace

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(tls_error, sizeof(tls_error), fmt, ap);
  va_end(ap);
}
void clear_error() { tls_error[0] = 0; }
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

}  // namespace dglb200

extern "C" {

int dglb200_abi_version(void) { return DGLB200_ABI_VERSION; }

const char* dglb200_last_error(void) { return dglb200::tls_error; }

int64_t dglb200_launch_count(void) { return dglb200::g_launches.load(); }

int dglb200_device_cc(void) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    dglb200::set_error("no usable CUDA device: %s", cudaGetErrorString(e));
    return -1;
  }
  int major = 0, minor = 0;
  DGLB200_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  DGLB200_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
  return major * 10 + minor;
}

}  // extern "C"
