// Shared helpers for libdgl_b200.so: error TLS, launch accounting, cache-hinted loads.
// Error convention mirrors the reference's API_BEGIN/API_END + DGLAPISetLastError
// (src/runtime/runtime_base.h:13-33, src/runtime/c_runtime_api.cc:130-148): entry points
// return 0 / -1 and leave a thread-local message.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "dgl_b200.h"

namespace dglb200 {

void set_error(const char* fmt, ...);
void clear_error();
void count_launch(int n = 1);

constexpr int kWarp = 32;
constexpr int kNumSMsB200 = 148;

inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// Bump allocator over the caller's workspace.
struct Workspace {
  char* base;
  size_t size, used;
  Workspace(void* p, size_t n) : base(static_cast<char*>(p)), size(n), used(0) {}
  template <typename T>
  T* take(size_t n) {
    size_t bytes = align_up(n * sizeof(T));
    if (base == nullptr || used + bytes > size) {
      used += bytes;  // keep counting so callers can report the needed size
      return nullptr;
    }
    T* r = reinterpret_cast<T*>(base + used);
    used += bytes;
    return r;
  }
};

#define DGLB200_REQUIRE(cond, ...)          \
  do {                                      \
    if (!(cond)) {                          \
      ::dglb200::set_error(__VA_ARGS__);    \
      return -1;                            \
    }                                       \
  } while (0)

#define DGLB200_CUDA(expr)                                                               \
  do {                                                                                   \
    cudaError_t e__ = (expr);                                                            \
    if (e__ != cudaSuccess) {                                                            \
      ::dglb200::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__),      \
                           __FILE__, __LINE__);                                          \
      return -1;                                                                         \
    }                                                                                    \
  } while (0)

// After a <<<>>> launch: count it and surface launch-configuration errors.
#define DGLB200_POST_LAUNCH(name)                                                        \
  do {                                                                                   \
    ::dglb200::count_launch();                                                           \
    cudaError_t e__ = cudaPeekAtLastError();                                             \
    if (e__ != cudaSuccess) {                                                            \
      (void)cudaGetLastError();                                                          \
      ::dglb200::set_error("launch of %s failed: %s (%s:%d)", name,                      \
                           cudaGetErrorString(e__), __FILE__, __LINE__);                 \
      return -1;                                                                         \
    }                                                                                    \
  } while (0)

// ---- device-side load/store helpers -----------------------------------------------------

// Structure arrays (column ids, edge ids) are read once per launch: bypass L1 so they do not
// evict feature rows that hub sources keep hot.
// (DGLB200_HOST_EMU: tests/native/cuda_emu compiles these sources for the host to run the kernels'
// logic without a GPU; the PTX below then reads as the plain C it stands for.)
__device__ __forceinline__ int32_t ld_stream_i32(const int32_t* p) {
#ifdef DGLB200_HOST_EMU
  return *p;
#else
  int32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
#endif
}
__device__ __forceinline__ int64_t ld_stream_i64(const int64_t* p) {
#ifdef DGLB200_HOST_EMU
  return *p;
#else
  int64_t v;
  asm volatile("ld.global.nc.L1::no_allocate.s64 %0, [%1];" : "=l"(v) : "l"(p));
  return v;
#endif
}

// base + id * row_bytes as ONE instruction (IMAD.WIDE.U32 with a 64-bit addend).  Written in PTX
// because the compiler otherwise splits a uniform base from the per-lane offset (3 instructions
// per gathered row in the SpMM inner loop).
__device__ __forceinline__ const char* row_ptr(const char* base, uint32_t id, uint32_t row_bytes) {
#ifdef DGLB200_HOST_EMU
  return base + (uint64_t)id * (uint64_t)row_bytes;
#else
  uint64_t r;
  asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(id), "r"(row_bytes), "l"(base));
  return reinterpret_cast<const char*>(r);
#endif
}

__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// A "chunk" is VEC consecutive elements of one feature row owned by one lane:
// 16 bytes on the vector path (4 x fp32 or 8 x bf16), one element on the scalar path.
template <typename T, int VEC>
struct Chunk;

template <>
struct Chunk<float, 4> {
  static __device__ __forceinline__ void load(const float* p, float (&v)[4]) {
    float4 t = __ldg(reinterpret_cast<const float4*>(p));
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  static __device__ __forceinline__ void store(float* p, const float (&v)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};
template <>
struct Chunk<float, 1> {
  static __device__ __forceinline__ void load(const float* p, float (&v)[1]) { v[0] = __ldg(p); }
  static __device__ __forceinline__ void store(float* p, const float (&v)[1]) { *p = v[0]; }
};
template <>
struct Chunk<__nv_bfloat16, 8> {
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&v)[8]) {
    uint4 t = __ldg(reinterpret_cast<const uint4*>(p));
    v[0] = bf16_lo(t.x); v[1] = bf16_hi(t.x); v[2] = bf16_lo(t.y); v[3] = bf16_hi(t.y);
    v[4] = bf16_lo(t.z); v[5] = bf16_hi(t.z); v[6] = bf16_lo(t.w); v[7] = bf16_hi(t.w);
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, const float (&v)[8]) {
    uint4 t;
    t.x = pack_bf16(v[0], v[1]); t.y = pack_bf16(v[2], v[3]);
    t.z = pack_bf16(v[4], v[5]); t.w = pack_bf16(v[6], v[7]);
    *reinterpret_cast<uint4*>(p) = t;
  }
};
template <>
struct Chunk<__nv_bfloat16, 1> {
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&v)[1]) {
    v[0] = __bfloat162float(*p);
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, const float (&v)[1]) {
    *p = __float2bfloat16_rn(v[0]);
  }
};

__device__ __forceinline__ float to_float(float v) { return v; }
__device__ __forceinline__ float to_float(__nv_bfloat16 v) { return __bfloat162float(v); }

}  // namespace dglb200
