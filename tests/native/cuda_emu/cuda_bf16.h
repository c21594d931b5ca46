// TEST INFRASTRUCTURE ONLY - host stand-in for <cuda_bf16.h> (see cuda_runtime.h in this directory):
// bfloat16 storage with round-to-nearest-even conversion, as the device intrinsics do.
#pragma once
#include "cuda_runtime.h"

struct __nv_bfloat16 {
  uint16_t bits;
  __nv_bfloat16() = default;
};
struct __nv_bfloat162 { __nv_bfloat16 x, y; };

inline uint16_t dglb200_emu_f2bf(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fff;            // NaN
  return (uint16_t)((u + 0x7fffu + ((u >> 16) & 1u)) >> 16);      // round to nearest even
}
inline __nv_bfloat16 __float2bfloat16_rn(float f) { __nv_bfloat16 b; b.bits = dglb200_emu_f2bf(f); return b; }
inline __nv_bfloat16 __float2bfloat16(float f) { return __float2bfloat16_rn(f); }
inline float __bfloat162float(__nv_bfloat16 b) { return __uint_as_float((uint32_t)b.bits << 16); }
inline __nv_bfloat162 __floats2bfloat162_rn(float lo, float hi) {
  __nv_bfloat162 v;
  v.x = __float2bfloat16_rn(lo);
  v.y = __float2bfloat16_rn(hi);
  return v;
}
