// Kernel family (3): per-edge u-op-v (SDDMM), per-destination edge softmax (+ backward) and the
// fused GAT aggregation.  See include/dgl_b200.h for the contracts.
//
// SDDMM is slot-parallel over the in-edge CSR (perfectly balanced whatever the degree skew):
// a group of LPE lanes owns a run of consecutive slots; it finds the row of its first slot by
// one binary search in indptr and then walks rows forward.  For a slot it reads the D-wide
// head segments of u[src] and v[dst] with 128-bit loads and (DOT) reduces with shuffles.
#include "common.cuh"

namespace dglb200 {

constexpr int kSddmmThreads = 256;
constexpr int kSlotsPerGroup = 8;

// dglb200_set_tuning("gat_head_slab", k): heads per pass of the fused GAT kernel (0 = one pass, the default)
static int& gat_head_slab() {
  static int v = 0;
  return v;
}
void set_gat_head_slab(int v) { gat_head_slab() = v; }

// dglb200_set_tuning("sddmm_dot_generic", 1): keep DOT on the generic slot-group kernel (tests compare the two)
static int& sddmm_dot_generic() {
  static int v = 0;
  return v;
}
void set_sddmm_dot_generic(int v) { sddmm_dot_generic() = v; }

template <typename T>
__device__ __forceinline__ void store_one(T* p, float v);
template <>
__device__ __forceinline__ void store_one<float>(float* p, float v) { *p = v; }
template <>
__device__ __forceinline__ void store_one<__nv_bfloat16>(__nv_bfloat16* p, float v) {
  *p = __float2bfloat16_rn(v);
}

template <typename T, int VEC, int LPE, int OP>
__global__ void __launch_bounds__(kSddmmThreads)
sddmm_kernel(int64_t n_rows, int64_t nnz, const int64_t* __restrict__ indptr,
             const int32_t* __restrict__ indices, const int32_t* __restrict__ eids,
             const T* __restrict__ u, const T* __restrict__ v, int64_t heads, int64_t dim,
             T* __restrict__ out) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t group = tid / LPE;
  const int gl = (int)(tid % LPE);
  const unsigned lane = threadIdx.x & 31;
  // mask of the lanes of this group inside the warp (groups never straddle warps: LPE | 32)
  const unsigned gmask = (LPE == 32) ? 0xffffffffu : (((1u << LPE) - 1u) << (lane / LPE * LPE));
  int64_t p = group * kSlotsPerGroup;
  const int64_t p_end = min(p + (int64_t)kSlotsPerGroup, nnz);
  if (p >= nnz) return;
  // row of the first slot: last r with indptr[r] <= p
  int64_t lo = 0, hi = n_rows;
  while (hi - lo > 1) {
    const int64_t mid = (lo + hi) >> 1;
    if (indptr[mid] <= p) lo = mid; else hi = mid;
  }
  int64_t r = lo;
  int64_t r_end = indptr[r + 1];
  const int64_t feat = heads * dim;
  for (; p < p_end; ++p) {
    while (p >= r_end) { ++r; r_end = indptr[r + 1]; }
    const int64_t s = indices[p];
    const int64_t e = eids ? (int64_t)eids[p] : p;
    const T* up = u + s * feat;
    const T* vp = v + r * feat;
    for (int64_t h = 0; h < heads; ++h) {
      float acc = 0.f;
      for (int64_t d = (int64_t)gl * VEC; d < dim; d += (int64_t)LPE * VEC) {
        float a[VEC], b[VEC];
        if (OP != DGLB200_SDDMM_COPY_V) Chunk<T, VEC>::load(up + h * dim + d, a);
        if (OP != DGLB200_SDDMM_COPY_U) Chunk<T, VEC>::load(vp + h * dim + d, b);
        if (OP == DGLB200_SDDMM_DOT) {
#pragma unroll
          for (int k = 0; k < VEC; ++k) acc = fmaf(a[k], b[k], acc);
        } else {
          float o[VEC];
#pragma unroll
          for (int k = 0; k < VEC; ++k)
            o[k] = (OP == DGLB200_SDDMM_ADD)      ? a[k] + b[k]
                   : (OP == DGLB200_SDDMM_MUL)    ? a[k] * b[k]
                   : (OP == DGLB200_SDDMM_COPY_U) ? a[k]
                                                  : b[k];
          Chunk<T, VEC>::store(out + e * feat + h * dim + d, o);
        }
      }
      if (OP == DGLB200_SDDMM_DOT) {
#pragma unroll
        for (int o = LPE / 2; o > 0; o >>= 1) acc += __shfl_xor_sync(gmask, acc, o);
        if (gl == 0) store_one<T>(out + e * heads + h, acc);
      }
    }
  }
}

// ---- DOT, warp per run of slots ----------------------------------------------------------------------
// The generic kernel above gives a slot to `dim / VEC` lanes and walks its heads one after the other: at 8 heads
// x 32 that is 8 dependent (load, reduce, store) rounds per slot, 187 warp instructions per edge, and ncu shows it
// latency-bound at 1.4 TB/s (profiles/r02_sddmm_dot_ncu_raw.csv).  Here a WARP owns kDotSlots consecutive slots and
// a lane owns K 16-byte chunks of the WHOLE feature row (chunk c = k * 32 + lane belongs to head c / LPH, LPH =
// dim / VEC lanes per head): the destination row stays in registers while the row lasts, the source rows of U
// slots are loaded back to back before any is used, the heads of a slot reduce side by side (log2(LPH) shuffles
// per register for LPH <= 32, a whole-warp reduce per head for LPH a multiple of 32), and a slot's results leave
// as one store instruction per register.  Lane-local sums and the butterfly order are those of the generic
// kernel: the results are bit-identical.
constexpr int kDotSlots = 64;
constexpr int kDotWarps = 8;

// LPH: lanes per head as a compile-time power of two <= 32 (shifts and unrolled butterflies), or 0 = a multiple of
// 32 given at run time in `lph_rt`.
template <typename T, int VEC, int K, int U, int LPH>
__global__ void __launch_bounds__(kDotWarps * 32)
sddmm_dot_warp_kernel(int64_t n_rows, int64_t nnz, const int64_t* __restrict__ indptr,
                      const int32_t* __restrict__ indices, const int32_t* __restrict__ eids,
                      const T* __restrict__ u, const T* __restrict__ v, int64_t heads, int lph_rt,
                      T* __restrict__ out) {
  const int lph = LPH > 0 ? LPH : lph_rt;
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * kDotWarps + (threadIdx.x >> 5);
  int64_t p = warp * kDotSlots;
  if (p >= nnz) return;
  const int64_t p_end = min(p + (int64_t)kDotSlots, nnz);
  const int64_t feat = heads * (int64_t)lph * VEC;
  const int chunks = (int)(heads * lph);
  int64_t lo = 0, hi = n_rows;                       // row of the first slot: last r with indptr[r] <= p
  while (hi - lo > 1) {
    const int64_t mid = (lo + hi) >> 1;
    if (indptr[mid] <= p) lo = mid; else hi = mid;
  }
  int64_t r = lo, r_end = indptr[r + 1];
  float vv[K][VEC];
  auto load_row = [&](const T* base, float (&dst)[K][VEC]) {
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const int c = k * 32 + lane;
      if (c < chunks) {
        Chunk<T, VEC>::load(base + (int64_t)c * VEC, dst[k]);
      } else {
#pragma unroll
        for (int t = 0; t < VEC; ++t) dst[k][t] = 0.f;
      }
    }
  };
  load_row(v + r * feat, vv);
  // `n` slots starting at p; called with the literal U for full batches so that the guards fold away
  auto batch = [&](const int n) {
    float uu[U][K][VEC];
#pragma unroll
    for (int j = 0; j < U; ++j)
      if (j < n) load_row(u + (int64_t)ld_stream_i32(indices + p + j) * feat, uu[j]);
#pragma unroll
    for (int j = 0; j < U; ++j) {
      if (j >= n) break;
      const int64_t q = p + j;
      if (q >= r_end) {                              // next non-empty row: its destination operand
        while (q >= r_end) { ++r; r_end = indptr[r + 1]; }
        load_row(v + r * feat, vv);
      }
      const int64_t e = eids ? (int64_t)eids[q] : q;
      if (LPH > 0) {
        T* orow = out + e * heads;
#pragma unroll
        for (int k = 0; k < K; ++k) {
          float acc = 0.f;
#pragma unroll
          for (int t = 0; t < VEC; ++t) acc = fmaf(uu[j][k][t], vv[k][t], acc);
#pragma unroll
          for (int o = LPH / 2; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
          const int c = k * 32 + lane;
          if (c < chunks && (c & (LPH - 1)) == 0) store_one<T>(orow + c / (LPH > 0 ? LPH : 1), acc);
        }
      } else {
        const int per_head = lph >> 5;               // registers of one head
        for (int h = 0; h < (int)heads; ++h) {
          float acc = 0.f;
#pragma unroll
          for (int k = 0; k < K; ++k) {
            if (k / per_head == h) {
#pragma unroll
              for (int t = 0; t < VEC; ++t) acc = fmaf(uu[j][k][t], vv[k][t], acc);
            }
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
          if (lane == 0) store_one<T>(out + e * heads + h, acc);
        }
      }
    }
    p += n;
  };
  while (p + U <= p_end) batch(U);
  if (p < p_end) batch((int)(p_end - p));
}

// Shapes the warp kernel takes: 16-byte chunks, a head = a power of two <= 32 of them or a multiple of 32, the
// row in at most 8 registers per lane.  Returns 0 when it launched, 1 when the generic kernel must do it.
template <typename T, int VEC>
int try_launch_sddmm_dot_warp(int64_t n_rows, int64_t nnz, const int64_t* indptr, const int32_t* indices,
                              const int32_t* eids, const T* u, const T* v, int64_t heads, int64_t dim,
                              T* out, cudaStream_t stream) {
  if (VEC == 1 || dim % VEC != 0 || nnz == 0) return 1;
  const int64_t lph = dim / VEC, chunks = heads * lph;
  const bool small = lph <= 32 && (lph & (lph - 1)) == 0;
  const bool large = lph > 32 && lph % 32 == 0;
  if (!(small || large) || chunks > 256 || lph < 2) return 1;
  const int k = (int)ceil_div(chunks, 32);
  const int64_t blocks = ceil_div(ceil_div(nnz, kDotSlots), kDotWarps);
  if (blocks >= (1ll << 31)) return 1;
#define DGLB200_DOT_CASE(KK, UU, LL)                                                                      \
  sddmm_dot_warp_kernel<T, VEC, KK, UU, LL><<<(unsigned)blocks, kDotWarps * 32, 0, stream>>>(              \
      n_rows, nnz, indptr, indices, eids, u, v, heads, (int)lph, out)
#define DGLB200_DOT_K(LL)                                                                                 \
  do {                                                                                                    \
    if (k == 1) DGLB200_DOT_CASE(1, 4, LL);                                                               \
    else if (k == 2) DGLB200_DOT_CASE(2, kBudget / 2, LL);                                                \
    else if (k <= 4) DGLB200_DOT_CASE(4, kBudget / 4, LL);                                                \
    else DGLB200_DOT_CASE(8, 1, LL);                                                                      \
  } while (0)
  constexpr int kBudget = VEC == 4 ? 8 : 4;          // chunk registers in flight per lane: U * K
  switch (small ? (int)lph : 0) {
    case 2: DGLB200_DOT_K(2); break;
    case 4: DGLB200_DOT_K(4); break;
    case 8: DGLB200_DOT_K(8); break;
    case 16: DGLB200_DOT_K(16); break;
    case 32: DGLB200_DOT_K(32); break;
    default: DGLB200_DOT_K(0); break;
  }
#undef DGLB200_DOT_K
#undef DGLB200_DOT_CASE
  DGLB200_POST_LAUNCH("sddmm_dot_warp_kernel");
  return 0;
}

template <typename T, int VEC, int OP>
int launch_sddmm_lpe(int64_t n_rows, int64_t nnz, const int64_t* indptr, const int32_t* indices,
                     const int32_t* eids, const T* u, const T* v, int64_t heads, int64_t dim,
                     T* out, cudaStream_t stream) {
  const int64_t per = ceil_div(dim, VEC);
  int lpe = 1;
  while (lpe < 32 && lpe < per) lpe <<= 1;
  const int64_t groups = ceil_div(nnz, kSlotsPerGroup);
  const int64_t blocks = ceil_div(groups * lpe, kSddmmThreads);
  DGLB200_REQUIRE(blocks < (1ll << 31), "sddmm: grid too large");
#define DGLB200_SDDMM_CASE(L)                                                                   \
  case L:                                                                                       \
    sddmm_kernel<T, VEC, L, OP><<<(unsigned)blocks, kSddmmThreads, 0, stream>>>(                 \
        n_rows, nnz, indptr, indices, eids, u, v, heads, dim, out);                              \
    break;
  switch (lpe) {
    DGLB200_SDDMM_CASE(1) DGLB200_SDDMM_CASE(2) DGLB200_SDDMM_CASE(4) DGLB200_SDDMM_CASE(8)
    DGLB200_SDDMM_CASE(16) DGLB200_SDDMM_CASE(32)
  }
#undef DGLB200_SDDMM_CASE
  DGLB200_POST_LAUNCH("sddmm_kernel");
  return 0;
}

template <typename T, int VECFULL>
int launch_sddmm(int64_t n_rows, int64_t nnz, const int64_t* indptr, const int32_t* indices,
                 const int32_t* eids, int op, const void* u_, const void* v_, int64_t heads,
                 int64_t dim, void* out_, cudaStream_t stream) {
  const T* u = static_cast<const T*>(u_);
  const T* v = static_cast<const T*>(v_);
  T* out = static_cast<T*>(out_);
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  const bool vec = dim % VECFULL == 0 && (u == nullptr || al(u)) && (v == nullptr || al(v)) &&
                   (op == DGLB200_SDDMM_DOT || al(out));
  if (op == DGLB200_SDDMM_DOT && vec && u != nullptr && v != nullptr && !sddmm_dot_generic()) {
    const int rc = try_launch_sddmm_dot_warp<T, VECFULL>(n_rows, nnz, indptr, indices, eids, u, v, heads, dim, out, stream);
    if (rc <= 0) return rc;
  }
#define DGLB200_SDDMM_OP(OPC)                                                                    \
  return vec ? launch_sddmm_lpe<T, VECFULL, OPC>(n_rows, nnz, indptr, indices, eids, u, v, heads, \
                                                 dim, out, stream)                               \
             : launch_sddmm_lpe<T, 1, OPC>(n_rows, nnz, indptr, indices, eids, u, v, heads, dim,  \
                                           out, stream)
  switch (op) {
    case DGLB200_SDDMM_DOT: DGLB200_SDDMM_OP(DGLB200_SDDMM_DOT);
    case DGLB200_SDDMM_ADD: DGLB200_SDDMM_OP(DGLB200_SDDMM_ADD);
    case DGLB200_SDDMM_MUL: DGLB200_SDDMM_OP(DGLB200_SDDMM_MUL);
    case DGLB200_SDDMM_COPY_U: DGLB200_SDDMM_OP(DGLB200_SDDMM_COPY_U);
    case DGLB200_SDDMM_COPY_V: DGLB200_SDDMM_OP(DGLB200_SDDMM_COPY_V);
  }
#undef DGLB200_SDDMM_OP
  set_error("sddmm: unknown op %d", op);
  return -1;
}

// ---- edge softmax ----------------------------------------------------------------------------
// One warp per destination row; heads are looped, lanes stride over the row's slots.

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Rows of H scores are contiguous (score[e, 0:H]): a lane owns whole slots and reads the H heads of a
// slot as ONE vector (H = 8: one 32-byte sector, every byte used; round 1 looped heads outermost and
// used 4 of every 32 bytes, three times).  A warp owns a destination row: rows of <= 32 slots keep
// their scores in registers (one read), longer rows make the max / sum / normalise passes over
// scores that stay in L1/L2; rows longer than kSoftmaxLongRow are left to the whole CTA (8 warps,
// shared-memory combine) so that a 22K-degree hub is not one warp's 700-iteration chain.
constexpr int kSoftmaxLongRow = 4096;
constexpr int kSoftmaxWarps = 8;

template <int H>
struct HeadVec {
  float v[H];
  __device__ __forceinline__ void load(const float* p) {
    if constexpr (H % 4 == 0) {
#pragma unroll
      for (int i = 0; i < H; i += 4) {
        const float4 t = *reinterpret_cast<const float4*>(p + i);
        v[i] = t.x; v[i + 1] = t.y; v[i + 2] = t.z; v[i + 3] = t.w;
      }
    } else if constexpr (H == 2) {
      const float2 t = *reinterpret_cast<const float2*>(p);
      v[0] = t.x; v[1] = t.y;
    } else {
#pragma unroll
      for (int i = 0; i < H; ++i) v[i] = p[i];
    }
  }
  __device__ __forceinline__ void store(float* p) const {
    if constexpr (H % 4 == 0) {
#pragma unroll
      for (int i = 0; i < H; i += 4) *reinterpret_cast<float4*>(p + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
    } else if constexpr (H == 2) {
      *reinterpret_cast<float2*>(p) = make_float2(v[0], v[1]);
    } else {
#pragma unroll
      for (int i = 0; i < H; ++i) p[i] = v[i];
    }
  }
};

// Combine a per-thread value over the `nthreads`-wide team (a warp, or the CTA through smem).
template <bool CTA, bool MAX>
__device__ __forceinline__ float team_reduce(float v, float* smem /*[kSoftmaxWarps + 1]*/) {
  v = MAX ? warp_max(v) : warp_sum(v);
  if constexpr (CTA) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();                      // smem free again
    if (lane == 0) smem[warp] = v;
    __syncthreads();
    float r = smem[0];
#pragma unroll
    for (int w = 1; w < kSoftmaxWarps; ++w) r = MAX ? fmaxf(r, smem[w]) : r + smem[w];
    v = r;
  }
  return v;
}

// One row by a team of lanes (tid in [0, nthreads)).  BWD == false: alpha = softmax(score);
// BWD == true: dscore = alpha * (dalpha - sum(alpha * dalpha)) with a = alpha, b = dalpha.
template <int H, bool CTA, bool BWD>
__device__ __forceinline__ void softmax_row(int64_t s, int64_t e, int tid, int nthreads,
                                            const int32_t* __restrict__ eids, const float* __restrict__ a,
                                            const float* __restrict__ b, float* __restrict__ out, float* smem) {
  auto eid_of = [&](int64_t p) { return eids ? (int64_t)eids[p] : p; };
  if (!CTA && e - s <= 32) {              // registers: one slot per lane
    const int64_t p = s + tid;
    const bool ok = p < e;
    const int64_t ed = ok ? eid_of(p) : 0;
    HeadVec<H> x, y;
    if (ok) x.load(a + ed * H);
    if (BWD && ok) y.load(b + ed * H);
#pragma unroll
    for (int h = 0; h < H; ++h) {
      if constexpr (!BWD) {
        const float m = warp_max(ok ? x.v[h] : -INFINITY);
        const float ex = ok ? expf(x.v[h] - m) : 0.f;
        const float z = warp_sum(ex);
        x.v[h] = ex / z;
      } else {
        const float dot = warp_sum(ok ? x.v[h] * y.v[h] : 0.f);
        x.v[h] = x.v[h] * (y.v[h] - dot);
      }
    }
    if (ok) x.store(out + ed * H);
    return;
  }
  float m[H], z[H];
#pragma unroll
  for (int h = 0; h < H; ++h) { m[h] = -INFINITY; z[h] = 0.f; }
  if constexpr (!BWD) {
    for (int64_t p = s + tid; p < e; p += nthreads) {
      HeadVec<H> x;
      x.load(a + eid_of(p) * H);
#pragma unroll
      for (int h = 0; h < H; ++h) m[h] = fmaxf(m[h], x.v[h]);
    }
#pragma unroll
    for (int h = 0; h < H; ++h) m[h] = team_reduce<CTA, true>(m[h], smem);
  }
  for (int64_t p = s + tid; p < e; p += nthreads) {
    const int64_t ed = eid_of(p);
    HeadVec<H> x, y;
    x.load(a + ed * H);
    if constexpr (BWD) y.load(b + ed * H);
#pragma unroll
    for (int h = 0; h < H; ++h) z[h] += BWD ? x.v[h] * y.v[h] : expf(x.v[h] - m[h]);
  }
#pragma unroll
  for (int h = 0; h < H; ++h) z[h] = team_reduce<CTA, false>(z[h], smem);
  for (int64_t p = s + tid; p < e; p += nthreads) {
    const int64_t ed = eid_of(p);
    HeadVec<H> x, y;
    x.load(a + ed * H);
    if constexpr (BWD) y.load(b + ed * H);
#pragma unroll
    for (int h = 0; h < H; ++h) x.v[h] = BWD ? x.v[h] * (y.v[h] - z[h]) : expf(x.v[h] - m[h]) / z[h];
    x.store(out + ed * H);
  }
}

template <int H, bool BWD>
__global__ void __launch_bounds__(kSoftmaxWarps * 32)
edge_softmax_rows_kernel(int64_t n_rows, const int64_t* __restrict__ indptr, const int32_t* __restrict__ eids,
                         const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out) {
  __shared__ float smem[kSoftmaxWarps + 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t row0 = (int64_t)blockIdx.x * kSoftmaxWarps;
  const int64_t r = row0 + warp;
  if (r < n_rows) {
    const int64_t s = indptr[r], e = indptr[r + 1];
    if (e > s && e - s <= kSoftmaxLongRow) softmax_row<H, false, BWD>(s, e, lane, 32, eids, a, b, out, smem);
  }
  // long rows of this CTA's window: the whole CTA, one row after another (uniform control flow)
  for (int w = 0; w < kSoftmaxWarps; ++w) {
    const int64_t rr = row0 + w;
    if (rr >= n_rows) break;
    const int64_t s = indptr[rr], e = indptr[rr + 1];
    if (e - s > kSoftmaxLongRow) softmax_row<H, true, BWD>(s, e, threadIdx.x, kSoftmaxWarps * 32, eids, a, b, out, smem);
  }
}

// generic head count (not 1, 2, 4, 8, 16): round 1's head-outermost kernel
__global__ void __launch_bounds__(256)
edge_softmax_kernel(int64_t n_rows, const int64_t* __restrict__ indptr,
                    const int32_t* __restrict__ eids, int64_t heads,
                    const float* __restrict__ score, float* __restrict__ alpha) {
  const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= n_rows) return;
  const int lane = threadIdx.x & 31;
  const int64_t s = indptr[r], e = indptr[r + 1];
  if (s == e) return;
  for (int64_t h = 0; h < heads; ++h) {
    float mx = -INFINITY;
    for (int64_t p = s + lane; p < e; p += 32) {
      const int64_t ed = eids ? (int64_t)eids[p] : p;
      mx = fmaxf(mx, score[ed * heads + h]);
    }
    mx = warp_max(mx);
    float sum = 0.f;
    for (int64_t p = s + lane; p < e; p += 32) {
      const int64_t ed = eids ? (int64_t)eids[p] : p;
      sum += expf(score[ed * heads + h] - mx);
    }
    sum = warp_sum(sum);
    for (int64_t p = s + lane; p < e; p += 32) {
      const int64_t ed = eids ? (int64_t)eids[p] : p;
      alpha[ed * heads + h] = expf(score[ed * heads + h] - mx) / sum;
    }
  }
}

__global__ void __launch_bounds__(256)
edge_softmax_backward_kernel(int64_t n_rows, const int64_t* __restrict__ indptr,
                             const int32_t* __restrict__ eids, int64_t heads,
                             const float* __restrict__ alpha, const float* __restrict__ dalpha,
                             float* __restrict__ dscore) {
  const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= n_rows) return;
  const int lane = threadIdx.x & 31;
  const int64_t s = indptr[r], e = indptr[r + 1];
  if (s == e) return;
  for (int64_t h = 0; h < heads; ++h) {
    float dot = 0.f;
    for (int64_t p = s + lane; p < e; p += 32) {
      const int64_t ed = eids ? (int64_t)eids[p] : p;
      dot += alpha[ed * heads + h] * dalpha[ed * heads + h];
    }
    dot = warp_sum(dot);
    for (int64_t p = s + lane; p < e; p += 32) {
      const int64_t ed = eids ? (int64_t)eids[p] : p;
      dscore[ed * heads + h] = alpha[ed * heads + h] * (dalpha[ed * heads + h] - dot);
    }
  }
}

template <bool BWD>
int launch_edge_softmax(int64_t n_rows, const int64_t* indptr, const int32_t* eids, int64_t heads,
                        const float* a, const float* b, float* out, cudaStream_t stream) {
  const bool al = ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(out) |
                    reinterpret_cast<uintptr_t>(b)) & 15) == 0;
  const unsigned blocks = (unsigned)ceil_div(n_rows, kSoftmaxWarps);
#define DGLB200_SOFTMAX_H(HH)                                                                          \
  if (heads == HH && (al || HH == 1)) {                                                                \
    edge_softmax_rows_kernel<HH, BWD><<<blocks, kSoftmaxWarps * 32, 0, stream>>>(n_rows, indptr, eids, a, b, out); \
    DGLB200_POST_LAUNCH("edge_softmax_rows_kernel");                                                   \
    return 0;                                                                                          \
  }
  DGLB200_SOFTMAX_H(1) DGLB200_SOFTMAX_H(2) DGLB200_SOFTMAX_H(4) DGLB200_SOFTMAX_H(8) DGLB200_SOFTMAX_H(16)
#undef DGLB200_SOFTMAX_H
  if constexpr (!BWD) {
    edge_softmax_kernel<<<(unsigned)ceil_div(n_rows * 32, 256), 256, 0, stream>>>(n_rows, indptr, eids, heads, a, out);
    DGLB200_POST_LAUNCH("edge_softmax_kernel");
  } else {
    edge_softmax_backward_kernel<<<(unsigned)ceil_div(n_rows * 32, 256), 256, 0, stream>>>(n_rows, indptr, eids, heads, a, b, out);
    DGLB200_POST_LAUNCH("edge_softmax_backward_kernel");
  }
  return 0;
}

// ---- fused GAT aggregation -------------------------------------------------------------------
// One launch for apply_edges(edge_attention) + update_all([src_mul_edge, copy_edge],[sum,sum]) +
// ft / z of examples/pytorch/gat/train.py:61-97.  A group of G lanes owns a destination row;
// lane chunk = VEC features of one head.  Pass 1 (STABLE only) finds the per-head row maximum of
// the scores, pass 2 accumulates a * ft[src] and z = sum a in slot order, pass 3 (optional)
// writes the per-edge attention by edge id.

__device__ __forceinline__ float gat_score(float as, float ad, float slope) {
  const float s = as + ad;
  return s >= 0.f ? s : s * slope;
}

template <int VEC, int G, int TILES>
__global__ void __launch_bounds__(128)
gat_aggregate_kernel(int64_t n_rows, const int64_t* __restrict__ indptr,
                     const int32_t* __restrict__ indices, const int32_t* __restrict__ eids,
                     const float* __restrict__ a_src, const float* __restrict__ a_dst,
                     const float* __restrict__ ft, int64_t heads, int64_t dim, int64_t ld_h, float slope,
                     int flavour, float* __restrict__ out, float* __restrict__ a_out,
                     float* __restrict__ z_out) {
  // `heads` = the heads of THIS pass (a head slab; the pointers are already shifted to its first head), ld_h = heads
  // of the whole operands: rows of ft / out are ld_h * dim apart, rows of a_src / a_dst / z / a_out ld_h apart.
  const int64_t ld_f = ld_h * dim;
  constexpr int RPW = 32 / G;
  const int lane = threadIdx.x & 31;
  const int gl = lane % G, grp = lane / G;
  const int64_t row = ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * RPW + grp;
  const int64_t feat = heads * dim;
  const bool rvalid = row < n_rows;
  const int64_t start = rvalid ? indptr[row] : 0;
  const int64_t end = rvalid ? indptr[row + 1] : 0;
  int64_t foff[TILES];
  int64_t head[TILES];
  bool fvalid[TILES];
  float adst[TILES], mx[TILES], z[TILES];
  float acc[TILES][VEC];
#pragma unroll
  for (int t = 0; t < TILES; ++t) {
    foff[t] = (int64_t)(t * G + gl) * VEC;
    fvalid[t] = rvalid && foff[t] < feat;
    head[t] = fvalid[t] ? foff[t] / dim : 0;
    adst[t] = fvalid[t] ? a_dst[row * ld_h + head[t]] : 0.f;
    mx[t] = 0.f;
    z[t] = 0.f;
#pragma unroll
    for (int v = 0; v < VEC; ++v) acc[t][v] = 0.f;
  }
  if (flavour == DGLB200_SOFTMAX_STABLE) {
#pragma unroll
    for (int t = 0; t < TILES; ++t) mx[t] = -INFINITY;
    // The dim / VEC lanes that share a head would all compute the same row maximum: they split the row's slots
    // instead (lane j of a head takes slots j, j + lph, ...) and combine with a butterfly - lph times less work in
    // this pass (8 x at 8 heads x 32), the same maximum bit for bit.
    const int64_t lph = dim / VEC;
    const bool split = dim % VEC == 0 && lph >= 2 && lph <= G && (lph & (lph - 1)) == 0;
    if (split) {
      const int64_t j = gl & (lph - 1);                   // (t * G + gl) % lph for every tile: lph | G
      for (int64_t p = start + j; p < end; p += 4 * lph) {
        int64_t s4[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) s4[u] = (p + u * lph < end) ? (int64_t)indices[p + u * lph] : -1;
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
          for (int t = 0; t < TILES; ++t)
            if (s4[u] >= 0 && fvalid[t]) mx[t] = fmaxf(mx[t], gat_score(a_src[s4[u] * ld_h + head[t]], adst[t], slope));
      }
    }
    for (int o = (int)(split ? lph >> 1 : 0); o > 0; o >>= 1) {
#pragma unroll
      for (int t = 0; t < TILES; ++t) mx[t] = fmaxf(mx[t], __shfl_xor_sync(0xffffffffu, mx[t], o));
    }
    for (int64_t p = split ? end : start; p < end; p += 4) {   // 4 independent id -> score chains per step
      int64_t s4[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) s4[u] = (p + u < end) ? (int64_t)indices[p + u] : -1;
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int t = 0; t < TILES; ++t)
          if (s4[u] >= 0 && fvalid[t]) mx[t] = fmaxf(mx[t], gat_score(a_src[s4[u] * ld_h + head[t]], adst[t], slope));
    }
  }
  // Main pass, U edges per step: all U column ids, then all U x TILES score + feature loads are issued
  // before anything is consumed (round 1 walked one edge at a time: every iteration waited a full
  // id -> row dependent-load chain, 53 ms at the C2-sized 8-head case against 32 ms for the three unfused
  // kernels).  Accumulation stays in slot order.
  // Full steps go through a copy of the body called with the literal U, so the `u < n` guards fold away and the
  // U x TILES loads really issue back to back (the same change took the SDDMM dot from 29.9 to 18.6 ms).
  constexpr int U = TILES >= 4 ? 2 : 4;
  auto step = [&](const int64_t p, const int n) {
    int64_t sidx[U];
    float as[U][TILES];
    float xv[U][TILES][VEC];
#pragma unroll
    for (int u = 0; u < U; ++u) sidx[u] = (u < n) ? (int64_t)indices[p + u] : 0;
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int t = 0; t < TILES; ++t) {
        as[u][t] = 0.f;
#pragma unroll
        for (int v = 0; v < VEC; ++v) xv[u][t][v] = 0.f;
        if (u < n && fvalid[t]) {
          as[u][t] = a_src[sidx[u] * ld_h + head[t]];
          Chunk<float, VEC>::load(ft + sidx[u] * ld_f + foff[t], xv[u][t]);
        }
      }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (u >= n) continue;
#pragma unroll
      for (int t = 0; t < TILES; ++t) {
        if (!fvalid[t]) continue;
        const float sc = gat_score(as[u][t], adst[t], slope);
        float a;
        if (flavour == DGLB200_SOFTMAX_STABLE) a = expf(sc - mx[t]);
        else a = fminf(fmaxf(expf(sc), -10.f), 10.f);
        z[t] += a;
#pragma unroll
        for (int v = 0; v < VEC; ++v) acc[t][v] = fmaf(a, xv[u][t][v], acc[t][v]);
      }
    }
  };
  {
    int64_t p = start;
    for (; p + U <= end; p += U) step(p, U);
    if (p < end) step(p, (int)(end - p));
  }
#pragma unroll
  for (int t = 0; t < TILES; ++t) {
    if (!fvalid[t]) continue;
    float o[VEC];
#pragma unroll
    for (int v = 0; v < VEC; ++v) o[v] = (end > start) ? acc[t][v] / z[t] : 0.f;
    Chunk<float, VEC>::store(out + row * ld_f + foff[t], o);
    if (z_out != nullptr && foff[t] % dim == 0) z_out[row * ld_h + head[t]] = z[t];
  }
  if (a_out != nullptr) {
    for (int64_t p = start; p < end; ++p) {
      const int64_t s = indices[p];
      const int64_t e = eids ? (int64_t)eids[p] : p;
#pragma unroll
      for (int t = 0; t < TILES; ++t) {
        if (!fvalid[t] || foff[t] % dim != 0) continue;  // one writer per head
        const float sc = gat_score(a_src[s * ld_h + head[t]], adst[t], slope);
        a_out[e * ld_h + head[t]] = (flavour == DGLB200_SOFTMAX_STABLE)
                                         ? expf(sc - mx[t]) / z[t]
                                         : fminf(fmaxf(expf(sc), -10.f), 10.f);
      }
    }
  }
}

template <int VEC>
int launch_gat(int64_t n_rows, const int64_t* indptr, const int32_t* indices, const int32_t* eids,
               const float* a_src, const float* a_dst, const float* ft, int64_t heads, int64_t dim,
               float slope, int flavour, float* out, float* a_out, float* z_out,
               cudaStream_t stream) {
  // Head slabs (dglb200_set_tuning("gat_head_slab", k)): a pass per slab of k heads keeps the rows it gathers
  // L2-resident, like the feature slabs of kernel (1); heads are independent (their own score, z and output columns),
  // so the passes write disjoint columns and the results are bit-identical.  OFF by default: at C2 size, 8 heads x 32
  // (239 MB operand), 4 passes of 2 heads took 38.8 ms against 22.7 ms for one pass - unlike the SpMM, every pass
  // re-does the per-edge work (column id, score, exp), and that, not the gather, is what the kernel is bound by.
  int64_t slab = heads;
  if (gat_head_slab() > 0) slab = gat_head_slab();
  if (slab < 1) slab = 1;
  if (slab > heads) slab = heads;
  if (VEC > 1 && (slab * dim) % VEC != 0) slab = heads;
  for (int64_t h0 = 0; h0 < heads; h0 += slab) {
    const int64_t hs = h0 + slab < heads ? slab : heads - h0;
    const int64_t chunks = ceil_div(hs * dim, VEC);
    DGLB200_REQUIRE(chunks <= 128, "gat_aggregate: heads*dim = %lld too wide for the fused kernel "
                    "(max %d); use sddmm + edge_softmax + spmm", (long long)(hs * dim), 128 * VEC);
#define DGLB200_GAT(G_, TL)                                                                      \
  {                                                                                              \
    const int64_t rows_per_block = 4 * (32 / G_);                                                \
    gat_aggregate_kernel<VEC, G_, TL><<<(unsigned)ceil_div(n_rows, rows_per_block), 128, 0,      \
                                        stream>>>(n_rows, indptr, indices, eids, a_src + h0,     \
                                                  a_dst + h0, ft + h0 * dim, hs, dim, heads,     \
                                                  slope, flavour, out + h0 * dim,                \
                                                  a_out ? a_out + h0 : nullptr,                  \
                                                  z_out ? z_out + h0 : nullptr);                 \
  }
    if (chunks <= 4) DGLB200_GAT(4, 1)
    else if (chunks <= 8) DGLB200_GAT(8, 1)
    else if (chunks <= 16) DGLB200_GAT(16, 1)
    else if (chunks <= 32) DGLB200_GAT(32, 1)
    else if (chunks <= 64) DGLB200_GAT(32, 2)
    else DGLB200_GAT(32, 4)
#undef DGLB200_GAT
    DGLB200_POST_LAUNCH("gat_aggregate_kernel");
  }
  return 0;
}

}  // namespace dglb200

using namespace dglb200;

extern "C" {

int dglb200_sddmm_csr(int64_t n_rows, int64_t nnz, const int64_t* indptr, const int32_t* indices,
                      const int32_t* eids, int op, int dtype, const void* u, const void* v,
                      int64_t heads, int64_t dim, void* out, void* stream) {
  clear_error();
  DGLB200_REQUIRE(n_rows >= 0 && nnz >= 0 && heads >= 1 && dim >= 1, "sddmm_csr: bad sizes");
  DGLB200_REQUIRE(n_rows < (1ll << 31) && nnz < (1ll << 31), "sddmm_csr: sizes must be < 2^31");
  DGLB200_REQUIRE(dtype == DGLB200_F32 || dtype == DGLB200_BF16, "sddmm_csr: unknown dtype %d", dtype);
  if (nnz == 0) return 0;
  DGLB200_REQUIRE(indptr && indices && out, "sddmm_csr: NULL pointer");
  DGLB200_REQUIRE(u != nullptr || op == DGLB200_SDDMM_COPY_V, "sddmm_csr: u is NULL");
  DGLB200_REQUIRE(v != nullptr || op == DGLB200_SDDMM_COPY_U, "sddmm_csr: v is NULL");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (op != DGLB200_SDDMM_DOT) {   // element-wise ops ignore head boundaries: one flat row of heads * dim
    dim *= heads;                  // elements, so (E, 8, 1) GAT scores move as 16-byte vectors too
    heads = 1;
  }
  return dtype == DGLB200_F32
             ? launch_sddmm<float, 4>(n_rows, nnz, indptr, indices, eids, op, u, v, heads, dim, out, st)
             : launch_sddmm<__nv_bfloat16, 8>(n_rows, nnz, indptr, indices, eids, op, u, v, heads, dim, out, st);
}

int dglb200_edge_softmax_csr(int64_t n_rows, int64_t nnz, const int64_t* indptr,
                             const int32_t* eids, int64_t heads, const float* score, float* alpha,
                             void* stream) {
  clear_error();
  DGLB200_REQUIRE(n_rows >= 0 && nnz >= 0 && heads >= 1, "edge_softmax_csr: bad sizes");
  if (nnz == 0 || n_rows == 0) return 0;
  DGLB200_REQUIRE(indptr && score && alpha, "edge_softmax_csr: NULL pointer");
  return launch_edge_softmax<false>(n_rows, indptr, eids, heads, score, nullptr, alpha, static_cast<cudaStream_t>(stream));
}

int dglb200_edge_softmax_backward_csr(int64_t n_rows, int64_t nnz, const int64_t* indptr,
                                      const int32_t* eids, int64_t heads, const float* alpha,
                                      const float* dalpha, float* dscore, void* stream) {
  clear_error();
  DGLB200_REQUIRE(n_rows >= 0 && nnz >= 0 && heads >= 1, "edge_softmax_backward_csr: bad sizes");
  if (nnz == 0 || n_rows == 0) return 0;
  DGLB200_REQUIRE(indptr && alpha && dalpha && dscore, "edge_softmax_backward_csr: NULL pointer");
  return launch_edge_softmax<true>(n_rows, indptr, eids, heads, alpha, dalpha, dscore, static_cast<cudaStream_t>(stream));
}

int dglb200_gat_aggregate_csr(int64_t n_rows, int64_t nnz, const int64_t* indptr,
                              const int32_t* indices, const int32_t* eids, const float* a_src,
                              const float* a_dst, const float* ft, int64_t heads, int64_t dim,
                              float negative_slope, int flavour, float* out, float* a_out,
                              float* z_out, void* stream) {
  clear_error();
  DGLB200_REQUIRE(n_rows >= 0 && nnz >= 0 && heads >= 1 && dim >= 1, "gat_aggregate_csr: bad sizes");
  DGLB200_REQUIRE(flavour == DGLB200_SOFTMAX_STABLE || flavour == DGLB200_SOFTMAX_GAT_EXAMPLE,
                  "gat_aggregate_csr: unknown flavour %d", flavour);
  if (n_rows == 0) return 0;
  DGLB200_REQUIRE(indptr && out && a_dst && (nnz == 0 || (indices && a_src && ft)),
                  "gat_aggregate_csr: NULL pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool vec = dim % 4 == 0 && ((reinterpret_cast<uintptr_t>(ft) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
  return vec ? launch_gat<4>(n_rows, indptr, indices, eids, a_src, a_dst, ft, heads, dim,
                             negative_slope, flavour, out, a_out, z_out, st)
             : launch_gat<1>(n_rows, indptr, indices, eids, a_src, a_dst, ft, heads, dim,
                             negative_slope, flavour, out, a_out, z_out, st);
}

}  // extern "C"
