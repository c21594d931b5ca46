// C-ABI entry points of the SpMM path: plan (long-row split table), forward, backward-x,
// and the slice fix-up kernel.  See include/dgl_b200.h for the contracts and
// spmm_kernel.cuh for the kernel design.
#include <mutex>
#include <string>

#include "scan.cuh"
#include "spmm_kernel.cuh"

namespace dglb200 {

// ---- long-row plan ---------------------------------------------------------------------------

__global__ void plan_flag_rows(const int64_t* indptr, int64_t n_rows, int chunk, int32_t* is_long,
                               int64_t* n_slices) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_rows) return;
  const int64_t d = indptr[r + 1] - indptr[r];
  const bool lg = d > chunk;
  is_long[r] = lg ? 1 : 0;
  n_slices[r] = lg ? (d + chunk - 1) / chunk : 0;
}

__global__ void plan_emit(const int32_t* is_long, const int32_t* long_pos, const int64_t* slice_pos,
                          int64_t n_rows, int64_t capacity, int32_t* long_rows, int64_t* chunk_ptr) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r > n_rows) return;
  if (r == n_rows) {  // terminator
    const int32_t nl = long_pos[n_rows];
    if (nl <= capacity) chunk_ptr[nl] = slice_pos[n_rows];
    return;
  }
  if (is_long[r]) {
    const int32_t k = long_pos[r];
    if (k < capacity) {
      long_rows[k] = (int32_t)r;
      chunk_ptr[k] = slice_pos[r];
    }
  }
}

// y[row, f] = ordered reduce of the row's slices.  One thread per (long row, feature).
template <typename T, int RED>
__global__ void spmm_fixup(const int32_t* long_rows, const int64_t* chunk_ptr, int n_long,
                           const float* partial, const int32_t* partial_arg, int64_t feat, int64_t y_ld,
                           T* y, int32_t* arg, int accumulate, const float* row_scale,
                           const float* bias, int relu) {
  const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int lr = blockIdx.y;
  if (f >= feat || lr >= n_long) return;
  const int64_t c0 = chunk_ptr[lr], c1 = chunk_ptr[lr + 1];
  float acc = (RED == DGLB200_RED_MAX) ? -INFINITY : 0.f;
  int32_t best = -1;
  for (int64_t c = c0; c < c1; ++c) {
    const float v = partial[c * feat + f];
    if (RED == DGLB200_RED_SUM) {
      acc += v;
    } else if (c == c0 || v > acc || (v != v && acc == acc)) {  // NaN is sticky, as in the kernel
      acc = v;
      if (partial_arg) best = partial_arg[c * feat + f];
    }
  }
  const int64_t row = long_rows[lr];
  if (RED == DGLB200_RED_SUM && accumulate) {
    if (sizeof(T) == 4) acc += reinterpret_cast<float*>(y)[row * y_ld + f];
    else acc += __bfloat162float(reinterpret_cast<__nv_bfloat16*>(y)[row * y_ld + f]);
  }
  if (RED == DGLB200_RED_MAX && arg) arg[row * feat + f] = best;
  if (RED == DGLB200_RED_SUM && (row_scale != nullptr || bias != nullptr || relu)) {
    acc = acc * (row_scale ? row_scale[row] : 1.f) + (bias ? bias[f] : 0.f);
    if (relu) acc = fmaxf(acc, 0.f);
  }
  if (sizeof(T) == 4) reinterpret_cast<float*>(y)[row * y_ld + f] = acc;
  else reinterpret_cast<__nv_bfloat16*>(y)[row * y_ld + f] = __float2bfloat16_rn(acc);
}

// out[p, :] = w[eids[p], :]: edge-id-ordered edge data -> CSR slot order (see
// dglb200_edge_data_to_slots).  One thread per (slot, UNIT-byte element); four independent gathers
// in flight per thread; id reads (L1-broadcast across a row's elements) and stores are coalesced.
template <typename V>
__global__ void __launch_bounds__(256)
edge_data_to_slots_kernel(int64_t nnz, const int32_t* __restrict__ eids, int64_t n_edges, int units,
                          const V* __restrict__ w, V* __restrict__ out, int32_t* __restrict__ bad) {
  constexpr int K = 4;
  const int64_t total = nnz * units;
  const int64_t base = ((int64_t)blockIdx.x * K) * blockDim.x + threadIdx.x;
  int64_t srcoff[K];
  V v[K];
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const int64_t i = base + (int64_t)k * blockDim.x;
    srcoff[k] = -1;
    if (i < total) {
      const int64_t p = i / units;
      const int32_t e = __ldg(eids + p);
      if (e < 0 || e >= n_edges) {
        if (bad != nullptr) atomicExch(bad, 1);
      } else {
        srcoff[k] = (int64_t)e * units + (i - p * units);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < K; ++k)
    if (srcoff[k] >= 0) v[k] = w[srcoff[k]];
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const int64_t i = base + (int64_t)k * blockDim.x;
    if (srcoff[k] >= 0) out[i] = v[k];
  }
}

void set_peer_push_tma(int v);   // peer.cu
void set_sddmm_dot_generic(int v);  // sddmm.cu
void set_gat_head_slab(int v);      // sddmm.cu

SpmmTuning& spmm_tuning() {
  static SpmmTuning t;
  return t;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// Bytes of L2 that one feature slab of the gathered operand may occupy (about half of the L2:
// the structure arrays and the output stream through the rest).
static size_t l2_slab_budget() {
  static size_t budget = 0;
  if (budget == 0) {
    int dev = 0, l2 = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, dev) != cudaSuccess || l2 <= 0) {
      (void)cudaGetLastError();
      l2 = 126 << 20;
    }
    budget = (size_t)l2 / 20 * 11;
  }
  return budget;
}

// Slab width (in 16-byte chunks) for a gather over `x_rows` source rows of `row_bytes` each with
// `nnz` gathers.  Whole row when the operand already fits in L2, when reuse is too low to matter
// (DRAM-bound: C5-like graphs) or when not even a 64-byte slab would fit.  Measured on C2
// (233K x 256 fp32, mean degree 492): 64 chunks 10.7 ms, 32: 9.5 ms, 16: 7.9 ms.
static int pick_slab_chunks(int64_t x_rows, int64_t nnz, size_t row_bytes) {
  if (x_rows <= 0 || nnz < 8 * x_rows) return 0;
  const size_t budget = l2_slab_budget();
  if ((size_t)x_rows * row_bytes <= budget) return 0;
  for (int sc = 32; sc >= 4; sc >>= 1)
    if ((size_t)x_rows * sc * 16 <= budget && (size_t)sc * 16 < row_bytes) return sc;
  return 0;
}

static int run_spmm(SpmmArgs a, int msg, int red, int dtype, const dglb200_spmm_plan_t* plan,
                    int64_t x_rows, int64_t nnz, cudaStream_t stream) {
  a.nnz = nnz;
  a.chunk = 1 << 30;
  a.n_long_rows = 0;
  a.n_chunks = 0;
  a.long_rows = nullptr;
  a.chunk_ptr = nullptr;
  a.partial = nullptr;
  a.partial_arg = nullptr;
  if (plan != nullptr && plan->n_long_rows > 0) {
    DGLB200_REQUIRE(plan->chunk >= 32 && plan->chunk % 32 == 0, "spmm: plan chunk %d invalid",
                    plan->chunk);
    DGLB200_REQUIRE(plan->long_rows && plan->chunk_ptr && plan->partial,
                    "spmm: plan with long rows needs long_rows, chunk_ptr and partial");
    const size_t need = (size_t)plan->n_chunks * (size_t)a.feat * sizeof(float);
    DGLB200_REQUIRE(plan->partial_bytes >= need, "spmm: plan partial buffer %zu < %zu bytes",
                    plan->partial_bytes, need);
    a.chunk = plan->chunk;
    a.n_long_rows = plan->n_long_rows;
    a.n_chunks = plan->n_chunks;
    a.long_rows = plan->long_rows;
    a.chunk_ptr = plan->chunk_ptr;
    a.partial = static_cast<float*>(plan->partial);
    a.partial_arg = (red == DGLB200_RED_MAX) ? plan->partial_arg : nullptr;
    DGLB200_REQUIRE(!(red == DGLB200_RED_MAX && a.arg != nullptr && a.partial_arg == nullptr),
                    "spmm: MAX with arg over a plan needs plan->partial_arg");
  }
  const int full_vec = dtype == DGLB200_F32 ? 4 : 8;
  if (a.x_ld == 0) a.x_ld = a.feat;
  if (a.y_ld == 0) a.y_ld = a.feat;
  // Vector (16-byte) path: gathered rows must start 16-byte aligned.  A feature width that is not
  // a multiple of the vector still qualifies for copy_src / scalar-weight messages when the rows of
  // x are padded (x_ld a multiple of the vector: the ragged last chunk reads the row's own padding);
  // its outputs are then stored element-wise.  ops.spmm pads 602-wide rows to 604 this way.
  const bool x_ok = aligned16(a.x) && a.x_ld % full_vec == 0;
  a.y_vec = aligned16(a.y) && a.y_ld % full_vec == 0;
  const bool ragged = a.feat % full_vec != 0;
  const bool ragged_ok = msg == DGLB200_MSG_COPY_SRC ||
                         (msg == DGLB200_MSG_SRC_MUL_EDGE && a.w_mode == kWScalar);
  bool vec = ragged ? (ragged_ok && x_ok && a.x_ld >= a.feat + 1) : (x_ok && a.y_vec);
  if (a.w != nullptr && (a.w_mode == kWFull || msg == DGLB200_MSG_COPY_EDGE)) vec = vec && aligned16(a.w);
  a.epi = (a.accumulate ? 1 : 0) | ((a.row_scale || a.bias || a.relu) ? 2 : 0);
  DGLB200_REQUIRE(!(a.accumulate && red != DGLB200_RED_SUM), "spmm: accumulate needs the SUM reducer");
  DGLB200_REQUIRE(!((a.row_scale || a.bias || a.relu) && red != DGLB200_RED_SUM),
                  "spmm: the row_scale / bias / relu epilogue needs the SUM reducer");
  if (a.x_halo != nullptr) {
    DGLB200_REQUIRE(dtype == DGLB200_F32 && msg == DGLB200_MSG_COPY_SRC && red == DGLB200_RED_SUM,
                    "spmm: a bf16 halo operand needs fp32 copy_src -> sum");
    DGLB200_REQUIRE(vec && !ragged && aligned16(a.x_halo) && a.feat % 8 == 0 && a.halo_first >= 0,
                    "spmm: a bf16 halo operand needs 16-byte aligned operands and feat a multiple of 8");
  }
  a.slab_chunks = 0;
  if (vec && msg != DGLB200_MSG_COPY_EDGE && x_rows > 0)
    a.slab_chunks = pick_slab_chunks(x_rows, nnz, (size_t)a.feat * (dtype == DGLB200_F32 ? 4 : 2));
  // Per-head weights (E, H) on H x D features: when the passes can be cut so that none straddles
  // two heads (a power-of-two number of 16-byte chunks dividing D), the scalar-weight kernel runs
  // with w[e, head of the pass]; otherwise the per-chunk / generic variants take over.
  a.w_slab_scalar = 0;
  // Which of the two wins is a matter of passes: a scalar-weight pass reads a whole 32-byte sector of the (E, H)
  // weights per edge for ONE float, and a head of 4 chunks means 8 such passes at 8 heads.  C2 size, 8 heads
  // (scripts/head_pass_probe.py): dim 16 - 10.2 ms scalar passes vs 7.7 ms chunk weights over the usual slabs;
  // dim 32 - 13.0 vs 15.3 ms.  So: scalar passes from 8 chunks per head up ("spmm_head_pass" 1 / 2 forces either).
  const int64_t head_chunks_all = a.w_group / full_vec;
  const bool scalar_passes = spmm_tuning().head_pass == 1 || (spmm_tuning().head_pass == 0 && head_chunks_all >= 8);
  if (vec && msg == DGLB200_MSG_SRC_MUL_EDGE && a.w_mode == kWGrouped && a.w_group % full_vec == 0 &&
      scalar_passes) {
    const int64_t head_chunks = a.w_group / full_vec;
    const int cap = a.slab_chunks > 0 ? a.slab_chunks : 64;
    int p = 64;
    while (p >= 4 && (p > cap || head_chunks % p != 0)) p >>= 1;
    if (p >= 4) {
      a.slab_chunks = p;
      a.w_slab_scalar = 1;
    }
  }
  int rc = dtype == DGLB200_F32 ? launch_spmm_f32(a, msg, red, vec, stream)
                                : launch_spmm_bf16(a, msg, red, vec, stream);
  if (rc != 0) return rc;
  if (a.n_long_rows > 0) {
    dim3 grid((unsigned)ceil_div(a.feat, 128), (unsigned)a.n_long_rows);
    DGLB200_REQUIRE(a.n_long_rows < 65536 * 1024, "spmm: too many long rows");
    // grid.y is limited to 65535: fold the overflow into several launches
    for (int base = 0; base < a.n_long_rows; base += 65535) {
      const int n = min(65535, a.n_long_rows - base);
      grid.y = (unsigned)n;
#define FIX(T, R)                                                                          \
  spmm_fixup<T, R><<<grid, 128, 0, stream>>>(a.long_rows + base, a.chunk_ptr + base, n,    \
                                             a.partial, a.partial_arg, a.feat, a.y_ld,     \
                                             static_cast<T*>(a.y), a.arg, a.accumulate, a.row_scale, a.bias, a.relu)
      if (dtype == DGLB200_F32) {
        if (red == DGLB200_RED_SUM) FIX(float, DGLB200_RED_SUM); else FIX(float, DGLB200_RED_MAX);
      } else {
        if (red == DGLB200_RED_SUM) FIX(__nv_bfloat16, DGLB200_RED_SUM);
        else FIX(__nv_bfloat16, DGLB200_RED_MAX);
      }
#undef FIX
      DGLB200_POST_LAUNCH("spmm_fixup");
    }
  }
  return 0;
}

// ---- scratch of the allocation-free entry point (dglb200_spmm_csr_static_scratch) -----------------------
// The long-row slices of an SpMM write fp32 partials that the fix-up kernel combines: normally the caller
// hands that scratch in (plan->partial).  Here it is a file-scope device array, so the call allocates
// nothing.  Every GPU that loads the library holds these DGLB200_STATIC_SCRATCH_BYTES until exit.
static __device__ __align__(256) float g_static_partial[DGLB200_STATIC_SCRATCH_BYTES / sizeof(float)];

constexpr int kMaxScratchDevices = 64;

// Device address of g_static_partial on the CURRENT device (in host code the array's name is only a
// host-side stand-in, not a device address): resolved once per device with that device current.
static int static_partial_address(float** out) {
  static std::mutex mu;
  static float* addr[kMaxScratchDevices] = {nullptr};
  int dev = -1;
  DGLB200_CUDA(cudaGetDevice(&dev));
  DGLB200_REQUIRE(dev >= 0 && dev < kMaxScratchDevices, "static scratch: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (addr[dev] == nullptr) {
    void* p = nullptr;
    DGLB200_CUDA(cudaGetSymbolAddress(&p, g_static_partial));
    DGLB200_REQUIRE(p != nullptr, "static scratch: cudaGetSymbolAddress returned NULL");
    addr[dev] = static_cast<float*>(p);
  }
  *out = addr[dev];
  return 0;
}

}  // namespace dglb200

using namespace dglb200;

extern "C" {

int dglb200_set_tuning(const char* key, int64_t value) {
  clear_error();
  DGLB200_REQUIRE(key != nullptr, "set_tuning: key is NULL");
  SpmmTuning& t = spmm_tuning();
  const std::string k(key);
  if (k == "spmm_ucap") t.ucap = (int)value;
  else if (k == "spmm_slab_chunks") t.slab_chunks = (int)value;
  else if (k == "spmm_threads") t.threads = (int)value;
  else if (k == "spmm_rows_per_group") t.rows_per_group = (int)value;
  else if (k == "spmm_head_pass") t.head_pass = (int)value;
  else if (k == "peer_push_tma") set_peer_push_tma((int)value);
  else if (k == "sddmm_dot_generic") set_sddmm_dot_generic((int)value);
  else if (k == "gat_head_slab") set_gat_head_slab((int)value);
  else {
    set_error("set_tuning: unknown key '%s'", key);
    return -1;
  }
  return 0;
}

int dglb200_spmm_plan_workspace_bytes(int64_t n_rows, size_t* bytes_host) {
  DGLB200_REQUIRE(bytes_host != nullptr && n_rows >= 0, "spmm_plan_workspace_bytes: bad args");
  size_t b = 0;
  b += align_up((size_t)(n_rows + 1) * sizeof(int32_t)) * 2;  // is_long, long_pos
  b += align_up((size_t)(n_rows + 1) * sizeof(int64_t)) * 2;  // n_slices, slice_pos
  b += scan_workspace_bytes(n_rows + 1, sizeof(int64_t)) * 2;
  *bytes_host = b + 1024;
  return 0;
}

int dglb200_spmm_plan(int64_t n_rows, const int64_t* indptr, int32_t chunk, int64_t capacity,
                      int32_t* long_rows, int64_t* chunk_ptr, int64_t* counts_host,
                      void* workspace, size_t workspace_bytes, void* stream_) {
  clear_error();
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  DGLB200_REQUIRE(n_rows >= 0 && n_rows < (1ll << 31), "spmm_plan: n_rows %lld out of range",
                  (long long)n_rows);
  DGLB200_REQUIRE(chunk >= 32 && chunk % 32 == 0, "spmm_plan: chunk must be a multiple of 32");
  DGLB200_REQUIRE(counts_host != nullptr, "spmm_plan: counts_host is NULL");
  counts_host[0] = counts_host[1] = 0;
  if (n_rows == 0) return 0;
  Workspace ws(workspace, workspace_bytes);
  int32_t* is_long = ws.take<int32_t>(n_rows + 1);
  int32_t* long_pos = ws.take<int32_t>(n_rows + 1);
  int64_t* n_slices = ws.take<int64_t>(n_rows + 1);
  int64_t* slice_pos = ws.take<int64_t>(n_rows + 1);
  DGLB200_REQUIRE(is_long && long_pos && n_slices && slice_pos, "spmm_plan: workspace too small");
  const unsigned blocks = (unsigned)ceil_div(n_rows + 1, 256);
  plan_flag_rows<<<blocks, 256, 0, stream>>>(indptr, n_rows, chunk, is_long, n_slices);
  DGLB200_POST_LAUNCH("plan_flag_rows");
  if (exclusive_scan<int32_t, int32_t>(is_long, long_pos, n_rows, 1, ws, stream) != 0) return -1;
  if (exclusive_scan<int64_t, int64_t>(n_slices, slice_pos, n_rows, 1, ws, stream) != 0) return -1;
  int32_t n_long = 0;
  int64_t n_sl = 0;
  DGLB200_CUDA(cudaMemcpyAsync(&n_long, long_pos + n_rows, sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
  DGLB200_CUDA(cudaMemcpyAsync(&n_sl, slice_pos + n_rows, sizeof(int64_t), cudaMemcpyDeviceToHost, stream));
  DGLB200_CUDA(cudaStreamSynchronize(stream));
  counts_host[0] = n_long;
  counts_host[1] = n_sl;
  if (n_long == 0) return 0;
  DGLB200_REQUIRE(long_rows != nullptr && chunk_ptr != nullptr && capacity >= n_long,
                  "spmm_plan: capacity %lld < %d long rows", (long long)capacity, n_long);
  plan_emit<<<blocks, 256, 0, stream>>>(is_long, long_pos, slice_pos, n_rows, capacity, long_rows,
                                        chunk_ptr);
  DGLB200_POST_LAUNCH("plan_emit");
  return 0;
}

int dglb200_spmm_csr(int64_t n_rows, int64_t x_rows, int64_t nnz, const int64_t* indptr,
                     const int32_t* indices, const int32_t* eids, int msg_op, int reduce_op,
                     int dtype, const void* x, int64_t feat, const void* w, int64_t w_width,
                     void* y, int32_t* arg, const dglb200_spmm_plan_t* plan, void* stream) {
  return dglb200_spmm_csr_strided(n_rows, x_rows, nnz, indptr, indices, eids, msg_op, reduce_op, dtype,
                                  x, feat, feat, w, w_width, y, feat, arg, plan, stream);
}

int dglb200_spmm_csr_strided(int64_t n_rows, int64_t x_rows, int64_t nnz, const int64_t* indptr,
                             const int32_t* indices, const int32_t* eids, int msg_op, int reduce_op,
                             int dtype, const void* x, int64_t x_ld, int64_t feat, const void* w,
                             int64_t w_width, void* y, int64_t y_ld, int32_t* arg,
                             const dglb200_spmm_plan_t* plan, void* stream) {
  return dglb200_spmm_csr_ext(n_rows, x_rows, nnz, indptr, indices, eids, msg_op, reduce_op, dtype, x, x_ld,
                              feat, w, w_width, y, y_ld, arg, plan, nullptr, stream);
}

int dglb200_spmm_csr_ext(int64_t n_rows, int64_t x_rows, int64_t nnz, const int64_t* indptr,
                         const int32_t* indices, const int32_t* eids, int msg_op, int reduce_op,
                         int dtype, const void* x, int64_t x_ld, int64_t feat, const void* w,
                         int64_t w_width, void* y, int64_t y_ld, int32_t* arg,
                         const dglb200_spmm_plan_t* plan, const dglb200_spmm_ext_t* ext, void* stream) {
  clear_error();
  DGLB200_REQUIRE(x_ld >= feat && y_ld >= feat, "spmm_csr: leading dimensions (%lld, %lld) < feat %lld",
                  (long long)x_ld, (long long)y_ld, (long long)feat);
  DGLB200_REQUIRE(n_rows >= 0 && nnz >= 0 && feat >= 1, "spmm_csr: bad sizes n_rows=%lld nnz=%lld feat=%lld",
                  (long long)n_rows, (long long)nnz, (long long)feat);
  DGLB200_REQUIRE(n_rows < (1ll << 31) && nnz < (1ll << 31) && x_rows < (1ll << 31),
                  "spmm_csr: sizes must be < 2^31");
  DGLB200_REQUIRE(dtype == DGLB200_F32 || dtype == DGLB200_BF16, "spmm_csr: unknown dtype %d", dtype);
  DGLB200_REQUIRE(msg_op >= DGLB200_MSG_COPY_SRC && msg_op <= DGLB200_MSG_COPY_EDGE,
                  "spmm_csr: unknown msg_op %d", msg_op);
  DGLB200_REQUIRE(reduce_op == DGLB200_RED_SUM || reduce_op == DGLB200_RED_MAX,
                  "spmm_csr: unknown reduce_op %d", reduce_op);
  if (n_rows == 0) return 0;
  DGLB200_REQUIRE(indptr != nullptr && y != nullptr, "spmm_csr: indptr / y is NULL");
  DGLB200_REQUIRE(nnz == 0 || msg_op == DGLB200_MSG_COPY_EDGE || (indices != nullptr && x != nullptr),
                  "spmm_csr: indices / x is NULL");
  SpmmArgs a{};
  a.n_rows = n_rows; a.indptr = indptr; a.indices = indices; a.eids = eids;
  a.x = x; a.w = w; a.y = y; a.arg = arg; a.feat = feat; a.x_ld = x_ld; a.y_ld = y_ld;
  a.w_width = 1; a.w_group = feat; a.w_mode = kWScalar;
  if (msg_op != DGLB200_MSG_COPY_SRC) {
    DGLB200_REQUIRE(nnz == 0 || w != nullptr, "spmm_csr: edge data w is NULL");
    if (msg_op == DGLB200_MSG_COPY_EDGE) {
      DGLB200_REQUIRE(w_width == feat, "spmm_csr: copy_edge needs w_width == feat (%lld != %lld)",
                      (long long)w_width, (long long)feat);
    } else {
      DGLB200_REQUIRE(w_width >= 1 && feat % w_width == 0,
                      "spmm_csr: w_width %lld must divide feat %lld", (long long)w_width, (long long)feat);
    }
    a.w_width = w_width;
    a.w_group = feat / w_width;
    a.w_mode = w_width == 1 ? kWScalar : (w_width == feat ? kWFull : kWGrouped);
  } else {
    a.w = nullptr;
  }
  if (reduce_op != DGLB200_RED_MAX) a.arg = nullptr;
  if (ext != nullptr) {
    a.accumulate = ext->accumulate != 0;
    a.x_halo = ext->x_halo;
    a.halo_first = ext->halo_first;
    a.row_scale = ext->row_scale;
    a.bias = ext->bias;
    a.relu = ext->relu != 0;
  }
  return run_spmm(a, msg_op, reduce_op, dtype, plan, x_rows, nnz, static_cast<cudaStream_t>(stream));
}

int dglb200_edge_data_to_slots(int64_t nnz, const int32_t* eids, int64_t n_edges, int64_t row_bytes,
                               const void* w, void* out, int32_t* bad_flag, void* stream_) {
  clear_error();
  DGLB200_REQUIRE(nnz >= 0 && nnz < (1ll << 31) && n_edges >= 0, "edge_data_to_slots: bad sizes");
  DGLB200_REQUIRE(row_bytes >= 2 && row_bytes % 2 == 0 && row_bytes <= 4096,
                  "edge_data_to_slots: row_bytes %lld must be even and in [2, 4096]", (long long)row_bytes);
  if (nnz == 0) return 0;
  DGLB200_REQUIRE(eids && w && out, "edge_data_to_slots: NULL pointer");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const bool u4 = row_bytes % 4 == 0 && (reinterpret_cast<uintptr_t>(w) & 3) == 0 &&
                  (reinterpret_cast<uintptr_t>(out) & 3) == 0;
  const int units = (int)(row_bytes / (u4 ? 4 : 2));
  const int64_t blocks = ceil_div(nnz * units, 256 * 4);
  DGLB200_REQUIRE(blocks < (1ll << 31), "edge_data_to_slots: grid too large");
  if (u4)
    edge_data_to_slots_kernel<uint32_t><<<(unsigned)blocks, 256, 0, stream>>>(
        nnz, eids, n_edges, units, static_cast<const uint32_t*>(w), static_cast<uint32_t*>(out), bad_flag);
  else
    edge_data_to_slots_kernel<uint16_t><<<(unsigned)blocks, 256, 0, stream>>>(
        nnz, eids, n_edges, units, static_cast<const uint16_t*>(w), static_cast<uint16_t*>(out), bad_flag);
  DGLB200_POST_LAUNCH("edge_data_to_slots_kernel");
  return 0;
}

int dglb200_spmm_csr_backward_x(int64_t n_rows, int64_t dy_rows, int64_t nnz, const int64_t* indptr,
                                const int32_t* indices, const int32_t* eids, int dtype,
                                const void* dy, int64_t feat, const void* w, int64_t w_width,
                                const int32_t* arg, void* dx, const dglb200_spmm_plan_t* plan,
                                void* stream) {
  clear_error();
  DGLB200_REQUIRE(n_rows >= 0 && nnz >= 0 && feat >= 1, "spmm_csr_backward_x: bad sizes");
  DGLB200_REQUIRE(n_rows < (1ll << 31) && nnz < (1ll << 31), "spmm_csr_backward_x: sizes must be < 2^31");
  DGLB200_REQUIRE(dtype == DGLB200_F32 || dtype == DGLB200_BF16, "spmm_csr_backward_x: unknown dtype");
  if (n_rows == 0) return 0;
  DGLB200_REQUIRE(indptr && dx && (nnz == 0 || (indices && dy)), "spmm_csr_backward_x: NULL pointer");
  SpmmArgs a{};
  a.n_rows = n_rows; a.indptr = indptr; a.indices = indices; a.eids = eids;
  a.x = dy; a.w = w; a.y = dx; a.arg = const_cast<int32_t*>(arg); a.feat = feat;
  a.w_width = 1; a.w_group = feat; a.w_mode = kWScalar;
  if (w != nullptr) {
    DGLB200_REQUIRE(w_width >= 1 && feat % w_width == 0, "spmm_csr_backward_x: w_width must divide feat");
    a.w_width = w_width;
    a.w_group = feat / w_width;
    a.w_mode = w_width == 1 ? kWScalar : (w_width == feat ? kWFull : kWGrouped);
  }
  int msg = (arg != nullptr) ? (int)kMsgMasked : (w != nullptr ? (int)DGLB200_MSG_SRC_MUL_EDGE : (int)DGLB200_MSG_COPY_SRC);
  return run_spmm(a, msg, DGLB200_RED_SUM, dtype, plan, dy_rows, nnz, static_cast<cudaStream_t>(stream));
}

int dglb200_spmm_csr_static_scratch(int64_t n_rows, int64_t x_rows, int64_t nnz, const int64_t* indptr,
                                    const int32_t* indices, const int32_t* eids, int msg_op, int reduce_op,
                                    int dtype, const void* x, int64_t feat, const void* w, int64_t w_width,
                                    void* y, const dglb200_spmm_plan_t* plan, void* stream) {
  clear_error();
  // no argmax output here: MAX with arg would need a second (int32) scratch of the same size
  if (plan == nullptr || plan->n_long_rows <= 0)     // no long rows: no scratch is used at all
    return dglb200_spmm_csr_strided(n_rows, x_rows, nnz, indptr, indices, eids, msg_op, reduce_op, dtype, x, feat,
                                    feat, w, w_width, y, feat, nullptr, nullptr, stream);
  DGLB200_REQUIRE(feat >= 1 && plan->n_chunks >= 0, "spmm_csr_static_scratch: bad sizes");
  const unsigned long long need = (unsigned long long)plan->n_chunks * (unsigned long long)feat * sizeof(float);
  DGLB200_REQUIRE(need <= (unsigned long long)DGLB200_STATIC_SCRATCH_BYTES,
                  "spmm_csr_static_scratch: %lld long-row slices x %lld features need %llu bytes of scratch, the "
                  "static scratch holds %llu (DGLB200_STATIC_SCRATCH_BYTES); use dglb200_spmm_csr with a "
                  "caller-provided plan->partial", (long long)plan->n_chunks, (long long)feat, need,
                  (unsigned long long)DGLB200_STATIC_SCRATCH_BYTES);
  float* scratch = nullptr;
  if (static_partial_address(&scratch) != 0) return -1;
  dglb200_spmm_plan_t local = *plan;                 // the caller's tables, our scratch
  local.partial = scratch;
  local.partial_arg = nullptr;
  local.partial_bytes = DGLB200_STATIC_SCRATCH_BYTES;
  // every slice writes all of its `feat` partials before the fix-up kernel (same stream) reads them: nothing
  // in the scratch has to start at zero, so the previous call's contents need no clearing
  return dglb200_spmm_csr_strided(n_rows, x_rows, nnz, indptr, indices, eids, msg_op, reduce_op, dtype, x, feat,
                                  feat, w, w_width, y, feat, nullptr, &local, stream);
}

}  // extern "C"
