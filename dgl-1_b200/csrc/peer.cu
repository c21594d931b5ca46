// Multi-GPU exchange over NVLink / NVSwitch peer memory: one kernel PUSHES this rank's rows into
// every peer's exchange buffer and then SIGNALS the peers through flag words in their memory; the
// consumer's stream waits on its own flag words.  No NCCL and no host round trip on the data path.
//
// The reference has no multi-GPU layer; this is the engine's own transport for
// the north-star's 1-D row partition (dgl_b200/dist.py).  Memory model of a push:
//   pusher CTA : stores rows into peer memory ... __threadfence_system() ... atomicAdd(local counter)
//   last CTA   : (observes every CTA's arrival) __threadfence_system(); st.volatile flag -> each peer
//   consumer   : ld.volatile own flag (spin) ... __threadfence_system() ... reads the rows
// i.e. release = fence.sys + relaxed store, acquire = relaxed load + fence.sys, chained through the
// device-scope arrival counter (fence cumulativity of the PTX memory model).
#include <type_traits>

#include "common.cuh"

namespace dglb200 {

constexpr int kMaxPeers = 16;
constexpr unsigned long long kWaitTimeoutNs = 20ull * 1000ull * 1000ull * 1000ull;   // 20 s, then err_flag

struct PushArgs {
  int n_dst;                        // destinations (peers) of this launch
  int64_t seg_ptr[kMaxPeers + 1];   // rows [seg_ptr[q], seg_ptr[q+1]) of `idx` go to destination q
  char* dst[kMaxPeers];             // first output row of destination q (peer memory)
  int32_t* signal[kMaxPeers];       // flag word of destination q to set once everything is visible (nullable)
  const int64_t* idx;               // row ids into x, concatenated by destination; NULL: row i of a segment = x row i
  int64_t x_rows;
  const char* x;
  int64_t x_ld_bytes;
  int64_t row_bytes;                // bytes of an x row that are moved
  int64_t dst_ld_bytes;
  const int32_t* wait_flags;        // local flag words that must all be >= wait_value before any store (nullable)
  int n_wait;
  int32_t wait_value;
  int32_t signal_value;
  int32_t* counter;                 // local CTA arrival counter (0 on entry; the kernel leaves it 0)
  int32_t* err_flag;                // set to 1 when a wait timed out
  int32_t* bad_flag;                // set to 1 on an out-of-range row id
  int rotate;                       // rotation of the destination order (the caller passes its rank)
};

__device__ __forceinline__ int32_t ld_flag(const int32_t* p) {
#ifdef DGLB200_HOST_EMU
  return __atomic_load_n(p, __ATOMIC_ACQUIRE);
#else
  int32_t v;
  asm volatile("ld.volatile.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
#endif
}
__device__ __forceinline__ void st_flag(int32_t* p, int32_t v) {
#ifdef DGLB200_HOST_EMU
  __atomic_store_n(p, v, __ATOMIC_RELEASE);
#else
  asm volatile("st.volatile.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
#endif
}
__device__ __forceinline__ void fence_system() {
#ifdef DGLB200_HOST_EMU
  __atomic_thread_fence(__ATOMIC_SEQ_CST);
#else
  __threadfence_system();
#endif
}

#ifndef DGLB200_HOST_EMU
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// Spin until flags[0..n) are all >= value (one thread).  Bounded: after kWaitTimeoutNs the error
// flag is raised and the caller proceeds (a wrong result that the checks catch, never a hung GPU).
__device__ __forceinline__ void wait_flags_ge(const int32_t* flags, int n, int32_t value, int32_t* err) {
  const unsigned long long t0 = global_ns();
  for (int i = 0; i < n; ++i) {
    unsigned backoff = 32;
    while (ld_flag(flags + i) - value < 0) {       // wrap-safe "flag < value"
      __nanosleep(backoff);
      if (backoff < 1024) backoff <<= 1;
      if (global_ns() - t0 > kWaitTimeoutNs) {
        if (err != nullptr) atomicExch(err, 1);
        return;
      }
    }
  }
  fence_system();
}
#endif

// One launch moves every destination's rows.  CTA b takes slice b of EVERY destination's segment,
// visiting the destinations in an order rotated by (b + rank): at any moment the CTAs of the grid are
// spread over all peers, so every NVLink direction of the switch is busy and no receiver is written
// by two senders' full grids at once.  Units are (row, 16-byte lane) pairs, four independent loads in
// flight per thread, stores posted to the peer.  CONVERT: fp32 rows are rounded to bf16 on the fly
// (half the bytes on the wire; the unit is then 32 bytes of input -> 16 bytes of output).
template <bool CONVERT>
__global__ void __launch_bounds__(256) peer_push_kernel(const PushArgs a) {
#ifndef DGLB200_HOST_EMU
  if (a.wait_flags != nullptr) {
    if (threadIdx.x == 0) wait_flags_ge(a.wait_flags, a.n_wait, a.wait_value, a.err_flag);
    __syncthreads();
  }
#endif
  constexpr int IN_UNIT = CONVERT ? 32 : 16;
  const int64_t lanes = a.row_bytes / IN_UNIT;
  constexpr int K = 4;
  for (int j = 0; j < a.n_dst; ++j) {
    const int q = (int)((j + (int64_t)blockIdx.x + a.rotate) % a.n_dst);
    const int64_t rows_q = a.seg_ptr[q + 1] - a.seg_ptr[q];
    // slice b of the segment
    const int64_t r_lo = rows_q * (int64_t)blockIdx.x / gridDim.x;
    const int64_t r_hi = rows_q * ((int64_t)blockIdx.x + 1) / gridDim.x;
    const int64_t total = (r_hi - r_lo) * lanes;
    const int64_t* idx = a.idx ? a.idx + a.seg_ptr[q] : nullptr;
    char* out = a.dst[q];
    for (int64_t u0 = threadIdx.x; u0 < total; u0 += (int64_t)blockDim.x * K) {
      uint4 v[K], v2[K];
      int64_t dst_off[K];
      bool ok[K];
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const int64_t u = u0 + (int64_t)k * blockDim.x;
        ok[k] = u < total;
        if (ok[k]) {
          const int64_t row = r_lo + u / lanes;
          const int64_t lane = u % lanes;
          const int64_t s = idx ? idx[row] : row;
          if (s < 0 || s >= a.x_rows) {
            atomicExch(a.bad_flag, 1);
            ok[k] = false;
          } else {
            const char* src = a.x + s * a.x_ld_bytes + lane * IN_UNIT;
            v[k] = *reinterpret_cast<const uint4*>(src);
            if constexpr (CONVERT) v2[k] = *reinterpret_cast<const uint4*>(src + 16);
            dst_off[k] = row * a.dst_ld_bytes + lane * 16;
          }
        }
      }
#pragma unroll
      for (int k = 0; k < K; ++k) {
        if (!ok[k]) continue;
        if constexpr (CONVERT) {
          uint4 t;
          t.x = pack_bf16(__uint_as_float(v[k].x), __uint_as_float(v[k].y));
          t.y = pack_bf16(__uint_as_float(v[k].z), __uint_as_float(v[k].w));
          t.z = pack_bf16(__uint_as_float(v2[k].x), __uint_as_float(v2[k].y));
          t.w = pack_bf16(__uint_as_float(v2[k].z), __uint_as_float(v2[k].w));
          *reinterpret_cast<uint4*>(out + dst_off[k]) = t;
        } else {
          *reinterpret_cast<uint4*>(out + dst_off[k]) = v[k];
        }
      }
    }
  }
  // ---- signal: every CTA's stores are visible system-wide before any flag is written ----------
  if (a.counter == nullptr) return;
  fence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const int32_t arrived = atomicAdd(a.counter, 1);
    if (arrived == (int32_t)gridDim.x - 1) {
      fence_system();
      atomicExch(a.counter, 0);
      for (int q = 0; q < a.n_dst; ++q)
        if (a.signal[q] != nullptr) st_flag(a.signal[q], a.signal_value);
    }
  }
}

struct SignalArgs {
  int n;
  int32_t* flag[kMaxPeers];
  int32_t value;
};

// Stream-ordered signal: everything enqueued before it on this stream has completed (kernel
// boundary), then `value` lands in each listed flag word (peer memory).
__global__ void peer_signal_kernel(const SignalArgs a) {
  fence_system();
  if ((int)threadIdx.x < a.n && a.flag[threadIdx.x] != nullptr) st_flag(a.flag[threadIdx.x], a.value);
}

// Stream-ordered wait: the kernels enqueued after it start only once flags[0..n) are all >= value.
__global__ void peer_wait_kernel(const int32_t* flags, int n, int32_t value, int32_t* err) {
#ifndef DGLB200_HOST_EMU
  if (threadIdx.x == 0) wait_flags_ge(flags, n, value, err);
#endif
}

#ifdef DGLB200_HOST_EMU
// The host emulation runs one launch at a time to completion, so a wait cannot spin INSIDE a launch
// (the signalling rank's launch could never start): the same wait is made on the host before the launch.
static void host_wait_flags(const int32_t* flags, int n, int32_t value, int32_t* err) {
  for (int i = 0; i < n; ++i) {
    long spins = 0;
    while (__atomic_load_n(flags + i, __ATOMIC_ACQUIRE) - value < 0) {
      std::this_thread::yield();
      if (++spins > 2000000000l) { if (err) *err = 1; return; }
    }
  }
}
#endif

}  // namespace dglb200

using namespace dglb200;

extern "C" int dglb200_peer_push(int n_dst, const int64_t* seg_ptr_host, void* const* dst_base_host,
                                 int32_t* const* signal_flags_host, const int64_t* idx, int64_t x_rows,
                                 const void* x, int64_t x_ld_bytes, int64_t row_bytes,
                                 int64_t dst_ld_bytes, int convert_to_bf16, const int32_t* wait_flags,
                                 int n_wait, int32_t wait_value, int32_t signal_value, int32_t* counter,
                                 int32_t* err_flag, int32_t* bad_flag, int rotate, int max_ctas,
                                 void* stream_) {
  clear_error();
  DGLB200_REQUIRE(n_dst >= 1 && n_dst <= kMaxPeers, "peer_push: n_dst %d out of range [1, %d]", n_dst, kMaxPeers);
  DGLB200_REQUIRE(seg_ptr_host && dst_base_host && x && bad_flag, "peer_push: NULL pointer");
  const int unit = convert_to_bf16 ? 32 : 16;
  DGLB200_REQUIRE(row_bytes >= unit && row_bytes % unit == 0, "peer_push: row_bytes %lld must be a multiple of %d",
                  (long long)row_bytes, unit);
  DGLB200_REQUIRE(x_ld_bytes >= row_bytes && x_ld_bytes % 16 == 0 && dst_ld_bytes % 16 == 0 &&
                      (reinterpret_cast<uintptr_t>(x) & 15) == 0,
                  "peer_push: rows must be 16-byte aligned");
  DGLB200_REQUIRE(dst_ld_bytes >= (convert_to_bf16 ? row_bytes / 2 : row_bytes), "peer_push: dst_ld_bytes too small");
  DGLB200_REQUIRE(signal_flags_host == nullptr || counter != nullptr, "peer_push: signalling needs a counter");
  PushArgs a{};
  a.n_dst = n_dst;
  int64_t max_rows = 0;
  for (int q = 0; q <= kMaxPeers; ++q) a.seg_ptr[q] = seg_ptr_host[q <= n_dst ? q : n_dst];
  for (int q = 0; q < kMaxPeers; ++q) {
    a.dst[q] = q < n_dst ? static_cast<char*>(dst_base_host[q]) : nullptr;
    a.signal[q] = (q < n_dst && signal_flags_host) ? signal_flags_host[q] : nullptr;
    if (q < n_dst) {
      const int64_t rows_q = a.seg_ptr[q + 1] - a.seg_ptr[q];
      DGLB200_REQUIRE(rows_q >= 0, "peer_push: seg_ptr must be non-decreasing");
      DGLB200_REQUIRE(rows_q == 0 || (a.dst[q] != nullptr && (reinterpret_cast<uintptr_t>(a.dst[q]) & 15) == 0),
                      "peer_push: destination %d is NULL or misaligned", q);
      if (rows_q > max_rows) max_rows = rows_q;
    }
  }
  a.idx = idx; a.x_rows = x_rows; a.x = static_cast<const char*>(x); a.x_ld_bytes = x_ld_bytes;
  a.row_bytes = row_bytes; a.dst_ld_bytes = dst_ld_bytes;
  a.wait_flags = wait_flags; a.n_wait = wait_flags ? n_wait : 0; a.wait_value = wait_value;
  a.signal_value = signal_value; a.counter = signal_flags_host ? counter : nullptr;
  a.err_flag = err_flag; a.bad_flag = bad_flag; a.rotate = rotate;
  // persistent grid: enough CTAs to keep the links busy, never more than the work
  const int64_t units = max_rows * (row_bytes / unit);
  int64_t blocks = ceil_div(units > 0 ? units : 1, 256 * 4);
  const int64_t cap = max_ctas > 0 ? max_ctas : 2 * kNumSMsB200;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
#ifdef DGLB200_HOST_EMU
  if (a.wait_flags) host_wait_flags(a.wait_flags, a.n_wait, a.wait_value, a.err_flag);
#endif
  if (convert_to_bf16) peer_push_kernel<true><<<(unsigned)blocks, 256, 0, stream>>>(a);
  else peer_push_kernel<false><<<(unsigned)blocks, 256, 0, stream>>>(a);
  DGLB200_POST_LAUNCH("peer_push_kernel");
  return 0;
}

extern "C" int dglb200_peer_signal(int n, int32_t* const* flags_host, int32_t value, void* stream_) {
  clear_error();
  DGLB200_REQUIRE(n >= 0 && n <= kMaxPeers && (n == 0 || flags_host), "peer_signal: bad arguments");
  if (n == 0) return 0;
  SignalArgs a{};
  a.n = n;
  a.value = value;
  for (int q = 0; q < kMaxPeers; ++q) a.flag[q] = q < n ? flags_host[q] : nullptr;
  peer_signal_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream_)>>>(a);
  DGLB200_POST_LAUNCH("peer_signal_kernel");
  return 0;
}

extern "C" int dglb200_peer_wait(const int32_t* flags, int n, int32_t value, int32_t* err_flag, void* stream_) {
  clear_error();
  DGLB200_REQUIRE(n >= 0 && (n == 0 || flags), "peer_wait: bad arguments");
  if (n == 0) return 0;
#ifdef DGLB200_HOST_EMU
  host_wait_flags(flags, n, value, err_flag);
#endif
  peer_wait_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream_)>>>(flags, n, value, err_flag);
  DGLB200_POST_LAUNCH("peer_wait_kernel");
  return 0;
}
