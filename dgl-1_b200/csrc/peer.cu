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

// ---- TMA variants: the REMOTE stores are bulk async copies (cp.async.bulk, shared -> peer global) -------
// Measured with the kernel above (C5, 2 and 8 ranks): a push CTA sustains only ~5 GB/s - every thread's
// posted NVLink stores sit in the SM's load/store path - so the link needs ~300 CTAs, and a SpMM that
// shares those SMs runs up to 3x slower while a push is in flight.  Here threads only do the LOCAL part
// (gather 16-byte units, optional bf16 rounding, st.shared into a ring of stages); one elected thread hands
// each filled stage to the TMA unit, which streams it to the peer on its own path.
constexpr int kTmaStages = 4;
constexpr int kTmaRows = 32;                       // rows per stage
constexpr int kTmaMaxRowBytes = 1024;              // output bytes per row (fp32 F <= 256, bf16 F <= 512)

#ifndef DGLB200_HOST_EMU
__device__ __forceinline__ void bulk_store_s2g(void* gdst, const void* ssrc, uint32_t bytes) {
  const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(ssrc);
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(saddr), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_proxy() { asm volatile("fence.proxy.async;" ::: "memory"); }

// release + signal shared by the TMA kernels (thread 0 of a CTA, after its bulk stores completed)
__device__ __forceinline__ void arrive_and_signal(const PushArgs& a) {
  if (a.counter == nullptr) return;
  fence_system();
  const int32_t arrived = atomicAdd(a.counter, 1);
  if (arrived == (int32_t)gridDim.x - 1) {
    fence_system();
    atomicExch(a.counter, 0);
    for (int q = 0; q < a.n_dst; ++q)
      if (a.signal[q] != nullptr) st_flag(a.signal[q], a.signal_value);
  }
}

template <bool CONVERT>
__global__ void __launch_bounds__(256) peer_push_tma_kernel(const PushArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  if (a.wait_flags != nullptr) {
    if (threadIdx.x == 0) wait_flags_ge(a.wait_flags, a.n_wait, a.wait_value, a.err_flag);
    __syncthreads();
  }
  constexpr int IN_UNIT = CONVERT ? 32 : 16;
  const int lanes = (int)(a.row_bytes / IN_UNIT);              // 16-byte output units per row
  const uint32_t out_row_bytes = (uint32_t)lanes * 16;
  const bool dense_dst = a.dst_ld_bytes == (int64_t)out_row_bytes;
  const uint32_t stage_bytes = kTmaRows * out_row_bytes;
  int stage = 0;
  int issued = 0;                                              // stages handed to the TMA so far (thread 0)
  for (int j = 0; j < a.n_dst; ++j) {
    const int q = (int)((j + (int64_t)blockIdx.x + a.rotate) % a.n_dst);
    const int64_t rows_q = a.seg_ptr[q + 1] - a.seg_ptr[q];
    const int64_t r_lo = rows_q * (int64_t)blockIdx.x / gridDim.x;
    const int64_t r_hi = rows_q * ((int64_t)blockIdx.x + 1) / gridDim.x;
    const int64_t* idx = a.idx ? a.idx + a.seg_ptr[q] : nullptr;
    for (int64_t r0 = r_lo; r0 < r_hi; r0 += kTmaRows) {
      const int nrows = (int)min((int64_t)kTmaRows, r_hi - r0);
      unsigned char* sbuf = smem_raw + (size_t)stage * stage_bytes;
      // the stage is free once the bulk store issued kTmaStages iterations ago has READ it
      if (threadIdx.x == 0 && issued >= kTmaStages) bulk_wait_read<kTmaStages - 1>();
      __syncthreads();
      const int total = nrows * lanes;
      constexpr int K = 4;            // independent gathers in flight per thread
      for (int u0 = threadIdx.x; u0 < total; u0 += (int)blockDim.x * K) {
        uint4 v[K], v2[K];
        int soff[K];
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const int u = u0 + k * (int)blockDim.x;
          soff[k] = -1;
          v[k] = make_uint4(0, 0, 0, 0);
          v2[k] = v[k];
          if (u < total) {
            const int rr = u / lanes, lane = u - rr * lanes;
            soff[k] = rr * (int)out_row_bytes + lane * 16;
            const int64_t srow = idx ? idx[r0 + rr] : r0 + rr;
            if (srow < 0 || srow >= a.x_rows) {
              atomicExch(a.bad_flag, 1);
            } else {
              const char* src = a.x + srow * a.x_ld_bytes + (int64_t)lane * IN_UNIT;
              v[k] = *reinterpret_cast<const uint4*>(src);
              if constexpr (CONVERT) v2[k] = *reinterpret_cast<const uint4*>(src + 16);
            }
          }
        }
#pragma unroll
        for (int k = 0; k < K; ++k) {
          if (soff[k] < 0) continue;
          uint4 t = v[k];
          if constexpr (CONVERT) {
            t.x = pack_bf16(__uint_as_float(v[k].x), __uint_as_float(v[k].y));
            t.y = pack_bf16(__uint_as_float(v[k].z), __uint_as_float(v[k].w));
            t.z = pack_bf16(__uint_as_float(v2[k].x), __uint_as_float(v2[k].y));
            t.w = pack_bf16(__uint_as_float(v2[k].z), __uint_as_float(v2[k].w));
          }
          *reinterpret_cast<uint4*>(sbuf + soff[k]) = t;
        }
      }
      fence_async_proxy();            // generic-proxy st.shared -> visible to the async proxy (TMA)
      __syncthreads();
      char* gdst = a.dst[q] + r0 * a.dst_ld_bytes;
      if (threadIdx.x == 0) {
        if (dense_dst) {
          bulk_store_s2g(gdst, sbuf, (uint32_t)nrows * out_row_bytes);
        } else {      // rows of a column block are dst_ld apart: one bulk store per row, one group per stage
          for (int rr = 0; rr < nrows; ++rr)
            bulk_store_s2g(gdst + (int64_t)rr * a.dst_ld_bytes, sbuf + (size_t)rr * out_row_bytes, out_row_bytes);
        }
        bulk_commit();
        ++issued;
      }
      stage = stage + 1 == kTmaStages ? 0 : stage + 1;
    }
  }
  // ---- signal: every bulk store has completed and is visible system-wide before any flag is written
  if (threadIdx.x == 0) {
    bulk_wait_all();
    fence_async_proxy();
    arrive_and_signal(a);
  }
}
#endif

// ---- fully asynchronous variant (no conversion): ONE warp per CTA drives both directions with the TMA ---
// Gather: each lane issues one bulk copy global -> shared for its row (cp.async.bulk ... mbarrier::
// complete_tx), kAsyncLoads stages of 32 rows in flight per warp, so the push no longer waits a DRAM round
// trip per stage like the thread-gather kernels; store: lane 0 hands each landed stage to a bulk copy
// shared -> peer global.  No register staging, no __syncthreads, and the SM's load/store path stays with
// the SpMM that runs beside it.
constexpr int kAsyncStages = 6;     // shared-memory stages
constexpr int kAsyncLoads = 4;      // stages of gathers in flight (the other two drain to the peer)

#ifndef DGLB200_HOST_EMU
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = (uint32_t)__cvta_generic_to_shared(bar);
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
  }
}
__device__ __forceinline__ void bulk_load_g2s(void* sdst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   (uint32_t)__cvta_generic_to_shared(sdst)),
               "l"(gsrc), "r"(bytes), "r"((uint32_t)__cvta_generic_to_shared(bar))
               : "memory");
}

__global__ void __launch_bounds__(32) peer_push_async_kernel(const PushArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t full[kAsyncStages];
  const int lane = threadIdx.x;
  if (lane == 0) {
    for (int s = 0; s < kAsyncStages; ++s) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_async_proxy();
    if (a.wait_flags != nullptr) wait_flags_ge(a.wait_flags, a.n_wait, a.wait_value, a.err_flag);
  }
  __syncwarp();
  const uint32_t row_bytes = (uint32_t)a.row_bytes;
  const bool dense_dst = a.dst_ld_bytes == a.row_bytes;
  const uint32_t stage_bytes = 32u * row_bytes;
  // the stages of this CTA as one flat sequence over all destinations: a producer cursor (loads issued) and a
  // consumer cursor (stores issued) walk the same (destination, row) sequence
  int64_t head = 0, tail = 0;
  int jq_h = 0, q_h = -1, jq_t = 0, q_t = -1;
  int64_t r_h = 0, r_hi_h = 0, r_t = 0, r_hi_t = 0;
  auto next_dst = [&](int& j, int& q, int64_t& r, int64_t& r_hi) -> bool {   // advance to a destination with rows left
    while (q < 0 || r >= r_hi) {
      if (j >= a.n_dst) return false;
      q = (int)((j + (int64_t)blockIdx.x + a.rotate) % a.n_dst);
      ++j;
      const int64_t rows_q = a.seg_ptr[q + 1] - a.seg_ptr[q];
      r = rows_q * (int64_t)blockIdx.x / gridDim.x;
      r_hi = rows_q * ((int64_t)blockIdx.x + 1) / gridDim.x;
    }
    return true;
  };
  bool more = next_dst(jq_h, q_h, r_h, r_hi_h);
  (void)next_dst(jq_t, q_t, r_t, r_hi_t);
  while (more || tail < head) {
    // ---- producer: keep kAsyncLoads stages of gathers in flight -------------------------------------
    if (more && head - tail < kAsyncLoads) {
      const int s = (int)(head % kAsyncStages);
      if (head >= kAsyncStages) {
        // the bulk store that last used this stage (number head - kAsyncStages) must have READ it: at most
        // kAsyncStages - kAsyncLoads younger stores may still be pending
        if (lane == 0) bulk_wait_read<kAsyncStages - kAsyncLoads>();
        __syncwarp();
      }
      const int nrows = (int)min((int64_t)32, r_hi_h - r_h);
      unsigned char* sbuf = smem_raw + (size_t)s * stage_bytes;
      const int64_t* idx = a.idx ? a.idx + a.seg_ptr[q_h] : nullptr;
      int64_t srow = -1;
      if (lane < nrows) {
        srow = idx ? idx[r_h + lane] : r_h + lane;
        if (srow < 0 || srow >= a.x_rows) {
          atomicExch(a.bad_flag, 1);
          srow = 0;                        // keep the byte count: read row 0 instead (flagged)
        }
      }
      if (lane == 0) mbar_expect_tx(&full[s], (uint32_t)nrows * row_bytes);
      __syncwarp();
      if (lane < nrows) bulk_load_g2s(sbuf + (size_t)lane * row_bytes, a.x + srow * a.x_ld_bytes, row_bytes, &full[s]);
      r_h += nrows;
      ++head;
      more = next_dst(jq_h, q_h, r_h, r_hi_h);
      if (more && head - tail < kAsyncLoads) continue;      // fill the pipeline before draining
    }
    // ---- consumer: the oldest stage has landed -> one bulk store to the peer ------------------------
    if (tail < head) {
      const int s = (int)(tail % kAsyncStages);
      mbar_wait(&full[s], (uint32_t)((tail / kAsyncStages) & 1));
      const int nrows = (int)min((int64_t)32, r_hi_t - r_t);
      unsigned char* sbuf = smem_raw + (size_t)s * stage_bytes;
      char* gdst = a.dst[q_t] + r_t * a.dst_ld_bytes;
      if (lane == 0) {
        if (dense_dst) {
          bulk_store_s2g(gdst, sbuf, (uint32_t)nrows * row_bytes);
        } else {
          for (int rr = 0; rr < nrows; ++rr)
            bulk_store_s2g(gdst + (int64_t)rr * a.dst_ld_bytes, sbuf + (size_t)rr * row_bytes, row_bytes);
        }
        bulk_commit();
      }
      __syncwarp();
      r_t += nrows;
      ++tail;
      (void)next_dst(jq_t, q_t, r_t, r_hi_t);
    }
  }
  if (lane == 0) {
    bulk_wait_all();
    fence_async_proxy();
    arrive_and_signal(a);
  }
}
#endif

// dglb200_set_tuning("peer_push_tma"): 0 = every thread stores to the peer; 1 = thread gather + TMA store;
// 2 = TMA both ways where the rows are not converted (fp32 wire), else 1
int& peer_push_tma() { static int v = 2; return v; }
void set_peer_push_tma(int v) { peer_push_tma() = v; }

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
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
#ifndef DGLB200_HOST_EMU
  const int64_t out_row_bytes = (row_bytes / unit) * 16;
  if (peer_push_tma() >= 2 && !convert_to_bf16 && row_bytes <= kTmaMaxRowBytes) {
    // TMA both ways: one warp per CTA; measured best at one CTA per SM (C5 x 8: 148 CTAs 10.39 ms/step,
    // 64: 11.16, 296: 11.76; thread stores: 10.79)
    int64_t nb = ceil_div(max_rows > 0 ? max_rows : 1, 32 * 8);
    const int64_t cap_a = max_ctas > 0 ? max_ctas : kNumSMsB200;
    if (nb > cap_a) nb = cap_a;
    if (nb < 1) nb = 1;
    const size_t smem = (size_t)kAsyncStages * 32 * (size_t)row_bytes;
    static bool attr_a = false;
    if (!attr_a) {
      DGLB200_CUDA(cudaFuncSetAttribute(peer_push_async_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        kAsyncStages * 32 * kTmaMaxRowBytes));
      attr_a = true;
    }
    peer_push_async_kernel<<<(unsigned)nb, 32, smem, stream>>>(a);
    DGLB200_POST_LAUNCH("peer_push_async_kernel");
    return 0;
  }
  if (peer_push_tma() >= 1 && out_row_bytes <= kTmaMaxRowBytes) {
    // thread gather (+ bf16 rounding) + TMA store; measured on the bf16 wire, C5 x 8: 296 CTAs 9.00 ms/step,
    // 148: 9.72, thread stores: 9.30
    const int64_t cap_t = max_ctas > 0 ? max_ctas : (convert_to_bf16 ? 2 : 1) * kNumSMsB200;
    if (blocks > cap_t) blocks = cap_t;
    if (blocks < 1) blocks = 1;
    const size_t smem = (size_t)kTmaStages * kTmaRows * (size_t)out_row_bytes;
    static bool attr_set = false;
    if (!attr_set) {
      DGLB200_CUDA(cudaFuncSetAttribute(peer_push_tma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        kTmaStages * kTmaRows * kTmaMaxRowBytes));
      DGLB200_CUDA(cudaFuncSetAttribute(peer_push_tma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        kTmaStages * kTmaRows * kTmaMaxRowBytes));
      attr_set = true;
    }
    if (convert_to_bf16) peer_push_tma_kernel<true><<<(unsigned)blocks, 256, smem, stream>>>(a);
    else peer_push_tma_kernel<false><<<(unsigned)blocks, 256, smem, stream>>>(a);
    DGLB200_POST_LAUNCH("peer_push_tma_kernel");
    return 0;
  }
#endif
  const int64_t cap = max_ctas > 0 ? max_ctas : 2 * kNumSMsB200;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
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
