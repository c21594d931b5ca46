// TEST INFRASTRUCTURE ONLY - a minimal host execution model for the CUDA sources under
// dgl-1_b200/csrc, so that `-m "not gpu"` tests can run the KERNEL SOURCE ITSELF (not a
// restatement) against the oracle in a container without a GPU.
//
// Model: a launch runs its blocks one after another; inside a block every CUDA thread is a host
// thread, so warp collectives (__shfl*_sync, __match_any_sync, __reduce_*_sync, __syncwarp) and
// __syncthreads() are real rendezvous between the participating lanes.  Lanes that have returned
// from the kernel drop out of later rendezvous (the kernels keep their collectives convergent, as
// full-mask collectives require on the hardware too).  __shared__ arrays are statics: one block is
// alive at a time.  Device pointers are host pointers; streams are ignored (everything is
// synchronous).  tests/native/cuda_emu/build.py rewrites `k<<<g, b, s, st>>>(args)` into
// dglb200_emu::launch(g, b, [&] { k(args); }) and compiles with g++ -std=c++20 -DDGLB200_HOST_EMU.
#pragma once
#include <atomic>
#include <barrier>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>
#include <type_traits>
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(...)
#define __shared__ static
#define __align__(n) __attribute__((aligned(n)))

// ---- vector types -----------------------------------------------------------------------------
struct dim3 {
  unsigned x, y, z;
  dim3(unsigned long long x_ = 1, unsigned long long y_ = 1, unsigned long long z_ = 1)
      : x((unsigned)x_), y((unsigned)y_), z((unsigned)z_) {}
};
struct float2 { float x, y; };
struct alignas(16) float4 { float x, y, z, w; };
struct alignas(16) uint4 { uint32_t x, y, z, w; };
struct uint2 { uint32_t x, y; };
inline float2 make_float2(float x, float y) { return float2{x, y}; }
inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return uint4{x, y, z, w}; }

// ---- runtime API subset ---------------------------------------------------------------------
typedef void* cudaStream_t;
typedef int cudaError_t;
enum { cudaSuccess = 0, cudaErrorEmu = 999 };
enum cudaMemcpyKind { cudaMemcpyHostToHost, cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyDefault };
enum cudaDeviceAttr { cudaDevAttrL2CacheSize, cudaDevAttrComputeCapabilityMajor, cudaDevAttrComputeCapabilityMinor, cudaDevAttrMultiProcessorCount };
struct cudaIpcMemHandle_t { char reserved[64]; };
enum { cudaIpcMemLazyEnablePeerAccess = 1 };
inline cudaError_t cudaGetLastError() { return cudaSuccess; }
inline cudaError_t cudaPeekAtLastError() { return cudaSuccess; }
inline const char* cudaGetErrorString(cudaError_t) { return "host emulation"; }
inline cudaError_t cudaGetDevice(int* d) { *d = 0; return cudaSuccess; }
inline cudaError_t cudaDeviceGetAttribute(int* v, cudaDeviceAttr a, int) {
  *v = a == cudaDevAttrL2CacheSize ? (126 << 20) : (a == cudaDevAttrComputeCapabilityMajor ? 10 : (a == cudaDevAttrMultiProcessorCount ? 148 : 0));
  return cudaSuccess;
}
inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t = nullptr) { memcpy(d, s, n); return cudaSuccess; }
inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { memcpy(d, s, n); return cudaSuccess; }
inline cudaError_t cudaMemsetAsync(void* d, int v, size_t n, cudaStream_t = nullptr) { memset(d, v, n); return cudaSuccess; }
inline cudaError_t cudaMemset(void* d, int v, size_t n) { memset(d, v, n); return cudaSuccess; }
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
inline cudaError_t cudaMalloc(void** p, size_t n) { *p = malloc(n ? n : 1); return *p ? cudaSuccess : cudaErrorEmu; }
template <typename T> inline cudaError_t cudaMalloc(T** p, size_t n) { return cudaMalloc(reinterpret_cast<void**>(p), n); }
inline cudaError_t cudaFree(void* p) { free(p); return cudaSuccess; }
template <typename T> inline cudaError_t cudaGetSymbolAddress(void** p, T& symbol) { *p = (void*)&symbol; return cudaSuccess; }
inline cudaError_t cudaIpcGetMemHandle(cudaIpcMemHandle_t*, void*) { return cudaErrorEmu; }
inline cudaError_t cudaIpcOpenMemHandle(void**, cudaIpcMemHandle_t, unsigned) { return cudaErrorEmu; }
inline cudaError_t cudaIpcCloseMemHandle(void*) { return cudaErrorEmu; }

// ---- execution state ------------------------------------------------------------------------
namespace dglb200_emu {
struct Warp {
  std::barrier<> bar;
  uint64_t slot[32];
  std::atomic<unsigned> exited{0};   // bit l set: lane l returned from the kernel
  explicit Warp(int lanes) : bar(lanes) { memset(slot, 0, sizeof(slot)); }
};
// DGLB200_EMU_STRICT=1: a warp collective whose MASK names a lane that has already returned from
// the kernel is undefined on the hardware; report it instead of letting the lane silently drop out.
inline bool strict() { static const bool s = getenv("DGLB200_EMU_STRICT") != nullptr; return s; }
inline void check_converged(Warp& w, unsigned mask, const char* what) {
  const unsigned gone = w.exited.load(std::memory_order_relaxed) & mask;
  if (strict() && gone != 0) {
    fprintf(stderr, "cuda_emu: %s with mask %08x names exited lane(s) %08x (divergent collective)\n", what, mask, gone);
    abort();
  }
}
struct Block {
  std::barrier<> bar;
  std::vector<std::unique_ptr<Warp>> warps;
  explicit Block(int threads) : bar(threads) {
    for (int w = 0; w * 32 < threads; ++w) {
      const int lanes = threads - w * 32 < 32 ? threads - w * 32 : 32;
      warps.emplace_back(new Warp(lanes));
    }
  }
};
inline Block*& cur_block() { static Block* b = nullptr; return b; }
}  // namespace dglb200_emu

extern thread_local dim3 threadIdx;
extern dim3 blockIdx, blockDim, gridDim;

namespace dglb200_emu {
inline int lane_id() { return (int)(threadIdx.x & 31u); }
inline Warp& warp() { return *cur_block()->warps[threadIdx.x >> 5]; }

// all lanes of the warp publish a value, then every lane reads what it needs
template <typename T, typename F>
inline auto exchange(unsigned mask, T v, F read) {
  static_assert(sizeof(T) <= 8, "shuffle of > 8 bytes");
  Warp& w = warp();
  check_converged(w, mask, "warp collective");
  uint64_t raw = 0;
  memcpy(&raw, &v, sizeof(T));
  w.slot[lane_id()] = raw;
  w.bar.arrive_and_wait();
  auto get = [&](int lane) { T r; memcpy(&r, &w.slot[lane], sizeof(T)); return r; };
  auto out = read(get);
  w.bar.arrive_and_wait();
  return out;
}

// Host threads that play the CUDA threads of a block; kept alive between launches (creating a few
// hundred threads per launch would dominate the small launches of the tests).  Leaked on purpose:
// the workers are detached and sleep on the condition variable until the process exits.
class Pool {
 public:
  static Pool& get() { static Pool* p = new Pool(); return *p; }
  void run(int n, const std::function<void(int)>& f) {
    std::unique_lock<std::mutex> lk(m_);
    while (n_workers_ < n) {
      const int id = n_workers_++;
      std::thread([this, id]() { work(id); }).detach();
    }
    job_ = &f;
    active_ = n;
    remaining_ = n;
    ++gen_;
    start_.notify_all();
    done_.wait(lk, [&] { return remaining_ == 0; });
    job_ = nullptr;
  }

 private:
  void work(int id) {
    uint64_t seen = 0;
    std::unique_lock<std::mutex> lk(m_);
    for (;;) {
      start_.wait(lk, [&] { return gen_ != seen; });
      seen = gen_;
      if (id >= active_) continue;
      const std::function<void(int)>* f = job_;
      lk.unlock();
      (*f)(id);
      lk.lock();
      if (--remaining_ == 0) done_.notify_one();
    }
  }
  std::mutex m_;
  std::condition_variable start_, done_;
  const std::function<void(int)>* job_ = nullptr;
  int n_workers_ = 0, active_ = 0, remaining_ = 0;
  uint64_t gen_ = 0;
};

inline std::mutex& launch_mutex() { static std::mutex m; return m; }

template <typename K>
void launch(dim3 grid, dim3 block, K kernel) {
  const int threads = (int)(block.x * block.y * block.z);
  const long long n_blocks = (long long)grid.x * grid.y * grid.z;
  if (threads <= 0 || n_blocks <= 0) return;
  // one launch at a time: the model's state (blockIdx, the current block, the pool) is global, and
  // tests may call the library from several host threads (ranks of a partition as threads)
  std::lock_guard<std::mutex> one_launch(launch_mutex());
  gridDim = grid;
  blockDim = block;
  // the blocks run one after another on the pool, each with fresh rendezvous objects (lanes that
  // exit early drop out of THEIR block's barriers only)
  std::barrier<> turn(threads);
  std::unique_ptr<Block> blk;
  Pool::get().run(threads, [&](int t) {
    threadIdx = dim3(t % block.x, (t / block.x) % block.y, t / (block.x * block.y));
    for (long long b = 0; b < n_blocks; ++b) {
      if (t == 0) {
        blockIdx = dim3(b % grid.x, (b / grid.x) % grid.y, b / ((long long)grid.x * grid.y));
        blk.reset(new Block(threads));
        cur_block() = blk.get();
      }
      turn.arrive_and_wait();
      kernel();
      blk->warps[t >> 5]->exited.fetch_or(1u << (t & 31), std::memory_order_relaxed);
      blk->warps[t >> 5]->bar.arrive_and_drop();
      blk->bar.arrive_and_drop();
      turn.arrive_and_wait();
    }
  });
  cur_block() = nullptr;
}
}  // namespace dglb200_emu

// ---- synchronisation + warp collectives -----------------------------------------------------------
inline void __syncthreads() { dglb200_emu::cur_block()->bar.arrive_and_wait(); }
inline void __syncwarp(unsigned mask = 0xffffffffu) {
  dglb200_emu::check_converged(dglb200_emu::warp(), mask, "__syncwarp");
  dglb200_emu::warp().bar.arrive_and_wait();
}

template <typename T>
inline T __shfl_sync(unsigned mask, T v, int src, int width = 32) {
  const int lane = dglb200_emu::lane_id();
  const int base = lane / width * width;
  return dglb200_emu::exchange(mask, v, [&](auto get) { return get(base + (((src % width) + width) % width)); });
}
template <typename T>
inline T __shfl_xor_sync(unsigned lanes, T v, int mask, int width = 32) {
  const int lane = dglb200_emu::lane_id();
  const int base = lane / width * width;
  const int peer = lane ^ mask;
  return dglb200_emu::exchange(lanes, v, [&](auto get) { return (peer >= base && peer < base + width) ? get(peer) : get(lane); });
}
template <typename T>
inline T __shfl_up_sync(unsigned mask, T v, unsigned delta, int width = 32) {
  const int lane = dglb200_emu::lane_id();
  const int base = lane / width * width;
  const int peer = lane - (int)delta;
  return dglb200_emu::exchange(mask, v, [&](auto get) { return peer >= base ? get(peer) : get(lane); });
}
template <typename T>
inline T __shfl_down_sync(unsigned mask, T v, unsigned delta, int width = 32) {
  const int lane = dglb200_emu::lane_id();
  const int base = lane / width * width;
  const int peer = lane + (int)delta;
  return dglb200_emu::exchange(mask, v, [&](auto get) { return peer < base + width ? get(peer) : get(lane); });
}
inline unsigned __ballot_sync(unsigned mask, int pred) {
  return dglb200_emu::exchange(mask, (int)(pred != 0), [&](auto get) {
    unsigned m = 0;
    for (int l = 0; l < 32; ++l) if ((mask >> l) & 1u) m |= (get(l) ? 1u : 0u) << l;
    return m;
  });
}
template <typename T>
inline unsigned __match_any_sync(unsigned mask, T v) {
  return dglb200_emu::exchange(mask, v, [&](auto get) {
    unsigned m = 0;
    for (int l = 0; l < 32; ++l) if (((mask >> l) & 1u) && get(l) == v) m |= 1u << l;
    return m;
  });
}
inline int __reduce_max_sync(unsigned mask, int v) {
  return dglb200_emu::exchange(mask, v, [&](auto get) { int r = v; for (int l = 0; l < 32; ++l) if ((mask >> l) & 1u) { int o = get(l); if (o > r) r = o; } return r; });
}
inline int __reduce_min_sync(unsigned mask, int v) {
  return dglb200_emu::exchange(mask, v, [&](auto get) { int r = v; for (int l = 0; l < 32; ++l) if ((mask >> l) & 1u) { int o = get(l); if (o < r) r = o; } return r; });
}
inline unsigned __activemask() { return 0xffffffffu; }

// ---- scalar intrinsics ------------------------------------------------------------------------
template <typename T> inline T __ldg(const T* p) { return *p; }
inline int __popc(unsigned v) { return __builtin_popcount(v); }
inline int __ffs(int v) { return __builtin_ffs(v); }
inline unsigned long long __umul64hi(unsigned long long a, unsigned long long b) { return (unsigned long long)(((unsigned __int128)a * b) >> 64); }
inline int __clz(int v) { return v == 0 ? 32 : __builtin_clz((unsigned)v); }
inline float __uint_as_float(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
inline uint32_t __float_as_uint(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
inline float2 __fadd2_rn(float2 a, float2 b) { return float2{a.x + b.x, a.y + b.y}; }
inline float2 __ffma2_rn(float2 a, float2 b, float2 c) { return float2{fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)}; }
template <typename A, typename B> inline auto min(A a, B b) -> typename std::common_type<A, B>::type { return a < b ? a : b; }
template <typename A, typename B> inline auto max(A a, B b) -> typename std::common_type<A, B>::type { return a > b ? a : b; }

// ---- atomics ----------------------------------------------------------------------------------
template <typename T> inline T atomicAdd(T* p, T v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
inline float atomicAdd(float* p, float v) {
  uint32_t* u = reinterpret_cast<uint32_t*>(p);
  uint32_t old = __atomic_load_n(u, __ATOMIC_SEQ_CST), nw;
  do { nw = __float_as_uint(__uint_as_float(old) + v); } while (!__atomic_compare_exchange_n(u, &old, nw, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST));
  return __uint_as_float(old);
}
template <typename T> inline T atomicExch(T* p, T v) { return __atomic_exchange_n(p, v, __ATOMIC_SEQ_CST); }
template <typename T> inline T atomicCAS(T* p, T expect, T v) {
  __atomic_compare_exchange_n(p, &expect, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST);
  return expect;    // the old value, as CUDA returns it
}
inline int atomicCAS(int* p, int expect, int v) { return atomicCAS<int>(p, expect, v); }
template <typename T> inline T atomicOr(T* p, T v) { return __atomic_fetch_or(p, v, __ATOMIC_SEQ_CST); }
inline int atomicExch(int* p, int v) { return __atomic_exchange_n(p, v, __ATOMIC_SEQ_CST); }
template <typename T> inline T atomicMax(T* p, T v) {
  T old = __atomic_load_n(p, __ATOMIC_SEQ_CST);
  while (old < v && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
  return old;
}
