// Host stand-ins for the CUDA built-ins the kNN iteration kernels use, so that their SOURCE TEXT
// (extracted from graph_learning_b200/csrc/knn_kernels.cuh by tests/test_knn_kernel_emulation.py) can be
// compiled with g++ and executed on the CPU: one OS thread per CUDA thread, CTAs one after another,
// __syncthreads = a 256-thread barrier, warp shuffles through a per-warp scratch line and a 32-thread
// barrier, atomics under one mutex.  Test infrastructure only; it checks logic (indexing, barriers,
// residual bookkeeping, double buffering), not performance and not the memory model of the GPU.
#pragma once
#include <atomic>
#include <barrier>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <mutex>
#include <system_error>
#include <thread>
#include <vector>

struct emu_dim3 { unsigned x = 1, y = 1, z = 1; };
static thread_local emu_dim3 threadIdx;
static emu_dim3 blockIdx, gridDim, blockDim;

struct float2 { float x, y; };
struct double2 { double x, y; };
struct uchar2 { unsigned char x, y; };
static inline uchar2 make_uchar2(unsigned char a, unsigned char b) { return uchar2{a, b}; }

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __shared__ static
#define __launch_bounds__(...)

template <typename T> static inline T __ldcs(const T *p) { return *p; }
template <typename T> static inline T __ldcg(const T *p) { return *p; }
template <typename T> static inline void __stcs(T *p, T v) { *p = v; }

static std::barrier<> *emu_cta_barrier = nullptr;
static std::vector<std::unique_ptr<std::barrier<>>> emu_warp_barrier;
static inline void __syncthreads() { emu_cta_barrier->arrive_and_wait(); }
static inline void __threadfence() { std::atomic_thread_fence(std::memory_order_seq_cst); }

static std::mutex emu_atomic_mutex;
template <typename T> static inline T atomicAdd(T *p, T v) {
  std::lock_guard<std::mutex> g(emu_atomic_mutex);
  const T old = *p;
  *p = old + v;
  return old;
}

template <typename V> static inline V __shfl_xor_sync(unsigned, V v, int o) {
  static V buf[64][32];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  buf[w][l] = v;
  emu_warp_barrier[w]->arrive_and_wait();
  const V r = buf[w][l ^ o];
  emu_warp_barrier[w]->arrive_and_wait();
  return r;
}

namespace gl {
template <typename V> static inline V warp_sum(V v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
}  // namespace gl

// run `kernel(args)` for a grid of `grid` CTAs of `threads` threads, CTAs sequentially.  Returns false when
// the OS refuses the threads (the caller skips the test): the pool is created behind a gate, so a failed
// creation releases the threads that already exist without running the kernel.
static bool emu_failed = false;
template <typename K, typename A> static void emu_launch(K kernel, const A &args, int grid, int threads) {
  if (emu_failed) return;
  gridDim.x = (unsigned)grid;
  blockDim.x = (unsigned)threads;
  emu_warp_barrier.clear();
  for (int w = 0; w < threads / 32; ++w) emu_warp_barrier.emplace_back(new std::barrier<>(32));
  for (int b = 0; b < grid && !emu_failed; ++b) {
    blockIdx.x = (unsigned)b;
    std::barrier<> bar(threads);
    emu_cta_barrier = &bar;
    std::atomic<int> gate{0};   // 0: wait, 1: run, 2: abort
    std::vector<std::thread> pool;
    pool.reserve(threads);
    try {
      for (int t = 0; t < threads; ++t)
        pool.emplace_back([&, t]() {
          while (gate.load(std::memory_order_acquire) == 0) std::this_thread::yield();
          if (gate.load(std::memory_order_acquire) != 1) return;
          threadIdx.x = (unsigned)t;
          kernel(args);
        });
      gate.store(1, std::memory_order_release);
    } catch (const std::system_error &) {
      emu_failed = true;
      gate.store(2, std::memory_order_release);
    }
    for (auto &th : pool) th.join();
  }
}
