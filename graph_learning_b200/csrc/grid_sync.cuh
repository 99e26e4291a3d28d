// grid_sync.cuh -- grid-wide barrier and fixed-order block sums of the resident (cooperatively launched)
// kernels: persistent.cu (K == 1) and persistent_batch.cu (K > 1).
#pragma once
#include "solver.cuh"

namespace gl {

__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void grid_barrier(unsigned *bar, unsigned &gen) {
  __syncthreads();
  if (threadIdx.x == 0) {
    gen += gridDim.x;
    __threadfence();
    atomicAdd(bar, 1u);
    while (ld_acquire_gpu(bar) < gen) {
    }
    __threadfence();
  }
  __syncthreads();
}

// fixed-order block sum of two values per thread; result valid in thread 0 (and broadcast via smem)
__device__ __forceinline__ void block_sum2(double &x, double &y, double (*s_red)[2]) {
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  x = warp_sum(x);
  y = warp_sum(y);
  __syncthreads();
  if (lane == 0) {
    s_red[wp][0] = x;
    s_red[wp][1] = y;
  }
  __syncthreads();
  x = y = 0.0;
#pragma unroll
  for (int q = 0; q < EDGE_THREADS / 32; ++q) {
    x += s_red[q][0];
    y += s_red[q][1];
  }
  __syncthreads();
}

}  // namespace gl
