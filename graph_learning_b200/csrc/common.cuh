// common.cuh -- shared helpers for the gl_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "../../include/gl_b200.h"

namespace gl {

constexpr int T = GL_TILE;            // tile edge (nodes)
constexpr int TT = T * T;             // elements per tile
constexpr int EDGE_THREADS = 256;     // 8 warps, 16 tile rows per warp
constexpr int ROWS_PER_WARP = T / (EDGE_THREADS / 32);

// ---- error plumbing -------------------------------------------------------------------------
extern thread_local char g_err[512];
extern std::atomic<long long> g_launches;

inline int fail(int code, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define GL_CUDA(expr)                                                                           \
  do {                                                                                          \
    cudaError_t _e = (expr);                                                                    \
    if (_e != cudaSuccess)                                                                      \
      return ::gl::fail(GL_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),      \
                        __FILE__, __LINE__);                                                    \
  } while (0)

#define GL_TRY(expr)            \
  do {                          \
    int _rc = (expr);           \
    if (_rc != GL_OK) return _rc; \
  } while (0)

// every kernel launch goes through this so gl_launch_count() is honest
#define GL_LAUNCH_CHECK()                                  \
  do {                                                     \
    ::gl::g_launches.fetch_add(1, std::memory_order_relaxed); \
    GL_CUDA(cudaGetLastError());                           \
  } while (0)

inline cudaStream_t as_stream(void *s) { return reinterpret_cast<cudaStream_t>(s); }

// ---- tile geometry ----------------------------------------------------------------------------
__host__ __device__ inline int tiles_per_side(int n) { return (n + T - 1) / T; }
__host__ __device__ inline int num_tiles(int n) {
  const int nt = tiles_per_side(n);
  return nt * (nt + 1) / 2;
}
// first tile index of tile-row I (tiles (I, I..nT-1))
__host__ __device__ inline int tile_row_start(int I, int nt) { return I * nt - I * (I - 1) / 2; }
__host__ __device__ inline int tile_index(int I, int J, int nt) { return tile_row_start(I, nt) + (J - I); }
// packed offset of edge (i, j), i < j   (operators.py:39-41 order)
__host__ __device__ inline int64_t packed_index(int64_t i, int64_t j, int64_t n) {
  return i * n - i * (i + 1) / 2 + (j - i - 1);
}

// ---- vector types: 16-byte vectors of the edge dtype ----------------------------------------
template <typename F> struct Vec;
template <> struct Vec<float> {
  using type = float4;
  static constexpr int N = 4;
};
template <> struct Vec<double> {
  using type = double2;
  static constexpr int N = 2;
};

template <typename F> __device__ __forceinline__ void vec_unpack(const typename Vec<F>::type &v, F *o);
template <> __device__ __forceinline__ void vec_unpack<float>(const float4 &v, float *o) {
  o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w;
}
template <> __device__ __forceinline__ void vec_unpack<double>(const double2 &v, double *o) {
  o[0] = v.x; o[1] = v.y;
}
template <typename F> __device__ __forceinline__ typename Vec<F>::type vec_pack(const F *o);
template <> __device__ __forceinline__ float4 vec_pack<float>(const float *o) {
  return make_float4(o[0], o[1], o[2], o[3]);
}
template <> __device__ __forceinline__ double2 vec_pack<double>(const double *o) {
  return make_double2(o[0], o[1]);
}

// streaming 128-bit accesses: the edge arrays are touched once per iteration and are far larger
// than L2 at the target size: cache-streaming (evict-first) loads and stores (ld/st.global.cs).
__device__ __forceinline__ float4 ld_stream(const float4 *p) { return __ldcs(p); }
__device__ __forceinline__ double2 ld_stream(const double2 *p) { return __ldcs(p); }
__device__ __forceinline__ void st_stream(float4 *p, const float4 &v) { __stcs(p, v); }
__device__ __forceinline__ void st_stream(double2 *p, const double2 &v) { __stcs(p, v); }

// ---- system-scope accesses for the peer (NVLink) exchange ------------------------------------------
__device__ __forceinline__ void st_release_sys(int *p, int v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
// Flag publication to all peers: ONE system-scope fence (orders everything this thread has observed,
// i.e. the rank's partials made visible by the preceding grid barrier / kernel boundary), then
// relaxed stores that do not wait for one another.  A st.release per peer costs a full NVLink
// round trip each (measured: 16.6 us for 8 peers against ~2 us this way).
__device__ __forceinline__ void st_relaxed_sys(int *p, int v) {
  asm volatile("st.relaxed.sys.global.s32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int *p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double ld_relaxed_sys(const double *p) {
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ int ld_volatile_int(const int *p) { return *reinterpret_cast<const volatile int *>(p); }

// ---- warp helpers ---------------------------------------------------------------------------------
template <typename V> __device__ __forceinline__ V warp_sum(V v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace gl
