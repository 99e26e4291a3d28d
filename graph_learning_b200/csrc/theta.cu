// theta.cu -- K2: the Kalofolias-Perraudin theta heuristic for a requested mean degree.
//
// Replaces get_theta without blocks, graph_learn/smooth_learning.py:201-219: each row of the SQUARE
// distance matrix is sorted with its zero diagonal included, b = sum of the first k sorted values,
//   theta_min_i = (k s_k^2 - b s_k)^-1/2,  theta_max_i = (k s_{k-1}^2 - b s_{k-1})^-1/2,
//   theta = sqrt(mean(theta_min) * mean(theta_max)).
// Only the k+1 smallest entries of a row matter, so instead of np.sort of the whole row (:202) one
// CTA gathers the row into shared memory (from tiled or packed Z -- no square matrix is built) and
// extracts its k+1 minima.  Bound: HBM/L2 (one gathered read of Z); arithmetic is float64.
#include <math_constants.h>

#include "solver.cuh"

namespace gl {

constexpr int TH_THREADS = 256;
constexpr int LAYOUT_DENSE = 2;

// Row i of the (virtual) square matrix streams through shared memory in chunks of `chunk` entries; the
// k+1 smallest seen so far are carried in front of the next chunk, so a row of any length is handled
// (one chunk up to ~50,000 float32 entries).  Tiled layout: only entries stored in tiles [tb, te) take
// part (a GPU shard; Z then points at the shard's first tile), the others count as +inf, and the zero
// diagonal belongs to the shard that owns the row's diagonal tile.  cand_out != nullptr: write the
// row's k+1 smallest (ascending, +inf padded) instead of the two statistics -- the per-shard half of
// the sharded form, merged by a second call on the gathered candidates (gl_theta_dense).
template <typename F>
__global__ void __launch_bounds__(TH_THREADS) theta_rows_kernel(const F *Z, int layout, int n, int nt, int k, int chunk,
                                                                int tb, int te, double *lo_out, double *hi_out,
                                                                F *cand_out, int64_t ld) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double *best_q = reinterpret_cast<double *>(smem_raw);          // [k + 1] sorted minima so far
  F *row = reinterpret_cast<F *>(smem_raw + sizeof(double) * ((k + 2) & ~1));  // [k + 1 carried | chunk]
  __shared__ double s_val[TH_THREADS / 32];
  __shared__ int s_idx[TH_THREADS / 32];
  const int i = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
  int carried = 0;
  for (int s0 = 0; s0 < n; s0 += chunk) {
    const int cnt = (n - s0 < chunk) ? n - s0 : chunk;
    for (int jj = tid; jj < cnt; jj += TH_THREADS) {
      const int j = s0 + jj;
      F v = (F)CUDART_INF;
      if (layout == LAYOUT_DENSE) {  // general (rows, n) matrix, row i as stored (block form, :220-235)
        v = Z[(int64_t)i * ld + j];
      } else if (layout == GL_LAYOUT_PACKED) {
        v = F(0);
        if (j != i) v = Z[packed_index(j < i ? j : i, j < i ? i : j, n)];
      } else {
        const int lo = j < i ? j : i, hi = j < i ? i : j;
        const int t = tile_index(lo / T, hi / T, nt);
        if (t >= tb && t < te) v = (j == i) ? F(0) : Z[(int64_t)(t - tb) * TT + (lo % T) * T + (hi % T)];
      }
      row[carried + jj] = v;
    }
    const int len = carried + cnt;
    __syncthreads();
    // k+1 rounds of arg-min extraction: round q yields sorted[q]
    for (int q = 0; q <= k; ++q) {
      double best = CUDART_INF;
      int bidx = -1;
      for (int j = tid; j < len; j += TH_THREADS) {
        const double v = (double)row[j];
        if (bidx < 0 || v < best) {
          best = v;
          bidx = j;
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bidx, o);
        if (oi >= 0 && (bidx < 0 || ov < best || (ov == best && oi < bidx))) {
          best = ov;
          bidx = oi;
        }
      }
      if (lane == 0) {
        s_val[wp] = best;
        s_idx[wp] = bidx;
      }
      __syncthreads();
      if (tid == 0) {
        double bv = s_val[0];
        int bi = s_idx[0];
        for (int w = 1; w < TH_THREADS / 32; ++w)
          if (s_idx[w] >= 0 && (bi < 0 || s_val[w] < bv || (s_val[w] == bv && s_idx[w] < bi))) {
            bv = s_val[w];
            bi = s_idx[w];
          }
        if (bi >= 0) row[bi] = (F)CUDART_INF;
        best_q[q] = bi >= 0 ? bv : CUDART_INF;
      }
      __syncthreads();
    }
    if (s0 + chunk < n) {  // carry the minima in front of the next chunk
      for (int q = tid; q <= k; q += TH_THREADS) row[q] = (F)best_q[q];
      carried = k + 1;
      __syncthreads();
    }
  }
  if (cand_out) {
    for (int q = tid; q <= k; q += TH_THREADS) cand_out[(int64_t)i * (k + 1) + q] = (F)best_q[q];
  } else if (tid == 0) {
    double b = 0.0;
    for (int q = 0; q < k; ++q) b += best_q[q];          // b = sum of sorted[0..k-1]   (:204)
    const double kk = (double)k, sk = best_q[k], sk1 = best_q[k - 1];
    lo_out[i] = 1.0 / sqrt(kk * sk * sk - b * sk);      // :205-208
    hi_out[i] = 1.0 / sqrt(kk * sk1 * sk1 - b * sk1);   // :209-212
  }
}

__global__ void __launch_bounds__(256) theta_mean_kernel(const double *lo, const double *hi, int n, double *out) {
  __shared__ double s[2][256];
  double a = 0.0, b = 0.0;
  for (int i = threadIdx.x; i < n; i += 256) {
    a += lo[i];
    b += hi[i];
  }
  s[0][threadIdx.x] = a;
  s[1][threadIdx.x] = b;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      s[0][threadIdx.x] += s[0][threadIdx.x + o];
      s[1][threadIdx.x] += s[1][threadIdx.x + o];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    out[0] = s[0][0] / n;
    out[1] = s[1][0] / n;
  }
}

}  // namespace gl
using namespace gl;

static int theta_launch(const void *Z, int layout, int rows, int n, int64_t ld, int k, int dtype, int tb, int te,
                        double *lo, double *hi, void *cand, cudaStream_t st) {
  const size_t esz = dtype == GL_F32 ? 4 : 8;
  const size_t head = sizeof(double) * (size_t)((k + 2) & ~1);
  const size_t budget = 200 * 1024;
  if (head + esz * (size_t)(k + 2) > budget)
    return fail(GL_E_ARG, "avg_degree %d is too large for the row-select kernel", k);
  size_t chunk = (budget - head) / esz - (size_t)(k + 1);
  if (chunk > (size_t)n) chunk = (size_t)n;
  const size_t smem = head + esz * (chunk + (size_t)(k + 1));
  const int nt = tiles_per_side(n);
  cudaError_t e;
  if (dtype == GL_F32) {
    e = cudaFuncSetAttribute(theta_rows_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess)
      theta_rows_kernel<float><<<rows, TH_THREADS, smem, st>>>((const float *)Z, layout, n, nt, k, (int)chunk, tb, te, lo, hi,
                                                               (float *)cand, ld);
  } else {
    e = cudaFuncSetAttribute(theta_rows_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess)
      theta_rows_kernel<double><<<rows, TH_THREADS, smem, st>>>((const double *)Z, layout, n, nt, k, (int)chunk, tb, te, lo, hi,
                                                                (double *)cand, ld);
  }
  if (e == cudaSuccess) {
    g_launches.fetch_add(1);
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) return fail(GL_E_CUDA, "theta row select failed: %s", cudaGetErrorString(e));
  return GL_OK;
}

static int theta_impl(const void *Z, int layout, int rows, int n, int64_t ld, int avg_degree, int dtype,
                      double *theta_host, void *stream) {
  if (n < 2 || rows < 1) return fail(GL_E_SHAPE, "need at least 2 nodes, got n=%d", n);
  if (avg_degree < 1 || avg_degree >= n)
    return fail(GL_E_ARG, "avg_degree must be in [1, n-1], got %d (n=%d)", avg_degree, n);
  if (!theta_host) return fail(GL_E_ARG, "null output");
  if (dtype != GL_F32 && dtype != GL_F64) return fail(GL_E_ARG, "dtype must be GL_F32 or GL_F64");
  cudaStream_t st = as_stream(stream);
  double *buf = nullptr;
  GL_CUDA(cudaMalloc((void **)&buf, sizeof(double) * (2 * (size_t)rows + 2)));
  int rc = theta_launch(Z, layout, rows, n, ld, avg_degree, dtype, 0, num_tiles(n), buf, buf + rows, nullptr, st);
  cudaError_t e = cudaSuccess;
  double h[2] = {0, 0};
  if (rc == GL_OK) {
    theta_mean_kernel<<<1, 256, 0, st>>>(buf, buf + rows, rows, buf + 2 * (size_t)rows);
    g_launches.fetch_add(1);
    e = cudaGetLastError();
  }
  if (rc == GL_OK && e == cudaSuccess)
    e = cudaMemcpyAsync(h, buf + 2 * (size_t)rows, sizeof(h), cudaMemcpyDeviceToHost, st);
  if (rc == GL_OK && e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(buf);
  if (rc != GL_OK) return rc;
  if (e != cudaSuccess) return fail(GL_E_CUDA, "gl_theta failed: %s", cudaGetErrorString(e));
  *theta_host = sqrt(h[0] * h[1]);  // :219
  return GL_OK;
}

extern "C" int gl_theta(const void *Z, int layout, int n, int avg_degree, int dtype, double *theta_host,
                        void *stream) {
  if (layout != GL_LAYOUT_PACKED && layout != GL_LAYOUT_TILED) return fail(GL_E_ARG, "bad layout");
  return theta_impl(Z, layout, n, n, 0, avg_degree, dtype, theta_host, stream);
}

extern "C" int gl_theta_dense(const void *M, int rows, int cols, int64_t ld, int avg_degree, int dtype,
                              double *theta_host, void *stream) {
  if (ld < cols) return fail(GL_E_SHAPE, "leading dimension %lld < cols %d", (long long)ld, cols);
  return theta_impl(M, LAYOUT_DENSE, rows, cols, ld, avg_degree, dtype, theta_host, stream);
}

extern "C" int gl_theta_partial(const void *Z_tiles, int n, int tile_begin, int tile_end, int avg_degree, int dtype,
                                void *cand_out, void *stream) {
  if (n < 2) return fail(GL_E_SHAPE, "need at least 2 nodes, got n=%d", n);
  if (avg_degree < 1 || avg_degree >= n)
    return fail(GL_E_ARG, "avg_degree must be in [1, n-1], got %d (n=%d)", avg_degree, n);
  if (dtype != GL_F32 && dtype != GL_F64) return fail(GL_E_ARG, "dtype must be GL_F32 or GL_F64");
  if (tile_begin < 0 || tile_end > num_tiles(n) || tile_end < tile_begin) return fail(GL_E_ARG, "bad tile range");
  if (!cand_out) return fail(GL_E_ARG, "null output");
  return theta_launch(Z_tiles, GL_LAYOUT_TILED, n, n, 0, avg_degree, dtype, tile_begin, tile_end, nullptr, nullptr,
                      cand_out, as_stream(stream));
}

// ---- allocation-free variant ---------------------------------------------------------------------
// gl_theta_noalloc: gl_theta for callers that may not call cudaMalloc (CUDA graph capture around the
// call site, serving loops, harnesses that forbid allocation).  No counterpart in the reference.
// The per-row statistics that gl_theta keeps in a cudaMalloc'd
// buffer live in a file-scope device array instead: 2 * GL_THETA_NOALLOC_MAX_NODES + 2 doubles (1 MiB)
// held by every device that loads the library until the process exits.
//
// * In host code `g_theta_scratch` is only a host-side stand-in for the symbol, not a device address:
//   the address a kernel may use comes from cudaGetSymbolAddress, taken once per device with that device
//   current (theta_scratch_ptr below) and passed to the kernels like any other pointer.
// * Every word the call reads from the scratch is written earlier in the same call, on the same stream
//   (row statistics by theta_rows_kernel for all n rows, the two means by theta_mean_kernel), and
//   nothing is accumulated into it, so no part of it has to start at zero and nothing is cleared.  The
//   array keeps the last call's values otherwise.
// * All calls on a device share the one array: calls that overlap on the GPU (different streams, or
//   different host threads) race.  Serialise them, or use gl_theta.
constexpr int THETA_NOALLOC_MAX_NODES = GL_THETA_NOALLOC_MAX_NODES;  // include/gl_b200.h
constexpr int THETA_NOALLOC_MAX_DEVICES = 64;
static __device__ __align__(256) double g_theta_scratch[2 * (size_t)THETA_NOALLOC_MAX_NODES + 2];

static int theta_scratch_ptr(double **out) {
  static std::atomic<double *> cache[THETA_NOALLOC_MAX_DEVICES];  // zero-initialised
  int dev = -1;
  GL_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= THETA_NOALLOC_MAX_DEVICES) return fail(GL_E_ARG, "device ordinal %d out of range", dev);
  double *p = cache[dev].load(std::memory_order_acquire);
  if (!p) {  // once per device, with that device current; concurrent first calls get the same address
    cudaError_t e = cudaGetSymbolAddress((void **)&p, g_theta_scratch);
    if (e != cudaSuccess || !p)
      return fail(GL_E_CUDA, "cudaGetSymbolAddress(theta scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
    cache[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return GL_OK;
}

extern "C" int gl_theta_noalloc(const void *Z, int layout, int n, int avg_degree, int dtype, double *theta_host,
                                void *stream) {
  if (layout != GL_LAYOUT_PACKED && layout != GL_LAYOUT_TILED) return fail(GL_E_ARG, "bad layout");
  if (n < 2) return fail(GL_E_SHAPE, "need at least 2 nodes, got n=%d", n);
  if (n > THETA_NOALLOC_MAX_NODES)
    return fail(GL_E_SHAPE, "gl_theta_noalloc handles at most %d nodes (static scratch), got n=%d; use gl_theta",
                THETA_NOALLOC_MAX_NODES, n);
  if (avg_degree < 1 || avg_degree >= n)
    return fail(GL_E_ARG, "avg_degree must be in [1, n-1], got %d (n=%d)", avg_degree, n);
  if (!Z || !theta_host) return fail(GL_E_ARG, "null argument");
  if (dtype != GL_F32 && dtype != GL_F64) return fail(GL_E_ARG, "dtype must be GL_F32 or GL_F64");
  double *buf = nullptr;
  int rc = theta_scratch_ptr(&buf);
  if (rc != GL_OK) return rc;
  cudaStream_t st = as_stream(stream);
  rc = theta_launch(Z, layout, n, n, 0, avg_degree, dtype, 0, num_tiles(n), buf, buf + n, nullptr, st);
  if (rc != GL_OK) return rc;
  theta_mean_kernel<<<1, 256, 0, st>>>(buf, buf + n, n, buf + 2 * (size_t)n);
  g_launches.fetch_add(1);
  cudaError_t e = cudaGetLastError();
  double h[2] = {0, 0};
  if (e == cudaSuccess) e = cudaMemcpyAsync(h, buf + 2 * (size_t)n, sizeof(h), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) return fail(GL_E_CUDA, "gl_theta_noalloc failed: %s", cudaGetErrorString(e));
  *theta_host = sqrt(h[0] * h[1]);  // smooth_learning.py:219
  return GL_OK;
}
