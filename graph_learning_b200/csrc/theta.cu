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

template <typename F>
__global__ void __launch_bounds__(TH_THREADS) theta_rows_kernel(const F *Z, int layout, int n, int nt, int k,
                                                                double *lo_out, double *hi_out, int64_t ld) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  F *row = reinterpret_cast<F *>(smem_raw);
  __shared__ double s_val[TH_THREADS / 32];
  __shared__ int s_idx[TH_THREADS / 32];
  __shared__ double s_sorted_prev, s_sorted_last, s_b;
  const int i = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
  for (int j = tid; j < n; j += TH_THREADS) {
    F v = F(0);
    if (layout == LAYOUT_DENSE) {  // general (rows, n) matrix, row i as stored (block form, :220-235)
      v = Z[(int64_t)i * ld + j];
    } else if (j != i) {
      const int lo = j < i ? j : i, hi = j < i ? i : j;
      if (layout == GL_LAYOUT_TILED) {
        const int I = lo / T, J = hi / T;
        v = Z[(int64_t)tile_index(I, J, nt) * TT + (lo % T) * T + (hi % T)];
      } else {
        v = Z[packed_index(lo, hi, n)];
      }
    }
    row[j] = v;
  }
  if (tid == 0) {
    s_b = 0.0;
    s_sorted_prev = 0.0;
    s_sorted_last = 0.0;
  }
  __syncthreads();
  // k+1 rounds of arg-min extraction: round q yields sorted[q]
  for (int q = 0; q <= k; ++q) {
    double best = CUDART_INF;
    int bidx = -1;
    for (int j = tid; j < n; j += TH_THREADS) {
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
      if (q < k) s_b += bv;  // b = sum of sorted[0..k-1]   (:204)
      s_sorted_prev = s_sorted_last;
      s_sorted_last = bv;
    }
    __syncthreads();
  }
  if (tid == 0) {
    const double kk = (double)k, b = s_b, sk = s_sorted_last, sk1 = s_sorted_prev;
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

static int theta_impl(const void *Z, int layout, int rows, int n, int64_t ld, int avg_degree, int dtype,
                      double *theta_host, void *stream) {
  if (n < 2 || rows < 1) return fail(GL_E_SHAPE, "need at least 2 nodes, got n=%d", n);
  if (avg_degree < 1 || avg_degree >= n)
    return fail(GL_E_ARG, "avg_degree must be in [1, n-1], got %d (n=%d)", avg_degree, n);
  if (!theta_host) return fail(GL_E_ARG, "null output");
  const size_t esz = dtype == GL_F32 ? 4 : 8;
  const size_t smem = esz * (size_t)n;
  if (smem > 200 * 1024)
    return fail(GL_E_SHAPE, "gl_theta holds one row of Z in shared memory: n=%d exceeds the %d-node limit "
                            "of this dtype; use the kNN-support path for larger graphs", n, (int)(200 * 1024 / esz));
  cudaStream_t st = as_stream(stream);
  double *buf = nullptr;
  GL_CUDA(cudaMalloc((void **)&buf, sizeof(double) * (2 * (size_t)rows + 2)));
  const int nt = tiles_per_side(n);
  int rc = GL_OK;
  cudaError_t e = cudaSuccess;
  if (dtype == GL_F32) {
    e = cudaFuncSetAttribute(theta_rows_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess)
      theta_rows_kernel<float><<<rows, TH_THREADS, smem, st>>>((const float *)Z, layout, n, nt, avg_degree, buf, buf + rows, ld);
  } else if (dtype == GL_F64) {
    e = cudaFuncSetAttribute(theta_rows_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess)
      theta_rows_kernel<double><<<rows, TH_THREADS, smem, st>>>((const double *)Z, layout, n, nt, avg_degree, buf, buf + rows, ld);
  } else {
    rc = fail(GL_E_ARG, "dtype must be GL_F32 or GL_F64");
  }
  if (rc == GL_OK && e == cudaSuccess) {
    g_launches.fetch_add(1);
    e = cudaGetLastError();
  }
  double h[2] = {0, 0};
  if (rc == GL_OK && e == cudaSuccess) {
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
