// layout.cu -- packed <-> tiled <-> square conversions and the stand-alone operator pair.
//
// "packed" is the reference's edge order (operators.py:39-41, scipy squareform); "tiled" is the
// resident layout of the solver (see include/gl_b200.h).  The operator kernels replace
// sum_squareform's CSR products (smooth_learning.py:13-44) and the dense-Laplacian forms
// (operators.py:44-94) for callers outside the solver loop.
#include <vector>

#include "solver.cuh"

namespace gl {

template <typename F>
__global__ void __launch_bounds__(256) pack_to_tiles_kernel(const F *packed, F *tiled, int n, int nt,
                                                            int tile_begin, const double *scale) {
  const int t = blockIdx.x + tile_begin, k = blockIdx.y;
  // decode (I, J) from t: small loop, nt <= a few thousand
  int I = 0, rem = t;
  while (rem >= nt - I) {
    rem -= nt - I;
    ++I;
  }
  const int J = I + rem;
  const int64_t E = (int64_t)n * (n - 1) / 2;
  const F *src = packed + (int64_t)k * E;
  F *dst = tiled + ((int64_t)k * gridDim.x + blockIdx.x) * TT;
  const double sc = scale ? scale[k] : 1.0;
  for (int idx = threadIdx.x; idx < TT; idx += blockDim.x) {
    const int r = idx / T, c = idx % T;
    const int gi = I * T + r, gj = J * T + c;
    F v = F(0);
    if (gj > gi && gj < n) {
      v = src[packed_index(gi, gj, n)];
      if (scale) v = (F)((double)v * sc);
    }
    dst[idx] = v;
  }
}

template <typename F>
__global__ void __launch_bounds__(256) tiles_to_packed_kernel(const F *tiled, F *packed, int n, int nt,
                                                              int tile_begin, double edge_tol) {
  const int t = blockIdx.x + tile_begin, k = blockIdx.y;
  int I = 0, rem = t;
  while (rem >= nt - I) {
    rem -= nt - I;
    ++I;
  }
  const int J = I + rem;
  const int64_t E = (int64_t)n * (n - 1) / 2;
  F *dst = packed + (int64_t)k * E;
  const F *src = tiled + ((int64_t)k * gridDim.x + blockIdx.x) * TT;
  for (int idx = threadIdx.x; idx < TT; idx += blockDim.x) {
    const int r = idx / T, c = idx % T;
    const int gi = I * T + r, gj = J * T + c;
    if (gj > gi && gj < n) {
      F v = src[idx];
      if ((double)v < edge_tol) v = F(0);  // smooth_learning.py:169
      dst[packed_index(gi, gj, n)] = v;
    }
  }
}

// tiled -> dense symmetric (n, n): 32x32 sub-blocks, mirror written through shared memory so both
// the (i,j) and the (j,i) stores are coalesced.
template <typename F>
__global__ void __launch_bounds__(256) tiles_to_square_kernel(const F *tiled, F *sq, int n, int nt,
                                                              double edge_tol) {
  __shared__ F s_blk[32][33];
  const int t = blockIdx.x;
  int I = 0, rem = t;
  while (rem >= nt - I) {
    rem -= nt - I;
    ++I;
  }
  const int J = I + rem;
  const F *src = tiled + (int64_t)t * TT;
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  for (int sb = 0; sb < 16; ++sb) {
    const int br = (sb / 4) * 32, bc = (sb % 4) * 32;
    if (I == J && br > bc) continue;  // strictly-lower sub-blocks of a diagonal tile: mirrors only
#pragma unroll
    for (int rr = 0; rr < 4; ++rr) {
      const int r = br + wp * 4 + rr, c = bc + lane;
      const int gi = I * T + r, gj = J * T + c;
      F v = F(0);
      if (gj > gi && gj < n) {
        v = src[r * T + c];
        if ((double)v < edge_tol) v = F(0);
      }
      s_blk[wp * 4 + rr][lane] = v;
      if (gi < n && gj < n && gj >= gi) sq[(int64_t)gi * n + gj] = v;  // includes the zero diagonal
    }
    __syncthreads();
#pragma unroll
    for (int rr = 0; rr < 4; ++rr) {
      // transposed: output row = column index of the block
      const int c = bc + wp * 4 + rr, r = br + lane;
      const int gi = I * T + r, gj = J * T + c;
      if (gi < n && gj < n && gj > gi) sq[(int64_t)gj * n + gi] = s_blk[lane][wp * 4 + rr];
    }
    __syncthreads();
  }
}

// ---- S w on packed input: row sums (warp per row) + column sums (thread per column) -----------
template <typename F>
__global__ void sumop_rows_kernel(const F *w, double *deg_acc, int n) {
  const int k = blockIdx.y;
  const int64_t E = (int64_t)n * (n - 1) / 2;
  const int row = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= n) return;
  const F *src = w + (int64_t)k * E + packed_index(row, row + 1, n) - (row + 1);
  double acc = 0.0;
  for (int j = row + 1 + lane; j < n; j += 32) acc += (double)src[j];
  acc = warp_sum(acc);
  if (lane == 0) deg_acc[(int64_t)k * n + row] = acc;
}
template <typename F, typename O>
__global__ void sumop_cols_kernel(const F *w, const double *deg_acc, O *deg, int n) {
  const int k = blockIdx.y;
  const int64_t E = (int64_t)n * (n - 1) / 2;
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const F *src = w + (int64_t)k * E;
  double acc = 0.0;
  for (int i = 0; i < j; ++i) acc += (double)src[packed_index(i, j, n)];
  deg[(int64_t)k * n + j] = (O)(acc + deg_acc[(int64_t)k * n + j]);
}

template <typename F>
__global__ void sumop_adj_kernel(const F *d, F *edge, int n) {
  const int k = blockIdx.y, i = blockIdx.x;
  const int64_t E = (int64_t)n * (n - 1) / 2;
  const F di = d[(int64_t)k * n + i];
  F *dst = edge + (int64_t)k * E + packed_index(i, i + 1, n) - (i + 1);
  for (int j = i + 1 + threadIdx.x; j < n; j += blockDim.x) dst[j] = di + d[(int64_t)k * n + j];
}

// Laplacian from packed weights: off-diagonal -w_ij, diagonal = degree (operators.py:44-70)
template <typename F>
__global__ void laplacian_kernel(const F *w, const F *deg, F *lap, int n) {
  const int k = blockIdx.y, i = blockIdx.x;
  const int64_t E = (int64_t)n * (n - 1) / 2;
  const F *src = w + (int64_t)k * E;
  F *dst = lap + ((int64_t)k * n + i) * n;
  for (int j = threadIdx.x; j < n; j += blockDim.x) {
    F v;
    if (j == i) v = deg[(int64_t)k * n + i];
    else if (j > i) v = -src[packed_index(i, j, n)];
    else v = -src[packed_index(j, i, n)];
    dst[j] = v;
  }
}
// adjoint: L -> L_ii + L_jj - L_ij - L_ji over i<j (operators.py:73-94)
template <typename F>
__global__ void laplacian_adj_kernel(const F *lap, F *edge, int n) {
  const int k = blockIdx.y, i = blockIdx.x;
  const int64_t E = (int64_t)n * (n - 1) / 2;
  const F *L = lap + (int64_t)k * n * n;
  const F lii = L[(int64_t)i * n + i];
  F *dst = edge + (int64_t)k * E + packed_index(i, i + 1, n) - (i + 1);
  for (int j = i + 1 + threadIdx.x; j < n; j += blockDim.x)
    dst[j] = lii + L[(int64_t)j * n + j] - L[(int64_t)i * n + j] - L[(int64_t)j * n + i];
}

int fill_tile_table(gl_solver *s) {
  std::vector<int2> h(s->ntl);
  int I = 0, rem = s->tile_begin;
  while (I < s->nt && rem >= s->nt - I) {
    rem -= s->nt - I;
    ++I;
  }
  int J = I + rem;
  for (int t = 0; t < s->ntl; ++t) {
    h[t] = make_int2(I, J);
    if (++J == s->nt) {
      ++I;
      J = I;
    }
  }
  if (s->ntl)
    GL_CUDA(cudaMemcpy(s->tile_ij, h.data(), sizeof(int2) * s->ntl, cudaMemcpyHostToDevice));
  return GL_OK;
}

}  // namespace gl

using namespace gl;

#define DISPATCH_DTYPE(dtype, CALL_F32, CALL_F64)                      \
  do {                                                                 \
    if ((dtype) == GL_F32) { CALL_F32; }                               \
    else if ((dtype) == GL_F64) { CALL_F64; }                          \
    else return fail(GL_E_ARG, "dtype must be GL_F32 or GL_F64");      \
  } while (0)

static int check_n(int n, int K) {
  if (n < 2) return fail(GL_E_SHAPE, "need at least 2 nodes, got n=%d", n);
  if (K < 1) return fail(GL_E_SHAPE, "need at least one graph, got K=%d", K);
  return GL_OK;
}

namespace gl {
// scale_dev: device array of K doubles or nullptr
int pack_to_tiles_impl(const void *packed, void *tiled, int n, int K, int dtype, int tb, int te,
                       const double *scale_dev, cudaStream_t st) {
  if (te <= tb) return GL_OK;
  dim3 grid(te - tb, K);
  const int nt = tiles_per_side(n);
  DISPATCH_DTYPE(dtype,
                 (pack_to_tiles_kernel<float><<<grid, 256, 0, st>>>((const float *)packed, (float *)tiled, n,
                                                                   nt, tb, scale_dev)),
                 (pack_to_tiles_kernel<double><<<grid, 256, 0, st>>>((const double *)packed,
                                                                    (double *)tiled, n, nt, tb, scale_dev)));
  GL_LAUNCH_CHECK();
  return GL_OK;
}
}  // namespace gl

extern "C" {

int gl_pack_to_tiles(const void *packed, void *tiled, int n, int K, int dtype, int tb, int te,
                     void *stream) {
  GL_TRY(check_n(n, K));
  if (tb < 0 || te > num_tiles(n) || te < tb) return fail(GL_E_ARG, "bad tile range [%d,%d)", tb, te);
  return pack_to_tiles_impl(packed, tiled, n, K, dtype, tb, te, nullptr, as_stream(stream));
}

int gl_tiles_to_packed(const void *tiled, void *packed, int n, int K, int dtype, int tb, int te,
                       void *stream) {
  GL_TRY(check_n(n, K));
  if (tb < 0 || te > num_tiles(n) || te < tb) return fail(GL_E_ARG, "bad tile range [%d,%d)", tb, te);
  if (te == tb) return GL_OK;
  dim3 grid(te - tb, K);
  const int nt = tiles_per_side(n);
  cudaStream_t st = as_stream(stream);
  const double keep = -INFINITY;
  DISPATCH_DTYPE(dtype,
                 (tiles_to_packed_kernel<float><<<grid, 256, 0, st>>>((const float *)tiled, (float *)packed,
                                                                     n, nt, tb, keep)),
                 (tiles_to_packed_kernel<double><<<grid, 256, 0, st>>>((const double *)tiled,
                                                                      (double *)packed, n, nt, tb, keep)));
  GL_LAUNCH_CHECK();
  return GL_OK;
}

int gl_tiles_to_square(const void *tiled, void *sq, int n, int dtype, double edge_tol, void *stream) {
  GL_TRY(check_n(n, 1));
  const int nt = tiles_per_side(n);
  cudaStream_t st = as_stream(stream);
  DISPATCH_DTYPE(dtype,
                 (tiles_to_square_kernel<float><<<num_tiles(n), 256, 0, st>>>((const float *)tiled,
                                                                             (float *)sq, n, nt, edge_tol)),
                 (tiles_to_square_kernel<double><<<num_tiles(n), 256, 0, st>>>((const double *)tiled,
                                                                              (double *)sq, n, nt, edge_tol)));
  GL_LAUNCH_CHECK();
  return GL_OK;
}

int gl_sum_op(const void *w, void *deg, int n, int K, int dtype, void *stream) {
  GL_TRY(check_n(n, K));
  cudaStream_t st = as_stream(stream);
  double *acc = nullptr;
  GL_CUDA(cudaMallocAsync(&acc, sizeof(double) * (size_t)n * K, st));
  dim3 g1((n + 7) / 8, K), g2((n + 127) / 128, K);
  if (dtype == GL_F32) {
    sumop_rows_kernel<float><<<g1, 256, 0, st>>>((const float *)w, acc, n);
    GL_LAUNCH_CHECK();
    sumop_cols_kernel<float, float><<<g2, 128, 0, st>>>((const float *)w, acc, (float *)deg, n);
    GL_LAUNCH_CHECK();
  } else if (dtype == GL_F64) {
    sumop_rows_kernel<double><<<g1, 256, 0, st>>>((const double *)w, acc, n);
    GL_LAUNCH_CHECK();
    sumop_cols_kernel<double, double><<<g2, 128, 0, st>>>((const double *)w, acc, (double *)deg, n);
    GL_LAUNCH_CHECK();
  } else {
    cudaFreeAsync(acc, st);
    return fail(GL_E_ARG, "dtype must be GL_F32 or GL_F64");
  }
  GL_CUDA(cudaFreeAsync(acc, st));
  return GL_OK;
}

int gl_sum_op_adj(const void *d, void *edge, int n, int K, int dtype, void *stream) {
  GL_TRY(check_n(n, K));
  cudaStream_t st = as_stream(stream);
  dim3 grid(n - 1, K);
  DISPATCH_DTYPE(dtype, (sumop_adj_kernel<float><<<grid, 256, 0, st>>>((const float *)d, (float *)edge, n)),
                 (sumop_adj_kernel<double><<<grid, 256, 0, st>>>((const double *)d, (double *)edge, n)));
  GL_LAUNCH_CHECK();
  return GL_OK;
}

int gl_laplacian_from_packed(const void *w, void *lap, int n, int K, int dtype, void *stream) {
  GL_TRY(check_n(n, K));
  cudaStream_t st = as_stream(stream);
  const size_t esz = dtype == GL_F32 ? 4 : 8;
  void *deg = nullptr;
  GL_CUDA(cudaMallocAsync(&deg, esz * (size_t)n * K, st));
  int rc = gl_sum_op(w, deg, n, K, dtype, stream);
  if (rc == GL_OK) {
    dim3 grid(n, K);
    if (dtype == GL_F32)
      laplacian_kernel<float><<<grid, 256, 0, st>>>((const float *)w, (const float *)deg, (float *)lap, n);
    else
      laplacian_kernel<double><<<grid, 256, 0, st>>>((const double *)w, (const double *)deg, (double *)lap, n);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (cudaGetLastError() != cudaSuccess) rc = fail(GL_E_CUDA, "laplacian kernel launch failed");
  }
  cudaFreeAsync(deg, st);
  return rc;
}

int gl_laplacian_adjoint(const void *lap, void *edge, int n, int K, int dtype, void *stream) {
  GL_TRY(check_n(n, K));
  cudaStream_t st = as_stream(stream);
  dim3 grid(n - 1, K);
  DISPATCH_DTYPE(dtype,
                 (laplacian_adj_kernel<float><<<grid, 256, 0, st>>>((const float *)lap, (float *)edge, n)),
                 (laplacian_adj_kernel<double><<<grid, 256, 0, st>>>((const double *)lap, (double *)edge, n)));
  GL_LAUNCH_CHECK();
  return GL_OK;
}

}  // extern "C"
