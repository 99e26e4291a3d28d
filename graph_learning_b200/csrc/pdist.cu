// pdist.cu -- K1 (exact form): pairwise squared distances between the COLUMNS of X by direct
// differences, written straight into tiled or packed edge order.
//
// Replaces `pdist(x.T) ** 2` (graph_learn/smooth_learning.py:278, clustering/kgraphs.py:96) and the
// broadcast-difference form of the GLMM M-step (clustering/glmm.py:67-68) via row_scale/col_shift:
//   y[t,i] = row_scale[t] * (X[t,i] - col_shift[i]),   Z_ij = sum_t (y[t,i] - y[t,j])^2.
// Direct differences do not cancel the way the Gram form ||x_i||^2+||x_j||^2-2<x_i,x_j> does on
// smooth signals (SURVEY 7.2 "Z precision"), so this kernel is the parity anchor for the
// tensor-core Gram kernel.  One CTA per 128x128 tile, 16x16 threads, 8x8 outputs per thread;
// fp32 mode accumulates 256-sample chunks in fp32 registers and folds them into float64 tile
// accumulators in shared memory.  Bound: FP32/FP64 FMA pipe (2 instructions per pair-sample).
#include "solver.cuh"

namespace gl {

constexpr int PD_KC = 16;       // samples per shared-memory stage
constexpr int PD_FLUSH = 16;    // stages per fp32 chunk (256 samples)

template <typename F>
__global__ void __launch_bounds__(256, 1)
pdist_kernel(const F *__restrict__ X, int64_t m, int n, int64_t ldx, const F *__restrict__ row_scale,
             const F *__restrict__ col_shift, F *__restrict__ Z, int layout, int tile_begin, int nt) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  F(*s_xi)[T] = reinterpret_cast<F(*)[T]>(smem_raw);
  F(*s_xj)[T] = reinterpret_cast<F(*)[T]>(smem_raw + sizeof(F) * PD_KC * T);
  double *s_out = reinterpret_cast<double *>(smem_raw + 2 * sizeof(F) * PD_KC * T);  // fp32 mode only

  const int t = blockIdx.x + tile_begin;
  int I = 0, rem = t;
  while (rem >= nt - I) {
    rem -= nt - I;
    ++I;
  }
  const int J = I + rem;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  constexpr bool F32 = sizeof(F) == 4;

  // loader mapping: 256 threads x 16 elements = 2 x (16 x 128)
  const int lcol = tid & (T - 1), lrow0 = tid >> 7;  // rows lrow0, lrow0+2, ...
  const int gi_l = I * T + lcol, gj_l = J * T + lcol;
  const F shi = (col_shift && gi_l < n) ? col_shift[gi_l] : F(0);
  const F shj = (col_shift && gj_l < n) ? col_shift[gj_l] : F(0);

  F acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = F(0);
  if (F32)
    for (int idx = tid; idx < TT; idx += 256) s_out[idx] = 0.0;

  // thread's output rows: ty*4+{0..3} and 64+ty*4+{0..3}; columns likewise with tx
  F pre_i[PD_KC / 2], pre_j[PD_KC / 2];
  auto fetch = [&](int64_t t0) {
#pragma unroll
    for (int q = 0; q < PD_KC / 2; ++q) {
      const int64_t tt = t0 + lrow0 + 2 * q;
      F vi = F(0), vj = F(0);
      if (tt < m) {
        const F sc = row_scale ? row_scale[tt] : F(1);
        if (gi_l < n) vi = sc * (X[tt * ldx + gi_l] - shi);
        if (gj_l < n) vj = sc * (X[tt * ldx + gj_l] - shj);
      }
      pre_i[q] = vi;
      pre_j[q] = vj;
    }
  };
  auto flush = [&]() {
    if (!F32) return;
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int r = (i >> 2) * 64 + ty * 4 + (i & 3), c = (j >> 2) * 64 + tx * 4 + (j & 3);
        s_out[r * T + c] += (double)acc[i][j];
        acc[i][j] = F(0);
      }
  };

  fetch(0);
  int stage = 0;
  for (int64_t t0 = 0; t0 < m; t0 += PD_KC) {
    __syncthreads();
#pragma unroll
    for (int q = 0; q < PD_KC / 2; ++q) {
      s_xi[lrow0 + 2 * q][lcol] = pre_i[q];
      s_xj[lrow0 + 2 * q][lcol] = pre_j[q];
    }
    __syncthreads();
    if (t0 + PD_KC < m) fetch(t0 + PD_KC);
#pragma unroll
    for (int kk = 0; kk < PD_KC; ++kk) {
      F xi[8], xj[8];
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          xi[h * 4 + e] = s_xi[kk][h * 64 + ty * 4 + e];
          xj[h * 4 + e] = s_xj[kk][h * 64 + tx * 4 + e];
        }
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const F df = xi[i] - xj[j];
          acc[i][j] += df * df;
        }
    }
    if (++stage == PD_FLUSH) {
      stage = 0;
      flush();
    }
  }
  flush();

  // ---- write the tile ---------------------------------------------------------------------------
  F *ztile = Z + (int64_t)blockIdx.x * TT;
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int r = (i >> 2) * 64 + ty * 4 + (i & 3), c = (j >> 2) * 64 + tx * 4 + (j & 3);
      const int gi = I * T + r, gj = J * T + c;
      const bool ok = gj > gi && gj < n;
      const F v = F32 ? (F)s_out[r * T + c] : acc[i][j];
      if (layout == GL_LAYOUT_TILED)
        ztile[r * T + c] = ok ? v : F(0);
      else if (ok)
        Z[packed_index(gi, gj, n)] = v;
    }
}

}  // namespace gl
using namespace gl;

extern "C" int gl_pdist_sq(const void *X, int64_t m, int n, int64_t ldx, const void *row_scale,
                           const void *col_shift, void *Z, int layout, int tile_begin, int tile_end,
                           int dtype, void *stream) {
  if (n < 2) return fail(GL_E_SHAPE, "need at least 2 nodes, got n=%d", n);
  if (m < 0 || ldx < n) return fail(GL_E_SHAPE, "bad signal matrix shape (m=%lld, ldx=%lld, n=%d)",
                                    (long long)m, (long long)ldx, n);
  if (layout != GL_LAYOUT_PACKED && layout != GL_LAYOUT_TILED) return fail(GL_E_ARG, "bad layout");
  if (layout == GL_LAYOUT_PACKED) {
    tile_begin = 0;
    tile_end = num_tiles(n);
  }
  if (tile_begin < 0 || tile_end > num_tiles(n) || tile_end < tile_begin)
    return fail(GL_E_ARG, "bad tile range [%d,%d)", tile_begin, tile_end);
  if (tile_end == tile_begin) return GL_OK;
  cudaStream_t st = as_stream(stream);
  const int nt = tiles_per_side(n);
  if (dtype == GL_F32) {
    const size_t smem = 2 * sizeof(float) * PD_KC * T + sizeof(double) * TT;
    GL_CUDA(cudaFuncSetAttribute(pdist_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    pdist_kernel<float><<<tile_end - tile_begin, 256, smem, st>>>(
        (const float *)X, m, n, ldx, (const float *)row_scale, (const float *)col_shift, (float *)Z, layout,
        tile_begin, nt);
  } else if (dtype == GL_F64) {
    const size_t smem = 2 * sizeof(double) * PD_KC * T;
    pdist_kernel<double><<<tile_end - tile_begin, 256, smem, st>>>(
        (const double *)X, m, n, ldx, (const double *)row_scale, (const double *)col_shift, (double *)Z,
        layout, tile_begin, nt);
  } else {
    return fail(GL_E_ARG, "dtype must be GL_F32 or GL_F64");
  }
  GL_LAUNCH_CHECK();
  return GL_OK;
}
