// knn_build.cu -- N3: k nearest neighbours between the COLUMNS of X, the front end of
// the kNN-support LogModel (config 5; Kalofolias & Perraudin 2019 restrict the support to a kNN
// graph).  The reference has no code for it.  Brute force and exact: squared distances by direct
// differences in fp32 (the arithmetic of pdist.cu, so d(i,j) == d(j,i) bit for bit), fused with the
// selection -- the n x n distance matrix is never written.
//
// One CTA owns 128 query nodes and streams all candidate nodes past them in tiles of 128:
//   1. 16x16 threads x 8x8 distances per thread, operands staged 16 features at a time;
//   2. the 128x128 tile goes to shared memory (self pairs and j >= n as +inf);
//   3. each warp scans 16 query rows: 4 coalesced reads per lane, a ballot against the row's current
//      k-th best tau -- almost always empty after the first tiles -- and a warp-parallel sorted
//      insertion (lane l holds entry l of the ascending list) for the few survivors.
// Ties keep the smaller node index first (candidates arrive in ascending order).
#include <math_constants.h>

#include "common.cuh"

namespace gl {

constexpr int KB_KC = 16;     // features per shared-memory stage
constexpr int KB_KMAX = 32;   // list length = warp width

struct KnnBuildSmem {
  float xi[KB_KC][T];
  float xj[KB_KC][T];
  float tile[T][T + 1];
  float best_d[T][KB_KMAX];
  int best_i[T][KB_KMAX];
  float tau[T];
};

__global__ void __launch_bounds__(256, 1)
knn_build_kernel(const float *__restrict__ X, int64_t m, int n, int64_t ldx, int k, int32_t *__restrict__ idx_out,
                 float *__restrict__ d2_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  KnnBuildSmem &S = *reinterpret_cast<KnnBuildSmem *>(smem_raw);
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4, lane = tid & 31, wq = tid >> 5;
  const int I = blockIdx.x, nt = tiles_per_side(n);
  const int lcol = tid & (T - 1), lrow0 = tid >> 7;
  const int gi_l = I * T + lcol;

  for (int x = tid; x < T * KB_KMAX; x += 256) {
    S.best_d[x / KB_KMAX][x % KB_KMAX] = CUDART_INF_F;
    S.best_i[x / KB_KMAX][x % KB_KMAX] = -1;
  }
  if (tid < T) S.tau[tid] = CUDART_INF_F;
  __syncthreads();

  for (int J = 0; J < nt; ++J) {
    const int gj_l = J * T + lcol;
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    for (int64_t t0 = 0; t0 < m; t0 += KB_KC) {
#pragma unroll
      for (int q = 0; q < KB_KC / 2; ++q) {
        const int64_t tt = t0 + lrow0 + 2 * q;
        float vi = 0.f, vj = 0.f;
        if (tt < m) {
          if (gi_l < n) vi = X[tt * ldx + gi_l];
          if (gj_l < n) vj = X[tt * ldx + gj_l];
        }
        S.xi[lrow0 + 2 * q][lcol] = vi;
        S.xj[lrow0 + 2 * q][lcol] = vj;
      }
      __syncthreads();
      const int kend = (int)min((int64_t)KB_KC, m - t0);
      for (int kk = 0; kk < kend; ++kk) {
        float xi[8], xj[8];
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            xi[h * 4 + e] = S.xi[kk][h * 64 + ty * 4 + e];
            xj[h * 4 + e] = S.xj[kk][h * 64 + tx * 4 + e];
          }
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float df = xi[i] - xj[j];
            acc[i][j] += df * df;
          }
      }
      __syncthreads();
    }
    // ---- tile to shared memory -----------------------------------------------------------------
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int r = (i >> 2) * 64 + ty * 4 + (i & 3), c = (j >> 2) * 64 + tx * 4 + (j & 3);
        const int gi = I * T + r, gj = J * T + c;
        S.tile[r][c] = (gj < n && gj != gi) ? acc[i][j] : CUDART_INF_F;
      }
    __syncthreads();
    // ---- selection: warp wq owns rows wq*16 .. wq*16+15 ---------------------------------------------
    for (int rr = 0; rr < T / 8; ++rr) {
      const int r = wq * (T / 8) + rr;
      float tau = S.tau[r];
      float v[4];
      unsigned any = 0;
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        v[t] = S.tile[r][lane + 32 * t];
        any |= __ballot_sync(0xffffffffu, v[t] < tau);
      }
      if (!any) continue;
      float bd = S.best_d[r][lane];
      int bi = S.best_i[r][lane];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        unsigned mask = __ballot_sync(0xffffffffu, v[t] < tau);
        while (mask) {
          const int L = __ffs(mask) - 1;
          const float cv = __shfl_sync(0xffffffffu, v[t], L);
          const int ci = J * T + L + 32 * t;
          // position = number of kept entries <= cv (ties: earlier index first)
          const int pos = __popc(__ballot_sync(0xffffffffu, lane < k && bd <= cv));
          const float pd = __shfl_up_sync(0xffffffffu, bd, 1);
          const int pi = __shfl_up_sync(0xffffffffu, bi, 1);
          if (lane == pos) {
            bd = cv;
            bi = ci;
          } else if (lane > pos) {
            bd = pd;
            bi = pi;
          }
          tau = __shfl_sync(0xffffffffu, bd, k - 1);
          mask &= mask - 1;
          mask &= __ballot_sync(0xffffffffu, v[t] < tau);
        }
      }
      S.best_d[r][lane] = bd;
      S.best_i[r][lane] = bi;
      if (lane == 0) S.tau[r] = tau;
    }
    __syncthreads();
  }
  // ---- results ------------------------------------------------------------------------------------
  for (int x = tid; x < T * k; x += 256) {
    const int r = x / k, l = x % k;
    const int gi = I * T + r;
    if (gi < n) {
      idx_out[(int64_t)gi * k + l] = S.best_i[r][l];
      d2_out[(int64_t)gi * k + l] = S.best_d[r][l];
    }
  }
}

}  // namespace gl
using namespace gl;

extern "C" int gl_knn_build(const void *X, int64_t m, int n, int64_t ldx, int k, int32_t *idx_out, void *d2_out,
                            void *stream) {
  if (n < 2) return fail(GL_E_SHAPE, "need at least 2 nodes, got n=%d", n);
  if (m < 1 || ldx < n) return fail(GL_E_SHAPE, "bad signal matrix shape (m=%lld, ldx=%lld, n=%d)",
                                    (long long)m, (long long)ldx, n);
  if (k < 1 || k > KB_KMAX) return fail(GL_E_ARG, "k must be in 1..%d, got %d", KB_KMAX, k);
  const size_t smem = sizeof(KnnBuildSmem);
  GL_CUDA(cudaFuncSetAttribute(knn_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  knn_build_kernel<<<tiles_per_side(n), 256, smem, as_stream(stream)>>>((const float *)X, m, n, ldx, k, idx_out,
                                                                        (float *)d2_out);
  GL_LAUNCH_CHECK();
  return GL_OK;
}
