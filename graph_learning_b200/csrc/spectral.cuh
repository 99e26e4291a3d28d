// spectral.cuh -- lambda_max of small symmetric PSD matrices inside a 256-thread CTA: the spectral norms
// of the batched stop rule (graph_learn/smooth_learning.py:158-159, np.linalg.norm(., ord=2) of (E, K) and
// (n, K) operands = sqrt(lambda_max) of their K x K Grams).  Shared by the per-iteration kernels
// (node_tail.cu) and the resident batched kernel (persistent_batch.cu).
#pragma once
#include <math_constants.h>

#include "solver.cuh"

namespace gl {

constexpr int NODE_THREADS = 256;

// lambda_max of four symmetric PSD K x K matrices at once (quarter of the block each) by repeated
// squaring B <- B^2 / tr(B^2) followed by a Rayleigh quotient with the original matrix.  With 26
// squarings the quotient's relative error is below 1e-8 for ANY spectrum (worst case over the
// gap is 1/(e*2^27)); sigma_max = sqrt(lambda_max), what np.linalg.norm(., ord=2) returns for
// the (E,K) / (n,K) operands at smooth_learning.py:158-159.
static __device__ void lambda_max4(const double *const *G, int K, double *scratch, double *out) {
  const int q = threadIdx.x / (NODE_THREADS / 4), lt = threadIdx.x % (NODE_THREADS / 4);
  constexpr int L = NODE_THREADS / 4;
  const int KK = K * K;
  const double *g = G[q];
  if (K == 1) {
    if (lt == 0) out[q] = g[0];
    __syncthreads();
    return;
  }
  // the squarings ping-pong between two buffers: in shared memory for K <= 16 (26 x 3 barriers with a
  // global round trip each measured ~50 us at K = 8), in the global scratch beyond that
  __shared__ double s_mat[4][2][256];
  double *B = KK <= 256 ? s_mat[q][0] : scratch + (int64_t)q * 2 * KK;
  double *C = KK <= 256 ? s_mat[q][1] : B + KK;
  __shared__ double s_tr[4];
  __shared__ int s_bad[4];
  if (lt == 0) {
    double tr = 0.0;
    for (int i = 0; i < K; ++i) tr += g[i * K + i];
    s_tr[q] = tr;
    s_bad[q] = !(tr > 0.0) || !isfinite(tr);
  }
  __syncthreads();
  if (!s_bad[q])
    for (int e = lt; e < KK; e += L) B[e] = g[e] / s_tr[q];
  __syncthreads();
  for (int it = 0; it < 26; ++it) {
    if (!s_bad[q])
      for (int e = lt; e < KK; e += L) {
        const int i = e / K, j = e % K;
        double acc = 0.0;
        for (int m = 0; m < K; ++m) acc += B[i * K + m] * B[m * K + j];
        C[e] = acc;
      }
    __syncthreads();
    if (lt == 0 && !s_bad[q]) {
      double tr = 0.0;
      for (int i = 0; i < K; ++i) tr += C[i * K + i];
      s_tr[q] = tr;
    }
    __syncthreads();
    if (!s_bad[q])
      for (int e = lt; e < KK; e += L) B[e] = C[e] / s_tr[q];
    __syncthreads();
  }
  if (lt == 0) {
    if (s_bad[q]) {
      double tr = 0.0;
      for (int i = 0; i < K; ++i) tr += g[i * K + i];
      out[q] = (tr == 0.0) ? 0.0 : CUDART_NAN;
    } else {
      int col = 0;
      for (int i = 1; i < K; ++i)
        if (B[i * K + i] > B[col * K + col]) col = i;
      double num = 0.0, den = 0.0;
      for (int i = 0; i < K; ++i) {
        double gv = 0.0;
        for (int j = 0; j < K; ++j) gv += g[i * K + j] * B[j * K + col];
        num += B[i * K + col] * gv;
        den += B[i * K + col] * B[i * K + col];
      }
      out[q] = num / den;
    }
  }
  __syncthreads();
}

}  // namespace gl
