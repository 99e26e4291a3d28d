// edge_pass.cu -- K3: the fused per-iteration sweep over the tiled edge vector.
//
// Replaces the edge-sized work of the reference loop body, graph_learn/smooth_learning.py:147-164,
// together with all four applications of the sum operator S / S^T (smooth_learning.py:13-44):
//
//   y  = (1-2 gamma beta) w - gamma (d_i + d_j)        :147   (S^T d is a gather of two node values)
//   p  = max(0, y - 2 gamma z)                         :150
//   q  = (1-2 gamma beta) p - gamma (pb_i + pb_j)      :153   (pb = dual prox output, known up front)
//   w' = w - y + q                                     :163
//   S p, S w'   accumulated per node                   :154, next iteration's :148
//   ||q - y||^2, ||w'||^2                              :158 (numerator; next iteration's denominator)
//
// One CTA owns one 128x128 tile (I,J) of the node-pair triangle: 8 warps x 16 rows, each lane
// owns 4 columns (one 16-byte vector per row for fp32, two for fp64).  Column partial sums live
// in registers for the whole tile, row partial sums are reduced across the warp with a
// reduce-scatter butterfly (31 shuffles per 16 rows x 2 quantities), and each tile issues exactly
// one float64 atomic per (node, quantity): 2 x 128 row atomics + 2 x 128 column atomics.
// Algorithmic traffic: read w, read z, write w = 12 B/edge fp32, 24 B/edge fp64.
#include "edge_tile.cuh"

namespace gl {

template <typename F, bool BATCH>
__global__ void __launch_bounds__(EDGE_THREADS, sizeof(F) == 4 ? 2 : 1) edge_pass_kernel(const EdgeArgs<F> A) {
  __shared__ EdgeSmem<F> sm;
  __shared__ int s_last;

  if (*reinterpret_cast<volatile int *>(A.flags + FLAG_DONE)) return;

  const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
  const int t = blockIdx.x, k = blockIdx.y;
  // peer mode: accumulate into the buffer of the current exchange parity
  const int seq = A.world > 1 ? ld_volatile_int(A.hdr + HDR_SEQ) : 0;
  double *const xb = A.xch + (seq & 1) * A.xch_stride;
  double *const g_sp = xb, *const g_sw = xb + (int64_t)gridDim.y * A.npad;
  double *const g_sum = xb + 2 * (int64_t)gridDim.y * A.npad;
  double r2 = 0.0, w2 = 0.0;
  edge_cta_tile<F, BATCH, false>(A, sm, t, k, g_sp, g_sw, r2, w2);
  if (BATCH) {
    // batch mode: only the DIAGONALS of Gram(q - y) and Gram(w') are kept up every iteration (the
    // graph's own squared norms); the node tail bounds the spectral-norm stop rule with them and asks
    // for the full K x K Grams (gram_edges_kernel) only when the bound cannot rule the stop out
    if (tid == 0) {
      const int K = gridDim.y;
      atomicAdd(g_sum + (int64_t)k * K + k, r2);
      atomicAdd(g_sum + (int64_t)K * K + (int64_t)k * K + k, w2);
    }
    return;
  }

  // ---- residual sums: fixed-order two-stage reduction (deterministic) -------------------------
  if (tid == 0) {
    A.cta_partial[2 * t] = r2;
    A.cta_partial[2 * t + 1] = w2;
    __threadfence();
    const int ticket = atomicAdd(A.flags + FLAG_EDGE_TICKET, 1);
    s_last = (ticket == (int)gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  r2 = w2 = 0.0;
  for (int i = tid; i < A.ntl; i += EDGE_THREADS) {
    r2 += __ldcg(A.cta_partial + 2 * i);
    w2 += __ldcg(A.cta_partial + 2 * i + 1);
  }
  r2 = warp_sum(r2);
  w2 = warp_sum(w2);
  __syncthreads();
  if (lane == 0) {
    sm.red[wp][0] = r2;
    sm.red[wp][1] = w2;
  }
  __syncthreads();
  if (tid == 0) {
    r2 = w2 = 0.0;
    for (int q = 0; q < EDGE_THREADS / 32; ++q) {
      r2 += sm.red[q][0];
      w2 += sm.red[q][1];
    }
    g_sum[0] += r2;  // xch slot is zeroed by the node tail; += keeps the sharded sum convention
    g_sum[1] += w2;
    A.flags[FLAG_EDGE_TICKET] = 0;
  }
}

// ---- initial degrees: d0 = S w0 and ||w0||^2 (smooth_learning.py:144) ---------------------------
// Not on the per-iteration path; one warp per tile row, plain shuffles.
template <typename F>
__global__ void __launch_bounds__(EDGE_THREADS) edge_init_kernel(const F *w, int64_t gstride,
                                                                 const int2 *tile_ij, double *xch,
                                                                 int64_t xch_stride, const int *hdr, int world,
                                                                 int n, int npad, int K) {
  double *const xb = xch + (world > 1 ? (ld_volatile_int(hdr + HDR_SEQ) & 1) : 0) * xch_stride;
  double *const sw = xb + (int64_t)K * npad;
  double *const gw = xb + 2 * (int64_t)K * npad + (int64_t)K * K;
  __shared__ double s_col[T];
  __shared__ double s_sq;
  const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
  const int t = blockIdx.x, k = blockIdx.y;
  const int2 ij = tile_ij[t];
  const int I = ij.x, J = ij.y;
  if (tid < T) s_col[tid] = 0.0;
  if (tid == 0) s_sq = 0.0;
  __syncthreads();
  const F *wt = w + (int64_t)k * gstride + (int64_t)t * TT;
  double colacc[4] = {0, 0, 0, 0}, sq = 0.0;
  for (int r = wp; r < T; r += EDGE_THREADS / 32) {
    const int gi = I * T + r;
    double rs = 0.0;
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      const int col = cc * 32 + lane, gj = J * T + col;
      const double v = (gj > gi && gj < n) ? (double)wt[r * T + col] : 0.0;
      rs += v;
      colacc[cc] += v;
      sq += v * v;
    }
    rs = warp_sum(rs);
    if (lane == 0 && gi < n) atomicAdd(sw + (int64_t)k * npad + gi, rs);
  }
#pragma unroll
  for (int cc = 0; cc < 4; ++cc) atomicAdd(&s_col[cc * 32 + lane], colacc[cc]);
  sq = warp_sum(sq);
  if (lane == 0) atomicAdd(&s_sq, sq);
  __syncthreads();
  if (tid < T && J * T + tid < n) atomicAdd(sw + (int64_t)k * npad + J * T + tid, s_col[tid]);
  // Gram diagonal only: the denominators of the cheap stop bound (K > 1) / of the stop rule (K == 1)
  if (tid == 0) atomicAdd(gw + (int64_t)k * K + k, s_sq);
}

template <typename F> static int edge_pass_t(gl_solver *s, cudaStream_t st) {
  EdgeArgs<F> A;
  A.w = static_cast<F *>(s->w);
  A.z = static_cast<const F *>(s->z);
  A.r = static_cast<F *>(s->r);
  A.gstride = s->gstride;
  A.a = static_cast<const F *>(s->a);
  A.b = static_cast<const F *>(s->b);
  A.xch = s->xch;
  A.xch_stride = s->xch_stride;
  A.cta_partial = s->cta_partial;
  A.hdr = reinterpret_cast<int *>(s->comm);
  A.world = s->world;
  A.rank = s->rank;
  A.peers = s->peers;
  A.tile_ij = s->tile_ij;
  A.flags = s->flags;
  A.n = s->n;
  A.npad = s->npad;
  A.ntl = s->ntl;
  A.c = (F)s->c;
  A.tgz = (F)(2.0 * s->gamma);
  if (s->ntl == 0) return GL_OK;
  dim3 grid(s->ntl, s->K);
  if (s->K == 1)
    edge_pass_kernel<F, false><<<grid, EDGE_THREADS, 0, st>>>(A);
  else
    edge_pass_kernel<F, true><<<grid, EDGE_THREADS, 0, st>>>(A);
  GL_LAUNCH_CHECK();
  return GL_OK;
}

int launch_edge_pass(gl_solver *s, cudaStream_t st) {
  return s->dtype == GL_F32 ? edge_pass_t<float>(s, st) : edge_pass_t<double>(s, st);
}

int launch_edge_init(gl_solver *s, cudaStream_t st) {
  if (s->ntl == 0) return GL_OK;
  dim3 grid(s->ntl, s->K);
  const int *hdr = reinterpret_cast<const int *>(s->comm);
  if (s->dtype == GL_F32)
    edge_init_kernel<float><<<grid, EDGE_THREADS, 0, st>>>(static_cast<const float *>(s->w), s->gstride,
                                                          s->tile_ij, s->xch, s->xch_stride, hdr, s->world,
                                                          s->n, s->npad, s->K);
  else
    edge_init_kernel<double><<<grid, EDGE_THREADS, 0, st>>>(static_cast<const double *>(s->w), s->gstride,
                                                           s->tile_ij, s->xch, s->xch_stride, hdr, s->world,
                                                           s->n, s->npad, s->K);
  GL_LAUNCH_CHECK();
  return GL_OK;
}

// peer mode: publish "my partials of the current exchange are complete" as a stand-alone launch.
// The node tail does the same on entry; this entry point exists for single-process emulation of
// several ranks on one stream, where every rank must have published before any node tail starts.
__global__ void signal_kernel(int *hdr, PeerView peers, int world, int rank, const int *flags) {
  if (ld_volatile_int(flags + FLAG_DONE)) return;  // nothing was produced: never publish early
  __threadfence_system();
  const int seq = ld_volatile_int(hdr + HDR_SEQ);
  for (int r = 0; r < world; ++r) st_relaxed_sys(peers.hdr[r] + rank, seq + 1);
}

int launch_signal(gl_solver *s, cudaStream_t st) {
  if (s->world <= 1) return GL_OK;
  signal_kernel<<<1, 1, 0, st>>>(reinterpret_cast<int *>(s->comm), s->peers, s->world, s->rank, s->flags);
  GL_LAUNCH_CHECK();
  return GL_OK;
}

}  // namespace gl
