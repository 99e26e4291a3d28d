// persistent_batch.cu -- whole iterations of a BATCH of K > 1 graphs in ONE resident kernel.
//
// The batched call of gsp_learn_graph_log_degrees (graph_learn/smooth_learning.py:125-174 on a (K, E)
// matrix: the M-step of GLMM / k-Graphs, clustering/glmm.py:18-85, kgraphs.py:90-121) iterates all K
// graphs in lock step under ONE stop rule on spectral norms (:158-159).  Per iteration the classic
// drivers launch edge_pass_kernel<batch> + node_tail_kernel + a gated gram_edges_kernel; at config 4
// (K = 8, n = 1,000: 16 MB of state, L2-resident) that is latency: three kernel boundaries around ~4 us
// of sweep.  Here the CTAs stay resident (cooperative launch) and loop over
//   E phase : (graph, tile) items, static stride assignment; degree partials by float64 atomics, the
//             item's (||q - y||^2, ||w'||^2) into a per-item slot (folded in fixed order later)
//   barrier
//   N phase : chunks of npc nodes x all K graphs (node_tail_kernel's batch mapping): node update and the
//             chunk's squared norms per graph of the dual residual and of d'
//   barrier
//   decide  : EVERY CTA folds the per-graph squared norms in the same fixed order and evaluates the same
//             rule, so the decision needs no broadcast.  The cheap bounds of node_tail.cu come first; only
//             when they cannot rule the stop out does the grid form the full K x K node Grams (one more
//             barrier) for the exact dual ratio, and only when that is below tol or on the last iteration
//             the exact K x K edge Grams (one more barrier) for the exact primal ratio.
// The sweep reads the iterate from one buffer and writes w' to a second one (the array that holds q - y on
// the per-iteration path): 12 B/edge instead of 16, and (w, w') is what the exact primal ratio needs.
// Same decisions as the per-iteration path; GL_B200_BATCH=0 restores it.
#include <math_constants.h>

#include <cstdlib>

#include "edge_tile.cuh"
#include "grid_sync.cuh"
#include "spectral.cuh"

namespace gl {

constexpr int BATCH_MAX_K = 16;   // K * K <= 256: every Gram lives in shared memory
constexpr int EG = 4;             // register block of the exact edge Grams

template <typename F> struct BatchArgs {
  EdgeArgs<F> E;            // E.w / E.r: the two iterate buffers (the sweep reads one and writes the other); world == 1
  double *d, *yb, *pb, *rb;
  double *dprev;            // [K][npad] d of the previous iterate (operand of the exact dual ratio)
  F *a, *b;
  double *item_partial;     // [2 parities][2][items]: (||w' - w - ...||^2 i.e. ||q - y||^2, ||w'||^2) of every (graph, tile) item
  double *diag_part;        // [2][K][nchunks]: per-chunk squared norms of (dual residual, d') per graph
  double *gram_part;        // [nchunks][2][K*K]: per-chunk node Grams, formed only on the iterations that need them
  double *egram_part;       // [grid][2][K*K]: per-CTA exact edge Grams (deciding iterations only)
  double *gprev, *resid;
  unsigned int *bar;
  long long *trace;         // nullable: [TRACE_CAP][TRACE_STAMPS] globaltimer stamps of CTA 0 (gl_solver_trace_enable)
  double gamma, four_ag, tol;
  int maxit, iters, K, npc, nchunks;
};

// CTA 0 / thread 0 stamps the phase boundaries of iteration `it` (tracing only): 0 sweep start, 1 sweep end,
// 2 past barrier 1, 3 node chunks done, 4 past barrier 2, 5 folds done, 6 decision taken (7 unused)
#define GL_BSTAMP(k)                                                                             \
  do {                                                                                           \
    if (P.trace && blockIdx.x == 0 && tid == 0) {                                                \
      long long t_;                                                                              \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                     \
      P.trace[(it % TRACE_CAP) * TRACE_STAMPS + (k)] = t_;                                       \
    }                                                                                            \
  } while (0)

// This CTA's share of Gram(R) and Gram(W), W = A.w the previous iterate, R = w' - W (= q - y, :163) with
// w' = A.w_out, over the tiled edge arrays of all K graphs (masked slots hold zeros in both buffers),
// written to out[2][K*K].
template <typename F>
__device__ void edge_gram_partial(const EdgeArgs<F> &A, int K, double *out, double *s_acc) {
  const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
  const int nb = (K + EG - 1) / EG, KK = K * K;
  const int64_t len = A.gstride;
  for (int which = 0; which < 2; ++which)
    for (int kb = 0; kb < nb; ++kb)
      for (int lb = kb; lb < nb; ++lb) {
        double acc[EG][EG];
#pragma unroll
        for (int i = 0; i < EG; ++i)
#pragma unroll
          for (int j = 0; j < EG; ++j) acc[i][j] = 0.0;
        for (int64_t e = (int64_t)blockIdx.x * EDGE_THREADS + tid; e < len; e += (int64_t)gridDim.x * EDGE_THREADS) {
          double xa[EG], xl[EG];
#pragma unroll
          for (int i = 0; i < EG; ++i) {
            const int k = kb * EG + i, l = lb * EG + i;
            xa[i] = xl[i] = 0.0;
            if (k < K) {
              const double wo = (double)__ldcg(A.w + (int64_t)k * len + e);
              xa[i] = which ? wo : (double)__ldcg(A.w_out + (int64_t)k * len + e) - wo;
            }
            if (l < K) {
              const double wo = (double)__ldcg(A.w + (int64_t)l * len + e);
              xl[i] = which ? wo : (double)__ldcg(A.w_out + (int64_t)l * len + e) - wo;
            }
          }
#pragma unroll
          for (int i = 0; i < EG; ++i)
#pragma unroll
            for (int j = 0; j < EG; ++j) acc[i][j] += xa[i] * xl[j];
        }
#pragma unroll
        for (int i = 0; i < EG; ++i)
#pragma unroll
          for (int j = 0; j < EG; ++j) {
            const double v = warp_sum(acc[i][j]);
            if (lane == 0) s_acc[wp * EG * EG + i * EG + j] = v;
          }
        __syncthreads();
        if (tid < EG * EG) {
          const int i = tid / EG, j = tid % EG, k = kb * EG + i, l = lb * EG + j;
          if (k < K && l < K) {
            double v = 0.0;
            for (int q = 0; q < EDGE_THREADS / 32; ++q) v += s_acc[q * EG * EG + tid];
            out[which * KK + k * K + l] = v;
            out[which * KK + l * K + k] = v;
          }
        }
        __syncthreads();
      }
}

// K x K Grams over the node axis of two (K, npad) arrays, by the whole grid: chunk owners write their chunk's
// partial Grams, one grid barrier, every CTA folds them in the same fixed order into out0 / out1 (shared
// memory, K*K each).  Only on the iterations whose decision needs spectral norms, and once when a launch ends.
template <typename F>
__device__ void node_gram_stage(const BatchArgs<F> &P, const double *x0, const double *x1, double *s_val, double *out0,
                                double *out1, unsigned &gen) {
  const int tid = threadIdx.x, K = P.K, KK = K * K, npad = P.E.npad;
  for (int c = blockIdx.x; c < P.nchunks; c += gridDim.x) {
    const int k = tid / P.npc, v = c * P.npc + tid % P.npc;
    const int64_t idx = (int64_t)k * npad + v;
    s_val[tid] = k < K ? __ldcg(x0 + idx) : 0.0;
    s_val[EDGE_THREADS + tid] = k < K ? __ldcg(x1 + idx) : 0.0;
    __syncthreads();
    for (int p = tid; p < 2 * KK; p += EDGE_THREADS) {
      const int which = p / KK, e = p % KK, kk = e / K, l = e % K;
      const double *xk = s_val + which * EDGE_THREADS + kk * P.npc, *xl = s_val + which * EDGE_THREADS + l * P.npc;
      double acc = 0.0;
      for (int j = 0; j < P.npc; ++j) acc += xk[j] * xl[j];
      P.gram_part[((int64_t)c * 2 + which) * KK + e] = acc;
    }
    __syncthreads();
  }
  grid_barrier(P.bar, gen);
  for (int p = tid; p < 2 * KK; p += EDGE_THREADS) {
    const int which = p / KK, e = p % KK;
    double acc = 0.0;
#pragma unroll 8
    for (int c = 0; c < P.nchunks; ++c) acc += __ldcg(P.gram_part + ((int64_t)c * 2 + which) * KK + e);
    (which == 0 ? out0 : out1)[e] = acc;
  }
  __syncthreads();
}

template <typename F>
__global__ void __launch_bounds__(EDGE_THREADS, sizeof(F) == 4 ? 2 : 1) persistent_batch_kernel(const BatchArgs<F> P) {
  __shared__ EdgeSmem<F> sm;
  __shared__ double s_grb[BATCH_MAX_K * BATCH_MAX_K];   // node Gram of the dual residual (:159 numerator)
  __shared__ double s_gx[BATCH_MAX_K * BATCH_MAX_K];    // node Gram of d of the previous iterate (its denominator) / of d' at launch end
  // per-graph squared norms, o = 2 k + which: [0, 2K) edge side (q - y, w'), [2K, 4K) node side (dual residual, d')
  __shared__ double s_sum[4 * BATCH_MAX_K];
  __shared__ double s_prev_w[BATCH_MAX_K], s_prev_d[BATCH_MAX_K];   // ||w_k||^2, ||d_k||^2 of the previous iterate
  __shared__ double s_grp[16][2];
  __shared__ double s_lam[4];
  __shared__ double s_lo, s_lo_dual;
  __shared__ int s_skip;
  EdgeArgs<F> A = P.E;
  const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
  int *flags = A.flags;
  if (ld_volatile_int(flags + FLAG_DONE)) return;  // set only by earlier launches: uniform
  int it = ld_volatile_int(flags + FLAG_IT);
  // The sweep reads the iterate from one buffer and writes w' to the other (12 B/edge: no q - y store); the
  // pair (w, w') is exactly what the deciding iterations need.  FLAG_WBUF says which buffer is current.
  int cur = ld_volatile_int(flags + FLAG_WBUF);
  const int K = P.K, KK = K * K;
  const int64_t total = (int64_t)K * A.npad;
  const int items = A.ntl * K;
  double *const xb = A.xch;
  // scratch aliased onto the column-partial area of the sweep (idle outside the E phase; >= 1,024 doubles):
  // [0, 512) node values of a chunk / folded edge Grams, [512, 640) warp partials of the edge Grams
  double *const s_val = reinterpret_cast<double *>(&sm.col[0][0][0]);
  double *const s_acc = s_val + 2 * EDGE_THREADS;
  if (tid < K) {
    s_prev_w[tid] = __ldcg(P.gprev + tid * K + tid);
    s_prev_d[tid] = __ldcg(P.gprev + KK + tid * K + tid);
  }
  __syncthreads();
  unsigned gen = 0;

  for (int step = 0; step < P.iters; ++step) {
    // item slots alternate by iteration: a fast CTA's next sweep must not overwrite what a slow CTA still folds
    double *const part = P.item_partial + (int64_t)(it & 1) * 2 * items;
    // ---------------- E phase --------------------------------------------------------------------
    GL_BSTAMP(0);
    A.w = cur ? P.E.r : P.E.w;
    A.w_out = cur ? P.E.w : P.E.r;
    for (int item = blockIdx.x; item < items; item += gridDim.x) {
      const int k = item / A.ntl, t = item - k * A.ntl;
      double r2 = 0.0, w2 = 0.0;
      edge_cta_tile<F, false, true, true>(A, sm, t, k, xb, xb + total, r2, w2);
      if (tid == 0) {
        part[item] = r2;            // [which][item]: the fold reads consecutive slots
        part[items + item] = w2;
      }
      __syncthreads();
    }
    GL_BSTAMP(1);
    grid_barrier(P.bar, gen);
    GL_BSTAMP(2);
    // ---------------- N phase: npc nodes x K graphs per chunk (node_tail_kernel's batch mapping) -----------
    for (int c = blockIdx.x; c < P.nchunks; c += gridDim.x) {
      const int k = tid / P.npc, v = c * P.npc + tid % P.npc;
      double rbv = 0.0, dn = 0.0;
      if (k < K) {
        const int64_t idx = (int64_t)k * A.npad + v;
        const double sp = __ldcg(xb + idx), sw = __ldcg(xb + total + idx);
        xb[idx] = 0.0;
        xb[total + idx] = 0.0;
        double dv = 0.0, ybn = 0.0, pbn = 0.0;
        if (v < A.n) {
          dv = __ldcg(P.d + idx);
          const double ybv = __ldcg(P.yb + idx);
          const double qb = __ldcg(P.pb + idx) + P.gamma * sp;   // :154
          rbv = qb - ybv;
          dn = (dv - ybv) + qb;                                  // :164
          ybn = dn + P.gamma * sw;                               // :148
          pbn = (ybn - sqrt(ybn * ybn + P.four_ag)) / 2.0;       // :151
        }
        P.dprev[idx] = dv;
        P.d[idx] = dn;
        P.yb[idx] = ybn;
        P.pb[idx] = pbn;
        P.rb[idx] = rbv;
        P.a[idx] = (F)(P.gamma * dn);
        P.b[idx] = (F)(P.gamma * pbn);
      }
      // the chunk's squared norms per graph: a graph's npc threads are consecutive (groups of <= 32 lanes)
      double q0 = rbv * rbv, q1 = dn * dn;
      const int width = P.npc < 32 ? P.npc : 32;
      for (int o = width >> 1; o > 0; o >>= 1) {
        q0 += __shfl_xor_sync(0xffffffffu, q0, o);
        q1 += __shfl_xor_sync(0xffffffffu, q1, o);
      }
      if ((tid & (width - 1)) == 0) {
        s_grp[tid / width][0] = q0;
        s_grp[tid / width][1] = q1;
      }
      __syncthreads();
      if (tid < 2 * K) {
        const int kk = tid >> 1, which = tid & 1, gpk = P.npc / width;
        double acc = 0.0;
        for (int g = 0; g < gpk; ++g) acc += s_grp[kk * gpk + g][which];
        P.diag_part[((int64_t)which * K + kk) * P.nchunks + c] = acc;   // [which][graph][chunk]: chunk-contiguous for the fold
      }
      __syncthreads();
    }
    GL_BSTAMP(3);
    grid_barrier(P.bar, gen);
    GL_BSTAMP(4);
    // ---------------- decision: identical on every CTA -----------------------------------------------
    // (Folding once in the last chunk CTA to finish -- a ticket before barrier 2 -- and letting every CTA read
    // the 4 K results was measured: 19.8 against 18.4 us per iteration at config 4; the serial ticket + fold in
    // front of the barrier costs more than the all-CTA fold behind it.  Not kept.)
    {  // 4 K sums, one warp each (round robin); the first load of every sum is issued before any is used
      const int nout = 4 * K;
      double acc[8];
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const int o = wp + 8 * r;
        const bool node = o >= 2 * K;
        const int oo = node ? o - 2 * K : o, kk = oo >> 1, which = oo & 1;
        acc[r] = 0.0;
        if (o < nout && lane < (node ? P.nchunks : A.ntl))
          acc[r] = node ? __ldcg(P.diag_part + ((int64_t)which * K + kk) * P.nchunks + lane)
                        : __ldcg(part + (int64_t)which * items + (int64_t)kk * A.ntl + lane);
      }
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const int o = wp + 8 * r;
        const bool node = o >= 2 * K;
        const int oo = node ? o - 2 * K : o, kk = oo >> 1, which = oo & 1;
        if (o < nout)
          for (int t = lane + 32; t < (node ? P.nchunks : A.ntl); t += 32)
            acc[r] += node ? __ldcg(P.diag_part + ((int64_t)which * K + kk) * P.nchunks + t)
                           : __ldcg(part + (int64_t)which * items + (int64_t)kk * A.ntl + t);
        const double v = warp_sum(acc[r]);
        if (lane == 0 && o < nout) s_sum[o] = v;
      }
    }
    __syncthreads();
    GL_BSTAMP(5);
    // K > 1: sigma_max(R)^2 >= max_k ||r_k||^2 and sigma_max(W)^2 <= sum_k ||w_k||^2 bound rel_primal from
    // below (likewise rel_dual): the stop is ruled out exactly while a bound is >= tol (node_tail.cu)
    const bool check = it >= 1;   // the first iteration never stops: smooth_learning.py:156-161
    bool skip = false, need = false, conv = false;
    double rd = CUDART_INF, rp = CUDART_INF;
    if (check) {
      if (tid == 0) {
        double mx = 0.0, tr = 0.0, mxd = 0.0, trd = 0.0;
        int bad = 0;
        for (int k = 0; k < K; ++k) {
          const double vr = s_sum[2 * k], u = s_prev_w[k], vd = s_sum[2 * K + 2 * k], ud = s_prev_d[k];
          bad |= (vr != vr) || (u != u) || (vd != vd) || (ud != ud);
          mx = vr > mx ? vr : mx;
          mxd = vd > mxd ? vd : mxd;
          tr += u;
          trd += ud;
        }
        s_lo = sqrt(mx) / sqrt(tr);
        s_lo_dual = sqrt(mxd) / sqrt(trd);
        bad |= (s_lo != s_lo) || (s_lo_dual != s_lo_dual);
        // 1e-3 slack covers the rounding of the float32 partial sums behind the primal bound
        const bool out = !bad && (s_lo * (1.0 - 1e-3) >= P.tol || s_lo_dual * (1.0 - 1e-3) >= P.tol);
        s_skip = out && it + 1 < P.maxit;   // the last iteration reports exact residuals
      }
      __syncthreads();
      skip = s_skip != 0;
      if (!skip) {   // uniform over the grid: exact dual ratio from the full node Grams
        node_gram_stage<F>(P, P.rb, P.dprev, s_val, s_grb, s_gx, gen);
        const double *G[4] = {s_grb, s_gx, s_grb, s_gx};
        lambda_max4(G, K, nullptr, s_lam);
        rd = sqrt(s_lam[2]) / sqrt(s_lam[3]);
        need = (rd < P.tol) || it + 1 >= P.maxit;   // the dual ratio is exact: it alone can still rule the stop out
      }
    }
    if (need) {   // uniform over the grid: exact spectral norms of the (E, K) operands
      edge_gram_partial<F>(A, K, P.egram_part + (int64_t)blockIdx.x * 2 * KK, s_acc);
      grid_barrier(P.bar, gen);
      for (int p = tid; p < 2 * KK; p += EDGE_THREADS) {
        double acc = 0.0;
#pragma unroll 8
        for (int c = 0; c < (int)gridDim.x; ++c) acc += __ldcg(P.egram_part + (int64_t)c * 2 * KK + p);
        s_val[p] = acc;
      }
      __syncthreads();
      const double *G[4] = {s_val, s_val + KK, s_val, s_val + KK};
      lambda_max4(G, K, nullptr, s_lam);
      rp = sqrt(s_lam[0]) / sqrt(s_lam[1]);
      conv = (rp < P.tol) && (rd < P.tol);   // NaN / inf compare false, like NumPy
    }
    __syncthreads();
    GL_BSTAMP(6);
    // rotate the denominators
    if (tid < K) {
      s_prev_w[tid] = s_sum[2 * tid + 1];
      s_prev_d[tid] = s_sum[2 * K + 2 * tid + 1];
    }
    ++it;
    cur ^= 1;
    const bool done = conv || it >= P.maxit;
    if (blockIdx.x == 0) {
      if (tid < K) {   // what the next launch (or the per-iteration kernels) start from
        P.gprev[tid * K + tid] = s_sum[2 * tid + 1];
        P.gprev[KK + tid * K + tid] = s_sum[2 * K + 2 * tid + 1];
      }
      if (tid == 0) {
        flags[FLAG_IT] = it;
        flags[FLAG_WBUF] = cur;
        if (check) {
          if (skip) {   // ruled out by the bounds (reported as such)
            P.resid[0] = s_lo;
            P.resid[1] = s_lo_dual;
          } else {
            P.resid[0] = need ? rp : s_lo;   // a lower bound unless this was a deciding iteration
            P.resid[1] = rd;
          }
        }
        if (conv) flags[FLAG_CONVERGED] = 1;
      }
    }
    if (done || step == P.iters - 1) {
      // the launch ends here: leave the full node Gram of d' behind (the per-iteration kernels' node tail
      // and gl_solver_node_tail read it as their denominator)
      node_gram_stage<F>(P, P.rb, P.d, s_val, s_grb, s_gx, gen);
      if (blockIdx.x == 0) {
        for (int e = tid; e < KK; e += EDGE_THREADS) P.gprev[KK + e] = s_gx[e];
        if (done && tid == 0) {
          __threadfence();
          flags[FLAG_DONE] = 1;
        }
      }
      break;
    }
    __syncthreads();
  }
}

// two-buffer mode: leave the iterate in the canonical buffer (s->w) for gl_solver_store / the stepwise calls
template <typename F>
__global__ void batch_finalize_kernel(const int *flags, F *w, const F *alt, int64_t len) {
  if (!flags[FLAG_WBUF]) return;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < len; e += (int64_t)gridDim.x * blockDim.x) w[e] = alt[e];
}

template <typename F> static int batch_grid_query(gl_solver *s) {
  int per_sm = 0, sms = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, persistent_batch_kernel<F>, EDGE_THREADS, 0) != cudaSuccess ||
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device) != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  if (per_sm > 2) per_sm = 2;   // egram_part is sized for two CTAs per SM (gl_solver_create)
  return per_sm * sms > 0 ? per_sm * sms : -1;
}

// K in 2..16 on the full tile range of one GPU, with the buffers gl_solver_create reserves for it
bool batch_applicable(gl_solver *s) {
  if (s->K < 2 || s->K > BATCH_MAX_K || s->world > 1 || s->ntl == 0) return false;
  if (s->tile_begin != 0 || s->tile_end != num_tiles(s->n)) return false;
  if (!s->batch_partial || !s->batch_node || !s->egram_part || !s->gram_part || !s->r) return false;
  if (s->batch_grid_max == 0)
    s->batch_grid_max = s->dtype == GL_F32 ? batch_grid_query<float>(s) : batch_grid_query<double>(s);
  return s->batch_grid_max > 0 && s->batch_grid_max <= s->egram_ctas;
}

template <typename F> static int batch_launch_t(gl_solver *s, int iters, cudaStream_t st) {
  BatchArgs<F> P;
  EdgeArgs<F> &A = P.E;
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
  A.world = 1;
  A.rank = 0;
  A.peers = s->peers;
  A.tile_ij = s->tile_ij;
  A.flags = s->flags;
  A.n = s->n;
  A.npad = s->npad;
  A.ntl = s->ntl;
  A.c = (F)s->c;
  A.tgz = (F)(2.0 * s->gamma);
  P.d = s->d; P.yb = s->yb; P.pb = s->pb; P.rb = s->rb;
  P.a = static_cast<F *>(s->a);
  P.b = static_cast<F *>(s->b);
  P.dprev = s->batch_node;
  P.diag_part = s->batch_node + (size_t)s->K * s->npad;
  P.item_partial = s->batch_partial;
  P.gram_part = s->gram_part;
  P.egram_part = s->egram_part;
  P.gprev = s->gprev;
  P.resid = s->resid;
  P.bar = s->grid_bar;
  P.trace = s->trace;
  P.gamma = s->gamma;
  P.four_ag = 4.0 * s->alpha * s->gamma;
  P.tol = s->tol;
  P.maxit = s->maxit;
  P.iters = iters;
  P.K = s->K;
  int kp = 1;
  while (kp < s->K) kp <<= 1;
  P.npc = EDGE_THREADS / kp;          // the chunking gram_part was sized for (gl_solver_create)
  P.nchunks = s->npad / P.npc;
  const int64_t items = (int64_t)s->ntl * s->K;
  int grid = s->batch_grid_max;
  const int64_t want = items > P.nchunks ? items : (int64_t)P.nchunks;
  if ((int64_t)grid > want) grid = (int)want;
  GL_CUDA(cudaMemsetAsync(s->grid_bar, 0, sizeof(unsigned int), st));
  void *args[] = {&P};
  GL_CUDA(cudaLaunchCooperativeKernel((void *)persistent_batch_kernel<F>, dim3(grid), dim3(EDGE_THREADS), args, 0, st));
  batch_finalize_kernel<F><<<296, 256, 0, st>>>(s->flags, A.w, A.r, (int64_t)s->K * s->gstride);
  g_launches.fetch_add(2, std::memory_order_relaxed);
  GL_CUDA(cudaGetLastError());
  GL_CUDA(cudaMemsetAsync(s->flags + FLAG_WBUF, 0, sizeof(int), st));
  return GL_OK;
}

// `iters` iterations of the batch (or fewer if the shared stop rule fires) in one cooperative launch.
int launch_persistent_batch(gl_solver *s, int iters, cudaStream_t st) {
  if (!batch_applicable(s)) return fail(GL_E_STATE, "batched persistent kernel is not applicable to this solver");
  return s->dtype == GL_F32 ? batch_launch_t<float>(s, iters, st) : batch_launch_t<double>(s, iters, st);
}

}  // namespace gl
