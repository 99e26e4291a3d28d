// node_tail.cu -- K4: the n-vector tail of every iteration, and the batch Gram kernel.
//
// Replaces the node-sized work of the reference loop body, graph_learn/smooth_learning.py:
//   qb = pb + gamma S p                       :154
//   rel_dual numerator / denominator          :159     (spectral norm for a (n,K) batch)
//   d  <- d - yb + qb                         :164
//   stop rule                                 :156-161, :166
// and prepares the NEXT iteration's dual prox, whose input is known as soon as S w' is:
//   yb' = d' + gamma S w'                     :148
//   pb' = (yb' - sqrt(yb'^2 + 4 alpha gamma))/2   :151
// Node arithmetic is float64 in both dtype modes (n values; negligible traffic).  The last CTA
// to finish evaluates the residual norms and sets the device-side stop flag, so a captured graph
// of iterations needs no host round trip (SURVEY 7.2 "stopping rule must match exactly").
#include <math_constants.h>

#include "spectral.cuh"

namespace gl {

struct NodeArgs {
  double *d, *yb, *pb, *rb;
  void *a, *b;
  void *ab;        // nullable: interleaved (a, b) pairs for the kNN slot kernel (one gather per far end)
  double *xch;     // [sp | sw | G_r | G_w]
  double *gprev;   // [Gram(W) | Gram(D)] of the previous iterate
  double *resid;   // rel_primal, rel_dual
  double *eig;     // scratch 4 x 2 x K*K
  double *node_partial;  // [gridDim.x][2] (K == 1)
  double *gram_part;     // batch mapping: [gridDim.x][2][K*K] per-CTA partial node Grams
  int npc;               // batch mapping: nodes per CTA (every CTA holds all K graphs of its nodes); 0 = flat mapping
  int *hdr;              // comm header (peer mode)
  int64_t xch_stride;
  int world, rank;
  PeerView peers;
  int *flags;
  int n, npad, K, f32, init, maxit;
  double gamma, four_ag, tol;
};

// Gram matrices over the node axis of x (K, npad): out[k*K+l] = sum_v x[k][v] x[l][v], for two
// arrays at once.  Fixed summation order (deterministic).
__device__ void node_gram2(const double *x0, const double *x1, int K, int n, int npad, double *out0,
                           double *out1) {
  __shared__ double s_part[2][NODE_THREADS];
  const int KK = K * K;
  const int tid = threadIdx.x;
  if (KK <= NODE_THREADS) {
    const int L = NODE_THREADS / KK;  // lanes per pair
    const int p = tid / L, l = tid % L;
    double a0 = 0.0, a1 = 0.0;
    if (p < KK) {
      const int k = p / K, m = p % K;
      for (int v = l; v < n; v += L) {
        a0 += __ldcg(x0 + (int64_t)k * npad + v) * __ldcg(x0 + (int64_t)m * npad + v);
        a1 += __ldcg(x1 + (int64_t)k * npad + v) * __ldcg(x1 + (int64_t)m * npad + v);
      }
    }
    s_part[0][tid] = a0;
    s_part[1][tid] = a1;
    __syncthreads();
    // tree over the L lanes of each pair (L is not a power of two in general: sequential halves)
    if (p < KK && l == 0) {
      double t0 = 0.0, t1 = 0.0;
      for (int i = 0; i < L; ++i) {
        t0 += s_part[0][tid + i];
        t1 += s_part[1][tid + i];
      }
      out0[p] = t0;
      out1[p] = t1;
    }
  } else {
    for (int p = tid; p < KK; p += NODE_THREADS) {
      const int k = p / K, m = p % K;
      double a0 = 0.0, a1 = 0.0;
      for (int v = 0; v < n; ++v) {
        a0 += __ldcg(x0 + (int64_t)k * npad + v) * __ldcg(x0 + (int64_t)m * npad + v);
        a1 += __ldcg(x1 + (int64_t)k * npad + v) * __ldcg(x1 + (int64_t)m * npad + v);
      }
      out0[p] = a0;
      out1[p] = a1;
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(NODE_THREADS) node_tail_kernel(const NodeArgs A) {
  __shared__ int s_last;
  __shared__ double s_lam[4];
  __shared__ double s_blk[NODE_THREADS / 32][2];
  if (!A.init && *reinterpret_cast<volatile int *>(A.flags + FLAG_DONE)) return;
  const int tid = threadIdx.x;
  double my_rb2 = 0.0, my_d2 = 0.0;
  const int64_t total = (int64_t)A.K * A.npad;
  // flat mapping: thread = entry of the (K, npad) arrays.  Batch mapping (K > 1): a CTA owns npc
  // consecutive nodes of ALL K graphs (thread = graph tid / npc, node tid % npc), so the K x K Grams
  // of the dual residual and of d' are formed per CTA from shared memory instead of by one CTA
  // walking all nodes serially (measured 71 us at K = 8, n = 1,000).
  int64_t idx = (int64_t)blockIdx.x * NODE_THREADS + tid;
  bool active = idx < total;
  if (A.npc > 0) {
    const int k = tid / A.npc;
    active = k < A.K;
    idx = (int64_t)k * A.npad + (int64_t)blockIdx.x * A.npc + tid % A.npc;
  }
  // ---- peer mode: wait until every rank has published its partials of this exchange, then read
  //      them straight out of the peers' memory over NVLink (fixed rank order -> bit-identical
  //      replicated state on all ranks).  Buffers alternate by exchange parity; the one NOT being
  //      read is cleared here for the next edge pass (all peers are past reading it).
  int seq = 0;
  int64_t rd_off = 0, clr_off = 0;
  if (A.world > 1) {
    seq = ld_volatile_int(A.hdr + HDR_SEQ);
    rd_off = (seq & 1) * A.xch_stride;
    clr_off = ((seq + 1) & 1) * A.xch_stride;
    // publish first: this kernel runs after the rank's edge pass has completed (stream order), so
    // its partials of exchange `seq` are final.  No fence or flag work inside the edge pass.
    if (blockIdx.x == 0 && tid < A.world) {
      __threadfence_system();
      st_relaxed_sys(A.peers.hdr[tid] + A.rank, seq + 1);
    }
    if (tid < A.world) {
      const long long t0 = clock64();
      while (ld_acquire_sys(A.hdr + tid) < seq + 1) {
        if (clock64() - t0 > 6000000000LL) {  // ~3 s: a peer died; fail instead of hanging the GPU
          A.flags[FLAG_ERROR] = 1;
          break;
        }
      }
    }
    __syncthreads();
  }
  double rb_keep = 0.0, d_keep = 0.0;
  if (active) {
    const int v = (int)(idx % A.npad);
    double sp, sw;
    if (A.world > 1) {
      sp = sw = 0.0;
      double vp[GL_MAX_PEERS], vw[GL_MAX_PEERS];
#pragma unroll
      for (int r = 0; r < GL_MAX_PEERS; ++r) {
        vp[r] = vw[r] = 0.0;
        if (r < A.world) {
          vp[r] = ld_relaxed_sys(A.peers.xch[r] + rd_off + idx);
          vw[r] = ld_relaxed_sys(A.peers.xch[r] + rd_off + total + idx);
        }
      }
#pragma unroll
      for (int r = 0; r < GL_MAX_PEERS; ++r) {
        sp += vp[r];
        sw += vw[r];
      }
      A.xch[clr_off + idx] = 0.0;
      A.xch[clr_off + total + idx] = 0.0;
    } else {
      sp = A.xch[idx];
      sw = A.xch[total + idx];
      A.xch[idx] = 0.0;
      A.xch[total + idx] = 0.0;
    }
    double dn = 0.0, rbv = 0.0, ybn = 0.0, pbn = 0.0;
    if (v < A.n) {
      if (A.init) {
        dn = sw;                                   // d0 = S w0            :144
      } else {
        const double dv = A.d[idx], ybv = A.yb[idx];
        const double qb = A.pb[idx] + A.gamma * sp;  //                    :154
        rbv = qb - ybv;
        dn = (dv - ybv) + qb;                       //                      :164
      }
      ybn = dn + A.gamma * sw;                       //                      :148
      pbn = (ybn - sqrt(ybn * ybn + A.four_ag)) / 2.0;  //                   :151
    }
    my_rb2 = rbv * rbv;
    my_d2 = dn * dn;
    rb_keep = rbv;
    d_keep = dn;
    A.d[idx] = dn;
    A.yb[idx] = ybn;
    A.pb[idx] = pbn;
    A.rb[idx] = rbv;
    if (A.f32) {
      static_cast<float *>(A.a)[idx] = (float)(A.gamma * dn);
      static_cast<float *>(A.b)[idx] = (float)(A.gamma * pbn);
      if (A.ab) static_cast<float2 *>(A.ab)[idx] = make_float2((float)(A.gamma * dn), (float)(A.gamma * pbn));
    } else {
      static_cast<double *>(A.a)[idx] = A.gamma * dn;
      static_cast<double *>(A.b)[idx] = A.gamma * pbn;
      if (A.ab) static_cast<double2 *>(A.ab)[idx] = make_double2(A.gamma * dn, A.gamma * pbn);
    }
  }
  if (A.npc > 0) {  // batch mapping: K x K partial Grams of this CTA's nodes, fixed summation order
    __shared__ double s_val[2][NODE_THREADS];
    s_val[0][tid] = rb_keep;
    s_val[1][tid] = d_keep;
    __syncthreads();
    const int KKb = A.K * A.K, npc = A.npc;
    for (int p = tid; p < 2 * KKb; p += NODE_THREADS) {
      const int which = p / KKb, e = p % KKb, k = e / A.K, l = e % A.K;
      const double *xk = s_val[which] + k * npc, *xl = s_val[which] + l * npc;
      double acc = 0.0;
      for (int j = 0; j < npc; ++j) acc += xk[j] * xl[j];
      A.gram_part[((int64_t)blockIdx.x * 2 + which) * KKb + e] = acc;
    }
  }
  if (A.K == 1) {  // single graph: per-CTA partial sums, combined in fixed order by the last CTA
    my_rb2 = warp_sum(my_rb2);
    my_d2 = warp_sum(my_d2);
    if ((tid & 31) == 0) {
      s_blk[tid >> 5][0] = my_rb2;
      s_blk[tid >> 5][1] = my_d2;
    }
    __syncthreads();
    if (tid == 0) {
      double r2 = 0.0, d2 = 0.0;
      for (int q = 0; q < NODE_THREADS / 32; ++q) {
        r2 += s_blk[q][0];
        d2 += s_blk[q][1];
      }
      A.node_partial[2 * blockIdx.x] = r2;
      A.node_partial[2 * blockIdx.x + 1] = d2;
    }
  }
  // ---- last CTA: residual norms, stop rule --------------------------------------------------
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const int ticket = atomicAdd(A.flags + FLAG_NODE_TICKET, 1);
    s_last = (ticket == (int)gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();

  const int KK = A.K * A.K;
  double *g_r = A.xch + 2 * total, *g_w = g_r + KK;   // edge-side Grams (summed over shards)
  double *g_rb = A.eig + 8 * (int64_t)KK, *g_d = g_rb + KK;  // node-side Grams, computed here
  double *g_clear = g_r;
  if (A.world > 1) {  // K == 1: sum the two edge-side scalars over the ranks into local scratch
    g_clear = A.xch + clr_off + 2 * total;
    g_r = A.node_partial + 2 * (int64_t)gridDim.x;
    g_w = g_r + 1;
    if (tid == 0) {
      double r2 = 0.0, w2 = 0.0;
      for (int r = 0; r < A.world; ++r) {
        r2 += ld_relaxed_sys(A.peers.xch[r] + rd_off + 2 * total);
        w2 += ld_relaxed_sys(A.peers.xch[r] + rd_off + 2 * total + 1);
      }
      g_r[0] = r2;
      g_w[0] = w2;
    }
    __syncthreads();
  }
  if (A.K == 1) {
    double r2 = 0.0, d2 = 0.0;
    for (int i = tid; i < (int)gridDim.x; i += NODE_THREADS) {
      r2 += __ldcg(A.node_partial + 2 * i);
      d2 += __ldcg(A.node_partial + 2 * i + 1);
    }
    r2 = warp_sum(r2);
    d2 = warp_sum(d2);
    __syncthreads();
    if ((tid & 31) == 0) {
      s_blk[tid >> 5][0] = r2;
      s_blk[tid >> 5][1] = d2;
    }
    __syncthreads();
    if (tid == 0) {
      r2 = d2 = 0.0;
      for (int q = 0; q < NODE_THREADS / 32; ++q) {
        r2 += s_blk[q][0];
        d2 += s_blk[q][1];
      }
      g_rb[0] = r2;
      g_d[0] = d2;
    }
    __syncthreads();
  } else if (A.npc > 0) {
    for (int p = tid; p < 2 * KK; p += NODE_THREADS) {
      const int which = p / KK, e = p % KK;
      double acc = 0.0;
      for (int c = 0; c < (int)gridDim.x; ++c) acc += __ldcg(A.gram_part + ((int64_t)c * 2 + which) * KK + e);
      (which == 0 ? g_rb : g_d)[e] = acc;
    }
    __syncthreads();
  } else {
    node_gram2(A.rb, A.d, A.K, A.n, A.npad, g_rb, g_d);
  }

  const int it = A.flags[FLAG_IT];
  // K > 1: the primal residual ratio needs the spectral norms of the (E, K) operands, i.e. full K x K
  // Grams over the edge axis.  Only their diagonals are maintained every iteration (edge pass); with
  //   sigma_max(R)^2 >= max_k ||r_k||^2   and   sigma_max(W)^2 <= sum_k ||w_k||^2
  // rel_primal >= s_lo, so the stop rule is RULED OUT exactly whenever s_lo >= tol (or the dual
  // ratio, exact here, is >= tol).  Otherwise FLAG_NEED hands the decision to gram_edges_kernel, which
  // forms the exact Grams and whose last CTA applies the rule: the decisions are those of the full rule.
  // The dual ratio is bounded the same way from the node-side Grams (exact and cheap, but their 26
  // squarings are not): the spectral norms are evaluated only when BOTH bounds fall below tol.
  const bool lazy = A.K > 1;
  __shared__ double s_lo, s_lo_dual;
  __shared__ int s_skip;
  if (!A.init && it >= 1) {
    if (tid == 0) s_skip = 0;
    if (lazy && tid == 0) {
      double mx = 0.0, tr = 0.0, mxd = 0.0, trd = 0.0;
      int bad = 0;
      for (int k = 0; k < A.K; ++k) {
        const double v = __ldcg(g_r + k * A.K + k), u = A.gprev[k * A.K + k];
        const double vd = g_rb[k * A.K + k], ud = A.gprev[KK + k * A.K + k];
        bad |= (v != v) || (u != u) || (vd != vd) || (ud != ud);
        mx = v > mx ? v : mx;
        mxd = vd > mxd ? vd : mxd;
        tr += u;
        trd += ud;
      }
      s_lo = sqrt(mx) / sqrt(tr);
      s_lo_dual = sqrt(mxd) / sqrt(trd);
      bad |= (s_lo != s_lo) || (s_lo_dual != s_lo_dual);
      // 1e-3 slack covers the rounding of the float32 partial sums behind the primal bound
      const bool out = !bad && (s_lo * (1.0 - 1e-3) >= A.tol || s_lo_dual * (1.0 - 1e-3) >= A.tol);
      s_skip = out && it + 1 < A.maxit;   // the last iteration reports exact residuals
    }
    __syncthreads();
    if (!s_skip) {
      const double *G[4] = {lazy ? g_rb : g_r, lazy ? A.gprev + KK : A.gprev, g_rb, A.gprev + KK};
      lambda_max4(G, A.K, A.eig, s_lam);
    }
  }
  __syncthreads();
  // rotate the denominators and clear the exchange slots
  for (int e = tid; e < KK; e += NODE_THREADS) {
    A.gprev[e] = __ldcg(g_w + e);
    A.gprev[KK + e] = g_d[e];
  }
  __syncthreads();
  for (int e = tid; e < 2 * KK; e += NODE_THREADS) g_clear[e] = 0.0;
  if (tid == 0) {
    A.flags[FLAG_NODE_TICKET] = 0;
    if (A.world > 1) {
      if (ld_volatile_int(A.flags + FLAG_ERROR)) A.flags[FLAG_DONE] = 1;
      A.hdr[HDR_SEQ] = seq + 1;  // this exchange is consumed
    }
    if (A.init) {
      A.flags[FLAG_IT] = 0;
      A.flags[FLAG_CONVERGED] = 0;
      A.flags[FLAG_NEED] = 0;
      A.flags[FLAG_DONE] = (A.maxit <= 0);
      A.resid[0] = CUDART_INF;
      A.resid[1] = CUDART_INF;
    } else {
      bool conv = false, need = false;
      if (it >= 1 && s_skip) {   // K > 1: ruled out by the bounds (reported as such)
        A.resid[0] = s_lo;
        A.resid[1] = s_lo_dual;
      } else if (it >= 1) {  // the first iteration never stops: smooth_learning.py:156-161
        const double rd = sqrt(s_lam[2]) / sqrt(s_lam[3]);
        A.resid[1] = rd;
        if (!lazy) {
          const double rp = sqrt(s_lam[0]) / sqrt(s_lam[1]);
          A.resid[0] = rp;
          conv = (rp < A.tol) && (rd < A.tol);  // NaN / inf compare false, like NumPy
        } else {
          // the dual ratio is exact here: it alone can still rule the stop out
          need = (rd < A.tol) || it + 1 >= A.maxit;
          if (!need) A.resid[0] = s_lo;              // a lower bound; exact on the deciding iterations
        }
      }
      A.flags[FLAG_IT] = it + 1;
      if (need) {
        A.flags[FLAG_NEED] = 1;
      } else {
        if (conv) A.flags[FLAG_CONVERGED] = 1;
        if (conv || it + 1 >= A.maxit) A.flags[FLAG_DONE] = 1;
      }
    }
  }
}

// K > 1, what the last CTA of gram_edges_kernel needs to decide (FLAG_NEED iterations): exact rel_primal
// from the spectral norms of Gram(q - y) and Gram(w) of the previous iterate, then the reference's stop
// rule (:158-166).
struct DecideArgs {
  double *g_r, *g_w, *resid, *eig;
  int *flags;
  int K, maxit;
  double tol;
};

int launch_node_tail(gl_solver *s, bool init, cudaStream_t st) {
  NodeArgs A;
  A.d = s->d; A.yb = s->yb; A.pb = s->pb; A.rb = s->rb;
  A.a = s->a; A.b = s->b; A.ab = s->ab;
  A.xch = s->xch; A.gprev = s->gprev; A.resid = s->resid; A.eig = s->eig_scratch;
  A.node_partial = s->node_partial;
  A.hdr = reinterpret_cast<int *>(s->comm);
  A.xch_stride = s->xch_stride;
  A.world = s->world;
  A.rank = s->rank;
  A.peers = s->peers;
  A.flags = s->flags;
  A.n = s->n; A.npad = s->npad; A.K = s->K; A.f32 = (s->dtype == GL_F32);
  A.init = init ? 1 : 0;
  A.maxit = s->maxit;
  A.gamma = s->gamma;
  A.four_ag = 4.0 * s->alpha * s->gamma;
  A.tol = s->tol;
  const int64_t total = (int64_t)s->K * s->npad;
  int grid = (int)((total + NODE_THREADS - 1) / NODE_THREADS);
  A.npc = 0;
  A.gram_part = s->gram_part;
  if (s->K > 1 && s->gram_part) {   // batch mapping (K <= 128: at least 2 nodes per CTA)
    int kp = 1;
    while (kp < s->K) kp <<= 1;
    A.npc = NODE_THREADS / kp;
    grid = s->npad / A.npc;
  }
  node_tail_kernel<<<grid, NODE_THREADS, 0, st>>>(A);
  GL_LAUNCH_CHECK();
  return GL_OK;
}

// ---- batch mode: Gram matrices of R = Q - Y and of the previous iterate W = W' - R over the local edge
// shard, only on the iterations whose stop decision needs them (FLAG_NEED, see node_tail_kernel).
// 8 x 8 register-blocked; grid = (chunks, graph-block pairs kb <= lb, {R, W}).
constexpr int GRAM_THREADS = NODE_THREADS;   // lambda_max4 in the deciding CTA is written for this block size
constexpr int GB = 8;

template <typename F>
__global__ void __launch_bounds__(GRAM_THREADS) gram_edges_kernel(const F *r, const F *w, int64_t gstride,
                                                                   int64_t len, int K, const DecideArgs D) {
  int *flags = D.flags;
  double *g_r = D.g_r, *g_w = D.g_w;
  if (!*reinterpret_cast<const volatile int *>(flags + FLAG_NEED)) return;
  __shared__ double s_acc[GRAM_THREADS / 32][GB * GB];
  __shared__ int s_last;
  __shared__ double s_lam[4];
  const int nb = (K + GB - 1) / GB;
  // decode pair index blockIdx.y -> (kb <= lb)
  int kb = 0, rem = blockIdx.y;
  while (rem >= nb - kb) {
    rem -= nb - kb;
    ++kb;
  }
  const int lb = kb + rem;
  const bool prev = blockIdx.z == 1;   // W of the previous iterate = w' - (q - y)   (:163)
  double *g = prev ? g_w : g_r;
  double acc[GB][GB];
#pragma unroll
  for (int i = 0; i < GB; ++i)
#pragma unroll
    for (int j = 0; j < GB; ++j) acc[i][j] = 0.0;
  const int64_t per = (len + gridDim.x - 1) / gridDim.x;
  const int64_t e0 = (int64_t)blockIdx.x * per, e1 = e0 + per < len ? e0 + per : len;
  for (int64_t e = e0 + threadIdx.x; e < e1; e += GRAM_THREADS) {
    double xa[GB], xb[GB];
#pragma unroll
    for (int i = 0; i < GB; ++i) {
      const int k = kb * GB + i, l = lb * GB + i;
      xa[i] = xb[i] = 0.0;
      if (k < K) xa[i] = prev ? (double)w[(int64_t)k * gstride + e] - (double)r[(int64_t)k * gstride + e]
                              : (double)r[(int64_t)k * gstride + e];
      if (l < K) xb[i] = prev ? (double)w[(int64_t)l * gstride + e] - (double)r[(int64_t)l * gstride + e]
                              : (double)r[(int64_t)l * gstride + e];
    }
#pragma unroll
    for (int i = 0; i < GB; ++i)
#pragma unroll
      for (int j = 0; j < GB; ++j) acc[i][j] += xa[i] * xb[j];
  }
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < GB; ++i)
#pragma unroll
    for (int j = 0; j < GB; ++j) {
      const double v = warp_sum(acc[i][j]);
      if (lane == 0) s_acc[wp][i * GB + j] = v;
    }
  __syncthreads();
  if (threadIdx.x < GB * GB) {
    const int i = threadIdx.x / GB, j = threadIdx.x % GB;
    const int k = kb * GB + i, l = lb * GB + j;
    if (k < K && l < K) {
      double v = 0.0;
      for (int q = 0; q < GRAM_THREADS / 32; ++q) v += s_acc[q][threadIdx.x];
      atomicAdd(g + (int64_t)k * K + l, v);
      if (kb != lb) atomicAdd(g + (int64_t)l * K + k, v);
    }
  }
  // ---- last CTA of the grid: the exact stop rule on the finished Grams ------------------------------
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const int ctas = (int)(gridDim.x * gridDim.y * gridDim.z);
    s_last = atomicAdd(flags + FLAG_EDGE_TICKET, 1) == ctas - 1;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const int KK = K * K;
  const double *G[4] = {g_r, g_w, g_r, g_w};
  lambda_max4(G, K, D.eig, s_lam);
  __syncthreads();
  for (int e = threadIdx.x; e < 2 * KK; e += GRAM_THREADS) g_r[e] = 0.0;   // g_w follows g_r: slots of the next pass
  if (threadIdx.x == 0) {
    const double rp = sqrt(s_lam[0]) / sqrt(s_lam[1]), rd = D.resid[1];
    D.resid[0] = rp;
    const bool conv = (rp < D.tol) && (rd < D.tol);
    flags[FLAG_EDGE_TICKET] = 0;
    flags[FLAG_NEED] = 0;
    if (conv) flags[FLAG_CONVERGED] = 1;
    if (conv || flags[FLAG_IT] >= D.maxit) flags[FLAG_DONE] = 1;
  }
}

int launch_gram_edges(gl_solver *s, cudaStream_t st) {
  if (s->K == 1 || s->ntl == 0) return GL_OK;
  const int nb = (s->K + GB - 1) / GB;
  const int64_t len = s->gstride;
  // 16 edges per thread: the kernel is latency-bound (a thread's loop is serial), so it wants many
  // short CTAs rather than a few long ones (8192-edge chunks measured 88 us at K = 8, n = 1,000)
  int chunks = (int)((len + 2047) / 2048);
  if (chunks > 2368) chunks = 2368;
  if (chunks < 1) chunks = 1;
  dim3 grid(chunks, nb * (nb + 1) / 2, 2);
  const int64_t total = (int64_t)s->K * s->npad;
  DecideArgs D;
  D.g_r = s->xch + 2 * total;
  D.g_w = D.g_r + (int64_t)s->K * s->K;
  D.resid = s->resid;
  D.eig = s->eig_scratch;
  D.flags = s->flags;
  D.K = s->K;
  D.maxit = s->maxit;
  D.tol = s->tol;
  if (s->dtype == GL_F32)
    gram_edges_kernel<float><<<grid, GRAM_THREADS, 0, st>>>(static_cast<const float *>(s->r),
                                                           static_cast<const float *>(s->w), s->gstride,
                                                           len, s->K, D);
  else
    gram_edges_kernel<double><<<grid, GRAM_THREADS, 0, st>>>(static_cast<const double *>(s->r),
                                                            static_cast<const double *>(s->w), s->gstride,
                                                            len, s->K, D);
  GL_LAUNCH_CHECK();
  return GL_OK;
}

}  // namespace gl
