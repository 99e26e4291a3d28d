// persistent.cu -- whole iterations in ONE resident kernel (single graph, K == 1).
//
// Same arithmetic as edge_pass.cu + node_tail.cu (graph_learn/smooth_learning.py:146-167), but the
// CTAs stay resident (cooperative launch, one or two per SM) and loop over
//   E phase : each CTA sweeps its tiles (static stride assignment), atomics into the exchange buffer
//   barrier
//   N phase : block 0 folds the CTAs' residual partials and (peer mode) publishes this rank's flag;
//             node chunks of 256 nodes are updated by the first CTAs, reading every rank's partials
//             straight from peer memory once all flags have arrived
//   barrier
//   decide  : every CTA evaluates the stop rule from the same partial sums (bit-identical), so no
//             third barrier and no broadcast is needed
// This removes the two kernel boundaries per iteration (launch gap + ramp-up/drain), which is what
// bounds the sharded case: at 8 GPUs the per-rank edge sweep is ~50 us.
#include <math_constants.h>

#include <cstdlib>

#include "edge_tile.cuh"
#include "grid_sync.cuh"

namespace gl {

template <typename F> struct PersistArgs {
  EdgeArgs<F> E;
  double *d, *yb, *pb, *rb;
  F *a, *b;
  double *node_partial;  // [nchunks][2] dual residual / ||d'||^2 partials, then [2] rank-summed edge sums
  double *gprev, *resid;
  unsigned int *bar;
  long long *trace;   // nullable: [TRACE_CAP][TRACE_STAMPS] globaltimer stamps of CTA 0
  double gamma, four_ag, tol;
  int maxit, iters, nchunks;
  int early_publish;   // peer mode: publish the degree partials before the fold (scalars behind HDR_SIG2)
};

__device__ __forceinline__ long long global_ns() {
  long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// CTA 0 / thread 0 stamps the phase boundaries of iteration `it` (tracing only)
#define GL_STAMP(k)                                                                              \
  do {                                                                                           \
    if (P.trace && blockIdx.x == 0 && tid == 0) P.trace[(it % TRACE_CAP) * TRACE_STAMPS + (k)] = global_ns(); \
  } while (0)

template <typename F>
__global__ void __launch_bounds__(EDGE_THREADS, sizeof(F) == 4 ? 2 : 1) persistent_kernel(const PersistArgs<F> P) {
  __shared__ EdgeSmem<F> sm;
  const EdgeArgs<F> &A = P.E;
  const int tid = threadIdx.x;
  int *flags = A.flags;
  if (ld_volatile_int(flags + FLAG_DONE)) return;  // set only by earlier launches: uniform
  int it = ld_volatile_int(flags + FLAG_IT);
  double gw_prev = __ldcg(P.gprev), gd_prev = __ldcg(P.gprev + 1);
  int seq = A.world > 1 ? ld_volatile_int(A.hdr + HDR_SEQ) : 0;
  unsigned gen = 0;
  const int64_t total = A.npad;  // K == 1

  for (int step = 0; step < P.iters; ++step) {
    const int64_t rd_off = (A.world > 1 ? (seq & 1) : 0) * A.xch_stride;
    const int64_t clr_off = (A.world > 1 ? ((seq + 1) & 1) : 0) * A.xch_stride;
    double *const xb = A.xch + rd_off;
    // ---------------- E phase --------------------------------------------------------------------
    GL_STAMP(0);
    double r2acc = 0.0, w2acc = 0.0;
    for (int t = blockIdx.x; t < A.ntl; t += gridDim.x) {
      double r2 = 0.0, w2 = 0.0;
      edge_cta_tile<F, false, true>(A, sm, t, 0, xb, xb + total, r2, w2);
      if (tid == 0) {
        r2acc += r2;
        w2acc += w2;
      }
      __syncthreads();
    }
    if (tid == 0) {
      A.cta_partial[2 * blockIdx.x] = r2acc;
      A.cta_partial[2 * blockIdx.x + 1] = w2acc;
    }
    GL_STAMP(1);
    grid_barrier(P.bar, gen);
    GL_STAMP(2);
    // ---------------- N phase --------------------------------------------------------------------
    // The degree partials are complete (barrier above): CTA 0 publishes them at once and goes straight to
    // its node chunk.  The two residual scalars are folded by a CTA that owns NO node chunk and travel
    // behind a SECOND flag (HDR_SIG2), which a peer only needs at the very end of its node phase -- the
    // fold is off every rank's critical path (measured on 2 GPUs: folding in CTA 0 after the publish was
    // slower than the old order, because CTA 0's own chunk then started a fence later).
    const bool early = A.world > 1 && P.early_publish && (int)gridDim.x > P.nchunks;
    const int fold_cta = early ? (int)gridDim.x - 1 : 0;
    if (blockIdx.x == 0 && early && tid < A.world) {  // one publishing thread per peer: fences and stores overlap
      __threadfence_system();
      st_relaxed_sys(A.peers.hdr[tid] + A.rank, seq + 1);
    }
    if ((int)blockIdx.x == fold_cta) {
      double r2 = 0.0, w2 = 0.0;
      for (int i = tid; i < (int)gridDim.x; i += EDGE_THREADS) {
        r2 += __ldcg(A.cta_partial + 2 * i);
        w2 += __ldcg(A.cta_partial + 2 * i + 1);
      }
      block_sum2(r2, w2, sm.red);
      if (tid == 0) {
        xb[2 * total] += r2;      // slot was cleared; += keeps the classic kernels' convention
        xb[2 * total + 1] += w2;
      }
      __syncthreads();
      if (A.world > 1 && tid < A.world) {
        __threadfence_system();
        st_relaxed_sys(A.peers.hdr[tid] + HDR_SIG2 + A.rank, seq + 1);
        if (!early) st_relaxed_sys(A.peers.hdr[tid] + A.rank, seq + 1);
      }
    }
    GL_STAMP(3);
    for (int c = blockIdx.x; c < P.nchunks; c += gridDim.x) {
      if (A.world > 1) {
        if (tid < A.world) {
          const long long t0 = clock64();
          while (ld_acquire_sys(A.hdr + tid) < seq + 1) {
            if (clock64() - t0 > 6000000000LL) {
              flags[FLAG_ERROR] = 1;
              break;
            }
          }
        }
        __syncthreads();
      }
      if (c == 0) GL_STAMP(4);
      const int64_t idx = (int64_t)c * EDGE_THREADS + tid;
      double my_rb2 = 0.0, my_d2 = 0.0;
      if (idx < total) {
        double sp = 0.0, sw = 0.0;
        if (A.world > 1) {
          // all peer loads in flight together (one NVLink round trip), summed in rank order
          double vp[GL_MAX_PEERS], vw[GL_MAX_PEERS];
#pragma unroll
          for (int r = 0; r < GL_MAX_PEERS; ++r) {
            vp[r] = vw[r] = 0.0;
            if (r < A.world && idx >= A.peers.first_node[r]) {   // below a rank's first row block its partials are zeros
              vp[r] = ld_relaxed_sys(A.peers.xch[r] + rd_off + idx);
              vw[r] = ld_relaxed_sys(A.peers.xch[r] + rd_off + total + idx);
            }
          }
#pragma unroll
          for (int r = 0; r < GL_MAX_PEERS; ++r) {
            sp += vp[r];
            sw += vw[r];
          }
        } else {
          sp = __ldcg(xb + idx);
          sw = __ldcg(xb + total + idx);
        }
        A.xch[clr_off + idx] = 0.0;
        A.xch[clr_off + total + idx] = 0.0;
        double dn = 0.0, rbv = 0.0, ybn = 0.0, pbn = 0.0;
        if (idx < A.n) {
          const double dv = __ldcg(P.d + idx), ybv = __ldcg(P.yb + idx);
          const double qb = __ldcg(P.pb + idx) + P.gamma * sp;   // :154
          rbv = qb - ybv;
          dn = (dv - ybv) + qb;                                  // :164
          ybn = dn + P.gamma * sw;                               // :148
          pbn = (ybn - sqrt(ybn * ybn + P.four_ag)) / 2.0;       // :151
        }
        my_rb2 = rbv * rbv;
        my_d2 = dn * dn;
        P.d[idx] = dn;
        P.yb[idx] = ybn;
        P.pb[idx] = pbn;
        P.rb[idx] = rbv;
        P.a[idx] = (F)(P.gamma * dn);
        P.b[idx] = (F)(P.gamma * pbn);
      }
      block_sum2(my_rb2, my_d2, sm.red);
      if (tid == 0) {
        P.node_partial[2 * c] = my_rb2;
        P.node_partial[2 * c + 1] = my_d2;
        if (c == 0) {  // edge-side sums over the ranks, then clear the slots of the next exchange
          double r2 = 0.0, w2 = 0.0;
          if (A.world > 1) {
            for (int r = 0; r < A.world; ++r) {   // the scalars travel behind their own flag (published after the fold)
              const long long t0 = clock64();
              while (ld_acquire_sys(A.hdr + HDR_SIG2 + r) < seq + 1) {
                if (clock64() - t0 > 6000000000LL) {
                  flags[FLAG_ERROR] = 1;
                  break;
                }
              }
            }
            for (int r = 0; r < A.world; ++r) {
              r2 += ld_relaxed_sys(A.peers.xch[r] + rd_off + 2 * total);
              w2 += ld_relaxed_sys(A.peers.xch[r] + rd_off + 2 * total + 1);
            }
          } else {
            r2 = xb[2 * total];
            w2 = xb[2 * total + 1];
          }
          P.node_partial[2 * P.nchunks] = r2;
          P.node_partial[2 * P.nchunks + 1] = w2;
          A.xch[clr_off + 2 * total] = 0.0;
          A.xch[clr_off + 2 * total + 1] = 0.0;
        }
      }
    }
    GL_STAMP(5);
    grid_barrier(P.bar, gen);
    GL_STAMP(6);
    // ---------------- decision: identical on every CTA -----------------------------------------------
    double rb2 = 0.0, d2 = 0.0;
    for (int i = tid; i < P.nchunks; i += EDGE_THREADS) {
      rb2 += __ldcg(P.node_partial + 2 * i);
      d2 += __ldcg(P.node_partial + 2 * i + 1);
    }
    block_sum2(rb2, d2, sm.red);
    const double g_r = __ldcg(P.node_partial + 2 * P.nchunks), g_w = __ldcg(P.node_partial + 2 * P.nchunks + 1);
    bool conv = false;
    double rp = CUDART_INF, rdl = CUDART_INF;
    if (it >= 1) {  // the first iteration never stops: smooth_learning.py:156-161
      rp = sqrt(g_r) / sqrt(gw_prev);
      rdl = sqrt(rb2) / sqrt(gd_prev);
      conv = (rp < P.tol) && (rdl < P.tol);
    }
    gw_prev = g_w;
    gd_prev = d2;
    GL_STAMP(7);
    const bool had_check = it >= 1;
    ++it;
    if (A.world > 1) ++seq;
    bool done = conv || it >= P.maxit;
    if (ld_volatile_int(flags + FLAG_ERROR)) done = true;  // written before the barrier above
    if (blockIdx.x == 0 && tid == 0) {
      flags[FLAG_IT] = it;
      if (had_check) {
        P.resid[0] = rp;
        P.resid[1] = rdl;
      }
      P.gprev[0] = gw_prev;
      P.gprev[1] = gd_prev;
      if (conv) flags[FLAG_CONVERGED] = 1;
      if (A.world > 1) A.hdr[HDR_SEQ] = seq;
      if (done) {
        __threadfence();
        flags[FLAG_DONE] = 1;
      }
    }
    if (done) break;
  }
}

// ---- small problems: every CTA owns at most ONE tile (n <= ~2,900) ---------------------------------------
// The whole problem is L2-resident there (config 2: 16 MB) and an iteration is latency: two grid barriers,
// the fold, the node chunk and the decision cost ~10 us against ~2 us of sweep.  With one tile per CTA the
// node phase is cheap enough to be done REDUNDANTLY by every CTA: after the single grid barrier (all degree
// partials complete) each CTA runs the node update of smooth_learning.py:148-164 over all n nodes itself,
// in the same fixed order, so every CTA holds bit-identical residual sums and takes the same stop decision
// -- no second barrier, no decision stage, no broadcast.  Only the owner of a 256-node chunk writes the
// node state; every CTA writes the (gamma d, gamma pb) values of its own tile's row and column block
// (identical values from every writer) and reads them back for the next sweep.
// Hazards are removed by buffering instead of barriers: the exchange buffer is TRIPLE buffered by
// iteration (sweep t + 1 adds into buffer (t + 1) % 3 while slower CTAs still read buffer t % 3, owners
// clear buffer (t + 2) % 3), the node state and the per-CTA residual partials are double buffered.
template <typename F> struct SmallArgs {
  EdgeArgs<F> E;                 // E.xch = exchange buffer 0, stride E.xch_stride, THREE buffers
  double *st[2][3];              // [parity][d, yb, pb]
  F *a, *b;
  double *cta_partial;           // [2][grid][2]
  double *gprev, *resid;
  unsigned int *bar;
  double gamma, four_ag, tol;
  int maxit, iters;
};

template <typename F>
__global__ void __launch_bounds__(EDGE_THREADS, sizeof(F) == 4 ? 2 : 1) persistent_small_kernel(const SmallArgs<F> P) {
  __shared__ EdgeSmem<F> sm;
  EdgeArgs<F> A = P.E;
  const int tid = threadIdx.x;
  int *flags = A.flags;
  if (ld_volatile_int(flags + FLAG_DONE)) return;
  int it = ld_volatile_int(flags + FLAG_IT);
  double gw_prev = __ldcg(P.gprev), gd_prev = __ldcg(P.gprev + 1);
  unsigned gen = 0;
  const int total = A.npad, grid = gridDim.x;
  const bool has_tile = (int)blockIdx.x < A.ntl;
  const int2 ij = has_tile ? A.tile_ij[blockIdx.x] : make_int2(-1, -1);
  A.a = P.a;
  A.b = P.b;
  if (it > 0) {   // continuing launch: the sweep's (a, b) follow from the node state of the last iteration
    const double *d = P.st[it & 1][0], *pb = P.st[it & 1][2];
    for (int j = tid; j < total; j += EDGE_THREADS)
      if (j / T == ij.x || j / T == ij.y) {
        P.a[j] = (F)(P.gamma * __ldcg(d + j));
        P.b[j] = (F)(P.gamma * __ldcg(pb + j));
      }
    __syncthreads();
  }
  for (int step = 0; step < P.iters; ++step) {
    const int par = it & 1;
    double *const xb = A.xch + (int64_t)(it % 3) * A.xch_stride;
    double *const xclr = A.xch + (int64_t)((it + 2) % 3) * A.xch_stride;
    double *const part = P.cta_partial + (int64_t)par * 2 * grid;
    // ---------------- E phase: this CTA's tile ------------------------------------------------------
    double r2 = 0.0, w2 = 0.0;
    if (has_tile) edge_cta_tile<F, false, true>(A, sm, blockIdx.x, 0, xb, xb + total, r2, w2);
    if (tid == 0) {
      part[2 * blockIdx.x] = r2;
      part[2 * blockIdx.x + 1] = w2;
    }
    grid_barrier(P.bar, gen);
    // ---------------- N phase, redundantly on every CTA ----------------------------------------------
    const double *d0 = P.st[par][0], *yb0 = P.st[par][1], *pb0 = P.st[par][2];
    double *d1 = P.st[par ^ 1][0], *yb1 = P.st[par ^ 1][1], *pb1 = P.st[par ^ 1][2];
    double my_rb2 = 0.0, my_d2 = 0.0;
    constexpr int NB = 8;   // nodes per thread in flight: all loads of a batch are issued before any store
    for (int j0 = tid; j0 < total; j0 += EDGE_THREADS * NB) {
      double sp[NB], sw[NB], dv[NB], ybv[NB], pbv[NB];
#pragma unroll
      for (int u = 0; u < NB; ++u) {
        const int j = j0 + u * EDGE_THREADS;
        const bool in = j < A.n;
        sp[u] = in ? __ldcg(xb + j) : 0.0;
        sw[u] = in ? __ldcg(xb + total + j) : 0.0;
        dv[u] = in ? __ldcg(d0 + j) : 0.0;
        ybv[u] = in ? __ldcg(yb0 + j) : 0.0;
        pbv[u] = in ? __ldcg(pb0 + j) : 0.0;
      }
#pragma unroll
      for (int u = 0; u < NB; ++u) {
        const int j = j0 + u * EDGE_THREADS;
        if (j >= total) break;
        double dn = 0.0, rbv = 0.0, ybn = 0.0, pbn = 0.0;
        if (j < A.n) {
          const double qb = pbv[u] + P.gamma * sp[u];            // :154
          rbv = qb - ybv[u];
          dn = (dv[u] - ybv[u]) + qb;                            // :164
          ybn = dn + P.gamma * sw[u];                            // :148
          pbn = (ybn - sqrt(ybn * ybn + P.four_ag)) / 2.0;       // :151
        }
        my_rb2 += rbv * rbv;
        my_d2 += dn * dn;
        if (j / T == ij.x || j / T == ij.y) {   // what this CTA's next sweep reads
          P.a[j] = (F)(P.gamma * dn);
          P.b[j] = (F)(P.gamma * pbn);
        }
        if (j / EDGE_THREADS == (int)blockIdx.x) {   // owner of the chunk: state of the next iteration, clear ahead
          d1[j] = dn;
          yb1[j] = ybn;
          pb1[j] = pbn;
          xclr[j] = 0.0;
          xclr[total + j] = 0.0;
        }
      }
    }
    block_sum2(my_rb2, my_d2, sm.red);
    double g_r = 0.0, g_w = 0.0;
    for (int i = tid; i < grid; i += EDGE_THREADS) {
      g_r += __ldcg(part + 2 * i);
      g_w += __ldcg(part + 2 * i + 1);
    }
    block_sum2(g_r, g_w, sm.red);
    bool conv = false;
    double rp = CUDART_INF, rdl = CUDART_INF;
    if (it >= 1) {  // the first iteration never stops: smooth_learning.py:156-161
      rp = sqrt(g_r) / sqrt(gw_prev);
      rdl = sqrt(my_rb2) / sqrt(gd_prev);
      conv = (rp < P.tol) && (rdl < P.tol);
    }
    gw_prev = g_w;
    gd_prev = my_d2;
    const bool had_check = it >= 1;
    ++it;
    const bool done = conv || it >= P.maxit;
    if (blockIdx.x == 0 && tid == 0) {
      flags[FLAG_IT] = it;
      if (had_check) {
        P.resid[0] = rp;
        P.resid[1] = rdl;
      }
      P.gprev[0] = gw_prev;
      P.gprev[1] = gd_prev;
      if (conv) flags[FLAG_CONVERGED] = 1;
      if (done) {
        __threadfence();
        flags[FLAG_DONE] = 1;
      }
    }
    if (done) break;
  }
}

// canonical node state <- the parity buffer the small kernel ended on (so that stepwise calls,
// gl_solver_node_state and a later gl_solver_run see one consistent state)
__global__ void small_finalize_kernel(const int *flags, double *d, double *yb, double *pb, const double *d1,
                                      const double *yb1, const double *pb1, int total) {
  if ((flags[FLAG_IT] & 1) == 0) return;
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < total; j += gridDim.x * blockDim.x) {
    d[j] = d1[j];
    yb[j] = yb1[j];
    pb[j] = pb1[j];
  }
}

template <typename F> static int small_launch_t(gl_solver *s, int iters, cudaStream_t st) {
  SmallArgs<F> P;
  EdgeArgs<F> &A = P.E;
  A.w = static_cast<F *>(s->w);
  A.z = static_cast<const F *>(s->z);
  A.r = nullptr;
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
  P.st[0][0] = s->d; P.st[0][1] = s->yb; P.st[0][2] = s->pb;
  P.st[1][0] = s->small_state; P.st[1][1] = s->small_state + s->npad; P.st[1][2] = s->small_state + 2 * (int64_t)s->npad;
  P.a = static_cast<F *>(s->a);
  P.b = static_cast<F *>(s->b);
  P.cta_partial = s->cta_partial;
  P.gprev = s->gprev;
  P.resid = s->resid;
  P.bar = s->grid_bar;
  P.gamma = s->gamma;
  P.four_ag = 4.0 * s->alpha * s->gamma;
  P.tol = s->tol;
  P.maxit = s->maxit;
  P.iters = iters;
  const int nchunks = (s->npad + EDGE_THREADS - 1) / EDGE_THREADS;
  const int grid = s->ntl > nchunks ? s->ntl : nchunks;
  GL_CUDA(cudaMemsetAsync(s->grid_bar, 0, sizeof(unsigned int), st));
  void *args[] = {&P};
  GL_CUDA(cudaLaunchCooperativeKernel((void *)persistent_small_kernel<F>, dim3(grid), dim3(EDGE_THREADS), args, 0, st));
  small_finalize_kernel<<<8, 256, 0, st>>>(s->flags, s->d, s->yb, s->pb, P.st[1][0], P.st[1][1], P.st[1][2], s->npad);
  g_launches.fetch_add(2, std::memory_order_relaxed);
  GL_LAUNCH_CHECK();
  return GL_OK;
}

// One tile per CTA and everything resident at once: is the redundant-node form applicable?
bool small_applicable(gl_solver *s) {
  if (s->K != 1 || s->world > 1 || s->ntl == 0 || s->trace) return false;
  if (s->tile_begin != 0 || s->tile_end != num_tiles(s->n)) return false;
  if (s->small_grid_max == 0) {
    int per_sm = 0, sms = 0;
    cudaError_t e = s->dtype == GL_F32
        ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, persistent_small_kernel<float>, EDGE_THREADS, 0)
        : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, persistent_small_kernel<double>, EDGE_THREADS, 0);
    if (e != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device) != cudaSuccess) {
      cudaGetLastError();
      s->small_grid_max = -1;
    } else {
      s->small_grid_max = per_sm * sms > 0 ? per_sm * sms : -1;
    }
  }
  return s->small_grid_max > 0 && s->ntl <= s->small_grid_max && s->small_state != nullptr;
}

int launch_persistent_small(gl_solver *s, int iters, cudaStream_t st) {
  return s->dtype == GL_F32 ? small_launch_t<float>(s, iters, st) : small_launch_t<double>(s, iters, st);
}

template <typename F> static int persistent_launch_t(gl_solver *s, int iters, cudaStream_t st) {
  PersistArgs<F> P;
  EdgeArgs<F> &A = P.E;
  A.w = static_cast<F *>(s->w);
  A.z = static_cast<const F *>(s->z);
  A.r = nullptr;
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
  P.d = s->d; P.yb = s->yb; P.pb = s->pb; P.rb = s->rb;
  P.a = static_cast<F *>(s->a);
  P.b = static_cast<F *>(s->b);
  P.node_partial = s->node_partial;
  P.gprev = s->gprev;
  P.resid = s->resid;
  P.bar = s->grid_bar;
  P.trace = s->trace;
  P.gamma = s->gamma;
  P.four_ag = 4.0 * s->alpha * s->gamma;
  P.tol = s->tol;
  P.maxit = s->maxit;
  P.iters = iters;
  P.nchunks = (s->npad + EDGE_THREADS - 1) / EDGE_THREADS;
  {
    const char *e = std::getenv("GL_B200_EARLY_PUBLISH");   // measurement / bisecting aid; default on
    P.early_publish = !(e && e[0] == '0');
  }
  if (s->persist_grid == 0) {
    int per_sm = 0, sms = 0;
    GL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, persistent_kernel<F>, EDGE_THREADS, 0));
    GL_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device));
    int grid = per_sm * sms;
    if (grid < 1) return fail(GL_E_CUDA, "persistent kernel does not fit on the device");
    const int want = s->ntl > P.nchunks ? s->ntl : P.nchunks;
    if (grid > want) grid = want;
    s->persist_grid = grid;
  }
  // (A grid shrunk to ceil(ntl / ceil(ntl / grid)) CTAs, so that every CTA sweeps the same number of tiles, was
  // measured in round 2: 69.4 vs 69.9 us per iteration on 8 GPUs, 111.6 vs 111.2 on 4 -- the barrier wait in the
  // phase trace comes from CTA-to-CTA variance, not from a partial last wave; not kept.)
  const int grid = s->persist_grid;
  GL_CUDA(cudaMemsetAsync(s->grid_bar, 0, sizeof(unsigned int), st));
  void *args[] = {&P};
  GL_CUDA(cudaLaunchCooperativeKernel((void *)persistent_kernel<F>, dim3(grid), dim3(EDGE_THREADS), args, 0, st));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return GL_OK;
}

// `iters` iterations (or fewer if the stop rule fires) in one cooperative launch.
int launch_persistent(gl_solver *s, int iters, cudaStream_t st) {
  if (s->K != 1 || s->ntl == 0) return fail(GL_E_STATE, "persistent kernel needs a single graph with local tiles");
  return s->dtype == GL_F32 ? persistent_launch_t<float>(s, iters, st) : persistent_launch_t<double>(s, iters, st);
}

}  // namespace gl
