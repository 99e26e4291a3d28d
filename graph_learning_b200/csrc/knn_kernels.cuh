// knn_kernels.cuh -- kernels of the kNN-support LogModel path (K6), shared by the one-shot solve
// (knn.cu) and the sharded, stepwise solver (knn_sharded.cu).  See knn.cu for the layout notes.
#pragma once
#include <cub/device/device_radix_sort.cuh>

#include <math_constants.h>

#include "solver.cuh"

namespace gl {

constexpr int KNN_THREADS = 256;
constexpr int KNN_UNROLL = 4;
template <typename F> struct Vec2;
template <> struct Vec2<float> { using type = float2; };
template <> struct Vec2<double> { using type = double2; };

static __global__ void knn_degree_count_kernel(const int32_t *ei, const int32_t *ej, int64_t nnz, int *cnt) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += (int64_t)gridDim.x * blockDim.x) {
    atomicAdd(cnt + ei[e], 1);
    atomicAdd(cnt + ej[e], 1);
  }
}
static __global__ void knn_max_kernel(const int *cnt, int n, int *out) {
  __shared__ int s[256];
  int m = 0;
  for (int i = threadIdx.x; i < n; i += 256) m = max(m, cnt[i]);
  s[threadIdx.x] = m;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) s[threadIdx.x] = max(s[threadIdx.x], s[threadIdx.x + o]);
    __syncthreads();
  }
  if (threadIdx.x == 0) *out = s[0];
}

// undirected list -> unsorted directed slots (key = source node, value = undirected edge id)
// Slots whose source lies outside [row_begin, row_end) (another shard's rows) get the sentinel key n,
// sort to the end and are dropped; *count receives the number of local slots.
// lab (nullable): node relabelling old -> new (single full-range shard: rows sorted by degree).
static __global__ void knn_expand_kernel(const int32_t *ei, const int32_t *ej, int64_t nnz, int32_t *key, int32_t *val,
                                  int row_begin, int row_end, int n, unsigned long long *count,
                                  const int32_t *lab) {
  unsigned long long mine = 0;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += (int64_t)gridDim.x * blockDim.x) {
    const int i = lab ? lab[ei[e]] : ei[e], j = lab ? lab[ej[e]] : ej[e];
    const bool oi = i >= row_begin && i < row_end, oj = j >= row_begin && j < row_end;
    key[2 * e] = oi ? i : n;
    val[2 * e] = (int32_t)e;
    key[2 * e + 1] = oj ? j : n;
    val[2 * e + 1] = (int32_t)e;
    mine += (oi ? 1 : 0) + (oj ? 1 : 0);
  }
  mine = warp_sum(mine);
  if ((threadIdx.x & 31) == 0 && mine) atomicAdd(count, mine);
}
// Degree-sorted relabelling (single full-range shard only).  A slice is as wide as its longest row, so
// rows of similar degree should share a slice: the nodes are renumbered in order of decreasing degree.
// The caller only ever sees per-EDGE results, so the renumbering stays internal (node vectors, dst).
static __global__ void knn_iota_kernel(int32_t *v, int n) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) v[i] = i;
}
static __global__ void knn_invert_kernel(const int32_t *order, int n, int32_t *lab) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) lab[order[i]] = i;
}
// Windowed variant (SELL-C-sigma): `arr` lists the nodes in a locality-preserving arrangement (the
// caller's numbering, or ascending order key); position p belongs to window p / sigma, and an
// ascending stable sort of these keys orders every window by decreasing degree while the windows stay
// where they are, so rows that share a slice are also close in the arrangement (their far ends fall
// into a few KB of the (a, b) array instead of all of it).  sigma == 0: one window (global sort).
static __global__ void knn_window_key_kernel(const int32_t *arr, const int *deg, int n, int sigma,
                                             unsigned long long *key) {
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
    const unsigned long long win = sigma > 0 ? (unsigned long long)(p / sigma) : 0ULL;
    key[p] = (win << 32) | (unsigned long long)(0x7fffffffu - (unsigned)deg[arr[p]]);
  }
}

// ---- sliced-ELL layout of the directed slots (SELL-32) ----------------------------------------------
// Rows (= source nodes, local index r = node - row_begin) are grouped into slices of 32 consecutive
// rows; slice s stores width[s] = max degree of its rows "columns" of 32 entries each, column-major:
//   entry(t, r) of slice s  at  slice_off[s] + t * 32 + (r & 31).
// Lane r of the warp that owns the slice walks its row with unit-stride (coalesced) warp accesses, keeps
// the row's degree sums in registers and STORES them: no shuffles, no atomics, deterministic sums.
// Padding entries hold dst = -1.

static __global__ void knn_row_count_kernel(const int32_t *src, int64_t nslots, int row_begin, int *cnt) {
  for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < nslots; s += (int64_t)gridDim.x * blockDim.x)
    atomicAdd(cnt + (src[s] - row_begin), 1);
}
// width of every slice (max row count) as the number of entries it occupies (width * 32)
static __global__ void knn_slice_size_kernel(const int *cnt, int rows, int nslices, long long *size) {
  const int s = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (s >= nslices) return;
  const int r = s * 32 + lane;
  int c = r < rows ? cnt[r] : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c = max(c, __shfl_xor_sync(0xffffffffu, c, o));
  if (lane == 0) size[s] = 32LL * c;
}
// scatter the source-sorted slots into the SELL arrays; eid_canon = undirected edge id on the copy whose
// source is the edge's first end (the copy gl_knn_store / the one-shot solve reads back), else -1
template <typename F>
__global__ void knn_sell_fill_kernel(const int32_t *ei, const int32_t *ej, const F *z, const F *w, int64_t nslots,
                                     const int32_t *src, const int32_t *eid, int row_begin, const int *row_ptr,
                                     const long long *slice_off, int32_t *dst_s, F *z_s, F *w_s, int32_t *eid_canon,
                                     const int32_t *lab) {
  for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < nslots; s += (int64_t)gridDim.x * blockDim.x) {
    const int32_t e = eid[s], me = src[s];
    const int32_t i = lab ? lab[ei[e]] : ei[e], j = lab ? lab[ej[e]] : ej[e];
    const int r = me - row_begin;
    const int t = (int)(s - row_ptr[r]);
    const long long pos = slice_off[r >> 5] + 32LL * t + (r & 31);
    dst_s[pos] = me == i ? j : i;
    z_s[pos] = z[e];
    w_s[pos] = w[e];
    eid_canon[pos] = me == i ? e : -1;
  }
}
template <typename F>
__global__ void knn_sell_writeback_kernel(const int32_t *eid_canon, const F *w_s, long long total, double edge_tol,
                                          F *w) {
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (long long)gridDim.x * blockDim.x) {
    const int32_t e = eid_canon[p];
    if (e >= 0) {
      const F v = w_s[p];
      w[e] = (double)v < edge_tol ? F(0) : v;
    }
  }
}

// MODE 0: iteration.  MODE 1: initial degrees S w0 and ||w0||^2.  One warp per slice (grid-stride), lane
// = row; KNN_UNROLL columns of the slice are loaded (then gathered) together before the first use.
// gsum receives HALF the slot sums (every undirected edge appears twice).
template <typename F, int MODE>
__global__ void __launch_bounds__(KNN_THREADS) knn_sell_kernel(const int32_t *__restrict__ dst,
                                                               const F *__restrict__ z, F *__restrict__ w,
                                                               const long long *__restrict__ slice_off, int nslices,
                                                               int row_begin,
                                                               const typename Vec2<F>::type *__restrict__ ab,
                                                               double *sp, double *sw, double *gsum,
                                                               const int *flags, F c, F tgz) {
  if (MODE == 0 && *reinterpret_cast<const volatile int *>(flags + FLAG_DONE)) return;
  __shared__ double s_red[KNN_THREADS / 32][2];
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  double res = 0.0, wsq = 0.0;
  const int warps = gridDim.x * (KNN_THREADS / 32);
  for (int s = blockIdx.x * (KNN_THREADS / 32) + wp; s < nslices; s += warps) {
    const long long base = slice_off[s] + lane, end = slice_off[s + 1];
    const int me = row_begin + s * 32 + lane;
    typename Vec2<F>::type abi;
    abi.x = abi.y = F(0);
    if (MODE == 0) abi = ab[me];
    double acc_p = 0.0, acc_w = 0.0;
    // two batches of KNN_UNROLL columns in flight: the next batch's (dst, w, z) loads are issued before
    // this batch's gathers are consumed
    int j[KNN_UNROLL], jn[KNN_UNROLL];
    F wo[KNN_UNROLL], zz[KNN_UNROLL], won[KNN_UNROLL], zzn[KNN_UNROLL];
    auto fetch = [&](long long p0, int *jj, F *ww, F *zv) {
#pragma unroll
      for (int u = 0; u < KNN_UNROLL; ++u) {
        const long long p = p0 + 32LL * u;
        const bool in = p < end;
        jj[u] = in ? dst[p] : -1;
        ww[u] = in ? w[p] : F(0);
        zv[u] = (MODE == 0 && in) ? z[p] : F(0);
      }
    };
    fetch(base, j, wo, zz);
    for (long long p0 = base; p0 < end; p0 += 32LL * KNN_UNROLL) {
      if (MODE == 1) {
#pragma unroll
        for (int u = 0; u < KNN_UNROLL; ++u)
          if (j[u] >= 0) {
            acc_w += (double)wo[u];
            wsq += (double)wo[u] * (double)wo[u];
          }
        fetch(p0 + 32LL * KNN_UNROLL, j, wo, zz);
        continue;
      }
      typename Vec2<F>::type abj[KNN_UNROLL];
#pragma unroll
      for (int u = 0; u < KNN_UNROLL; ++u) abj[u] = ab[j[u] < 0 ? me : j[u]];   // far end: one gather
      fetch(p0 + 32LL * KNN_UNROLL, jn, won, zzn);
#pragma unroll
      for (int u = 0; u < KNN_UNROLL; ++u)
        if (j[u] >= 0) {
          const F y = c * wo[u] - (abi.x + abj[u].x);
          F pv = y - tgz * zz[u];
          pv = pv > F(0) ? pv : F(0);
          const F q = c * pv - (abi.y + abj[u].y);
          const F r = q - y;
          const F wn = (wo[u] - y) + q;
          w[p0 + 32LL * u] = wn;
          acc_p += (double)pv;
          acc_w += (double)wn;
          res += (double)r * (double)r;
          wsq += (double)wn * (double)wn;
        }
#pragma unroll
      for (int u = 0; u < KNN_UNROLL; ++u) {
        j[u] = jn[u];
        wo[u] = won[u];
        zz[u] = zzn[u];
      }
    }
    if (MODE == 0) sp[me] = acc_p;   // every row is owned by exactly one lane: plain stores
    sw[me] = acc_w;
  }
  res = warp_sum(res);
  wsq = warp_sum(wsq);
  if (lane == 0) {
    s_red[wp][0] = res;
    s_red[wp][1] = wsq;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double r2 = 0.0, w2 = 0.0;
    for (int q = 0; q < KNN_THREADS / 32; ++q) {
      r2 += s_red[q][0];
      w2 += s_red[q][1];
    }
    if (MODE == 0) atomicAdd(gsum, 0.5 * r2);
    atomicAdd(gsum + 1, 0.5 * w2);
  }
}

// Whole iteration in ONE launch (single shard).  Because a lane owns ALL slots of its row, the row's
// degree sums S p and S w' are complete inside that lane when its slice ends, so the node update of
// smooth_learning.py:148-164 for that node follows at once: no (S p, S w') round trip through memory, no
// node kernel, no second launch.  The other rows still gather the OLD (gamma d, gamma pb) pairs during
// this sweep, so the pairs are double-buffered (ab_in / ab_out swap every iteration).  Residual sums:
// per-CTA partials, combined in a fixed order by the last CTA (ticket), which applies the stop rule.
template <typename F> struct KnnFusedArgs {
  const int32_t *dst;
  const F *z;
  F *w;
  const long long *slice_off;
  int nslices, n, maxit;
  const typename Vec2<F>::type *ab_in;
  typename Vec2<F>::type *ab_out;
  double *d, *yb, *pb;
  double *part;          // [gridDim.x][4]
  double *gprev, *resid;
  int *flags;
  F c, tgz;
  double gamma, four_ag, tol;
  // ---- PEER mode (row-sharded over several GPUs, one process each; see knn_fused_kernel) --------------
  int world, rank, row_begin;
  typename Vec2<F>::type *ab_out_peer[GL_MAX_PEERS];   // every rank's OUT buffer of this iteration (peer-mapped)
  double *scal_peer[GL_MAX_PEERS];                     // every rank's scalar block [2][GL_MAX_PEERS][4]
  int *hdr_peer[GL_MAX_PEERS];                         // every rank's header: sig[r] = last sequence rank r published
  int *hdr;                                            // own header (sig[], HDR_SEQ)
  const int32_t *need;                                 // nullable [n]: bit r set = rank r's slots reference this node
  const double *scal;                                  // own scalar block (peers write their partials here)
  double *ring;                                        // [2][2] denominators of the previous two iterates
};

// (Evict-first __ldcs/__stcs slot accesses were measured on B200 in round 2 and lost: 0.154 against
// 0.085 ms per iteration at n = 1e6 under the Morton order, profiles/r2_knn_row_order_stream.jsonl.)
//
// PEER = true: the same launch on a row shard [row_begin, row_begin + 32 nslices) of a multi-GPU solve, with
// the exchange done by the kernel itself over NVLink peer memory instead of two NCCL collectives:
//   * the new (gamma d, gamma pb) pair of a row is stored into EVERY rank's out buffer (8-byte remote
//     stores, unit stride across a warp), so after the sweep every rank holds the full array;
//   * the last CTA writes this rank's four residual partials into every rank's scalar block and then
//     publishes a sequence flag to every rank (st after a system fence);
//   * the next launch starts by waiting for all ranks' flags, then EVERY rank evaluates the stop rule for
//     the iteration just completed from the partials summed in rank order (identical inputs, identical
//     decision, no collective).  The decision therefore lags one launch, but a launch that decides "stop"
//     returns before touching w, so the weights and the iteration count are exactly the reference's.
// Buffer reuse is safe without further flags: a rank writes into a peer's out buffer only after that peer
// has published the end of the previous iteration, which was the last reader of that buffer.
// U = columns of a slice in flight per batch (two batches are in flight), MINB = resident CTAs per SM the
// register budget is sized for: (4, 4) = 64 registers.  Other (U, MINB) pairs were measured in round 2 and
// none was faster (DESIGN K6); only this instantiation is built.
template <typename F, bool PEER = false, int U = KNN_UNROLL, int MINB = (sizeof(F) == 4 ? 4 : 2)>
__global__ void __launch_bounds__(KNN_THREADS, MINB) knn_fused_kernel(const KnnFusedArgs<F> A) {
  if (*reinterpret_cast<const volatile int *>(A.flags + FLAG_DONE)) return;
  __shared__ double s_red[KNN_THREADS / 32][4];
  __shared__ int s_last;
  int seq = 0, L = 0;
  if constexpr (PEER) {
    __shared__ int s_stop;
    seq = ld_volatile_int(A.hdr + HDR_SEQ);
    if ((int)threadIdx.x < A.world) {
      const long long t0 = clock64();
      while (ld_acquire_sys(A.hdr + threadIdx.x) < seq) {
        if (clock64() - t0 > 6000000000LL) {
          A.flags[FLAG_ERROR] = 1;
          break;
        }
      }
    }
    __syncthreads();
    L = ld_volatile_int(A.flags + FLAG_IT);   // iterations completed so far (written by the previous launch)
    if (threadIdx.x == 0) {
      int stop = ld_volatile_int(A.flags + FLAG_ERROR) ? 1 : 0;
      if (L >= 1) {   // stop rule for iteration L - 1 (smooth_learning.py:156-166)
        const double *sc = A.scal + ((L - 1) & 1) * GL_MAX_PEERS * 4;
        double t[4] = {0.0, 0.0, 0.0, 0.0};
        for (int r = 0; r < A.world; ++r)
          for (int q = 0; q < 4; ++q) t[q] += __ldcg(sc + 4 * r + q);
        const double g_r = 0.5 * t[0], g_w = 0.5 * t[1];   // every undirected edge appears in two slots
        bool conv = false;
        double rp = 0.0, rd = 0.0;
        if (L - 1 >= 1) {   // the first iteration never stops: :156-161
          rp = sqrt(g_r) / sqrt(__ldcg(A.ring + 2 * (L & 1)));
          rd = sqrt(t[2]) / sqrt(__ldcg(A.ring + 2 * (L & 1) + 1));
          conv = (rp < A.tol) && (rd < A.tol);   // NaN / inf compare false, like NumPy
        }
        if (conv || L >= A.maxit) stop = 1;
        if (blockIdx.x == 0) {
          A.ring[2 * ((L + 1) & 1)] = g_w;       // denominators of the next check (:157-159)
          A.ring[2 * ((L + 1) & 1) + 1] = t[3];
          if (L - 1 >= 1) {
            A.resid[0] = rp;
            A.resid[1] = rd;
          }
          if (conv) A.flags[FLAG_CONVERGED] = 1;
        }
      }
      if (stop && blockIdx.x == 0) {
        __threadfence();
        A.flags[FLAG_DONE] = 1;
      }
      s_stop = stop;
    }
    __syncthreads();
    if (s_stop) return;
  }
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  const int32_t *__restrict__ dst = A.dst;
  const F *__restrict__ z = A.z;
  F *__restrict__ w = A.w;
  const typename Vec2<F>::type *__restrict__ ab = A.ab_in;
  const F c = A.c, tgz = A.tgz;
  double res = 0.0, wsq = 0.0, rb2 = 0.0, d2 = 0.0;
  const int warps = gridDim.x * (KNN_THREADS / 32);
  for (int s = blockIdx.x * (KNN_THREADS / 32) + wp; s < A.nslices; s += warps) {
    const long long base = A.slice_off[s] + lane, end = A.slice_off[s + 1];
    const int me = (PEER ? A.row_begin : 0) + s * 32 + lane;
    const typename Vec2<F>::type abi = ab[me];
    const bool live = me < A.n;
    // node state of this row: issued now, consumed after the slice
    // (pb is not stored: it is a function of yb, recomputed below -- 16 MB less traffic per iteration at n = 1e6)
    const double dv = live ? A.d[me] : 0.0, ybv = live ? A.yb[me] : 0.0;
    // per-row partial sums in the edge dtype (a row holds a few dozen slots), folded into float64 when
    // the row ends -- the dense tile routine does the same over 64 edges (edge_tile.cuh)
    F acc_p = F(0), acc_w = F(0), row_r = F(0), row_w = F(0);
    int j[U], jn[U];
    F wo[U], zz[U], won[U], zzn[U];
    auto fetch = [&](long long p0, int *jj, F *ww, F *zv) {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long p = p0 + 32LL * u;
        const bool in = p < end;
        jj[u] = in ? dst[p] : -1;
        ww[u] = in ? w[p] : F(0);
        zv[u] = in ? z[p] : F(0);
      }
    };
    fetch(base, j, wo, zz);
    for (long long p0 = base; p0 < end; p0 += 32LL * U) {
      typename Vec2<F>::type abj[U];
#pragma unroll
      for (int u = 0; u < U; ++u) abj[u] = ab[j[u] < 0 ? me : j[u]];   // far end: one gather
      fetch(p0 + 32LL * U, jn, won, zzn);
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (j[u] >= 0) {
          const F y = c * wo[u] - (abi.x + abj[u].x);
          F pv = y - tgz * zz[u];
          pv = pv > F(0) ? pv : F(0);
          const F q = c * pv - (abi.y + abj[u].y);
          const F r = q - y;
          const F wn = (wo[u] - y) + q;
          w[p0 + 32LL * u] = wn;
          acc_p += pv;
          acc_w += wn;
          row_r += r * r;
          row_w += wn * wn;
        }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        j[u] = jn[u];
        wo[u] = won[u];
        zz[u] = zzn[u];
      }
    }
    res += (double)row_r;
    wsq += (double)row_w;
    // ---- node update of row `me` (the arithmetic of knn_node_kernel) -----------------------------------
    typename Vec2<F>::type pr;
    pr.x = pr.y = F(0);
    if (live) {
      const double pbv = (ybv - sqrt(ybv * ybv + A.four_ag)) / 2.0;   // :151 of the previous iteration, bit-identical
      const double qb = pbv + A.gamma * (double)acc_p;            //                      :154
      const double rbv = qb - ybv;
      const double dn = (dv - ybv) + qb;                          //                      :164
      const double ybn = dn + A.gamma * (double)acc_w;            //                      :148
      const double pbn = (ybn - sqrt(ybn * ybn + A.four_ag)) / 2.0;   //                  :151
      rb2 += rbv * rbv;
      d2 += dn * dn;
      A.d[me] = dn;
      A.yb[me] = ybn;
      pr.x = (F)(A.gamma * dn);
      pr.y = (F)(A.gamma * pbn);
    }
    A.ab_out[me] = pr;
    if constexpr (PEER) {
      // only the ranks whose slots have this node as a far end need its pair (with spatially compact
      // shards that is a thin boundary layer: ~10 % of the rows at n = 1e6 on 8 ranks)
      const int nd = A.need ? (live ? A.need[me] : 0) : -1;
#pragma unroll
      for (int r = 0; r < GL_MAX_PEERS; ++r)
        if (r < A.world && r != A.rank && ((nd >> r) & 1)) A.ab_out_peer[r][me] = pr;
    }
  }
  res = warp_sum(res);
  wsq = warp_sum(wsq);
  rb2 = warp_sum(rb2);
  d2 = warp_sum(d2);
  if (lane == 0) {
    s_red[wp][0] = res;
    s_red[wp][1] = wsq;
    s_red[wp][2] = rb2;
    s_red[wp][3] = d2;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double t[4] = {0.0, 0.0, 0.0, 0.0};
    for (int q = 0; q < KNN_THREADS / 32; ++q)
      for (int i = 0; i < 4; ++i) t[i] += s_red[q][i];
    for (int i = 0; i < 4; ++i) A.part[4 * blockIdx.x + i] = t[i];
    if constexpr (PEER) __threadfence_system();   // this CTA's remote pair stores before its ticket
    else __threadfence();
    const int ticket = atomicAdd(A.flags + FLAG_EDGE_TICKET, 1);
    s_last = (ticket == (int)gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // ---- last CTA: fixed-order sums over the CTAs, stop rule (smooth_learning.py:156-166) --------------
  double t[4] = {0.0, 0.0, 0.0, 0.0};
  for (int i = threadIdx.x; i < (int)gridDim.x; i += KNN_THREADS)
    for (int q = 0; q < 4; ++q) t[q] += __ldcg(A.part + 4 * i + q);
  for (int q = 0; q < 4; ++q) t[q] = warp_sum(t[q]);
  __syncthreads();
  if (lane == 0)
    for (int q = 0; q < 4; ++q) s_red[wp][q] = t[q];
  __syncthreads();
  if (threadIdx.x != 0) return;
  for (int q = 0; q < 4; ++q) t[q] = 0.0;
  for (int v = 0; v < KNN_THREADS / 32; ++v)
    for (int q = 0; q < 4; ++q) t[q] += s_red[v][q];
  if constexpr (PEER) {
    // this rank's partials into every rank's scalar block, then the sequence flag (release at system scope)
    for (int r = 0; r < A.world; ++r)
      for (int q = 0; q < 4; ++q) A.scal_peer[r][((L & 1) * GL_MAX_PEERS + A.rank) * 4 + q] = t[q];
    A.flags[FLAG_EDGE_TICKET] = 0;
    A.flags[FLAG_IT] = L + 1;
    A.hdr[HDR_SEQ] = seq + 1;
    __threadfence_system();
    for (int r = 0; r < A.world; ++r) st_relaxed_sys(A.hdr_peer[r] + A.rank, seq + 1);
    return;
  }
  const double g_r = 0.5 * t[0], g_w = 0.5 * t[1];   // every undirected edge appears in two slots
  A.flags[FLAG_EDGE_TICKET] = 0;
  const int it = A.flags[FLAG_IT];
  bool conv = false;
  if (it >= 1) {  // the first iteration never stops: smooth_learning.py:156-161
    const double rp = sqrt(g_r) / sqrt(A.gprev[0]);
    const double rd = sqrt(t[2]) / sqrt(A.gprev[1]);
    A.resid[0] = rp;
    A.resid[1] = rd;
    conv = (rp < A.tol) && (rd < A.tol);  // NaN / inf compare false, like NumPy
  }
  A.flags[FLAG_IT] = it + 1;
  if (conv) A.flags[FLAG_CONVERGED] = 1;
  if (conv || it + 1 >= A.maxit) A.flags[FLAG_DONE] = 1;
  A.gprev[0] = g_w;   // denominators of the next check: the iterate before the update (:157-159)
  A.gprev[1] = t[3];
}

// Node phase of the kNN path (single graph, single GPU): the arithmetic of node_tail.cu for K == 1
// (smooth_learning.py:148-166) as a grid-stride kernel -- a few hundred CTAs however large n is,
// several independent nodes in flight per thread -- because one CTA per 256 nodes (3,907 CTAs at
// n = 1e6, each paying a fence, a barrier and a ticket) measured 123 us per launch, latency-bound.
// Only the interleaved (a, b) pairs the slot kernel gathers are written, not a and b separately.
constexpr int KNN_NODE_THREADS = 256;
constexpr int KNN_NODE_UNROLL = 4;
struct KnnNodeArgs {
  double *d, *yb, *pb, *rb, *xch, *gprev, *resid, *node_partial;
  void *ab;
  int *flags;
  int n, npad, init, maxit;
  int row_begin, row_end;   // nodes this rank updates ([0, npad) on a single GPU)
  int fused;                // 1: the last CTA also applies the stop rule (single GPU)
  double *scal;             // !fused: [||q-y||^2, ||w'||^2, ||qb-yb||^2, ||d'||^2] partials of this rank
  double gamma, four_ag, tol;
};

template <typename F>
__global__ void __launch_bounds__(KNN_NODE_THREADS) knn_node_kernel(const KnnNodeArgs A) {
  __shared__ int s_last;
  __shared__ double s_blk[KNN_NODE_THREADS / 32][2];
  if (!A.init && *reinterpret_cast<volatile int *>(A.flags + FLAG_DONE)) return;
  const int tid = threadIdx.x;
  double my_rb2 = 0.0, my_d2 = 0.0;
  typename Vec2<F>::type *ab = static_cast<typename Vec2<F>::type *>(A.ab);
  // KNN_NODE_UNROLL independent nodes per thread: all loads first (the stores below could alias them
  // as far as the compiler knows, so it would not hoist the next node's loads by itself)
  const int64_t gstride = (int64_t)gridDim.x * KNN_NODE_THREADS;
  for (int64_t idx0 = A.row_begin + (int64_t)blockIdx.x * KNN_NODE_THREADS + tid; idx0 < A.row_end;
       idx0 += gstride * KNN_NODE_UNROLL) {
    double sp[KNN_NODE_UNROLL], sw[KNN_NODE_UNROLL], dv[KNN_NODE_UNROLL], ybv[KNN_NODE_UNROLL], pbv[KNN_NODE_UNROLL];
#pragma unroll
    for (int u = 0; u < KNN_NODE_UNROLL; ++u) {
      const int64_t idx = idx0 + u * gstride;
      const bool in = idx < A.row_end;
      sp[u] = in ? A.xch[idx] : 0.0;
      sw[u] = in ? A.xch[A.npad + idx] : 0.0;
      const bool live = in && idx < A.n && !A.init;
      dv[u] = live ? A.d[idx] : 0.0;
      ybv[u] = live ? A.yb[idx] : 0.0;
      pbv[u] = live ? A.pb[idx] : 0.0;
    }
#pragma unroll
    for (int u = 0; u < KNN_NODE_UNROLL; ++u) {
      const int64_t idx = idx0 + u * gstride;
      if (idx >= A.row_end) continue;
      double dn = 0.0, rbv = 0.0, ybn = 0.0, pbn = 0.0;
      if (idx < A.n) {
        if (A.init) {
          dn = sw[u];                                       // d0 = S w0            :144
        } else {
          const double qb = pbv[u] + A.gamma * sp[u];       //                      :154
          rbv = qb - ybv[u];
          dn = (dv[u] - ybv[u]) + qb;                       //                      :164
        }
        ybn = dn + A.gamma * sw[u];                         //                      :148
        pbn = (ybn - sqrt(ybn * ybn + A.four_ag)) / 2.0;    //                      :151
      }
      my_rb2 += rbv * rbv;
      my_d2 += dn * dn;
      A.d[idx] = dn;
      A.yb[idx] = ybn;
      A.pb[idx] = pbn;
      A.rb[idx] = rbv;
      typename Vec2<F>::type pr;
      pr.x = (F)(A.gamma * dn);
      pr.y = (F)(A.gamma * pbn);
      ab[idx] = pr;
    }
  }
  my_rb2 = warp_sum(my_rb2);
  my_d2 = warp_sum(my_d2);
  if ((tid & 31) == 0) {
    s_blk[tid >> 5][0] = my_rb2;
    s_blk[tid >> 5][1] = my_d2;
  }
  __syncthreads();
  if (tid == 0) {
    double r2 = 0.0, d2 = 0.0;
    for (int q = 0; q < KNN_NODE_THREADS / 32; ++q) {
      r2 += s_blk[q][0];
      d2 += s_blk[q][1];
    }
    A.node_partial[2 * blockIdx.x] = r2;
    A.node_partial[2 * blockIdx.x + 1] = d2;
    __threadfence();
    const int ticket = atomicAdd(A.flags + FLAG_NODE_TICKET, 1);
    s_last = (ticket == (int)gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // ---- last CTA: fixed-order sums, stop rule (smooth_learning.py:156-166) -------------------------
  double r2 = 0.0, d2 = 0.0;
  for (int i = tid; i < (int)gridDim.x; i += KNN_NODE_THREADS) {
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
  if (tid != 0) return;
  r2 = d2 = 0.0;
  for (int q = 0; q < KNN_NODE_THREADS / 32; ++q) {
    r2 += s_blk[q][0];
    d2 += s_blk[q][1];
  }
  double *g = A.xch + 2 * (int64_t)A.npad;   // [ ||q - y||^2 , ||w'||^2 ] from the slot kernel
  const double g_r = __ldcg(g), g_w = __ldcg(g + 1);
  g[0] = 0.0;
  g[1] = 0.0;
  A.flags[FLAG_NODE_TICKET] = 0;
  if (!A.fused) {  // sharded: hand the partial sums to the exchange; knn_decide_kernel applies the rule
    A.scal[0] = g_r;
    A.scal[1] = g_w;
    A.scal[2] = r2;
    A.scal[3] = d2;
    return;
  }
  if (A.init) {
    A.flags[FLAG_IT] = 0;
    A.flags[FLAG_CONVERGED] = 0;
    A.flags[FLAG_DONE] = (A.maxit <= 0);
    A.resid[0] = CUDART_INF;
    A.resid[1] = CUDART_INF;
  } else {
    const int it = A.flags[FLAG_IT];
    bool conv = false;
    if (it >= 1) {  // the first iteration never stops: smooth_learning.py:156-161
      const double rp = sqrt(g_r) / sqrt(A.gprev[0]);
      const double rd = sqrt(r2) / sqrt(A.gprev[1]);
      A.resid[0] = rp;
      A.resid[1] = rd;
      conv = (rp < A.tol) && (rd < A.tol);  // NaN / inf compare false, like NumPy
    }
    A.flags[FLAG_IT] = it + 1;
    if (conv) A.flags[FLAG_CONVERGED] = 1;
    if (conv || it + 1 >= A.maxit) A.flags[FLAG_DONE] = 1;
  }
  A.gprev[0] = g_w;   // denominators of the next check: the iterate before the update (:157-159)
  A.gprev[1] = d2;
}

// Stop rule from the rank-summed scalars (identical inputs, identical decision on every rank).
static __global__ void knn_decide_kernel(const double *scal, double *gprev, double *resid, int *flags, int init, int maxit,
                                  double tol) {
  if (!init && flags[FLAG_DONE]) return;
  const double g_r = scal[0], g_w = scal[1], r2 = scal[2], d2 = scal[3];
  if (init) {
    flags[FLAG_IT] = 0;
    flags[FLAG_CONVERGED] = 0;
    flags[FLAG_DONE] = (maxit <= 0);
    resid[0] = CUDART_INF;
    resid[1] = CUDART_INF;
  } else {
    const int it = flags[FLAG_IT];
    bool conv = false;
    if (it >= 1) {
      const double rp = sqrt(g_r) / sqrt(gprev[0]);
      const double rd = sqrt(r2) / sqrt(gprev[1]);
      resid[0] = rp;
      resid[1] = rd;
      conv = (rp < tol) && (rd < tol);
    }
    flags[FLAG_IT] = it + 1;
    if (conv) flags[FLAG_CONVERGED] = 1;
    if (conv || it + 1 >= maxit) flags[FLAG_DONE] = 1;
  }
  gprev[0] = g_w;
  gprev[1] = d2;
}

template <typename F> static int launch_knn_node(gl_solver *s, bool init, cudaStream_t st) {
  KnnNodeArgs A;
  A.d = s->d; A.yb = s->yb; A.pb = s->pb; A.rb = s->rb;
  A.xch = s->xch; A.gprev = s->gprev; A.resid = s->resid; A.node_partial = s->node_partial;
  A.ab = s->ab;
  A.flags = s->flags;
  A.n = s->n; A.npad = s->npad; A.init = init ? 1 : 0; A.maxit = s->maxit;
  A.row_begin = 0; A.row_end = s->npad; A.fused = 1; A.scal = nullptr;
  A.gamma = s->gamma;
  A.four_ag = 4.0 * s->alpha * s->gamma;
  A.tol = s->tol;
  int grid = (s->npad + KNN_NODE_THREADS - 1) / KNN_NODE_THREADS;
  if (grid > 148 * 2) grid = 148 * 2;
  knn_node_kernel<F><<<grid, KNN_NODE_THREADS, 0, st>>>(A);
  GL_LAUNCH_CHECK();
  return GL_OK;
}


}  // namespace gl
