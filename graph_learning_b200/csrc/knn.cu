// knn.cu -- K6: the LogModel iteration restricted to an explicit edge list (config 5).
//
// The reference has NO code for this; it is the same forward-backward-forward
// iteration as graph_learn/smooth_learning.py:146-167 with the sum operator S restricted to the
// columns of the support, and step size gamma = step / (2 beta + sqrt(2 d_max)) (signless-Laplacian
// bound on ||S_mask||_2; equals smooth_learning.py:135-137 on the complete graph).  Parity is
// pinned against oracle/logmodel_oracle.py::solve_knn and, with the complete edge list, against the
// dense solver.
//
// Layout: every undirected edge (i, j) is stored TWICE, as directed slots i->j and j->i, sorted by
// source node, each slot carrying its own copy of w and z.  Both copies apply the same update
// (a_i + a_j is commutative, so they stay bit-identical), and each slot contributes only to the
// degree sums of its SOURCE node: consecutive slots share the source, a warp reduces its 32 slots by
// segments with shuffles and issues one float64 atomic per (warp, source node) instead of four per
// edge, and nothing is scattered to the far end of an edge.  Traffic: 20 B per slot = 40 B per
// undirected edge per iteration in fp32 (src, dst, w, z, w'); SURVEY 8(d)'s algorithmic figure of
// 20 B per stored edge is what roofline.achieved is quoted on.  The node vectors (a, b: 4 B x n) are
// L2-resident gathers.  The node tail is the dense solver's (node_tail.cu) on a workspace that owns
// no tiles.
#include <cub/device/device_radix_sort.cuh>

#include <math_constants.h>

#include "solver.cuh"

namespace gl {

constexpr int KNN_THREADS = 256;
constexpr int KNN_UNROLL = 4;
template <typename F> struct Vec2;
template <> struct Vec2<float> { using type = float2; };
template <> struct Vec2<double> { using type = double2; };

__global__ void knn_degree_count_kernel(const int32_t *ei, const int32_t *ej, int64_t nnz, int *cnt) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += (int64_t)gridDim.x * blockDim.x) {
    atomicAdd(cnt + ei[e], 1);
    atomicAdd(cnt + ej[e], 1);
  }
}
__global__ void knn_max_kernel(const int *cnt, int n, int *out) {
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
__global__ void knn_expand_kernel(const int32_t *ei, const int32_t *ej, int64_t nnz, int32_t *key, int32_t *val) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += (int64_t)gridDim.x * blockDim.x) {
    key[2 * e] = ei[e];
    val[2 * e] = (int32_t)e;
    key[2 * e + 1] = ej[e];
    val[2 * e + 1] = (int32_t)e;
  }
}
// sorted slots: fill in the far end and this slot's copies of z and w
template <typename F>
__global__ void knn_fill_kernel(const int32_t *ei, const int32_t *ej, const F *z, const F *w, int64_t nslots,
                                const int32_t *src, const int32_t *eid, int32_t *dst, F *zs, F *ws) {
  for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < nslots; s += (int64_t)gridDim.x * blockDim.x) {
    const int32_t e = eid[s], i = ei[e], j = ej[e];
    dst[s] = src[s] == i ? j : i;
    zs[s] = z[e];
    ws[s] = w[e];
  }
}
// result back in the caller's edge order, from the slot whose source is the edge's first end; the
// edge_tol threshold of smooth_learning.py:169 applied on the way
template <typename F>
__global__ void knn_writeback_kernel(const int32_t *ei, const int32_t *src, const int32_t *eid, const F *ws,
                                     int64_t nslots, double edge_tol, F *w) {
  for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < nslots; s += (int64_t)gridDim.x * blockDim.x) {
    const int32_t e = eid[s];
    if (src[s] == ei[e]) {
      const F v = ws[s];
      w[e] = (double)v < edge_tol ? F(0) : v;
    }
  }
}

// MODE 0: iteration.  MODE 1: initial degrees S w0 and ||w0||^2.  One thread per directed slot,
// warps own 32 consecutive slots; gsum receives HALF the slot sums (every edge appears twice).
template <typename F, int MODE>
__global__ void __launch_bounds__(KNN_THREADS) knn_slot_kernel(const int32_t *__restrict__ src,
                                                               const int32_t *__restrict__ dst,
                                                               const F *__restrict__ z, F *__restrict__ w,
                                                               int64_t nslots,
                                                               const typename Vec2<F>::type *__restrict__ ab,
                                                               double *sp, double *sw,
                                                               double *gsum, const int *flags, F c, F tgz) {
  if (MODE == 0 && *reinterpret_cast<const volatile int *>(flags + FLAG_DONE)) return;
  __shared__ double s_red[KNN_THREADS / 32][2];
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  double res = 0.0, wsq = 0.0;
  // Each warp takes KNN_UNROLL groups of 32 consecutive slots per step and issues all their loads
  // (then all their gathers) before the first use: one group per step leaves the kernel
  // latency-bound (measured 1.25 TB/s moved at 1.1e7 edges).
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * KNN_UNROLL;
  for (int64_t s0 = ((int64_t)blockIdx.x * (KNN_THREADS / 32) + wp) * 32 * KNN_UNROLL; s0 < nslots; s0 += stride) {
    int i[KNN_UNROLL], j[KNN_UNROLL];
    F wo[KNN_UNROLL], zz[KNN_UNROLL];
#pragma unroll
    for (int u = 0; u < KNN_UNROLL; ++u) {
      const int64_t s = s0 + u * 32 + lane;
      const bool valid = s < nslots;
      i[u] = valid ? src[s] : -1;
      wo[u] = valid ? w[s] : F(0);
      if (MODE == 0) {
        j[u] = valid ? dst[s] : 0;
        zz[u] = valid ? z[s] : F(0);
      }
    }
    typename Vec2<F>::type abi[KNN_UNROLL], abj[KNN_UNROLL];
    if (MODE == 0) {
#pragma unroll
      for (int u = 0; u < KNN_UNROLL; ++u) {
        abi[u] = ab[i[u] < 0 ? 0 : i[u]];   // (a, b) of both ends: one gather each
        abj[u] = ab[j[u]];
      }
    }
#pragma unroll
    for (int u = 0; u < KNN_UNROLL; ++u) {
      const int64_t s = s0 + u * 32 + lane;
      const bool valid = s < nslots;
      F vp = F(0), vw = F(0);   // segment sums in the edge dtype (<= 32 terms), float64 from the atomic on
      if (valid) {
        if (MODE == 1) {
          vw = wo[u];
          wsq += (double)vw * (double)vw;
        } else {
          const F y = c * wo[u] - (abi[u].x + abj[u].x);
          F p = y - tgz * zz[u];
          p = p > F(0) ? p : F(0);
          const F q = c * p - (abi[u].y + abj[u].y);
          const F r = q - y;
          const F wn = (wo[u] - y) + q;
          w[s] = wn;
          vp = p;
          vw = wn;
          res += (double)r * (double)r;
          wsq += (double)wn * (double)wn;
        }
      }
      // segmented suffix sums over runs of equal source (slots are sorted by source)
      const int iu = i[u];
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const int io = __shfl_down_sync(0xffffffffu, iu, off);
        const F po = __shfl_down_sync(0xffffffffu, vp, off);
        const F wo2 = __shfl_down_sync(0xffffffffu, vw, off);
        if (lane + off < 32 && io == iu) {
          vp += po;
          vw += wo2;
        }
      }
      const int iprev = __shfl_up_sync(0xffffffffu, iu, 1);
      if (valid && (lane == 0 || iprev != iu)) {
        if (MODE == 0 && vp != F(0)) atomicAdd(sp + iu, (double)vp);
        atomicAdd(sw + iu, (double)vw);
      }
    }
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

// Node phase of the kNN path (single graph, single GPU): the arithmetic of node_tail.cu for K == 1
// (smooth_learning.py:148-166) as a grid-stride kernel -- a few hundred CTAs however large n is,
// several independent nodes in flight per thread -- because one CTA per 256 nodes (3,907 CTAs at
// n = 1e6, each paying a fence, a barrier and a ticket) measured 123 us per launch, latency-bound.
// Only the interleaved (a, b) pairs the slot kernel gathers are written, not a and b separately.
constexpr int KNN_NODE_THREADS = 256;
struct KnnNodeArgs {
  double *d, *yb, *pb, *rb, *xch, *gprev, *resid, *node_partial;
  void *ab;
  int *flags;
  int n, npad, init, maxit;
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
  for (int64_t idx = (int64_t)blockIdx.x * KNN_NODE_THREADS + tid; idx < A.npad;
       idx += (int64_t)gridDim.x * KNN_NODE_THREADS) {
    const double sp = A.xch[idx], sw = A.xch[A.npad + idx];
    A.xch[idx] = 0.0;
    A.xch[A.npad + idx] = 0.0;
    double dn = 0.0, rbv = 0.0, ybn = 0.0, pbn = 0.0;
    if (idx < A.n) {
      if (A.init) {
        dn = sw;                                          // d0 = S w0            :144
      } else {
        const double dv = A.d[idx], ybv = A.yb[idx];
        const double qb = A.pb[idx] + A.gamma * sp;       //                      :154
        rbv = qb - ybv;
        dn = (dv - ybv) + qb;                             //                      :164
      }
      ybn = dn + A.gamma * sw;                            //                      :148
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

template <typename F> static int launch_knn_node(gl_solver *s, bool init, cudaStream_t st) {
  KnnNodeArgs A;
  A.d = s->d; A.yb = s->yb; A.pb = s->pb; A.rb = s->rb;
  A.xch = s->xch; A.gprev = s->gprev; A.resid = s->resid; A.node_partial = s->node_partial;
  A.ab = s->ab;
  A.flags = s->flags;
  A.n = s->n; A.npad = s->npad; A.init = init ? 1 : 0; A.maxit = s->maxit;
  A.gamma = s->gamma;
  A.four_ag = 4.0 * s->alpha * s->gamma;
  A.tol = s->tol;
  int grid = (s->npad + KNN_NODE_THREADS - 1) / KNN_NODE_THREADS;
  if (grid > 148 * 4) grid = 148 * 4;
  knn_node_kernel<F><<<grid, KNN_NODE_THREADS, 0, st>>>(A);
  GL_LAUNCH_CHECK();
  return GL_OK;
}

template <typename F>
static int knn_solve_t(const int32_t *ei, const int32_t *ej, const F *z, int64_t nnz, int n, double alpha,
                       double beta, int maxit, double tol, double step_size, double edge_tol, F *w, int *n_iter,
                       double *rel_primal, double *rel_dual, cudaStream_t st) {
  gl_solver *s = nullptr;
  GL_TRY(gl_solver_create(&s, n, 1, sizeof(F) == 4 ? GL_F32 : GL_F64, 0, 0));
  if (cudaMalloc(&s->ab, 2 * sizeof(F) * (size_t)s->npad) != cudaSuccess) {
    gl_solver_destroy(s);
    return fail(GL_E_ALLOC, "kNN solve: out of device memory");
  }
  cudaMemsetAsync(s->ab, 0, 2 * sizeof(F) * (size_t)s->npad, st);
  int rc = GL_OK;
  int *cnt = nullptr;
  int *h_flags = nullptr;
  cudaGraphExec_t gexec = nullptr;
  // directed, source-sorted slots
  const int64_t nslots = 2 * nnz;
  int32_t *key_in = nullptr, *val_in = nullptr, *src = nullptr, *eid = nullptr, *dst = nullptr;
  F *zs = nullptr, *ws = nullptr;
  void *sort_tmp = nullptr;
  auto done = [&](int code) {
    for (void *q : {(void *)key_in, (void *)val_in, (void *)src, (void *)eid, (void *)dst, (void *)zs, (void *)ws, sort_tmp})
      if (q) cudaFree(q);
    if (cnt) cudaFree(cnt);
    if (h_flags) cudaFreeHost(h_flags);
    if (gexec) cudaGraphExecDestroy(gexec);
    gl_solver_destroy(s);
    return code;
  };
#define KNN_CUDA(expr)                                                                      \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess)                                                                  \
      return done(fail(GL_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),   \
                       __FILE__, __LINE__));                                                \
  } while (0)
  if (nslots >= (int64_t)1 << 31) return done(fail(GL_E_SHAPE, "edge list too long for 32-bit slot ids"));
  const int grid = (int)std::min<int64_t>((nnz + KNN_THREADS - 1) / KNN_THREADS, 148 * 8);
  const int sgrid = (int)std::min<int64_t>((nslots + KNN_THREADS - 1) / KNN_THREADS, 148 * 8);
  // d_max of the support
  KNN_CUDA(cudaMalloc((void **)&cnt, sizeof(int) * ((size_t)n + 1)));
  KNN_CUDA(cudaMemsetAsync(cnt, 0, sizeof(int) * ((size_t)n + 1), st));
  int dmax = 0;
  if (nnz > 0) {
    knn_degree_count_kernel<<<grid, KNN_THREADS, 0, st>>>(ei, ej, nnz, cnt);
    g_launches.fetch_add(1);
    knn_max_kernel<<<1, 256, 0, st>>>(cnt, n, cnt + n);
    g_launches.fetch_add(1);
    KNN_CUDA(cudaGetLastError());
    KNN_CUDA(cudaMemcpyAsync(&dmax, cnt + n, sizeof(int), cudaMemcpyDeviceToHost, st));
    KNN_CUDA(cudaStreamSynchronize(st));
  }
  if (nnz > 0) {
    const size_t ib = sizeof(int32_t) * (size_t)nslots, fb = sizeof(F) * (size_t)nslots;
    KNN_CUDA(cudaMalloc((void **)&key_in, ib));
    KNN_CUDA(cudaMalloc((void **)&val_in, ib));
    KNN_CUDA(cudaMalloc((void **)&src, ib));
    KNN_CUDA(cudaMalloc((void **)&eid, ib));
    KNN_CUDA(cudaMalloc((void **)&dst, ib));
    KNN_CUDA(cudaMalloc((void **)&zs, fb));
    KNN_CUDA(cudaMalloc((void **)&ws, fb));
    knn_expand_kernel<<<grid, KNN_THREADS, 0, st>>>(ei, ej, nnz, key_in, val_in);
    g_launches.fetch_add(1);
    int bits = 1;
    while (bits < 31 && ((int64_t)1 << bits) < (int64_t)n) ++bits;
    size_t tmp_bytes = 0;
    KNN_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, key_in, src, val_in, eid, (int)nslots, 0, bits, st));
    KNN_CUDA(cudaMalloc(&sort_tmp, tmp_bytes ? tmp_bytes : 16));
    KNN_CUDA(cub::DeviceRadixSort::SortPairs(sort_tmp, tmp_bytes, key_in, src, val_in, eid, (int)nslots, 0, bits, st));
    knn_fill_kernel<F><<<sgrid, KNN_THREADS, 0, st>>>(ei, ej, z, w, nslots, src, eid, dst, zs, ws);
    g_launches.fetch_add(1);
    KNN_CUDA(cudaGetLastError());
  }
  s->alpha = alpha; s->beta = beta; s->step = step_size; s->maxit = maxit; s->tol = tol;
  s->gamma = step_size / (2.0 * beta + std::sqrt(2.0 * (double)dmax));
  s->c = 1.0 - 2.0 * s->gamma * beta;
  double *sp = s->xch, *sw = s->xch + s->npad, *gsum = s->xch + 2 * (int64_t)s->npad;
  const F c = (F)s->c, tgz = (F)(2.0 * s->gamma);
  KNN_CUDA(cudaMemsetAsync(s->xch, 0, sizeof(double) * (size_t)s->xch_len, st));
  KNN_CUDA(cudaMemsetAsync(s->flags, 0, sizeof(int) * FLAG_COUNT, st));
  if (nnz > 0) {
    knn_slot_kernel<F, 1><<<sgrid, KNN_THREADS, 0, st>>>(src, dst, zs, ws, nslots, (const typename Vec2<F>::type *)s->ab,
                                                        sp, sw, gsum, s->flags, c, tgz);
    g_launches.fetch_add(1);
    KNN_CUDA(cudaGetLastError());
  }
  rc = launch_knn_node<F>(s, true, st);
  if (rc != GL_OK) return done(rc);
  s->begun = true;

  // iteration graph: `chunk` iterations per replay, stop flag polled between replays
  const int chunk = maxit < 16 ? (maxit > 0 ? maxit : 1) : 16;
  cudaStream_t cs;
  KNN_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
  cudaEvent_t ev;
  cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
  cudaEventRecord(ev, st);
  cudaStreamWaitEvent(cs, ev, 0);
  cudaGraph_t graph = nullptr;
  const long long before = g_launches.load();
  cudaError_t ce = cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal);
  if (ce == cudaSuccess) {
    for (int i = 0; i < chunk && rc == GL_OK; ++i) {
      if (nnz > 0)
        knn_slot_kernel<F, 0><<<sgrid, KNN_THREADS, 0, cs>>>(src, dst, zs, ws, nslots, (const typename Vec2<F>::type *)s->ab,
                                                            sp, sw, gsum, s->flags, c, tgz);
      rc = launch_knn_node<F>(s, false, cs);
    }
    ce = cudaStreamEndCapture(cs, &graph);
  }
  g_launches.store(before);
  if (ce == cudaSuccess && rc == GL_OK) ce = cudaGraphInstantiate(&gexec, graph, 0);
  if (graph) cudaGraphDestroy(graph);
  if (ce == cudaSuccess && rc == GL_OK) ce = cudaMallocHost((void **)&h_flags, sizeof(int) * FLAG_COUNT);
  if (ce == cudaSuccess && rc == GL_OK) {
    const int per_iter = (nnz > 0 ? 1 : 0) + 1;
    for (int it = 0; it < maxit; it += chunk) {
      ce = cudaGraphLaunch(gexec, cs);
      g_launches.fetch_add((long long)per_iter * chunk);
      if (ce == cudaSuccess)
        ce = cudaMemcpyAsync(h_flags, s->flags, sizeof(int) * FLAG_COUNT, cudaMemcpyDeviceToHost, cs);
      if (ce == cudaSuccess) ce = cudaStreamSynchronize(cs);
      if (ce != cudaSuccess || h_flags[FLAG_DONE]) break;
    }
  }
  if (ce == cudaSuccess) ce = cudaStreamSynchronize(cs);
  cudaEventDestroy(ev);
  cudaStreamDestroy(cs);
  if (rc != GL_OK) return done(rc);
  if (ce != cudaSuccess) return done(fail(GL_E_CUDA, "kNN solve failed: %s", cudaGetErrorString(ce)));
  rc = gl_solver_status(s, n_iter, rel_primal, rel_dual, nullptr, st);
  if (rc == GL_OK && nnz > 0) {
    knn_writeback_kernel<F><<<sgrid, KNN_THREADS, 0, st>>>(ei, src, eid, ws, nslots, edge_tol, w);
    g_launches.fetch_add(1);
    KNN_CUDA(cudaGetLastError());
    KNN_CUDA(cudaStreamSynchronize(st));
  }
#undef KNN_CUDA
  return done(rc);
}

}  // namespace gl
using namespace gl;

extern "C" int gl_knn_logmodel_solve(const int32_t *ei, const int32_t *ej, const void *z, int64_t nnz, int n,
                                     int dtype, double alpha, double beta, int maxit, double tol,
                                     double step_size, double edge_tol, void *w, int *n_iter,
                                     double *rel_primal, double *rel_dual, void *stream) {
  if (n < 2) return fail(GL_E_SHAPE, "need at least 2 nodes, got n=%d", n);
  if (nnz < 0) return fail(GL_E_SHAPE, "negative edge count");
  cudaStream_t st = as_stream(stream);
  if (dtype == GL_F32)
    return knn_solve_t<float>(ei, ej, (const float *)z, nnz, n, alpha, beta, maxit, tol, step_size, edge_tol,
                              (float *)w, n_iter, rel_primal, rel_dual, st);
  if (dtype == GL_F64)
    return knn_solve_t<double>(ei, ej, (const double *)z, nnz, n, alpha, beta, maxit, tol, step_size, edge_tol,
                               (double *)w, n_iter, rel_primal, rel_dual, st);
  return fail(GL_E_ARG, "dtype must be GL_F32 or GL_F64");
}
