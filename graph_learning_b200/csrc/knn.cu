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
#include "knn_kernels.cuh"

namespace gl {

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
  unsigned long long *slot_count = nullptr;
  auto done = [&](int code) {
    for (void *q : {(void *)key_in, (void *)val_in, (void *)src, (void *)eid, (void *)dst, (void *)zs, (void *)ws, sort_tmp, (void *)slot_count})
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
    KNN_CUDA(cudaMalloc((void **)&slot_count, sizeof(unsigned long long)));
    KNN_CUDA(cudaMemsetAsync(slot_count, 0, sizeof(unsigned long long), st));
    knn_expand_kernel<<<grid, KNN_THREADS, 0, st>>>(ei, ej, nnz, key_in, val_in, 0, n, n, slot_count);
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
