// knn.cu -- K6: the LogModel iteration restricted to an explicit edge list (config 5).
//
// The reference has NO code for this; it is the same forward-backward-forward
// iteration as graph_learn/smooth_learning.py:146-167 with the sum operator S restricted to the
// columns of the support, and step size gamma = step / (2 beta + sqrt(2 d_max)) (signless-Laplacian
// bound on ||S_mask||_2; equals smooth_learning.py:135-137 on the complete graph).  Parity is
// pinned against oracle/logmodel_oracle.py::solve_knn and, with the complete edge list, against the
// dense solver.
//
// Layout: every undirected edge (i, j) is stored TWICE, as directed slots i->j and j->i, each slot
// carrying its own copy of w and z, grouped by source node in a sliced-ELL format (SELL-32, see
// knn_kernels.cuh).  Both copies apply the same update (a_i + a_j is commutative, so they stay
// bit-identical), and each slot contributes only to the degree sums of its SOURCE node, which one
// lane owns: no atomics, no shuffles, nothing scattered to the far end of an edge, deterministic
// sums.  Traffic: 16 B per slot (dst, w, z, w') = 32 B per undirected edge per iteration in fp32
// plus padding; SURVEY 8(d)'s algorithmic figure of 20 B per stored edge is what the roofline is
// quoted on.  The node vectors (interleaved (a, b) pairs, 8 B x n) are L2-resident gathers.
// This file holds the one-shot entry point; the stepwise / sharded solver is knn_sharded.cu.
#include "knn_kernels.cuh"

namespace gl {

// One-shot solve = the stepwise solver on a single shard [0, npad) in its one-launch form (slot sweep,
// node update and stop rule in knn_fused_kernel), replayed from a CUDA graph, stop flag polled between
// replays.
static int knn_solve(const int32_t *ei, const int32_t *ej, const void *z, int64_t nnz, int n, int dtype, double alpha,
                     double beta, int maxit, double tol, double step_size, double edge_tol, void *w, int *n_iter,
                     double *rel_primal, double *rel_dual, const int32_t *order_key, int sort_window, int layout,
                     cudaStream_t st) {
  int dmax = 0;
  GL_TRY(gl_knn_max_degree(ei, ej, nnz, n, &dmax, st));
  gl_knn_solver *k = nullptr;
  const int npad = tiles_per_side(n) * T;
  GL_TRY(gl_knn_create_ordered(&k, ei, ej, z, w, nnz, n, dtype, 0, npad, npad, order_key, sort_window, layout, st));
  {
    const int frc = gl_knn_set_fused(k, 2);   // whole iteration in one launch
    if (frc != GL_OK) {
      gl_knn_destroy(k);
      return frc;
    }
  }
  int *h_flags = nullptr;
  cudaGraphExec_t gexec = nullptr;
  cudaStream_t cs = nullptr;
  auto done = [&](int code) {
    if (h_flags) cudaFreeHost(h_flags);
    if (gexec) cudaGraphExecDestroy(gexec);
    if (cs) cudaStreamDestroy(cs);
    gl_knn_destroy(k);
    return code;
  };
  int rc = gl_knn_begin(k, alpha, beta, step_size, maxit, tol, dmax, st);
  if (rc != GL_OK) return done(rc);
  // iterations per graph replay; even, because the two (a, b) buffers swap roles every launch and the
  // captured pointers must line up again at the start of every replay (launches past maxit are no-ops)
  int chunk = maxit < 32 ? (maxit > 0 ? maxit : 1) : 32;
  chunk += chunk & 1;
  cudaError_t ce = cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking);
  cudaEvent_t ev = nullptr;
  if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
  if (ce == cudaSuccess) ce = cudaEventRecord(ev, st);
  if (ce == cudaSuccess) ce = cudaStreamWaitEvent(cs, ev, 0);
  cudaGraph_t graph = nullptr;
  const long long before = g_launches.load();
  if (ce == cudaSuccess) ce = cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal);
  if (ce == cudaSuccess) {
    for (int i = 0; i < chunk && rc == GL_OK; ++i) {
      rc = gl_knn_edge_step(k, cs);
      if (rc == GL_OK) rc = gl_knn_node_step(k, cs);
    }
    ce = cudaStreamEndCapture(cs, &graph);
  }
  const long long per_chunk = g_launches.load() - before;   // launches recorded once, replayed many times
  g_launches.store(before);
  if (ce == cudaSuccess && rc == GL_OK) ce = cudaGraphInstantiate(&gexec, graph, 0);
  if (graph) cudaGraphDestroy(graph);
  if (ce == cudaSuccess && rc == GL_OK) ce = cudaMallocHost((void **)&h_flags, sizeof(int) * 2 * FLAG_COUNT);
  cudaEvent_t evs[2] = {nullptr, nullptr};
  for (auto &e2 : evs)
    if (ce == cudaSuccess && rc == GL_OK) ce = cudaEventCreateWithFlags(&e2, cudaEventDisableTiming);
  if (ce == cudaSuccess && rc == GL_OK) {
    // pipelined replay (as gl_solver_run): replay g + 1 is enqueued before the flags of replay g are read,
    // so the device never idles while the host polls; iterations past the stop are device-side no-ops
    const int max_graphs = (maxit + chunk - 1) / chunk;
    const int *dflags = gl_knn_flags_ptr(k);
    int launched = 0, checked = 0;
    bool fin = maxit <= 0;
    while (!fin && ce == cudaSuccess) {
      while (launched < max_graphs && launched - checked < 2 && ce == cudaSuccess) {
        ce = cudaGraphLaunch(gexec, cs);
        g_launches.fetch_add(per_chunk);
        if (ce == cudaSuccess)
          ce = cudaMemcpyAsync(h_flags + (launched & 1) * FLAG_COUNT, dflags, sizeof(int) * FLAG_COUNT,
                               cudaMemcpyDeviceToHost, cs);
        if (ce == cudaSuccess) ce = cudaEventRecord(evs[launched & 1], cs);
        ++launched;
      }
      if (checked >= launched || ce != cudaSuccess) break;
      ce = cudaEventSynchronize(evs[checked & 1]);
      fin = h_flags[(checked & 1) * FLAG_COUNT + FLAG_DONE] != 0;
      ++checked;
    }
  }
  for (auto e2 : evs)
    if (e2) cudaEventDestroy(e2);
  if (ce == cudaSuccess && cs) ce = cudaStreamSynchronize(cs);
  if (ev) cudaEventDestroy(ev);
  if (rc != GL_OK) return done(rc);
  if (ce != cudaSuccess) return done(fail(GL_E_CUDA, "kNN solve failed: %s", cudaGetErrorString(ce)));
  rc = gl_knn_status(k, n_iter, rel_primal, rel_dual, nullptr, st);
  if (rc == GL_OK && nnz > 0) {
    // untouched entries cannot occur on a single shard: every edge has its first end in [0, n)
    rc = gl_knn_store(k, w, edge_tol, st);
    if (rc == GL_OK && cudaStreamSynchronize(st) != cudaSuccess) rc = fail(GL_E_CUDA, "kNN store failed");
  }
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
  if (dtype != GL_F32 && dtype != GL_F64) return fail(GL_E_ARG, "dtype must be GL_F32 or GL_F64");
  return knn_solve(ei, ej, z, nnz, n, dtype, alpha, beta, maxit, tol, step_size, edge_tol, w, n_iter, rel_primal,
                   rel_dual, nullptr, 0, 0, as_stream(stream));
}

extern "C" int gl_knn_logmodel_solve_ordered(const int32_t *ei, const int32_t *ej, const void *z, int64_t nnz, int n,
                                             int dtype, double alpha, double beta, int maxit, double tol,
                                             double step_size, double edge_tol, void *w, int *n_iter,
                                             double *rel_primal, double *rel_dual, const int32_t *order_key,
                                             int sort_window, int layout, void *stream) {
  if (n < 2) return fail(GL_E_SHAPE, "need at least 2 nodes, got n=%d", n);
  if (nnz < 0) return fail(GL_E_SHAPE, "negative edge count");
  if (dtype != GL_F32 && dtype != GL_F64) return fail(GL_E_ARG, "dtype must be GL_F32 or GL_F64");
  return knn_solve(ei, ej, z, nnz, n, dtype, alpha, beta, maxit, tol, step_size, edge_tol, w, n_iter, rel_primal,
                   rel_dual, order_key, sort_window, layout, as_stream(stream));
}
