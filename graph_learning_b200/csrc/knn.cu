// knn.cu -- K6: the LogModel iteration restricted to an explicit edge list (config 5).
//
// The reference has NO code for this; it is the same forward-backward-forward
// iteration as graph_learn/smooth_learning.py:146-167 with the sum operator S restricted to the
// columns of the support, and step size gamma = step / (2 beta + sqrt(2 d_max)) (signless-Laplacian
// bound on ||S_mask||_2; equals smooth_learning.py:135-137 on the complete graph).  Parity is
// pinned against oracle/logmodel_oracle.py::solve_knn and, with the complete edge list, against the
// dense solver.  Algorithmic traffic: 20 B per stored edge per iteration in fp32 (i, j, w, z, w').
// The node tail is the dense solver's (node_tail.cu) on a workspace that owns no tiles.
#include "solver.cuh"

namespace gl {

constexpr int KNN_THREADS = 256;

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

// MODE 0: iteration.  MODE 1: initial degrees S w0 and ||w0||^2.
template <typename F, int MODE>
__global__ void __launch_bounds__(KNN_THREADS) knn_edge_kernel(const int32_t *__restrict__ ei,
                                                               const int32_t *__restrict__ ej,
                                                               const F *__restrict__ z, F *__restrict__ w,
                                                               int64_t nnz, const F *__restrict__ a,
                                                               const F *__restrict__ b, double *sp, double *sw,
                                                               double *gsum, const int *flags, F c, F tgz) {
  if (MODE == 0 && *reinterpret_cast<const volatile int *>(flags + FLAG_DONE)) return;
  __shared__ double s_red[KNN_THREADS / 32][2];
  double res = 0.0, wsq = 0.0;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += (int64_t)gridDim.x * blockDim.x) {
    const int i = ei[e], j = ej[e];
    const F wo = w[e];
    if (MODE == 1) {
      atomicAdd(sw + i, (double)wo);
      atomicAdd(sw + j, (double)wo);
      wsq += (double)wo * (double)wo;
    } else {
      const F y = c * wo - (a[i] + a[j]);
      F p = y - tgz * z[e];
      p = p > F(0) ? p : F(0);
      const F q = c * p - (b[i] + b[j]);
      const F r = q - y;
      const F wn = (wo - y) + q;
      w[e] = wn;
      if (p != F(0)) {
        atomicAdd(sp + i, (double)p);
        atomicAdd(sp + j, (double)p);
      }
      atomicAdd(sw + i, (double)wn);
      atomicAdd(sw + j, (double)wn);
      res += (double)r * (double)r;
      wsq += (double)wn * (double)wn;
    }
  }
  res = warp_sum(res);
  wsq = warp_sum(wsq);
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
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
    if (MODE == 0) atomicAdd(gsum, r2);
    atomicAdd(gsum + 1, w2);
  }
}

template <typename F> __global__ void knn_threshold_kernel(F *w, int64_t nnz, double edge_tol) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += (int64_t)gridDim.x * blockDim.x)
    if ((double)w[e] < edge_tol) w[e] = F(0);
}

template <typename F>
static int knn_solve_t(const int32_t *ei, const int32_t *ej, const F *z, int64_t nnz, int n, double alpha,
                       double beta, int maxit, double tol, double step_size, double edge_tol, F *w, int *n_iter,
                       double *rel_primal, double *rel_dual, cudaStream_t st) {
  gl_solver *s = nullptr;
  GL_TRY(gl_solver_create(&s, n, 1, sizeof(F) == 4 ? GL_F32 : GL_F64, 0, 0));
  int rc = GL_OK;
  int *cnt = nullptr;
  int *h_flags = nullptr;
  cudaGraphExec_t gexec = nullptr;
  auto done = [&](int code) {
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
  const int grid = (int)std::min<int64_t>((nnz + KNN_THREADS - 1) / KNN_THREADS, 148 * 8);
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
  s->alpha = alpha; s->beta = beta; s->step = step_size; s->maxit = maxit; s->tol = tol;
  s->gamma = step_size / (2.0 * beta + std::sqrt(2.0 * (double)dmax));
  s->c = 1.0 - 2.0 * s->gamma * beta;
  double *sp = s->xch, *sw = s->xch + s->npad, *gsum = s->xch + 2 * (int64_t)s->npad;
  const F c = (F)s->c, tgz = (F)(2.0 * s->gamma);
  KNN_CUDA(cudaMemsetAsync(s->xch, 0, sizeof(double) * (size_t)s->xch_len, st));
  KNN_CUDA(cudaMemsetAsync(s->flags, 0, sizeof(int) * FLAG_COUNT, st));
  if (nnz > 0) {
    knn_edge_kernel<F, 1><<<grid, KNN_THREADS, 0, st>>>(ei, ej, z, w, nnz, (const F *)s->a, (const F *)s->b, sp,
                                                       sw, gsum, s->flags, c, tgz);
    g_launches.fetch_add(1);
    KNN_CUDA(cudaGetLastError());
  }
  rc = launch_node_tail(s, true, st);
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
        knn_edge_kernel<F, 0><<<grid, KNN_THREADS, 0, cs>>>(ei, ej, z, w, nnz, (const F *)s->a, (const F *)s->b,
                                                           sp, sw, gsum, s->flags, c, tgz);
      rc = launch_node_tail(s, false, cs);
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
    knn_threshold_kernel<F><<<grid, KNN_THREADS, 0, st>>>(w, nnz, edge_tol);
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
