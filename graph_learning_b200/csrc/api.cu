// api.cu -- C-ABI entry points of the solver (include/gl_b200.h) and library plumbing.
//
// gl_solver_* mirrors the phases of gsp_learn_graph_log_degrees (graph_learn/smooth_learning.py):
//   create/load  <- :125-133 (shape handling, operator set-up -- no operator is materialised here)
//   begin        <- :135-144 (step size, initial point, d = S w)
//   edge_pass + node_tail (per iteration) <- :146-167
//   store        <- :169-174 (threshold, output layout)
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "solver.cuh"

namespace gl {
thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};
int pack_to_tiles_impl(const void *packed, void *tiled, int n, int K, int dtype, int tb, int te,
                       const double *scale_dev, cudaStream_t st);
}  // namespace gl
using namespace gl;

namespace {

template <typename F>
__global__ void scale_rows_kernel(F *z, int64_t gstride, const double *scale) {
  const int k = blockIdx.y;
  const double sc = scale[k];
  F *p = z + (int64_t)k * gstride;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < gstride;
       i += (int64_t)gridDim.x * blockDim.x)
    p[i] = (F)((double)p[i] * sc);
}

template <typename F>
__global__ void __launch_bounds__(256) store_packed_kernel(const F *tiled, int64_t gstride, F *packed,
                                                          const int2 *tile_ij, int n, double edge_tol) {
  const int t = blockIdx.x, k = blockIdx.y;
  const int2 ij = tile_ij[t];
  const int64_t E = (int64_t)n * (n - 1) / 2;
  F *dst = packed + (int64_t)k * E;
  const F *src = tiled + (int64_t)k * gstride + (int64_t)t * TT;
  for (int idx = threadIdx.x; idx < TT; idx += blockDim.x) {
    const int r = idx / T, c = idx % T;
    const int gi = ij.x * T + r, gj = ij.y * T + c;
    if (gj > gi && gj < n) {
      F v = src[idx];
      if ((double)v < edge_tol) v = F(0);  // smooth_learning.py:169
      dst[packed_index(gi, gj, n)] = v;
    }
  }
}

void free_all(gl_solver *s) {
  if (s->graph_exec) cudaGraphExecDestroy(s->graph_exec);
  for (void *p : s->ipc_opened)
    if (p) cudaIpcCloseMemHandle(p);
  void *ptrs[] = {s->w, s->z, s->r, s->tile_ij, s->d, s->yb, s->pb, s->rb, s->a, s->b,
                  s->comm, s->cta_partial, s->node_partial, s->gprev, s->resid, s->eig_scratch, s->flags, s->grid_bar, s->trace, s->ab, s->gram_part,
                  s->small_state, s->batch_partial, s->batch_node, s->egram_part};
  for (void *p : ptrs)
    if (p) cudaFree(p);
}

int alloc_zero(void **p, size_t bytes) {
  if (bytes == 0) bytes = 16;
  cudaError_t e = cudaMalloc(p, bytes);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(GL_E_ALLOC, "cudaMalloc(%zu bytes) failed: %s", bytes, cudaGetErrorString(e));
  }
  GL_CUDA(cudaMemset(*p, 0, bytes));
  return GL_OK;
}

}  // namespace

extern "C" {

int gl_version(void) { return 100; }
const char *gl_last_error(void) { return g_err; }
int64_t gl_launch_count(void) { return (int64_t)g_launches.load(); }

int64_t gl_num_edges(int n) { return (int64_t)n * (n - 1) / 2; }
int gl_nodes_from_edges(int64_t e) {
  if (e < 1) return -1;
  const int n = (int)std::llround((1.0 + std::sqrt(1.0 + 8.0 * (double)e)) / 2.0);  // :133
  return ((int64_t)n * (n - 1) / 2 == e) ? n : -1;
}
int gl_num_tiles(int n) { return num_tiles(n); }
int64_t gl_tiled_elems(int n) { return (int64_t)num_tiles(n) * TT; }

int gl_shard_tiles(int n, int rank, int world, int *tile_begin, int *tile_end) {
  if (n < 2 || world < 1 || rank < 0 || rank >= world) return fail(GL_E_ARG, "bad shard arguments");
  const int nt = tiles_per_side(n), total = num_tiles(n);
  // weights: off-diagonal tile 2, diagonal tile 1
  const int64_t wsum = 2LL * total - nt;
  auto boundary = [&](int r) -> int {
    if (r <= 0) return 0;
    if (r >= world) return total;
    const int64_t target = wsum * r / world;
    int64_t acc = 0;
    int t = 0;
    for (int I = 0; I < nt && acc < target; ++I)
      for (int J = I; J < nt && acc < target; ++J) {
        acc += (I == J) ? 1 : 2;
        ++t;
      }
    return t;
  };
  *tile_begin = boundary(rank);
  *tile_end = boundary(rank + 1);
  return GL_OK;
}

int gl_solver_create(gl_solver **out, int n, int K, int dtype, int tile_begin, int tile_end) {
  if (!out) return fail(GL_E_ARG, "null output pointer");
  *out = nullptr;
  if (n < 2) return fail(GL_E_SHAPE, "need at least 2 nodes, got n=%d", n);
  if (K < 1) return fail(GL_E_SHAPE, "need at least one graph, got K=%d", K);
  if (dtype != GL_F32 && dtype != GL_F64) return fail(GL_E_ARG, "dtype must be GL_F32 or GL_F64");
  if (tile_begin < 0 || tile_end > num_tiles(n) || tile_end < tile_begin)
    return fail(GL_E_ARG, "bad tile range [%d,%d) for n=%d (%d tiles)", tile_begin, tile_end, n,
                num_tiles(n));
  // K > 1 shares ONE stop decision through spectral norms of (E, K) matrices (smooth_learning.py:158-159);
  // the exact rule forms K x K edge Grams from the local tiles only, so a batch cannot be tile-sharded
  // (config 4 is 16 MB of state: it runs on one GPU, "replicas only" when more are offered).
  if (K > 1 && (tile_begin != 0 || tile_end != num_tiles(n)))
    return fail(GL_E_ARG, "a batch of K=%d graphs cannot be tile-sharded (shared spectral stop rule); "
                          "use the full tile range on one GPU", K);
  gl_solver *s = new (std::nothrow) gl_solver();
  if (!s) return fail(GL_E_ALLOC, "host allocation failed");
  GL_CUDA(cudaGetDevice(&s->device));
  s->n = n; s->K = K; s->dtype = dtype;
  s->nt = tiles_per_side(n); s->npad = s->nt * T;
  s->tile_begin = tile_begin; s->tile_end = tile_end; s->ntl = tile_end - tile_begin;
  s->esz = dtype == GL_F32 ? 4 : 8;
  s->gstride = (int64_t)s->ntl * TT;
  const size_t ebytes = (size_t)K * (size_t)s->gstride * s->esz;
  const size_t nk = (size_t)K * s->npad;
  const size_t KK = (size_t)K * K;
  s->xch_len = (int64_t)(2 * nk + 2 * KK);
  int rc = GL_OK;
  if (rc == GL_OK) rc = alloc_zero(&s->w, ebytes);
  if (rc == GL_OK) rc = alloc_zero(&s->z, ebytes);
  if (rc == GL_OK && K > 1) rc = alloc_zero(&s->r, ebytes);
  if (rc == GL_OK) rc = alloc_zero((void **)&s->tile_ij, sizeof(int2) * (size_t)s->ntl);
  if (rc == GL_OK) rc = alloc_zero((void **)&s->d, 8 * nk);
  if (rc == GL_OK) rc = alloc_zero((void **)&s->yb, 8 * nk);
  if (rc == GL_OK) rc = alloc_zero((void **)&s->pb, 8 * nk);
  if (rc == GL_OK) rc = alloc_zero((void **)&s->rb, 8 * nk);
  if (rc == GL_OK) rc = alloc_zero(&s->a, s->esz * nk);
  if (rc == GL_OK) rc = alloc_zero(&s->b, s->esz * nk);
  s->xch_stride = (s->xch_len + 31) / 32 * 32;
  if (rc == GL_OK) rc = alloc_zero((void **)&s->comm, 8 * (size_t)(COMM_HDR + 3 * s->xch_stride));   // 3rd buffer: small kernel
  if (rc == GL_OK && K == 1 && tile_begin == 0 && tile_end == num_tiles(n) && s->ntl <= 1024)
    rc = alloc_zero((void **)&s->small_state, 8 * 3 * (size_t)s->npad);
  if (rc == GL_OK) s->xch = s->comm + COMM_HDR;
  if (rc == GL_OK) rc = alloc_zero((void **)&s->cta_partial, 16 * (size_t)((s->ntl > 1024 ? s->ntl : 1024) + s->npad / 256 + 8));
  if (rc == GL_OK) rc = alloc_zero((void **)&s->node_partial, 16 * (nk / 256 + 2));
  if (rc == GL_OK) rc = alloc_zero((void **)&s->gprev, 8 * 2 * KK);
  if (rc == GL_OK) rc = alloc_zero((void **)&s->resid, 16);
  if (rc == GL_OK) rc = alloc_zero((void **)&s->eig_scratch, 8 * 10 * KK + 8 * K);
  if (rc == GL_OK && K > 1 && K <= 128) {  // batch node tail: one CTA per 256/K' nodes, all K graphs
    int kp = 1;
    while (kp < K) kp <<= 1;
    const size_t ctas = (size_t)s->npad / (256 / kp);
    rc = alloc_zero((void **)&s->gram_part, sizeof(double) * ctas * 2 * KK);
  }
  if (rc == GL_OK && K > 1 && K <= 16 && s->gram_part) {  // resident batched kernel (persistent_batch.cu)
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device) == cudaSuccess && sms > 0) {
      s->egram_ctas = 2 * sms;
      rc = alloc_zero((void **)&s->batch_partial, sizeof(double) * 2 * 2 * (size_t)s->ntl * K);
      int kp2 = 1;
      while (kp2 < K) kp2 <<= 1;
      const size_t chunks = (size_t)s->npad / (256 / kp2);
      if (rc == GL_OK) rc = alloc_zero((void **)&s->batch_node, sizeof(double) * (nk + chunks * 2 * K));
      if (rc == GL_OK) rc = alloc_zero((void **)&s->egram_part, sizeof(double) * (size_t)s->egram_ctas * 2 * KK);
    } else {
      cudaGetLastError();
    }
  }
  if (rc == GL_OK) rc = alloc_zero((void **)&s->flags, sizeof(int) * FLAG_COUNT);
  if (rc == GL_OK) rc = alloc_zero((void **)&s->grid_bar, 64);
  if (rc == GL_OK) rc = fill_tile_table(s);
  if (rc != GL_OK) {
    free_all(s);
    delete s;
    return rc;
  }
  *out = s;
  return GL_OK;
}

int gl_solver_destroy(gl_solver *s) {
  if (!s) return GL_OK;
  free_all(s);
  delete s;
  return GL_OK;
}

void *gl_solver_z_tiles(gl_solver *s, int k) {
  return s ? static_cast<char *>(s->z) + (size_t)k * s->gstride * s->esz : nullptr;
}
void *gl_solver_w_tiles(gl_solver *s, int k) {
  return s ? static_cast<char *>(s->w) + (size_t)k * s->gstride * s->esz : nullptr;
}

static int upload_scale(gl_solver *s, const double *z_scale_host, double **dev, cudaStream_t st) {
  *dev = nullptr;
  if (!z_scale_host) return GL_OK;
  double *p = s->eig_scratch + 10 * (size_t)s->K * s->K;
  GL_CUDA(cudaMemcpyAsync(p, z_scale_host, sizeof(double) * s->K, cudaMemcpyHostToDevice, st));
  *dev = p;
  return GL_OK;
}

int gl_solver_load(gl_solver *s, const void *z_packed, const double *z_scale_host, const void *w_packed,
                   void *stream) {
  if (!s) return fail(GL_E_ARG, "null solver");
  if (!z_packed) return fail(GL_E_ARG, "null distances");
  cudaStream_t st = as_stream(stream);
  double *sc = nullptr;
  GL_TRY(upload_scale(s, z_scale_host, &sc, st));
  GL_TRY(pack_to_tiles_impl(z_packed, s->z, s->n, s->K, s->dtype, s->tile_begin, s->tile_end, sc, st));
  if (w_packed)
    GL_TRY(pack_to_tiles_impl(w_packed, s->w, s->n, s->K, s->dtype, s->tile_begin, s->tile_end, nullptr, st));
  else
    GL_CUDA(cudaMemsetAsync(s->w, 0, (size_t)s->K * s->gstride * s->esz, st));
  s->begun = false;
  return GL_OK;
}

int gl_shard_rows(int n, int tile_begin, int tile_end, int *row_begin, int *row_end, int64_t *edge_begin,
                  int64_t *edge_end) {
  if (n < 2) return fail(GL_E_SHAPE, "need at least 2 nodes, got n=%d", n);
  if (tile_begin < 0 || tile_end > num_tiles(n) || tile_end < tile_begin) return fail(GL_E_ARG, "bad tile range");
  const int nt = tiles_per_side(n);
  auto tile_row = [&](int t) {
    int I = 0;
    while (I < nt && t >= nt - I) {
      t -= nt - I;
      ++I;
    }
    return I;
  };
  int r0 = 0, r1 = 0;
  if (tile_end > tile_begin) {
    r0 = tile_row(tile_begin) * T;
    r1 = (tile_row(tile_end - 1) + 1) * T;
    if (r1 > n) r1 = n;
  }
  auto base = [&](int64_t i) { return i * (n - 1) - i * (i - 1) / 2; };  // packed offset of row i
  if (row_begin) *row_begin = r0;
  if (row_end) *row_end = r1;
  if (edge_begin) *edge_begin = base(r0);
  if (edge_end) *edge_end = base(r1 < n ? r1 : n - 1);
  return GL_OK;
}

int gl_solver_load_rows(gl_solver *s, const void *z_rows, const double *z_scale_host, const void *w_rows,
                        void *stream) {
  if (!s) return fail(GL_E_ARG, "null solver");
  if (s->K != 1) return fail(GL_E_ARG, "gl_solver_load_rows serves a single graph");
  int64_t e0 = 0, e1 = 0;
  GL_TRY(gl_shard_rows(s->n, s->tile_begin, s->tile_end, nullptr, nullptr, &e0, &e1));
  if (e1 == e0) {   // a shard whose tiles hold no edge (e.g. the last node alone in its diagonal tile)
    const size_t bytes = (size_t)s->gstride * s->esz;
    GL_CUDA(cudaMemsetAsync(s->z, 0, bytes, as_stream(stream)));
    GL_CUDA(cudaMemsetAsync(s->w, 0, bytes, as_stream(stream)));
    s->begun = false;
    return GL_OK;
  }
  if (!z_rows) return fail(GL_E_ARG, "null distances");
  // the kernels index the packed vector from its start; only elements of the shard's rows are touched
  const char *zb = static_cast<const char *>(z_rows) - (size_t)e0 * s->esz;
  const char *wb = w_rows ? static_cast<const char *>(w_rows) - (size_t)e0 * s->esz : nullptr;
  return gl_solver_load(s, zb, z_scale_host, wb, stream);
}

int gl_solver_scale_z(gl_solver *s, const double *z_scale_host, void *stream) {
  if (!s) return fail(GL_E_ARG, "null solver");
  if (!z_scale_host || s->ntl == 0) return GL_OK;
  cudaStream_t st = as_stream(stream);
  double *sc = nullptr;
  GL_TRY(upload_scale(s, z_scale_host, &sc, st));
  dim3 grid(592, s->K);
  if (s->dtype == GL_F32)
    scale_rows_kernel<float><<<grid, 256, 0, st>>>((float *)s->z, s->gstride, sc);
  else
    scale_rows_kernel<double><<<grid, 256, 0, st>>>((double *)s->z, s->gstride, sc);
  GL_LAUNCH_CHECK();
  return GL_OK;
}

int gl_solver_begin(gl_solver *s, double alpha, double beta, double step_size, int maxit, double tol,
                    void *stream) {
  if (!s) return fail(GL_E_ARG, "null solver");
  cudaStream_t st = as_stream(stream);
  s->alpha = alpha; s->beta = beta; s->step = step_size; s->maxit = maxit; s->tol = tol;
  // smooth_learning.py:135-137 -- sqrt(2(n-1)) is the exact 2-norm of S
  s->gamma = step_size / (2.0 * beta + std::sqrt(2.0 * (double)(s->n - 1)));
  s->c = 1.0 - 2.0 * s->gamma * beta;
  s->begun = false;
  if (s->graph_exec) {  // parameters are baked into the captured launches
    cudaGraphExecDestroy(s->graph_exec);
    s->graph_exec = nullptr;
  }
  // peer mode: the buffer of the coming exchange was cleared by the previous node tail and the
  // other one may still be read by a peer -- never touch them from the host side
  if (s->world <= 1) GL_CUDA(cudaMemsetAsync(s->xch, 0, sizeof(double) * (size_t)(3 * s->xch_stride), st));
  GL_CUDA(cudaMemsetAsync(s->flags, 0, sizeof(int) * FLAG_COUNT, st));
  GL_TRY(launch_edge_init(s, st));   // local partial of S w0 and the squared norm of every graph's w0
  return GL_OK;
}

int gl_solver_begin_finish(gl_solver *s, void *stream) {
  if (!s) return fail(GL_E_ARG, "null solver");
  GL_TRY(launch_node_tail(s, true, as_stream(stream)));
  s->begun = true;
  return GL_OK;
}

int gl_solver_edge_pass(gl_solver *s, void *stream) {
  if (!s || !s->begun) return fail(GL_E_STATE, "gl_solver_begin/begin_finish must be called first");
  cudaStream_t st = as_stream(stream);
  return launch_edge_pass(s, st);
}

int gl_solver_node_tail(gl_solver *s, void *stream) {
  if (!s || !s->begun) return fail(GL_E_STATE, "gl_solver_begin/begin_finish must be called first");
  cudaStream_t st = as_stream(stream);
  GL_TRY(launch_node_tail(s, false, st));
  return launch_gram_edges(s, st);    // K > 1: a no-op unless the node tail asked for the exact norms
}

void *gl_solver_comm_ptr(gl_solver *s, int64_t *n_bytes) {
  if (!s) return nullptr;
  if (n_bytes) *n_bytes = 8 * (int64_t)(COMM_HDR + 2 * s->xch_stride);
  return s->comm;
}

int gl_solver_comm_handle(gl_solver *s, void *handle64) {
  if (!s || !handle64) return fail(GL_E_ARG, "null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  GL_CUDA(cudaIpcGetMemHandle(&h, s->comm));
  std::memcpy(handle64, &h, sizeof(h));
  return GL_OK;
}

int gl_solver_comm_connect_ptrs(gl_solver *s, int rank, int world, void *const *comm_ptrs) {
  if (!s) return fail(GL_E_ARG, "null solver");
  if (world < 1 || world > GL_MAX_PEERS || rank < 0 || rank >= world)
    return fail(GL_E_ARG, "peer exchange supports 1..%d ranks (got rank %d of %d)", GL_MAX_PEERS, rank, world);
  if (world > 1 && s->K != 1) return fail(GL_E_ARG, "peer exchange needs a single graph (K == 1)");
  if (s->graph_exec) {
    cudaGraphExecDestroy(s->graph_exec);
    s->graph_exec = nullptr;
  }
  s->peers = PeerView{};
  for (int r = 0; r < world; ++r) {
    double *base = (r == rank) ? s->comm : static_cast<double *>(comm_ptrs[r]);
    if (!base) return fail(GL_E_ARG, "null comm pointer for rank %d", r);
    s->peers.hdr[r] = reinterpret_cast<int *>(base);
    s->peers.xch[r] = base + COMM_HDR;
  }
  s->world = world;
  s->rank = rank;
  s->begun = false;
  return GL_OK;
}

int gl_solver_comm_first_nodes(gl_solver *s, const int *first_node) {
  if (!s) return fail(GL_E_ARG, "null solver");
  if (s->graph_exec) {   // the table is baked into captured launches
    cudaGraphExecDestroy(s->graph_exec);
    s->graph_exec = nullptr;
  }
  for (int r = 0; r < GL_MAX_PEERS; ++r) {
    int v = (first_node && r < s->world) ? first_node[r] : 0;
    if (v < 0 || v > s->npad) return fail(GL_E_ARG, "first_node[%d] = %d out of range", r, v);
    s->peers.first_node[r] = v;
  }
  return GL_OK;
}

int gl_solver_comm_connect(gl_solver *s, int rank, int world, const void *handles) {
  if (!s || !handles) return fail(GL_E_ARG, "null argument");
  if (world < 1 || world > GL_MAX_PEERS || rank < 0 || rank >= world)
    return fail(GL_E_ARG, "peer exchange supports 1..%d ranks (got rank %d of %d)", GL_MAX_PEERS, rank, world);
  void *ptrs[GL_MAX_PEERS] = {};
  for (int r = 0; r < world; ++r) {
    if (r == rank) continue;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, static_cast<const char *>(handles) + 64 * r, sizeof(h));
    if (s->ipc_opened[r]) {
      cudaIpcCloseMemHandle(s->ipc_opened[r]);
      s->ipc_opened[r] = nullptr;
    }
    GL_CUDA(cudaIpcOpenMemHandle(&ptrs[r], h, cudaIpcMemLazyEnablePeerAccess));
    s->ipc_opened[r] = ptrs[r];
  }
  return gl_solver_comm_connect_ptrs(s, rank, world, ptrs);
}

int gl_solver_comm_publish(gl_solver *s, void *stream) {
  if (!s) return fail(GL_E_ARG, "null solver");
  return launch_signal(s, as_stream(stream));
}

void *gl_solver_exchange_ptr(gl_solver *s, int64_t *n_doubles) {
  if (!s) return nullptr;
  if (n_doubles) *n_doubles = s->xch_len;
  return s->xch;
}

int gl_solver_run(gl_solver *s, int poll_every, void *stream) {
  if (!s || !s->begun) return fail(GL_E_STATE, "gl_solver_begin/begin_finish must be called first");
  cudaStream_t user = as_stream(stream);
  const char *env = std::getenv("GL_B200_PERSISTENT");
  const char *env_batch = std::getenv("GL_B200_BATCH");
  const bool batch = s->K > 1 && s->ntl > 0 && !(env && env[0] == '0') && !(env_batch && env_batch[0] == '0') &&
                     batch_applicable(s);
  const bool persistent = (s->K == 1 && s->ntl > 0 && !(env && env[0] == '0')) || batch;
  const char *env_small = std::getenv("GL_B200_SMALL");
  const bool small = persistent && !batch && !(env_small && env_small[0] == '0') && small_applicable(s);
  if (poll_every <= 0) poll_every = persistent ? 50 : (((int64_t)s->ntl * s->K >= 4096) ? 8 : 32);
  if (poll_every > s->maxit) poll_every = s->maxit > 0 ? s->maxit : 1;

  // private stream: the caller's stream may be the legacy default stream, which cannot be captured
  cudaStream_t st;
  cudaEvent_t ev;
  GL_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  GL_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  int rc = GL_OK;
  int *h_flags = nullptr;
  cudaEvent_t evs[2] = {nullptr, nullptr};
  auto cleanup = [&]() {
    if (h_flags) cudaFreeHost(h_flags);
    for (auto e : evs)
      if (e) cudaEventDestroy(e);
    cudaEventDestroy(ev);
    cudaStreamDestroy(st);
  };
#define RUN_CUDA(expr)                                                                           \
  do {                                                                                           \
    cudaError_t _e = (expr);                                                                     \
    if (_e != cudaSuccess) {                                                                     \
      rc = fail(GL_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__,     \
                __LINE__);                                                                       \
      cleanup();                                                                                 \
      return rc;                                                                                 \
    }                                                                                            \
  } while (0)
  RUN_CUDA(cudaEventRecord(ev, user));
  RUN_CUDA(cudaStreamWaitEvent(st, ev, 0));
  RUN_CUDA(cudaMallocHost((void **)&h_flags, sizeof(int) * 2 * FLAG_COUNT));
  RUN_CUDA(cudaEventCreateWithFlags(&evs[0], cudaEventDisableTiming));
  RUN_CUDA(cudaEventCreateWithFlags(&evs[1], cudaEventDisableTiming));

  const int per_iter = (s->ntl ? 1 : 0) + ((s->K > 1 && s->ntl) ? 1 : 0) + 1;
  if (!persistent && (!s->graph_exec || s->graph_iters != poll_every)) {
    if (s->graph_exec) {
      cudaGraphExecDestroy(s->graph_exec);
      s->graph_exec = nullptr;
    }
    cudaGraph_t graph = nullptr;
    const long long before = g_launches.load();
    RUN_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    for (int i = 0; i < poll_every && rc == GL_OK; ++i) {
      rc = launch_edge_pass(s, st);
      if (rc == GL_OK) rc = launch_node_tail(s, false, st);
      if (rc == GL_OK) rc = launch_gram_edges(s, st);
    }
    cudaError_t ce = cudaStreamEndCapture(st, &graph);
    g_launches.store(before);  // capture records, it does not launch
    if (rc != GL_OK) {
      if (graph) cudaGraphDestroy(graph);
      cleanup();
      return rc;
    }
    RUN_CUDA(ce);
    ce = cudaGraphInstantiate(&s->graph_exec, graph, 0);
    cudaGraphDestroy(graph);
    RUN_CUDA(ce);
    s->graph_iters = poll_every;
  }
  // pipelined replay: chunk g+1 is enqueued before the flags of chunk g are inspected; iterations
  // past the stop are device-side no-ops, so over-enqueueing costs only empty launches.
  const int max_graphs = (s->maxit + poll_every - 1) / poll_every;
  int launched = 0, checked = 0;
  bool done = s->maxit <= 0;
  while (!done) {
    while (launched < max_graphs && launched - checked < 2) {
      if (persistent) {
        // one resident kernel runs the whole chunk (small problems: the one-barrier redundant-node form;
        // K > 1: the batched form with the shared spectral stop rule)
        rc = batch   ? launch_persistent_batch(s, poll_every, st)
             : small ? launch_persistent_small(s, poll_every, st)
                     : launch_persistent(s, poll_every, st);
        if (rc != GL_OK) {
          cleanup();
          return rc;
        }
      } else {
        RUN_CUDA(cudaGraphLaunch(s->graph_exec, st));
        g_launches.fetch_add((long long)per_iter * poll_every);
      }
      RUN_CUDA(cudaMemcpyAsync(h_flags + (launched & 1) * FLAG_COUNT, s->flags, sizeof(int) * FLAG_COUNT,
                               cudaMemcpyDeviceToHost, st));
      RUN_CUDA(cudaEventRecord(evs[launched & 1], st));
      ++launched;
    }
    if (checked >= launched) break;
    RUN_CUDA(cudaEventSynchronize(evs[checked & 1]));
    done = h_flags[(checked & 1) * FLAG_COUNT + FLAG_DONE] != 0;
    ++checked;
  }
  RUN_CUDA(cudaStreamSynchronize(st));
  // make later work on the caller's stream ordered after ours (already complete)
  cleanup();
#undef RUN_CUDA
  return GL_OK;
}

int gl_solver_status(gl_solver *s, int *n_iter, double *rel_primal, double *rel_dual, int *converged,
                     void *stream) {
  if (!s) return fail(GL_E_ARG, "null solver");
  cudaStream_t st = as_stream(stream);
  int hf[FLAG_COUNT];
  double hr[2];
  GL_CUDA(cudaMemcpyAsync(hf, s->flags, sizeof(hf), cudaMemcpyDeviceToHost, st));
  GL_CUDA(cudaMemcpyAsync(hr, s->resid, sizeof(hr), cudaMemcpyDeviceToHost, st));
  GL_CUDA(cudaStreamSynchronize(st));
  if (hf[FLAG_ERROR]) return fail(GL_E_STATE, "peer exchange timed out waiting for another rank");
  if (n_iter) *n_iter = hf[FLAG_IT];
  if (converged) *converged = hf[FLAG_CONVERGED];
  if (rel_primal) *rel_primal = hr[0];
  if (rel_dual) *rel_dual = hr[1];
  return GL_OK;
}

int gl_solver_store(gl_solver *s, void *w_packed, double edge_tol, void *stream) {
  if (!s) return fail(GL_E_ARG, "null solver");
  if (s->ntl == 0) return GL_OK;
  cudaStream_t st = as_stream(stream);
  dim3 grid(s->ntl, s->K);
  if (s->dtype == GL_F32)
    store_packed_kernel<float><<<grid, 256, 0, st>>>((const float *)s->w, s->gstride, (float *)w_packed,
                                                    s->tile_ij, s->n, edge_tol);
  else
    store_packed_kernel<double><<<grid, 256, 0, st>>>((const double *)s->w, s->gstride, (double *)w_packed,
                                                     s->tile_ij, s->n, edge_tol);
  GL_LAUNCH_CHECK();
  return GL_OK;
}

int gl_solver_trace_enable(gl_solver *s) {
  if (!s) return fail(GL_E_ARG, "null solver");
  if (!s->trace) GL_TRY(alloc_zero((void **)&s->trace, sizeof(long long) * TRACE_CAP * TRACE_STAMPS));
  return GL_OK;
}

int gl_solver_trace_read(gl_solver *s, int64_t *stamps_host, int max_iters) {
  if (!s || !s->trace) return fail(GL_E_STATE, "tracing is not enabled on this solver");
  if (max_iters > TRACE_CAP) max_iters = TRACE_CAP;
  if (max_iters <= 0) return GL_OK;
  GL_CUDA(cudaMemcpy(stamps_host, s->trace, sizeof(long long) * TRACE_STAMPS * max_iters, cudaMemcpyDeviceToHost));
  return GL_OK;
}

int gl_solver_node_state(gl_solver *s, double *d_dev, void *stream) {
  if (!s) return fail(GL_E_ARG, "null solver");
  GL_CUDA(cudaMemcpy2DAsync(d_dev, sizeof(double) * s->n, s->d, sizeof(double) * s->npad,
                            sizeof(double) * s->n, s->K, cudaMemcpyDeviceToDevice, as_stream(stream)));
  return GL_OK;
}

int gl_logmodel_solve(const void *z_packed, const double *z_scale_host, int n, int K, int dtype,
                      double alpha, double beta, const void *edge_init_packed, int maxit, double tol,
                      double step_size, double edge_tol, void *w_packed, int *n_iter, double *rel_primal,
                      double *rel_dual, void *stream) {
  gl_solver *s = nullptr;
  GL_TRY(gl_solver_create(&s, n, K, dtype, 0, num_tiles(n)));
  int rc = gl_solver_load(s, z_packed, z_scale_host, edge_init_packed, stream);
  if (rc == GL_OK) rc = gl_solver_begin(s, alpha, beta, step_size, maxit, tol, stream);
  if (rc == GL_OK) rc = gl_solver_begin_finish(s, stream);
  if (rc == GL_OK) rc = gl_solver_run(s, 0, stream);
  if (rc == GL_OK) rc = gl_solver_status(s, n_iter, rel_primal, rel_dual, nullptr, stream);
  if (rc == GL_OK) rc = gl_solver_store(s, w_packed, edge_tol, stream);
  if (rc == GL_OK && cudaStreamSynchronize(as_stream(stream)) != cudaSuccess)
    rc = fail(GL_E_CUDA, "stream synchronize failed after solve");
  gl_solver_destroy(s);
  return rc;
}

int gl_logmodel_solve_host(const void *z_host, const double *z_scale_host, int n, int K, int dtype,
                           double alpha, double beta, const void *edge_init_host, int maxit, double tol,
                           double step_size, double edge_tol, void *w_host, int *n_iter,
                           double *rel_primal, double *rel_dual) {
  if (n < 2) return fail(GL_E_SHAPE, "need at least 2 nodes, got n=%d", n);
  if (K < 1) return fail(GL_E_SHAPE, "need at least one graph, got K=%d", K);
  if (dtype != GL_F32 && dtype != GL_F64) return fail(GL_E_ARG, "dtype must be GL_F32 or GL_F64");
  const size_t bytes = (size_t)K * (size_t)gl_num_edges(n) * (dtype == GL_F32 ? 4 : 8);
  void *zd = nullptr, *wd = nullptr;
  cudaStream_t st = nullptr;
  int rc = GL_OK;
  auto cu = [&](cudaError_t e, const char *what) {
    if (e != cudaSuccess && rc == GL_OK) rc = fail(GL_E_CUDA, "%s failed: %s", what, cudaGetErrorString(e));
  };
  cu(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking), "cudaStreamCreate");
  if (rc == GL_OK) cu(cudaMalloc(&zd, bytes), "cudaMalloc(Z)");
  if (rc == GL_OK) cu(cudaMalloc(&wd, bytes), "cudaMalloc(W)");
  if (rc == GL_OK) cu(cudaMemcpyAsync(zd, z_host, bytes, cudaMemcpyHostToDevice, st), "H2D(Z)");
  if (rc == GL_OK && edge_init_host)
    cu(cudaMemcpyAsync(wd, edge_init_host, bytes, cudaMemcpyHostToDevice, st), "H2D(edge_init)");
  if (rc == GL_OK)
    rc = gl_logmodel_solve(zd, z_scale_host, n, K, dtype, alpha, beta, edge_init_host ? wd : nullptr, maxit,
                           tol, step_size, edge_tol, wd, n_iter, rel_primal, rel_dual, st);
  if (rc == GL_OK) cu(cudaMemcpyAsync(w_host, wd, bytes, cudaMemcpyDeviceToHost, st), "D2H(W)");
  if (rc == GL_OK) cu(cudaStreamSynchronize(st), "cudaStreamSynchronize");
  if (zd) cudaFree(zd);
  if (wd) cudaFree(wd);
  if (st) cudaStreamDestroy(st);
  return rc;
}

}  // extern "C"
