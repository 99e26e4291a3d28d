// knn_sharded.cu -- stepwise kNN-support solver for row-sharded multi-GPU runs (config 5, SURVEY 8(e):
// "kNN variant: shard the edge list by contiguous ranges of the row index").
//
// Every rank owns the nodes [row_begin, row_end) and ALL directed slots whose source is one of them
// (knn_kernels.cuh), so the degree sums of its nodes are complete locally: nothing is reduced across
// ranks except four residual scalars.  What a rank lacks is the (a, b) pair of far-end nodes it does
// not own; after its node step every rank's slice of the interleaved `ab` array is all-gathered
// (8 B x n in fp32; by the caller: NCCL in-place all-gather, graph_learning_b200/sharded.py).
// Per iteration:  edge_step -> node_step -> [all-gather ab | all-reduce 4 scalars] -> decide.
// The node state (d, yb, pb) of foreign rows is never touched.
//
// Peer mode (round 2, gl_knn_comm_connect): the exchange moves INTO the one-launch iteration kernel
// (knn_fused_kernel<F, PEER = true>): a row's new pair is stored straight into the out buffer of every
// rank that references that node (gl_knn_comm_need), the residual partials and a sequence flag follow
// through peer-mapped memory, and the next launch evaluates the stop rule for the iteration before it.
// One launch per iteration per rank, no collective; the ab buffers, the scalar block and the flags live
// in one CUDA-IPC exported "comm block" per rank.
#include <cstdlib>
#include <cstring>

#include <cub/device/device_scan.cuh>

#include "knn_kernels.cuh"

struct gl_knn_solver {
  gl_solver *s = nullptr;     // node-state workspace (no tiles)
  int dtype = GL_F32;
  int64_t nnz = 0, nslots = 0;
  int row_begin = 0, row_end = 0;
  int64_t ab_len = 0;          // entries of the ab array (>= npad; world x rows per rank)
  // SELL-32 arrays of the own rows' directed slots (knn_kernels.cuh)
  int32_t *dst = nullptr, *eid_canon = nullptr;
  void *zs = nullptr, *ws = nullptr;
  long long *slice_off = nullptr;   // [nslices + 1]
  int nslices = 0;
  long long total = 0;              // entries incl. padding
  int fused = 0;                    // 1: single shard, the node kernel applies the stop rule itself
                                    // 2: single shard, whole iteration in one launch (knn_fused_kernel)
                                    // 3: row shard in PEER mode: one launch per iteration, exchange over peer memory
  // comm block (one allocation, CUDA-IPC exportable): [header 512 B | scalar block 2 x 8 x 4 doubles |
  // (a, b) buffer 0 | (a, b) buffer 1]; s->ab = buffer 0, ab2 = buffer 1
  char *comm = nullptr;
  size_t ab_bytes = 0;
  int world = 1, rank = 0;
  char *peer_comm[GL_MAX_PEERS] = {};
  void *ipc_opened[GL_MAX_PEERS] = {};
  double *ring = nullptr;           // fused == 3: [2][2] denominators of the previous two iterates
  int32_t *need = nullptr;          // fused == 3, optional: per node, bit r = rank r references it (gl_knn_comm_need)
  void *ab2 = nullptr;              // second (a, b) buffer, roles swap every iteration (fused >= 2)
  double *part = nullptr;           // fused == 2: [CTAs][4] residual partials
  int parity = 0;                   // fused == 2: which buffer the next iteration READS
  double *scal = nullptr;      // [4] partial sums for the exchange
};

namespace gl {

template <typename F>
static int knn_create_t(gl_knn_solver *k, const int32_t *ei, const int32_t *ej, const F *z, const F *w0,
                        const int32_t *okey, int window, cudaStream_t st) {
  const int64_t nnz = k->nnz, all = 2 * nnz;
  const int n = k->s->n;
  if (nnz == 0) return GL_OK;
  int32_t *key_in = nullptr, *val_in = nullptr, *src_all = nullptr, *eid_all = nullptr;
  int32_t *lab = nullptr, *deg = nullptr, *deg_sorted = nullptr, *ids = nullptr, *order = nullptr;
  int32_t *okey_sorted = nullptr, *arr = nullptr;
  unsigned long long *count = nullptr, *wkey = nullptr, *wkey_sorted = nullptr;
  void *tmp = nullptr, *tmp0 = nullptr;
  F *w_zero = nullptr;
  int rc = GL_OK;
  auto cleanup = [&]() {
    for (void *q : {(void *)key_in, (void *)val_in, (void *)src_all, (void *)eid_all, (void *)count, tmp, (void *)w_zero,
                    (void *)lab, (void *)deg, (void *)deg_sorted, (void *)ids, (void *)order, tmp0,
                    (void *)okey_sorted, (void *)arr, (void *)wkey, (void *)wkey_sorted})
      if (q) cudaFree(q);
  };
#define KS_CUDA(expr)                                                                                      \
  do {                                                                                                     \
    cudaError_t _e = (expr);                                                                               \
    if (_e != cudaSuccess) {                                                                               \
      cleanup();                                                                                           \
      return fail(GL_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
    }                                                                                                      \
  } while (0)
  const size_t ib = sizeof(int32_t) * (size_t)all;
  KS_CUDA(cudaMalloc((void **)&key_in, ib));
  KS_CUDA(cudaMalloc((void **)&val_in, ib));
  KS_CUDA(cudaMalloc((void **)&src_all, ib));
  KS_CUDA(cudaMalloc((void **)&eid_all, ib));
  KS_CUDA(cudaMalloc((void **)&count, sizeof(unsigned long long)));
  KS_CUDA(cudaMemsetAsync(count, 0, sizeof(unsigned long long), st));
  const int grid = (int)std::min<int64_t>((nnz + KNN_THREADS - 1) / KNN_THREADS, 148 * 8);
  if (k->row_begin == 0 && k->row_end >= k->s->npad) {
    // single shard owning every row: renumber the nodes by decreasing degree (SELL padding -> ~0)
    const size_t nb = sizeof(int32_t) * ((size_t)n + 1);
    KS_CUDA(cudaMalloc((void **)&lab, nb));
    KS_CUDA(cudaMalloc((void **)&deg, nb));
    KS_CUDA(cudaMalloc((void **)&deg_sorted, nb));
    KS_CUDA(cudaMalloc((void **)&ids, nb));
    KS_CUDA(cudaMalloc((void **)&order, nb));
    KS_CUDA(cudaMemsetAsync(deg, 0, nb, st));
    const int ngrid = std::min((n + KNN_THREADS - 1) / KNN_THREADS, 148 * 8);
    knn_degree_count_kernel<<<grid, KNN_THREADS, 0, st>>>(ei, ej, nnz, deg);
    knn_iota_kernel<<<ngrid, KNN_THREADS, 0, st>>>(ids, n);
    size_t t0b = 0;
    if (!okey && window == 0) {
      KS_CUDA(cub::DeviceRadixSort::SortPairsDescending(nullptr, t0b, deg, deg_sorted, ids, order, n, 0, 32, st));
      KS_CUDA(cudaMalloc(&tmp0, t0b ? t0b : 16));
      KS_CUDA(cub::DeviceRadixSort::SortPairsDescending(tmp0, t0b, deg, deg_sorted, ids, order, n, 0, 32, st));
    } else {
      // locality-preserving variants: arrange the nodes by the caller's order key (stable; none = the
      // caller's numbering), then sort by decreasing degree inside windows of `window` positions
      // (0: one window = global sort with the arrangement as tie order; < 0: arrangement only)
      const int32_t *arrangement = ids;
      if (okey) {
        KS_CUDA(cudaMalloc((void **)&okey_sorted, nb));
        KS_CUDA(cudaMalloc((void **)&arr, nb));
        KS_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, t0b, okey, okey_sorted, ids, arr, n, 0, 32, st));
        KS_CUDA(cudaMalloc(&tmp0, t0b ? t0b : 16));
        KS_CUDA(cub::DeviceRadixSort::SortPairs(tmp0, t0b, okey, okey_sorted, ids, arr, n, 0, 32, st));
        KS_CUDA(cudaFree(tmp0));
        tmp0 = nullptr;
        arrangement = arr;
      }
      if (window < 0) {
        KS_CUDA(cudaMemcpyAsync(order, arrangement, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToDevice, st));
      } else {
        const size_t kb = sizeof(unsigned long long) * ((size_t)n + 1);
        KS_CUDA(cudaMalloc((void **)&wkey, kb));
        KS_CUDA(cudaMalloc((void **)&wkey_sorted, kb));
        knn_window_key_kernel<<<ngrid, KNN_THREADS, 0, st>>>(arrangement, deg, n, window, wkey);
        g_launches.fetch_add(1);
        KS_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, t0b, wkey, wkey_sorted, arrangement, order, n, 0, 64, st));
        KS_CUDA(cudaMalloc(&tmp0, t0b ? t0b : 16));
        KS_CUDA(cub::DeviceRadixSort::SortPairs(tmp0, t0b, wkey, wkey_sorted, arrangement, order, n, 0, 64, st));
      }
    }
    knn_invert_kernel<<<ngrid, KNN_THREADS, 0, st>>>(order, n, lab);
    g_launches.fetch_add(3);
    KS_CUDA(cudaGetLastError());
  }
  knn_expand_kernel<<<grid, KNN_THREADS, 0, st>>>(ei, ej, nnz, key_in, val_in, k->row_begin, k->row_end, n, count, lab);
  g_launches.fetch_add(1);
  int bits = 1;
  while (bits < 31 && ((int64_t)1 << bits) <= (int64_t)n) ++bits;   // the sentinel key n must fit
  size_t tmp_bytes = 0;
  KS_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, key_in, src_all, val_in, eid_all, (int)all, 0, bits, st));
  KS_CUDA(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 16));
  KS_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, key_in, src_all, val_in, eid_all, (int)all, 0, bits, st));
  unsigned long long h_count = 0;
  KS_CUDA(cudaMemcpyAsync(&h_count, count, sizeof(h_count), cudaMemcpyDeviceToHost, st));
  KS_CUDA(cudaStreamSynchronize(st));
  k->nslots = (int64_t)h_count;
  // ---- SELL-32 build: row counts -> row_ptr, slice sizes -> slice_off, scatter -----------------------
  const int rows = k->row_end - k->row_begin;
  k->nslices = (rows + 31) / 32;
  int *cnt = nullptr, *row_ptr = nullptr;
  long long *size = nullptr;
  void *scan_tmp = nullptr;
  auto cleanup2 = [&]() {
    for (void *q : {(void *)cnt, (void *)row_ptr, (void *)size, scan_tmp})
      if (q) cudaFree(q);
  };
#define KS2_CUDA(expr)                                                                                     \
  do {                                                                                                     \
    cudaError_t _e = (expr);                                                                               \
    if (_e != cudaSuccess) {                                                                               \
      cleanup2();                                                                                          \
      cleanup();                                                                                           \
      return fail(GL_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
    }                                                                                                      \
  } while (0)
  KS2_CUDA(cudaMalloc((void **)&cnt, sizeof(int) * (size_t)(rows + 1)));
  KS2_CUDA(cudaMalloc((void **)&row_ptr, sizeof(int) * (size_t)(rows + 1)));
  KS2_CUDA(cudaMalloc((void **)&size, sizeof(long long) * (size_t)(k->nslices + 1)));
  KS2_CUDA(cudaMalloc((void **)&k->slice_off, sizeof(long long) * (size_t)(k->nslices + 1)));
  KS2_CUDA(cudaMemsetAsync(cnt, 0, sizeof(int) * (size_t)(rows + 1), st));
  KS2_CUDA(cudaMemsetAsync(size, 0, sizeof(long long) * (size_t)(k->nslices + 1), st));
  if (k->nslots > 0) {
    const int sgrid = (int)std::min<int64_t>((k->nslots + KNN_THREADS - 1) / KNN_THREADS, 148 * 8);
    knn_row_count_kernel<<<sgrid, KNN_THREADS, 0, st>>>(src_all, k->nslots, k->row_begin, cnt);
    g_launches.fetch_add(1);
  }
  size_t b1 = 0, b2 = 0;
  KS2_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, b1, cnt, row_ptr, rows + 1, st));
  KS2_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, b2, size, k->slice_off, k->nslices + 1, st));
  KS2_CUDA(cudaMalloc(&scan_tmp, std::max(b1, b2) + 16));
  KS2_CUDA(cub::DeviceScan::ExclusiveSum(scan_tmp, b1, cnt, row_ptr, rows + 1, st));
  if (k->nslices > 0) {
    knn_slice_size_kernel<<<(k->nslices + 7) / 8, 256, 0, st>>>(cnt, rows, k->nslices, size);
    g_launches.fetch_add(1);
  }
  KS2_CUDA(cub::DeviceScan::ExclusiveSum(scan_tmp, b2, size, k->slice_off, k->nslices + 1, st));
  KS2_CUDA(cudaMemcpyAsync(&k->total, k->slice_off + k->nslices, sizeof(long long), cudaMemcpyDeviceToHost, st));
  KS2_CUDA(cudaStreamSynchronize(st));
  if (k->total > 0) {
    const size_t lb = sizeof(int32_t) * (size_t)k->total, fb = sizeof(F) * (size_t)k->total;
    KS2_CUDA(cudaMalloc((void **)&k->dst, lb));
    KS2_CUDA(cudaMalloc((void **)&k->eid_canon, lb));
    KS2_CUDA(cudaMalloc(&k->zs, fb));
    KS2_CUDA(cudaMalloc(&k->ws, fb));
    KS2_CUDA(cudaMemsetAsync(k->dst, 0xFF, lb, st));          // -1: padding
    KS2_CUDA(cudaMemsetAsync(k->eid_canon, 0xFF, lb, st));
    KS2_CUDA(cudaMemsetAsync(k->zs, 0, fb, st));
    KS2_CUDA(cudaMemsetAsync(k->ws, 0, fb, st));
    const F *w_init = w0;
    if (!w_init) {
      if (!w_zero) {
        KS2_CUDA(cudaMalloc((void **)&w_zero, sizeof(F) * (size_t)nnz));
        KS2_CUDA(cudaMemsetAsync(w_zero, 0, sizeof(F) * (size_t)nnz, st));
      }
      w_init = w_zero;
    }
    const int sgrid = (int)std::min<int64_t>((k->nslots + KNN_THREADS - 1) / KNN_THREADS, 148 * 8);
    knn_sell_fill_kernel<F><<<sgrid, KNN_THREADS, 0, st>>>(ei, ej, z, w_init, k->nslots, src_all, eid_all, k->row_begin,
                                                           row_ptr, k->slice_off, k->dst, (F *)k->zs, (F *)k->ws,
                                                           k->eid_canon, lab);
    g_launches.fetch_add(1);
    KS2_CUDA(cudaGetLastError());
  }
  KS2_CUDA(cudaStreamSynchronize(st));
  cleanup2();
#undef KS2_CUDA
#undef KS_CUDA
  cleanup();
  return rc;
}

template <typename F> static int knn_slot_launch(gl_knn_solver *k, bool init, cudaStream_t st) {
  gl_solver *s = k->s;
  if (k->nslices == 0) return GL_OK;
  double *sp = s->xch, *sw = s->xch + s->npad, *gsum = s->xch + 2 * (int64_t)s->npad;
  const int grid = std::max(1, std::min((k->nslices + KNN_THREADS / 32 - 1) / (KNN_THREADS / 32), 148 * 8));
  const F c = (F)s->c, tgz = (F)(2.0 * s->gamma);
  using V = typename Vec2<F>::type;
  if (init)
    knn_sell_kernel<F, 1><<<grid, KNN_THREADS, 0, st>>>(k->dst, (const F *)k->zs, (F *)k->ws, k->slice_off, k->nslices,
                                                       k->row_begin, (const V *)s->ab, sp, sw, gsum, s->flags, c, tgz);
  else
    knn_sell_kernel<F, 0><<<grid, KNN_THREADS, 0, st>>>(k->dst, (const F *)k->zs, (F *)k->ws, k->slice_off, k->nslices,
                                                       k->row_begin, (const V *)s->ab, sp, sw, gsum, s->flags, c, tgz);
  GL_LAUNCH_CHECK();
  return GL_OK;
}

constexpr int KNN_FUSED_MAX_GRID = 148 * 8;
constexpr size_t KNN_COMM_HDR = 512, KNN_COMM_SCAL = sizeof(double) * 2 * GL_MAX_PEERS * 4;

template <typename F> static int knn_fused_launch(gl_knn_solver *k, cudaStream_t st) {
  gl_solver *s = k->s;
  using V = typename Vec2<F>::type;
  KnnFusedArgs<F> A;
  A.dst = k->dst;
  A.z = (const F *)k->zs;
  A.w = (F *)k->ws;
  A.slice_off = k->slice_off;
  A.nslices = k->nslices;
  A.n = s->n;
  A.maxit = s->maxit;
  A.ab_in = (const V *)(k->parity ? k->ab2 : s->ab);
  A.ab_out = (V *)(k->parity ? s->ab : k->ab2);
  A.d = s->d; A.yb = s->yb; A.pb = s->pb;
  A.part = k->part;
  A.gprev = s->gprev;
  A.resid = s->resid;
  A.flags = s->flags;
  A.c = (F)s->c;
  A.tgz = (F)(2.0 * s->gamma);
  A.gamma = s->gamma;
  A.four_ag = 4.0 * s->alpha * s->gamma;
  A.tol = s->tol;
  static const int per_sm = [] {
    const char *e = getenv("GL_KNN_CTAS_PER_SM");   // measurement aid; default = the resident count (one wave)
    return e ? atoi(e) : 0;
  }();
  const int resident = 148 * (per_sm > 0 ? per_sm : (sizeof(F) == 4 ? 4 : 2));
  const int want = (k->nslices + KNN_THREADS / 32 - 1) / (KNN_THREADS / 32);
  const int grid = std::max(1, std::min(want, std::min(KNN_FUSED_MAX_GRID, resident)));
  A.world = k->world;
  A.rank = k->rank;
  A.row_begin = k->row_begin;
  A.hdr = reinterpret_cast<int *>(k->comm);
  A.scal = reinterpret_cast<const double *>(k->comm + KNN_COMM_HDR);
  A.ring = k->ring;
  A.need = k->need;
  for (int r = 0; r < GL_MAX_PEERS; ++r) {
    char *base = (r == k->rank || r >= k->world) ? k->comm : k->peer_comm[r];
    A.hdr_peer[r] = reinterpret_cast<int *>(base);
    A.scal_peer[r] = reinterpret_cast<double *>(base + KNN_COMM_HDR);
    A.ab_out_peer[r] = reinterpret_cast<V *>(base + KNN_COMM_HDR + KNN_COMM_SCAL + (k->parity ? 0 : k->ab_bytes));
  }
  // (U, MINB) = (4, 4) is the measured best at n = 1e6, k = 20 on B200 (ms per iteration, tools/knn_probe.py,
  // profiles/r2_knn_probe.jsonl): (4, 4) 0.085 | (6, 3) 0.087 | (8, 2) 0.095 | (8, 3, spills) 0.097 | (2, 6) 0.107
  if (k->fused == 3)
    knn_fused_kernel<F, true><<<grid, KNN_THREADS, 0, st>>>(A);
  else
    knn_fused_kernel<F, false><<<grid, KNN_THREADS, 0, st>>>(A);
  GL_LAUNCH_CHECK();
  k->parity ^= 1;
  return GL_OK;
}

template <typename F> static int knn_node_launch(gl_knn_solver *k, bool init, cudaStream_t st) {
  gl_solver *s = k->s;
  KnnNodeArgs A;
  A.d = s->d; A.yb = s->yb; A.pb = s->pb; A.rb = s->rb;
  A.xch = s->xch; A.gprev = s->gprev; A.resid = s->resid; A.node_partial = s->node_partial;
  A.ab = s->ab;
  A.flags = s->flags;
  A.n = s->n; A.npad = s->npad; A.init = init ? 1 : 0; A.maxit = s->maxit;
  A.row_begin = k->row_begin; A.row_end = k->row_end; A.scal = k->scal;
  A.fused = k->fused == 3 ? 0 : k->fused;   // peer mode: the init node step hands its partials to the exchange
  A.gamma = s->gamma;
  A.four_ag = 4.0 * s->alpha * s->gamma;
  A.tol = s->tol;
  const int rows = k->row_end - k->row_begin;
  int grid = (rows + KNN_NODE_THREADS - 1) / KNN_NODE_THREADS;
  grid = std::max(1, std::min(grid, 148 * 2));
  knn_node_kernel<F><<<grid, KNN_NODE_THREADS, 0, st>>>(A);
  GL_LAUNCH_CHECK();
  return GL_OK;
}

}  // namespace gl
using namespace gl;

extern "C" int gl_knn_max_degree(const int32_t *ei, const int32_t *ej, int64_t nnz, int n, int *dmax_host,
                                 void *stream) {
  if (n < 2 || nnz < 0 || !dmax_host) return fail(GL_E_ARG, "bad arguments");
  *dmax_host = 0;
  if (nnz == 0) return GL_OK;
  cudaStream_t st = as_stream(stream);
  int *cnt = nullptr;
  GL_CUDA(cudaMalloc((void **)&cnt, sizeof(int) * ((size_t)n + 1)));
  cudaMemsetAsync(cnt, 0, sizeof(int) * ((size_t)n + 1), st);
  const int grid = (int)std::min<int64_t>((nnz + KNN_THREADS - 1) / KNN_THREADS, 148 * 8);
  knn_degree_count_kernel<<<grid, KNN_THREADS, 0, st>>>(ei, ej, nnz, cnt);
  knn_max_kernel<<<1, 256, 0, st>>>(cnt, n, cnt + n);
  g_launches.fetch_add(2);
  cudaError_t e = cudaMemcpyAsync(dmax_host, cnt + n, sizeof(int), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(cnt);
  if (e != cudaSuccess) return fail(GL_E_CUDA, "gl_knn_max_degree: %s", cudaGetErrorString(e));
  return GL_OK;
}

extern "C" int gl_knn_create(gl_knn_solver **out, const int32_t *ei, const int32_t *ej, const void *z, const void *w0,
                             int64_t nnz, int n, int dtype, int row_begin, int row_end, int64_t ab_len, void *stream) {
  return gl_knn_create_ordered(out, ei, ej, z, w0, nnz, n, dtype, row_begin, row_end, ab_len, nullptr, 0, 0, stream);
}

extern "C" int gl_knn_create_ordered(gl_knn_solver **out, const int32_t *ei, const int32_t *ej, const void *z,
                                     const void *w0, int64_t nnz, int n, int dtype, int row_begin, int row_end,
                                     int64_t ab_len, const int32_t *order_key, int sort_window, int layout,
                                     void *stream) {
  if (!out) return fail(GL_E_ARG, "null output");
  if (layout != 0)
    return fail(GL_E_ARG, "layout must be 0 (directed slots); the blocked single-copy layout of round 1 measured slower "
                          "on B200 (profiles/r2_knn_row_order.jsonl) and was removed");
  *out = nullptr;
  if (n < 2) return fail(GL_E_SHAPE, "need at least 2 nodes, got n=%d", n);
  if (nnz < 0 || 2 * nnz >= ((int64_t)1 << 31)) return fail(GL_E_SHAPE, "bad edge count %lld", (long long)nnz);
  if (dtype != GL_F32 && dtype != GL_F64) return fail(GL_E_ARG, "dtype must be GL_F32 or GL_F64");
  if (row_begin < 0 || row_end < row_begin) return fail(GL_E_ARG, "bad row range [%d,%d)", row_begin, row_end);
  gl_knn_solver *k = new gl_knn_solver();
  int rc = gl_solver_create(&k->s, n, 1, dtype, 0, 0);
  if (rc != GL_OK) {
    delete k;
    return rc;
  }
  k->dtype = dtype;
  k->nnz = nnz;
  k->row_begin = row_begin;
  k->row_end = std::min(row_end, k->s->npad);
  k->ab_len = std::max<int64_t>(ab_len, k->s->npad);
  const size_t esz = dtype == GL_F32 ? 4 : 8;
  cudaStream_t st = as_stream(stream);
  k->ab_bytes = 2 * esz * (size_t)k->ab_len;
  const size_t comm_bytes = KNN_COMM_HDR + KNN_COMM_SCAL + 2 * k->ab_bytes;
  cudaError_t e = cudaMalloc((void **)&k->comm, comm_bytes);
  if (e == cudaSuccess) e = cudaMemsetAsync(k->comm, 0, comm_bytes, st);
  if (e == cudaSuccess) {
    k->s->ab = k->comm + KNN_COMM_HDR + KNN_COMM_SCAL;
    k->ab2 = k->comm + KNN_COMM_HDR + KNN_COMM_SCAL + k->ab_bytes;
  }
  if (e == cudaSuccess) e = cudaMalloc((void **)&k->scal, 4 * sizeof(double));
  if (e == cudaSuccess) e = cudaMemsetAsync(k->scal, 0, 4 * sizeof(double), st);
  if (e != cudaSuccess) {
    gl_knn_destroy(k);
    return fail(GL_E_ALLOC, "gl_knn_create: %s", cudaGetErrorString(e));
  }
  rc = dtype == GL_F32 ? knn_create_t<float>(k, ei, ej, (const float *)z, (const float *)w0, order_key, sort_window, st)
                       : knn_create_t<double>(k, ei, ej, (const double *)z, (const double *)w0, order_key, sort_window, st);
  if (rc != GL_OK) {
    gl_knn_destroy(k);
    return rc;
  }
  *out = k;
  return GL_OK;
}

extern "C" int gl_knn_destroy(gl_knn_solver *k) {
  if (!k) return GL_OK;
  for (void *q : k->ipc_opened)
    if (q) cudaIpcCloseMemHandle(q);
  for (void *q : {(void *)k->dst, (void *)k->eid_canon, k->zs, k->ws, (void *)k->slice_off, (void *)k->scal,
                  (void *)k->part, (void *)k->ring, (void *)k->need, (void *)k->comm})
    if (q) cudaFree(q);
  if (k->s) {
    k->s->ab = nullptr;   // lives inside the comm block freed above
    gl_solver_destroy(k->s);
  }
  delete k;
  return GL_OK;
}

extern "C" int64_t gl_knn_local_slots(gl_knn_solver *k) { return k ? k->nslots : -1; }
extern "C" int64_t gl_knn_padded_slots(gl_knn_solver *k) { return k ? (int64_t)k->total : -1; }

extern "C" int gl_knn_begin(gl_knn_solver *k, double alpha, double beta, double step_size, int maxit, double tol,
                            int dmax, void *stream) {
  if (!k) return fail(GL_E_ARG, "null solver");
  gl_solver *s = k->s;
  cudaStream_t st = as_stream(stream);
  s->alpha = alpha; s->beta = beta; s->step = step_size; s->maxit = maxit; s->tol = tol;
  s->gamma = step_size / (2.0 * beta + std::sqrt(2.0 * (double)dmax));
  s->c = 1.0 - 2.0 * s->gamma * beta;
  GL_CUDA(cudaMemsetAsync(s->xch, 0, sizeof(double) * (size_t)s->xch_len, st));
  GL_CUDA(cudaMemsetAsync(s->flags, 0, sizeof(int) * FLAG_COUNT, st));
  if (k->dtype == GL_F32) {
    GL_TRY(knn_slot_launch<float>(k, true, st));
    GL_TRY(knn_node_launch<float>(k, true, st));
  } else {
    GL_TRY(knn_slot_launch<double>(k, true, st));
    GL_TRY(knn_node_launch<double>(k, true, st));
  }
  s->begun = true;
  k->parity = 0;  // the init node step wrote buffer 0
  return GL_OK;   // caller: exchange ab / scalars, then gl_knn_decide(k, 1, stream)
}

extern "C" int gl_knn_edge_step(gl_knn_solver *k, void *stream) {
  if (!k || !k->s->begun) return fail(GL_E_STATE, "gl_knn_begin first");
  if (k->fused >= 2)
    return k->dtype == GL_F32 ? knn_fused_launch<float>(k, as_stream(stream))
                              : knn_fused_launch<double>(k, as_stream(stream));
  return k->dtype == GL_F32 ? knn_slot_launch<float>(k, false, as_stream(stream))
                            : knn_slot_launch<double>(k, false, as_stream(stream));
}

extern "C" int gl_knn_node_step(gl_knn_solver *k, void *stream) {
  if (!k || !k->s->begun) return fail(GL_E_STATE, "gl_knn_begin first");
  if (k->fused >= 2) return GL_OK;   // the node update ran inside the edge step
  return k->dtype == GL_F32 ? knn_node_launch<float>(k, false, as_stream(stream))
                            : knn_node_launch<double>(k, false, as_stream(stream));
}

extern "C" int gl_knn_decide(gl_knn_solver *k, int init, void *stream) {
  if (!k || !k->s->begun) return fail(GL_E_STATE, "gl_knn_begin first");
  gl_solver *s = k->s;
  if (k->fused == 1 || k->fused == 2) return GL_OK;   // single shard: the node kernel has applied the stop rule
  if (k->fused == 3 && !init) return GL_OK;           // peer mode: every launch decides for the previous iteration
  knn_decide_kernel<<<1, 1, 0, as_stream(stream)>>>(k->scal, s->gprev, s->resid, s->flags, init, s->maxit, s->tol);
  GL_LAUNCH_CHECK();
  if (k->fused == 3)   // denominators of the first check: ring slot of "iteration -1"
    GL_CUDA(cudaMemcpyAsync(k->ring + 2, s->gprev, 2 * sizeof(double), cudaMemcpyDeviceToDevice, as_stream(stream)));
  return GL_OK;
}

extern "C" void *gl_knn_ab_ptr(gl_knn_solver *k, int64_t *n_entries) {
  if (!k) return nullptr;
  if (n_entries) *n_entries = k->ab_len;
  return k->s->ab;
}

extern "C" void *gl_knn_scalars_ptr(gl_knn_solver *k) { return k ? k->scal : nullptr; }
extern "C" const int *gl_knn_flags_ptr(gl_knn_solver *k) { return k ? k->s->flags : nullptr; }

extern "C" int gl_knn_status(gl_knn_solver *k, int *n_iter, double *rel_primal, double *rel_dual, int *converged,
                             void *stream) {
  if (!k) return fail(GL_E_ARG, "null solver");
  return gl_solver_status(k->s, n_iter, rel_primal, rel_dual, converged, stream);
}

extern "C" int gl_knn_store(gl_knn_solver *k, void *w_out, double edge_tol, void *stream) {
  if (!k) return fail(GL_E_ARG, "null solver");
  cudaStream_t st = as_stream(stream);
  if (k->total == 0) return GL_OK;
  const int grid = (int)std::min<long long>((k->total + KNN_THREADS - 1) / KNN_THREADS, 148 * 8);
  if (k->dtype == GL_F32)
    knn_sell_writeback_kernel<float><<<grid, KNN_THREADS, 0, st>>>(k->eid_canon, (const float *)k->ws, k->total, edge_tol,
                                                                  (float *)w_out);
  else
    knn_sell_writeback_kernel<double><<<grid, KNN_THREADS, 0, st>>>(k->eid_canon, (const double *)k->ws, k->total,
                                                                   edge_tol, (double *)w_out);
  GL_LAUNCH_CHECK();
  return GL_OK;
}

// single shard with the stop rule fused into the node kernel (what the one-shot solve uses)
extern "C" int gl_knn_set_fused(gl_knn_solver *k, int fused) {
  if (!k) return fail(GL_E_ARG, "null solver");
  if (fused < 0 || fused > 2) return fail(GL_E_ARG, "fused must be 0, 1 or 2");
  if (k->world > 1) return fail(GL_E_STATE, "a peer-connected shard runs only the one-launch peer iteration");
  if (fused == 2) {
    if (k->row_begin != 0 || k->row_end != k->s->npad)
      return fail(GL_E_STATE, "the one-launch iteration needs a single shard owning all rows");
    if (k->s->begun) return fail(GL_E_STATE, "choose the mode before gl_knn_begin");
    if (!k->part) GL_CUDA(cudaMalloc((void **)&k->part, sizeof(double) * 4 * KNN_FUSED_MAX_GRID));
  }
  k->fused = fused;
  return GL_OK;
}

// ---- peer (NVLink) exchange of the row-sharded solve ----------------------------------------------------
extern "C" int gl_knn_comm_handle(gl_knn_solver *k, void *handle64) {
  if (!k || !handle64) return fail(GL_E_ARG, "null argument");
  cudaIpcMemHandle_t h;
  GL_CUDA(cudaIpcGetMemHandle(&h, k->comm));
  std::memcpy(handle64, &h, sizeof(h));
  return GL_OK;
}

extern "C" int gl_knn_comm_connect(gl_knn_solver *k, int rank, int world, const void *handles) {
  if (!k || !handles) return fail(GL_E_ARG, "null argument");
  if (world < 1 || world > GL_MAX_PEERS || rank < 0 || rank >= world)
    return fail(GL_E_ARG, "peer exchange supports 1..%d ranks (got rank %d of %d)", GL_MAX_PEERS, rank, world);
  if (k->s->begun) return fail(GL_E_STATE, "connect before gl_knn_begin");
  for (int r = 0; r < world; ++r) {
    if (r == rank) continue;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, static_cast<const char *>(handles) + 64 * r, sizeof(h));
    if (k->ipc_opened[r]) {
      cudaIpcCloseMemHandle(k->ipc_opened[r]);
      k->ipc_opened[r] = nullptr;
    }
    void *p = nullptr;
    GL_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    k->ipc_opened[r] = p;
    k->peer_comm[r] = static_cast<char *>(p);
  }
  k->world = world;
  k->rank = rank;
  if (!k->part) GL_CUDA(cudaMalloc((void **)&k->part, sizeof(double) * 4 * KNN_FUSED_MAX_GRID));
  if (!k->ring) {
    GL_CUDA(cudaMalloc((void **)&k->ring, sizeof(double) * 4));
    GL_CUDA(cudaMemset(k->ring, 0, sizeof(double) * 4));
  }
  k->fused = 3;
  return GL_OK;
}

// Far ends of this shard's slots as a 0/1 mark per node (the caller sums rank r's marks shifted to bit r
// over the ranks -- any all-reduce -- and hands the result back with gl_knn_comm_need).
static __global__ void knn_mark_far_ends_kernel(const int32_t *dst, long long total, int32_t *mark) {
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (long long)gridDim.x * blockDim.x) {
    const int32_t j = dst[p];
    if (j >= 0) mark[j] = 1;
  }
}

extern "C" int gl_knn_mark_far_ends(gl_knn_solver *k, int32_t *mark_dev, void *stream) {
  if (!k || !mark_dev) return fail(GL_E_ARG, "null argument");
  cudaStream_t st = as_stream(stream);
  GL_CUDA(cudaMemsetAsync(mark_dev, 0, sizeof(int32_t) * (size_t)k->s->n, st));
  if (k->total > 0) {
    const int grid = (int)std::min<long long>((k->total + KNN_THREADS - 1) / KNN_THREADS, 148 * 8);
    knn_mark_far_ends_kernel<<<grid, KNN_THREADS, 0, st>>>(k->dst, k->total, mark_dev);
    GL_LAUNCH_CHECK();
  }
  return GL_OK;
}

extern "C" int gl_knn_comm_need(gl_knn_solver *k, const int32_t *need_dev, void *stream) {
  if (!k) return fail(GL_E_ARG, "null solver");
  if (k->fused != 3) return fail(GL_E_STATE, "gl_knn_comm_connect first");
  if (!need_dev) {   // back to pushing every pair to every rank
    if (k->need) cudaFree(k->need);
    k->need = nullptr;
    return GL_OK;
  }
  if (!k->need) GL_CUDA(cudaMalloc((void **)&k->need, sizeof(int32_t) * (size_t)k->s->n));
  GL_CUDA(cudaMemcpyAsync(k->need, need_dev, sizeof(int32_t) * (size_t)k->s->n, cudaMemcpyDeviceToDevice, as_stream(stream)));
  return GL_OK;
}
