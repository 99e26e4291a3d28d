/* gl_b200.h -- C-ABI of the B200-native LogModel hot path (libgl_b200.so).
 *
 * The reference (LTS4/graph-learning) is pure Python: it has no FFI layer.  The boundary it
 * offers for this path is a set of Python call surfaces; every entry point below sits directly
 * underneath one of them and names the reference lines it replaces (paths relative to the
 * reference tree).  INTEGRATION.md shows the ctypes stub a reference maintainer would add.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes, no torch / C++ types.
 *   - Every pointer named *_dev is a CUDA device pointer on the current device; *_host is host
 *     memory.  `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *   - dtype: GL_F32 = 0, GL_F64 = 1.  Edge arrays (w, z) are stored and computed in that type;
 *     node vectors (degrees, dual variable) and all residual sums are always float64.
 *   - Return value 0 = success.  Non-zero = error code (GL_E_*); gl_last_error() returns the
 *     message of the calling thread's last failure.  The Python layer maps GL_E_SHAPE to
 *     ValueError (the reference raises ValueError for ndim > 2, smooth_learning.py:130-131)
 *     and everything else to RuntimeError.
 *   - "packed" = the reference's edge order: row-major strict upper triangle,
 *     k(i,j) = i*n - i*(i+1)/2 + (j-i-1)  (operators.py:39-41, scipy squareform).
 *   - "tiled" = this library's resident layout: 128x128 tiles (I<=J) of the upper triangle of
 *     the node-pair grid, tiles in row-major triangle order, each tile row-major and contiguous
 *     (64 KiB fp32).  Invalid slots (i>=j or j>=n) hold 0.  gl_tiled_elems(n) gives the size.
 *   - There is NO CPU fallback: every compute entry point fails with GL_E_CUDA when no device
 *     is usable.
 */
#ifndef GL_B200_H
#define GL_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GL_F32 0
#define GL_F64 1

#define GL_OK 0
#define GL_E_SHAPE 1   /* bad sizes / ndim  -> ValueError */
#define GL_E_ARG 2     /* bad argument value */
#define GL_E_CUDA 3    /* CUDA runtime error (message has the cudaError string) */
#define GL_E_ALLOC 4
#define GL_E_STATE 5   /* call order / workspace state */

#define GL_TILE 128

#define GL_LAYOUT_PACKED 0
#define GL_LAYOUT_TILED 1

typedef struct gl_solver gl_solver; /* opaque solver workspace (one per n, K, dtype, shard) */

/* ---- library --------------------------------------------------------------------------- */
int gl_version(void);
const char *gl_last_error(void);
/* Number of kernels this library has launched in this process (all threads).  bench.py reads
 * it before/after the timed region to fill "gpu_launches".  Kernels replayed from a CUDA graph
 * are counted per replay. */
int64_t gl_launch_count(void);

/* ---- sizes / layout -------------------------------------------------------------------- */
int64_t gl_num_edges(int n);        /* n(n-1)/2 */
int gl_nodes_from_edges(int64_t e); /* inverse, smooth_learning.py:133; -1 if not triangular */
int gl_num_tiles(int n);            /* nT(nT+1)/2, nT = ceil(n/128) */
int64_t gl_tiled_elems(int n);      /* gl_num_tiles(n) * 128 * 128 */
/* Work-balanced contiguous tile range of shard `rank` out of `world` (diagonal tiles weigh
 * half).  SURVEY 8(e): each GPU owns a contiguous edge shard. */
int gl_shard_tiles(int n, int rank, int world, int *tile_begin, int *tile_end);

/* Node rows [row_begin, row_end) touched by tiles [tile_begin, tile_end), and the range
 * [edge_begin, edge_end) of the packed vector that holds those rows (contiguous: the packed order is
 * row-major).  A rank of a sharded solve needs only that slice of Z on its device
 * (gl_solver_load_rows). */
int gl_shard_rows(int n, int tile_begin, int tile_end, int *row_begin, int *row_end,
                  int64_t *edge_begin, int64_t *edge_end);

/* packed <-> tiled for K graphs stored (K, E) / (K, tiled_elems).  Tiles [tile_begin,tile_end)
 * only; `tiled_dev` points at the first tile of the shard. */
int gl_pack_to_tiles(const void *packed_dev, void *tiled_dev, int n, int K, int dtype,
                     int tile_begin, int tile_end, void *stream);
int gl_tiles_to_packed(const void *tiled_dev, void *packed_dev, int n, int K, int dtype,
                       int tile_begin, int tile_end, void *stream);
/* tiled -> dense symmetric (n,n) with zero diagonal: scipy squareform(w), the return shape of
 * the 1-D call (smooth_learning.py:171-172).  Entries below edge_tol are written as 0
 * (smooth_learning.py:169); pass -INFINITY to keep everything. */
int gl_tiles_to_square(const void *tiled_dev, void *square_dev, int n, int dtype, double edge_tol,
                       void *stream);

/* ---- edge <-> degree operator pair (smooth_learning.py:13-44, sum_squareform) ----------- */
/* S w : (K,E) packed -> (K,n) weighted degrees.  Replaces `sum_op @ edge_w`
 * (smooth_learning.py:144,148,154). */
int gl_sum_op(const void *w_packed_dev, void *deg_dev, int n, int K, int dtype, void *stream);
/* S^T d : (K,n) -> (K,E) packed, d_i + d_j per edge.  Replaces `sum_op_t @ d`
 * (smooth_learning.py:147,153). */
int gl_sum_op_adj(const void *d_dev, void *edge_packed_dev, int n, int K, int dtype, void *stream);
/* Dense-Laplacian forms (operators.py:44-94).  lap (K,n,n), w (K,E) packed. */
int gl_laplacian_from_packed(const void *w_packed_dev, void *lap_dev, int n, int K, int dtype,
                             void *stream);                 /* operators.py:44-70 */
int gl_laplacian_adjoint(const void *lap_dev, void *edge_packed_dev, int n, int K, int dtype,
                         void *stream);                     /* operators.py:73-94 */

/* ---- pairwise squared distances (smooth_learning.py:278: pdist(x.T)**2) ------------------ */
/* X is (m, n) row-major with leading dimension ldx (elements), nodes are COLUMNS.
 * row_scale (nullable, (m,)) and col_shift (nullable, (n,)) give
 *   y[t, i] = row_scale[t] * (X[t, i] - col_shift[i])
 * i.e. the responsibility-weighted, mean-centred signals of the GLMM M-step
 * (clustering/glmm.py:61-68); Z_ij = sum_t (y[t,i] - y[t,j])^2.
 * Output layout GL_LAYOUT_PACKED ((E,)) or GL_LAYOUT_TILED (tiles [tile_begin,tile_end)). */
int gl_pdist_sq(const void *X_dev, int64_t m, int n, int64_t ldx, const void *row_scale_dev,
                const void *col_shift_dev, void *Z_dev, int layout, int tile_begin, int tile_end,
                int dtype, void *stream);

/* Tensor-core form of the same quantity, float32 only (north_star (1)): the Gram
 * matrix <y_i, y_j> runs as a tcgen05 / TMA GEMM on error-compensated bf16 slices of y
 * (`terms` = 3: y0+y1+y2, 6 MMAs per k-step, float32-grade; 2: y0+y1, 3 MMAs, ~16 bits),
 * accumulated in TMEM over `chunk` samples at a time (<= 0: default 256) and folded into registers
 * by the epilogue warps; the norm epilogue |y_i|^2 + |y_j|^2 - 2<y_i,y_j> (norms in float64,
 * result clamped at 0 like a distance) is fused and writes either layout directly.
 * `variant`: 0 = precise (default): exact power-of-two-quantised leading slice (terms = 3),
 *                separate TMEM accumulators for the leading and the small cross terms, float64
 *                folds, 128-wide units, 32-sample k-blocks (SWIZZLE_64B);
 *            1 = same with 64-sample k-blocks (SWIZZLE_128B);
 *            2 / 3 = one fp32 accumulator, 256- / 128-wide units, 32-sample k-blocks (faster; the
 *                tensor core's truncating accumulation leaves a bias of ~1e-5 |y|^2);
 *            4 / 5 = same with 64-sample k-blocks (4 needs terms = 2);
 *            6 = the precise scheme as a CTA pair: tcgen05.mma.cta_group::2 (M = 256) on 2-CTA clusters,
 *                each CTA loading its own rows and half of the shared operand (terms = 3);
 *            7 = TWO fp16 slices (terms = 2) of per-node power-of-two normalised signals: exact 8-bit
 *                leading slice + 11-bit remainder, four products per k-step, rescaled exactly in the
 *                epilogue; float32-grade like variant 0 and 1.36x faster (what the Python front end uses);
 *            8 = variant 7 on CTA pairs (terms = 2; bit-identical to 7, measured slightly slower).
 * `workspace_dev` (any alignment, gl_gram_workspace_bytes(m, n, terms) bytes) receives the split,
 * transposed operand planes.  Needs an sm_100 device. */
int64_t gl_gram_workspace_bytes(int64_t m, int n, int terms);
int gl_pdist_sq_gram(const void *X_dev, int64_t m, int n, int64_t ldx, const void *row_scale_dev,
                     const void *col_shift_dev, void *Z_dev, int layout, int tile_begin, int tile_end,
                     int terms, int chunk, int variant, void *workspace_dev, int64_t workspace_bytes,
                     void *stream);

/* ---- theta heuristic (smooth_learning.py:201-219, get_theta without blocks) --------------- */
/* Z in either layout (full triangle).  Writes theta to *theta_host (synchronises).  Rows of any
 * length stream through shared memory in chunks (one chunk up to ~50,000 float32 entries). */
int gl_theta(const void *Z_dev, int layout, int n, int avg_degree, int dtype, double *theta_host,
             void *stream);
/* Sharded form: Z_tiles_dev points at the first tile of the range [tile_begin, tile_end) a rank holds
 * (e.g. gl_solver_z_tiles); cand_out_dev (n, avg_degree + 1) of `dtype` receives, per node, the
 * avg_degree + 1 smallest entries of its row found in those tiles (ascending, +inf padded; the zero
 * diagonal counts in the shard owning the row's diagonal tile).  The ranks gather their candidate
 * lists side by side, (n, world * (avg_degree + 1)), and gl_theta_dense on that matrix returns the
 * theta of the full matrix (the statistic only reads the avg_degree + 1 smallest entries of a row). */
int gl_theta_partial(const void *Z_tiles_dev, int n, int tile_begin, int tile_end, int avg_degree,
                     int dtype, void *cand_out_dev, void *stream);
/* Same statistic on a general dense (rows, cols) matrix with leading dimension ld, rows used as
 * stored (no implicit diagonal): the per-block-pair call of the block form
 * (smooth_learning.py:220-235, get_theta(sq_pdists[slice1, slice2], ...)). */
int gl_theta_dense(const void *M_dev, int rows, int cols, int64_t ld, int avg_degree, int dtype,
                   double *theta_host, void *stream);
/* Allocation-free gl_theta.  No counterpart in the reference: added
 * for callers that may not call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same arguments and result as gl_theta, but the per-row statistics
 * live in a file-scope static device array of 2 * GL_THETA_NOALLOC_MAX_NODES + 2 doubles (1 MiB, held by
 * every device that loads the library until exit) instead of a cudaMalloc'd buffer; the call makes no
 * cudaMalloc / cudaFree and calls no allocating library routine.  n > GL_THETA_NOALLOC_MAX_NODES is
 * refused with GL_E_SHAPE.  All calls on a device share the one array (it keeps the last call's
 * contents; every word a call reads it has written itself, so nothing needs clearing): calls that
 * overlap on the GPU race -- serialise them or use gl_theta. */
#define GL_THETA_NOALLOC_MAX_NODES 65536
int gl_theta_noalloc(const void *Z_dev, int layout, int n, int avg_degree, int dtype, double *theta_host,
                     void *stream);

/* ---- the solver (smooth_learning.py:47-174, gsp_learn_graph_log_degrees) ----------------- */
/* Workspace for K graphs on n nodes owning tiles [tile_begin, tile_end) of the triangle
 * (0, gl_num_tiles(n) for a single GPU).  Allocates w, z (tiled) and all node vectors.  A batch
 * (K > 1) must own the full range: its shared stop rule needs spectral norms over all edges
 * (smooth_learning.py:158-159), so it is not tile-sharded (GL_E_ARG). */
int gl_solver_create(gl_solver **out, int n, int K, int dtype, int tile_begin, int tile_end);
int gl_solver_destroy(gl_solver *s);
/* Device pointers of the resident tiled state of graph k (for zero-copy producers/consumers). */
void *gl_solver_z_tiles(gl_solver *s, int k);
void *gl_solver_w_tiles(gl_solver *s, int k);
/* Load (theta-scaled) distances / initial weights from packed (K,E) device arrays.
 * z_scale: nullable host array of K doubles multiplied into Z on load (theta per graph,
 * smooth_learning.py:289, glmm.py:79, kgraphs.py:102).  w_packed_dev NULL = zeros
 * (smooth_learning.py:140-143). */
int gl_solver_load(gl_solver *s, const void *z_packed_dev, const double *z_scale_host,
                   const void *w_packed_dev, void *stream);
/* gl_solver_load for a tile shard when only the shard's slice of the packed vector is on the device
 * (single graph): z_rows_dev / w_rows_dev point at packed element edge_begin of gl_shard_rows. */
int gl_solver_load_rows(gl_solver *s, const void *z_rows_dev, const double *z_scale_host,
                        const void *w_rows_dev, void *stream);
/* Same, when Z already sits in the solver's tiled buffer (gl_pdist_sq wrote it there). */
int gl_solver_scale_z(gl_solver *s, const double *z_scale_host, void *stream);
/* Set parameters and initialise the dual state d = S w (smooth_learning.py:135-144).
 * For a sharded solver this leaves the LOCAL partial degrees in the exchange buffer; the
 * caller all-reduces gl_solver_exchange_ptr() and then calls gl_solver_begin_finish(). */
int gl_solver_begin(gl_solver *s, double alpha, double beta, double step_size, int maxit,
                    double tol, void *stream);
int gl_solver_begin_finish(gl_solver *s, void *stream);
/* One iteration = edge pass, [all-reduce of the exchange buffer by the caller when sharded],
 * node tail.  Both are stream-ordered kernel launches, capturable in a CUDA graph; they are
 * no-ops on the device once the stop rule (smooth_learning.py:166) or maxit has fired. */
int gl_solver_edge_pass(gl_solver *s, void *stream);
int gl_solver_node_tail(gl_solver *s, void *stream);
/* Exchange buffer: float64 [K*2*npad degrees partials | K*2 residual sums], to be summed over
 * shards between edge pass and node tail.  *n_doubles receives its length. */
void *gl_solver_exchange_ptr(gl_solver *s, int64_t *n_doubles);
/* Peer (NVLink) exchange, single graph: instead of an all-reduce by the caller, the node tail of
 * every rank reads all ranks' exchange buffers directly from peer-mapped memory, synchronised by
 * sequence flags each rank's node tail publishes on entry.  Each rank exports the
 * 64-byte CUDA IPC handle of its comm block, the caller all-gathers them (any transport) and
 * connects; or passes already-mapped device pointers (same process / symmetric memory).  After
 * connecting with world > 1 the caller performs NO collective between edge pass and node tail,
 * and begin_finish directly follows begin.  world == 1 restores the plain mode. */
int gl_solver_comm_handle(gl_solver *s, void *handle64_out);
int gl_solver_comm_connect(gl_solver *s, int rank, int world, const void *handles /* world x 64 B */);
/* Optional hint after connecting: first_node[r] (world ints) = the first node rank r's tiles touch (its
 * first tile row x 128).  Rank r's degree partials below that node are exact zeros, so the persistent
 * kernel's node phase skips those peer reads (about a quarter of the NVLink traffic of the exchange on
 * 8 ranks; sums unchanged bit for bit).  NULL = read everything. */
int gl_solver_comm_first_nodes(gl_solver *s, const int *first_node);
void *gl_solver_comm_ptr(gl_solver *s, int64_t *n_bytes);
int gl_solver_comm_connect_ptrs(gl_solver *s, int rank, int world, void *const *comm_ptrs);
/* Publish this rank's flag as a stand-alone launch (the node tail also does it on entry).  Only
 * needed when several ranks are emulated on ONE stream: publish all, then run the node tails. */
int gl_solver_comm_publish(gl_solver *s, void *stream);
/* Single-GPU driver: runs iterations from an internally captured CUDA graph until the device
 * stop flag is set; polls every `poll_every` iterations (<=0: default).  Synchronises. */
int gl_solver_run(gl_solver *s, int poll_every, void *stream);
/* Read back iteration count and the last relative residuals (synchronises the stream).  For K > 1
 * rel_primal is exact on the iteration that ended the solve (stop rule or maxit); in between it is the
 * lower bound max_k ||q_k - y_k|| / ||W||_F the node tail uses to rule the stop out without forming
 * the K x K edge Grams (the stop decisions are those of the exact rule, smooth_learning.py:158-166). */
int gl_solver_status(gl_solver *s, int *n_iter, double *rel_primal, double *rel_dual,
                     int *converged, void *stream);
/* Optional phase tracing of the persistent iteration kernel (K == 1): after enabling, CTA 0 stamps
 * %globaltimer (ns) at 8 phase boundaries of every iteration: [E start, E done, after barrier 1,
 * after fold+publish, after peer-flag wait, node chunk done, after barrier 2, after decision].
 * gl_solver_trace_read copies the first max_iters x 8 stamps (iteration index modulo 4096). */
int gl_solver_trace_enable(gl_solver *s);
int gl_solver_trace_read(gl_solver *s, int64_t *stamps_host, int max_iters);
/* Results: packed (K,E) with the edge_tol threshold applied (smooth_learning.py:169); or the
 * raw node state for tests. */
int gl_solver_store(gl_solver *s, void *w_packed_dev, double edge_tol, void *stream);
int gl_solver_node_state(gl_solver *s, double *d_dev /*(K,n)*/, void *stream);

/* ---- one-shot conveniences (what the Python function calls) ------------------------------ */
/* gsp_learn_graph_log_degrees on device arrays: distances (K,E) packed -> weights (K,E). */
int gl_logmodel_solve(const void *z_packed_dev, const double *z_scale_host, int n, int K, int dtype,
                      double alpha, double beta, const void *edge_init_packed_dev, int maxit,
                      double tol, double step_size, double edge_tol, void *w_packed_dev,
                      int *n_iter, double *rel_primal, double *rel_dual, void *stream);
/* Same with HOST buffers (pageable or pinned): H2D of Z (and edge_init), solve, D2H of W.  This is
 * the entry point the reference-side binding of INTEGRATION.md calls (integration/_gl_b200.py, driven by
 * tests/test_gpu_boundary.py against the reference's goldens).  bench.py's "e2e" times the Python
 * drop-in function instead, which returns the reference's dense (n, n) shape for a 1-D call. */
int gl_logmodel_solve_host(const void *z_packed_host, const double *z_scale_host, int n, int K,
                           int dtype, double alpha, double beta,
                           const void *edge_init_packed_host, int maxit, double tol,
                           double step_size, double edge_tol, void *w_packed_host, int *n_iter,
                           double *rel_primal, double *rel_dual);

/* ---- kNN-restricted support (config 5; no reference code) ---------------- */
/* Edge list (ei < ej), nnz entries, z = theta-scaled squared distances of those edges.
 * gamma = step / (2 beta + sqrt(2 d_max)).  w_dev in: initial (or zeros), out: result. */
int gl_knn_logmodel_solve(const int32_t *ei_dev, const int32_t *ej_dev, const void *z_dev,
                          int64_t nnz, int n, int dtype, double alpha, double beta, int maxit,
                          double tol, double step_size, double edge_tol, void *w_dev, int *n_iter,
                          double *rel_primal, double *rel_dual, void *stream);

/* Stepwise form of the same solver for row-sharded multi-GPU runs ("shard the edge list
 * by contiguous ranges of the row index").  A rank owns the nodes [row_begin, row_end) and every
 * directed slot i->j with i among them, so its degree sums are complete locally.  Per iteration:
 *   gl_knn_edge_step -> gl_knn_node_step -> caller all-gathers the rank's slice of the interleaved
 *   (gamma d, gamma pb) array (gl_knn_ab_ptr: ab_len entries of 2 x edge dtype; rank r's slice
 *   starts at entry r's row_begin) and all-reduces the 4 float64 scalars of gl_knn_scalars_ptr
 *   -> gl_knn_decide (stop rule, identical on every rank).
 * gl_knn_begin performs the initial degrees and node step; the caller exchanges likewise and calls
 * gl_knn_decide(k, 1, ...).  The edge list passed to gl_knn_create may be the full list or only the
 * edges touching the rank's rows; it must stay alive until gl_knn_destroy (gl_knn_store reads it).
 * gl_knn_store writes the weights of the edges whose FIRST end the rank owns, in that list's order.
 * d_max for the step size is global: gl_knn_max_degree on the full list (or any other way). */
typedef struct gl_knn_solver gl_knn_solver;
int gl_knn_max_degree(const int32_t *ei_dev, const int32_t *ej_dev, int64_t nnz, int n, int *dmax_host,
                      void *stream);
int gl_knn_create(gl_knn_solver **out, const int32_t *ei_dev, const int32_t *ej_dev, const void *z_dev,
                  const void *w0_dev /* nullable: zeros */, int64_t nnz, int n, int dtype, int row_begin,
                  int row_end, int64_t ab_len, void *stream);
/* Same with a choice of the INTERNAL row order of a single full-range shard (ignored otherwise; the
 * caller never sees it: results are per edge of the caller's list).  By default the rows are sorted
 * by decreasing degree over the whole node set (no SELL padding, but neighbouring rows are unrelated
 * nodes, so every far-end gather of a warp touches its own sector).  order_key_dev (nullable, n
 * int32 >= 0, e.g. Morton codes of the points the kNN lists were built from): the nodes are first
 * arranged by ascending key (stable; null = the caller's numbering).  sort_window > 0: degree sort
 * inside windows of that many consecutive positions of the arrangement (SELL-C-sigma: padding stays
 * small and the far ends of a CTA's rows stay within a few KB of the (a, b) array); 0: one window
 * (global sort, the arrangement only breaks ties); < 0: the arrangement as it is.
 * layout must be 0 (every edge as two directed slots, the form above); a blocked single-copy layout
 * was built in round 1, measured slower on B200 in round 2 and removed. */
int gl_knn_create_ordered(gl_knn_solver **out, const int32_t *ei_dev, const int32_t *ej_dev,
                          const void *z_dev, const void *w0_dev, int64_t nnz, int n, int dtype,
                          int row_begin, int row_end, int64_t ab_len, const int32_t *order_key_dev,
                          int sort_window, int layout, void *stream);
int gl_knn_logmodel_solve_ordered(const int32_t *ei_dev, const int32_t *ej_dev, const void *z_dev,
                                  int64_t nnz, int n, int dtype, double alpha, double beta,
                                  int maxit, double tol, double step_size, double edge_tol,
                                  void *w_dev, int *n_iter, double *rel_primal, double *rel_dual,
                                  const int32_t *order_key_dev, int sort_window, int layout,
                                  void *stream);
int gl_knn_destroy(gl_knn_solver *k);
int64_t gl_knn_local_slots(gl_knn_solver *k);
/* entries of the SELL arrays including padding (gl_knn_local_slots counts the live ones) */
int64_t gl_knn_padded_slots(gl_knn_solver *k);
/* Single shard only.  1: the node step applies the stop rule itself (gl_knn_decide becomes a no-op).
 * 2: the whole iteration -- slot sweep, node update, stop rule -- runs inside gl_knn_edge_step as ONE
 * launch (gl_knn_node_step and gl_knn_decide become no-ops); choose before gl_knn_begin. */
int gl_knn_set_fused(gl_knn_solver *k, int fused);
/* Peer (NVLink) exchange for the row-sharded solve: each rank exports the 64-byte CUDA IPC handle of its
 * comm block (header, scalar block, both (a, b) buffers), the caller all-gathers the handles and
 * connects.  From then on gl_knn_edge_step is the WHOLE iteration in one launch: slot sweep, node update
 * of the own rows with the new (a, b) pairs stored into every rank's buffer, residual partials and a
 * sequence flag pushed to every rank; the next launch waits for all flags and applies the stop rule to
 * the iteration just completed (identical on every rank; a launch that decides "stop" leaves w
 * untouched, so iteration count and weights are those of the reference rule).  gl_knn_node_step and
 * gl_knn_decide(k, 0) become no-ops; gl_knn_begin + the caller's one initial all-gather / all-reduce +
 * gl_knn_decide(k, 1) stay as above.  Enqueue maxit + 1 launches (the last one only decides). */
int gl_knn_comm_handle(gl_knn_solver *k, void *handle64_out);
int gl_knn_comm_connect(gl_knn_solver *k, int rank, int world, const void *handles /* world x 64 B */);
/* Optional traffic cut for the peer exchange: gl_knn_mark_far_ends writes 1 at every node this shard's
 * slots reference (mark_dev: n int32); the caller combines the ranks' marks into a bit set per node
 * (bit r = rank r references it; e.g. an all-reduce SUM of mark << rank) and passes it to
 * gl_knn_comm_need on every rank.  A node's new pair is then stored only into the ranks that read it --
 * with spatially compact shards a thin boundary layer instead of every row.  NULL restores the full push. */
int gl_knn_mark_far_ends(gl_knn_solver *k, int32_t *mark_dev, void *stream);
int gl_knn_comm_need(gl_knn_solver *k, const int32_t *need_dev, void *stream);
int gl_knn_begin(gl_knn_solver *k, double alpha, double beta, double step_size, int maxit, double tol,
                 int dmax, void *stream);
int gl_knn_edge_step(gl_knn_solver *k, void *stream);
int gl_knn_node_step(gl_knn_solver *k, void *stream);
int gl_knn_decide(gl_knn_solver *k, int init, void *stream);
void *gl_knn_ab_ptr(gl_knn_solver *k, int64_t *n_entries);
void *gl_knn_scalars_ptr(gl_knn_solver *k);
int gl_knn_status(gl_knn_solver *k, int *n_iter, double *rel_primal, double *rel_dual, int *converged,
                  void *stream);
/* Device pointer of the shard's 8 status words (int): [0] = solve finished (stop rule or maxit),
 * [1] = iterations completed, [2] = stop rule fired, [5] = peer exchange timed out.  For drivers that
 * poll without blocking the stream (asynchronous copy into pinned memory between graph replays). */
const int *gl_knn_flags_ptr(gl_knn_solver *k);
int gl_knn_store(gl_knn_solver *k, void *w_out_dev, double edge_tol, void *stream);

/* k nearest neighbours between the COLUMNS of X (m, n) float32 (no reference code):
 * exact brute force on squared Euclidean distances, fused distance + selection (the n x n matrix
 * is never written).  idx_out (n, k) int32 and d2_out (n, k) float32, ascending per row, self
 * excluded, ties by smaller index; unused entries (n - 1 < k) hold -1 / +inf.  1 <= k <= 32.
 * theta for a requested degree then follows from gl_theta_dense on [0 | d2] (n, k + 1): the
 * reference statistic (smooth_learning.py:201-219) only reads the k + 1 smallest entries of a row. */
int gl_knn_build(const void *X_dev, int64_t m, int n, int64_t ldx, int k, int32_t *idx_out_dev,
                 void *d2_out_dev, void *stream);

/* Assignment step of the batched callers: per-sample quadratic forms against K
 * square matrices A_k (K, n, n), e.g. the learned Laplacians,
 *     Q[s, k] = y^T (A_k + diag_add I) y,   y = X[s, :] - shift[k, :]   (shift nullable = 0)
 * replacing the dense contractions np.einsum("ni,kij,nj->nk") of reference kgraphs.py:87-88 and the
 * eigenvector projection of glmm.py:253-274 (which equals Q - (lambda_0 + regul) D^2 with
 * diag_add = regul, U[k] = the first eigenvector and D[s, k] = U[k] . y).  Evaluated over the non-zero
 * entries of A_k only (diagonal + symmetrised upper triangle), exact for any square matrix.
 * gl_quadform_create compacts the matrices (synchronises once for the entry count); the handle owns
 * its device arrays.  gl_quadform_apply: X (m, n) row-major with leading dimension ldx, Q_out (m, K);
 * U (K, n) and D_out (m, K) go together (both nullable); labels_out (m) int32 = argmin_k Q[s, k], first
 * minimum on ties as np.argmin (kgraphs.py:113), nullable.  All arrays of the handle's dtype. */
typedef struct gl_quadform gl_quadform;
int gl_quadform_create(gl_quadform **out, const void *A_dev, int K, int n, int dtype, double diag_add,
                       void *stream);
int gl_quadform_destroy(gl_quadform *q);
int64_t gl_quadform_nnz(const gl_quadform *q);
int gl_quadform_apply(gl_quadform *q, const void *X_dev, int64_t m, int64_t ldx, const void *shift_dev,
                      const void *U_dev, void *Q_out_dev, void *D_out_dev, int32_t *labels_out_dev,
                      void *stream);

#ifdef __cplusplus
}
#endif
#endif /* GL_B200_H */
