// solver.cuh -- workspace of the LogModel primal-dual solver (opaque `gl_solver` of the C-ABI).
#pragma once
#include "common.cuh"

// Device flag words (int): see node_tail / edge_pass.
enum {
  FLAG_DONE = 0,        // stop rule or maxit fired: later launches are no-ops
  FLAG_IT = 1,          // iterations completed
  FLAG_CONVERGED = 2,   // stop rule (smooth_learning.py:166) fired
  FLAG_EDGE_TICKET = 3, // last-CTA-done counter of the edge pass
  FLAG_NODE_TICKET = 4, // last-CTA-done counter of the node tail
  FLAG_ERROR = 5,       // peer exchange timed out
  FLAG_NEED = 6,        // K > 1: the node tail could not rule the stop out from the cheap bound; the
                        // edge Gram + decide kernels evaluate the exact spectral norms this iteration
  FLAG_WBUF = 7,        // K > 1, resident kernel: which buffer holds the current iterate (0 = w, 1 = r used as second w)
  FLAG_COUNT = 8
};

// Peer (NVLink) exchange: every rank owns one "comm block" = [header | xch buffer 0 | xch buffer 1].
// Header ints: sig[r] (r < 8) = last exchange sequence number rank r has published to this rank;
// HDR_SEQ = number of exchanges this rank has consumed (buffer parity = seq & 1).
constexpr int GL_MAX_PEERS = 8;
constexpr int TRACE_CAP = 4096, TRACE_STAMPS = 8;
constexpr int COMM_HDR = 64;  // doubles (512 B)
constexpr int HDR_SEQ = 16;
constexpr int HDR_SIG2 = 8;   // sig2[r] (r < 8): last sequence number whose residual scalars rank r has published
struct PeerView {
  double *xch[GL_MAX_PEERS];  // base of every rank's xch buffer 0 (peer-mapped)
  int *hdr[GL_MAX_PEERS];     // base of every rank's header
  int first_node[GL_MAX_PEERS];  // rank r's tiles touch no node below this one: its partials there are exact zeros
                                 // and need not be read (0 = read everything; gl_solver_comm_first_nodes)
};

struct gl_solver {
  int n = 0, K = 0, dtype = 0;
  int nt = 0, npad = 0;
  int tile_begin = 0, tile_end = 0, ntl = 0;  // local tile range / count
  size_t esz = 4;
  int64_t gstride = 0;  // elements per graph in the local tiled buffers
  int device = 0;

  // edge state (edge dtype), [K][ntl*TT]
  void *w = nullptr, *z = nullptr, *r = nullptr /* K>1 only: second iterate buffer of the resident batched kernel; q - y of the last pass on the per-iteration path */;
  int2 *tile_ij = nullptr;  // [ntl] global (I, J) of each local tile
  // node state
  double *d = nullptr, *yb = nullptr, *pb = nullptr, *rb = nullptr;  // [K][npad]
  void *a = nullptr, *b = nullptr;  // [K][npad] edge dtype: gamma*d, gamma*pb
  void *ab = nullptr;               // kNN path only: [npad] interleaved (a, b) pairs
  // exchange buffer (float64): [sp K*npad | sw K*npad | G_r K*K | G_w K*K]
  double *comm = nullptr;  // the comm block; xch = comm + COMM_HDR
  double *xch = nullptr;
  int64_t xch_len = 0, xch_stride = 0;
  int world = 1, rank = 0;  // > 1 ranks: peer mode (exchange over mapped peer memory, K == 1)
  PeerView peers = {};
  void *ipc_opened[GL_MAX_PEERS] = {};
  double *cta_partial = nullptr;  // [ntl][2] per-CTA residual sums of the edge pass (K == 1)
  double *node_partial = nullptr; // [node CTAs][2] per-CTA dual residual sums of the node tail (K == 1)
  double *gram_part = nullptr;    // K > 1: [node CTAs][2][K*K] per-CTA partial node Grams (batch mapping)
  double *gprev = nullptr;        // [2][K*K]: Gram(W), Gram(D) of the previous iterate (denominators)
  double *resid = nullptr;        // [2]: rel_primal, rel_dual of the last checked iteration
  double *eig_scratch = nullptr;  // [4][2][K*K] repeated-squaring buffers
  int *flags = nullptr;           // [FLAG_COUNT]
  unsigned int *grid_bar = nullptr;  // grid barrier counter of the persistent kernel
  int persist_grid = 0;           // resident CTAs of the persistent kernel (0 = not sized yet)
  double *small_state = nullptr;  // K == 1, full range, small n: second parity of [d | yb | pb] (persistent_small_kernel)
  int small_grid_max = 0;         // resident CTAs of persistent_small_kernel (0 = not sized yet, -1 = unusable)
  double *batch_partial = nullptr; // K > 1: [2][ntl * K][2] per-item residual sums of persistent_batch_kernel
  double *batch_node = nullptr;   // K > 1: [K*npad] d of the previous iterate | [chunks][2][K] per-chunk squared norms
  double *egram_part = nullptr;   // K > 1: [egram_ctas][2][K*K] per-CTA exact edge Grams (deciding iterations)
  int egram_ctas = 0;             // CTAs egram_part was sized for (two per SM)
  int batch_grid_max = 0;         // resident CTAs of persistent_batch_kernel (0 = not sized yet, -1 = unusable)
  long long *trace = nullptr;     // optional phase timestamps of the persistent kernel: [TRACE_CAP][TRACE_STAMPS] ns

  // parameters of the current solve
  double alpha = 1, beta = 1, step = 0.5, gamma = 0, c = 0, tol = 1e-5;
  int maxit = 1000;
  bool begun = false;

  // captured iteration graph (single-GPU driver)
  cudaGraphExec_t graph_exec = nullptr;
  int graph_iters = 0;
  cudaStream_t graph_stream_tag = nullptr;
};

namespace gl {
int launch_edge_pass(gl_solver *s, cudaStream_t st);
int launch_edge_init(gl_solver *s, cudaStream_t st);
int launch_signal(gl_solver *s, cudaStream_t st);
int launch_gram_edges(gl_solver *s, cudaStream_t st);
int launch_node_tail(gl_solver *s, bool init, cudaStream_t st);
int launch_persistent(gl_solver *s, int iters, cudaStream_t st);
int launch_persistent_small(gl_solver *s, int iters, cudaStream_t st);
bool small_applicable(gl_solver *s);
int launch_persistent_batch(gl_solver *s, int iters, cudaStream_t st);
bool batch_applicable(gl_solver *s);
int fill_tile_table(gl_solver *s);
}  // namespace gl
