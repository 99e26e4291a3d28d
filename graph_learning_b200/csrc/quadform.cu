// quadform.cu -- N2 / N1: the assignment step of the batched callers.
//
//   k-Graphs   (reference kgraphs.py:87-88, :112-113): smoothness[s, k] = x_s^T L_k x_s, labels = argmin_k
//   GLMM E-step (reference glmm.py:253-274): sum_{j>=1} (lambda_j + regul) (v_j^T y)^2 with y = x_s - mu_k,
//              which equals y^T (L_k + regul I) y - (lambda_0 + regul) (v_0^T y)^2 (orthonormal v_j).
//
// The reference contracts the samples against DENSE (n, n) Laplacians (2 m n^2 K flops).  The learned
// Laplacians are sparse (a few edges per node survive the solver's threshold), so the quadratic form
// is evaluated over the non-zero entries only: q = sum_i A_ii y_i^2 + sum_{i<j} (A_ij + A_ji) y_i y_j.
// That identity holds for ANY square matrix, so the kernel does not assume a Laplacian.
//
//   gl_quadform_create : compacts K dense matrices into entry lists (i, j, value), deterministic order
//                        (row-major; warp-ballot ordered compaction, one cub scan).
//   gl_quadform_apply  : one CTA (1024 threads) per block of S samples.  The block's signals are staged
//                        transposed in shared memory (node-major, padded: lane = sample reads are
//                        conflict-free).  Entries are stored row by row, the diagonal first and flagged,
//                        so a warp streams a contiguous, entry-balanced range of rows as
//                            q += y_i * (A_ii y_i + sum_j (A_ij + A_ji) y_j)
//                        with ONE 8-byte warp-uniform entry load and ONE gather per entry (y_i is gathered
//                        once per row).  Partial sums are combined across warps in a fixed order and the
//                        argmin over k is fused into the tail.  X is read exactly once; the kernel is
//                        bound by the L1/shared-memory pipe (about 2 wavefronts per entry per 32 samples).
#include <cub/device/device_scan.cuh>

#include "common.cuh"

struct gl_quadform {
  int K, n, dtype;
  int64_t nnz;
  int32_t *rowoff;   // (K*n + 1) entry offsets; graph k owns [rowoff[k*n], rowoff[(k+1)*n])
  void *ent;         // (nnz) packed entries {column | row-start flag, value}, row-major, diagonal first
  int32_t *bounds;   // (K, QF_MAX_SLOTS + 1) entry offsets of row-aligned, entry-balanced stream shares
};

namespace gl {

constexpr int QF_THREADS = 1024;
constexpr int QF_WARPS = QF_THREADS / 32;
constexpr int QF_UNROLL = 4;
constexpr int QF_STAGE = 8;                   // row loads in flight per lane while staging
constexpr size_t QF_SMEM_CAP = 224 * 1024;   // one CTA per SM (227 KB is the hardware limit)
constexpr unsigned QF_ROW_START = 0x80000000u;
constexpr int QF_MAX_SLOTS = QF_WARPS * 8;    // S = 4: 8 slots per warp; coarser S uses every (8/R)-th bound

// packed entry: column index (bit 31 = first entry of its row, i.e. the diagonal) and value
template <typename F> struct QfEntry;
template <> struct __align__(8) QfEntry<float> {
  unsigned j;
  float v;
};
template <> struct __align__(16) QfEntry<double> {
  unsigned j, pad;
  double v;
};
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait1() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait0() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ QfEntry<float> qf_load(const QfEntry<float> *p) {
  const int2 t = __ldg(reinterpret_cast<const int2 *>(p));
  QfEntry<float> e;
  e.j = (unsigned)t.x;
  e.v = __int_as_float(t.y);
  return e;
}
__device__ __forceinline__ QfEntry<double> qf_load(const QfEntry<double> *p) {
  const int4 t = __ldg(reinterpret_cast<const int4 *>(p));
  QfEntry<double> e;
  e.j = (unsigned)t.x;
  e.pad = 0;
  e.v = __hiloint2double(t.w, t.z);
  return e;
}

// ---- compaction -------------------------------------------------------------------------------------
// one warp per (k, i) row: 1 (the diagonal, always kept) + #{j > i : A_ij + A_ji != 0}
template <typename F>
__global__ void qf_count_kernel(const F *__restrict__ A, int K, int n, int32_t *__restrict__ cnt) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= K * n) return;
  const int k = row / n, i = row % n;
  const F *Ak = A + (size_t)k * n * n;
  int c = 0;
  for (int j = i + 1 + lane; j < n; j += 32) c += (Ak[(size_t)i * n + j] + Ak[(size_t)j * n + i]) != F(0);
  c = warp_sum(c);
  if (lane == 0) cnt[row] = c + 1;
}

template <typename F>
__global__ void qf_fill_kernel(const F *__restrict__ A, int K, int n, F diag_add, const int32_t *__restrict__ rowoff,
                               QfEntry<F> *__restrict__ ent) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= K * n) return;
  const int k = row / n, i = row % n;
  const F *Ak = A + (size_t)k * n * n;
  int pos = rowoff[row];
  if (lane == 0) {
    QfEntry<F> e{};
    e.j = (unsigned)i | QF_ROW_START;
    e.v = Ak[(size_t)i * n + i] + diag_add;
    ent[pos] = e;
  }
  ++pos;
  for (int j0 = i + 1; j0 < n; j0 += 32) {
    const int j = j0 + lane;
    F v = F(0);
    if (j < n) v = Ak[(size_t)i * n + j] + Ak[(size_t)j * n + i];
    const unsigned keep = __ballot_sync(0xffffffffu, v != F(0));
    if (v != F(0)) {
      const int p = pos + __popc(keep & ((1u << lane) - 1u));
      QfEntry<F> e{};
      e.j = (unsigned)j;
      e.v = v;
      ent[p] = e;
    }
    pos += __popc(keep);
  }
}

// share t of graph k starts at the first row whose offset reaches t/QF_MAX_SLOTS of the graph's entries
__global__ void qf_bounds_kernel(const int32_t *__restrict__ rowoff, int K, int n, int32_t *__restrict__ bounds) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= K * (QF_MAX_SLOTS + 1)) return;
  const int k = x / (QF_MAX_SLOTS + 1), t = x % (QF_MAX_SLOTS + 1);
  const int32_t *ro = rowoff + (size_t)k * n;
  const int e0 = ro[0], len = ro[n] - e0;
  const int target = (int)((int64_t)len * t / QF_MAX_SLOTS);
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (ro[mid] - e0 < target) lo = mid + 1; else hi = mid;
  }
  bounds[x] = t == QF_MAX_SLOTS ? ro[n] : ro[lo];
}

// ---- evaluation ---------------------------------------------------------------------------------------
struct QfArgs {
  const void *X;
  int64_t m, ldx;
  int n, K;
  const int32_t *bounds;
  const void *ent;
  const void *shift;   // (K, n) or null
  const void *U;       // (K, n) or null
  void *Q;             // (m, K)
  void *D;             // (m, K) or null
  int32_t *labels;     // (m) or null
};

// S samples per CTA; a warp holds R = 32 / S slots of S samples each, every slot streams its own rows.
template <typename F, int S, bool SHIFT>
__global__ void __launch_bounds__(QF_THREADS, 1) quadform_kernel(QfArgs a) {
  constexpr int R = 32 / S;
  constexpr int SP = S + 1;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  F *xs = reinterpret_cast<F *>(smem_raw);                 // [n][SP]
  F *mus = xs + (size_t)a.n * SP;                          // [n]
  F *us = mus + a.n;                                       // [n]
  double *red = reinterpret_cast<double *>(us + a.n + (a.n & 1));   // [2][QF_WARPS * R][S]
  constexpr int EPP = 16 / (int)sizeof(QfEntry<F>);   // entries per 16-byte piece
  constexpr int RING = S * EPP;                       // entries per ring buffer (one piece per lane)
  QfEntry<F> *ring_base = reinterpret_cast<QfEntry<F> *>(                              // [slots][2][RING], 16-byte aligned
      (reinterpret_cast<uintptr_t>(red + 2 * QF_WARPS * R * S) + 15) & ~uintptr_t(15));
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int s = lane % S, sub = lane / S;
  const unsigned slot_mask = S == 32 ? 0xffffffffu : (((1u << (S & 31)) - 1u) << (sub * S));
  const int n = a.n, K = a.K;
  const int64_t s0 = (int64_t)blockIdx.x * S;
  const F *X = static_cast<const F *>(a.X);
  const QfEntry<F> *ent = static_cast<const QfEntry<F> *>(a.ent);

  // staging: a warp takes whole sample rows, QF_STAGE coalesced 128-byte loads in flight per lane;
  // the transposed stores have stride SP = S + 1 words (odd): conflict-free
  for (int r = warp; r < S; r += QF_WARPS) {
    const int64_t gs = s0 + r;
    if (gs < a.m) {
      const F *row = X + gs * a.ldx;
      int i = lane;
      for (; i + (QF_STAGE - 1) * 32 < n; i += QF_STAGE * 32) {
        F v[QF_STAGE];
#pragma unroll
        for (int u = 0; u < QF_STAGE; ++u) v[u] = __ldcs(row + i + 32 * u);
#pragma unroll
        for (int u = 0; u < QF_STAGE; ++u) xs[(size_t)(i + 32 * u) * SP + r] = v[u];
      }
      for (; i < n; i += 32) xs[(size_t)i * SP + r] = __ldcs(row + i);
    } else {
      for (int i = lane; i < n; i += 32) xs[(size_t)i * SP + r] = F(0);
    }
  }
  const bool dots = a.U != nullptr;

  double best = 0.0;
  int best_k = 0;
  const int slot = warp * R + sub, nslots = QF_WARPS * R;
  constexpr int BSTEP = QF_MAX_SLOTS / (QF_WARPS * R);
  int nb0 = __ldg(a.bounds + slot * BSTEP), nb1 = __ldg(a.bounds + (slot + 1) * BSTEP);
  for (int k = 0; k < K; ++k) {
    __syncthreads();   // xs ready (k = 0); previous graph's mus / us / red consumed
    if (SHIFT) {
      const F *mu = static_cast<const F *>(a.shift) + (size_t)k * n;
      for (int i = tid; i < n; i += QF_THREADS) mus[i] = mu[i];
    }
    if (dots) {
      const F *u = static_cast<const F *>(a.U) + (size_t)k * n;
      for (int i = tid; i < n; i += QF_THREADS) us[i] = u[i];
    }
    __syncthreads();
    // this slot's contiguous, row-aligned share of graph k's entry stream (precomputed bounds; the
    // next graph's pair is fetched now so that its latency is hidden)
    const int ebeg = nb0, eend = nb1;
    if (k + 1 < K) {
      nb0 = __ldg(a.bounds + (size_t)(k + 1) * (QF_MAX_SLOTS + 1) + slot * BSTEP);
      nb1 = __ldg(a.bounds + (size_t)(k + 1) * (QF_MAX_SLOTS + 1) + (slot + 1) * BSTEP);
    }
    const int abeg = ebeg & ~(EPP - 1);                       // 16-byte aligned start
    const int nbuf = eend > abeg ? (eend - abeg + RING - 1) / RING : 0;
    QfEntry<F> *ring = ring_base + (size_t)slot * 2 * RING;
    auto issue = [&](int bfr) {
      const int pe = abeg + bfr * RING + s * EPP;             // first entry of this lane's piece
      if (bfr < nbuf && pe < eend) cp_async16(ring + (bfr & 1) * RING + s * EPP, ent + pe);
      cp_async_commit();
    };
    F yi = F(0), acc[2] = {F(0), F(0)};
    issue(0);
    for (int bfr = 0; bfr < nbuf; ++bfr) {
      issue(bfr + 1);
      cp_async_wait1();
      __syncwarp(slot_mask);
      const QfEntry<F> *buf = ring + (bfr & 1) * RING;
      const int base = abeg + bfr * RING;
      int e = ebeg > base ? ebeg : base;
      const int hi = eend < base + RING ? eend : base + RING;
      for (; e + QF_UNROLL <= hi; e += QF_UNROLL) {
        QfEntry<F> en[QF_UNROLL];
        F g[QF_UNROLL];
#pragma unroll
        for (int u = 0; u < QF_UNROLL; ++u) en[u] = buf[e - base + u];
#pragma unroll
        for (int u = 0; u < QF_UNROLL; ++u) {
          const unsigned j = en[u].j & ~QF_ROW_START;
          g[u] = xs[(size_t)j * SP + s];
          if (SHIFT) g[u] -= mus[j];
        }
#pragma unroll
        for (int u = 0; u < QF_UNROLL; ++u) {
          yi = (en[u].j & QF_ROW_START) ? g[u] : yi;   // diagonal entry opens its row: y_i by select, no branch
          acc[u & 1] += (en[u].v * g[u]) * yi;
        }
      }
      for (; e < hi; ++e) {
        const QfEntry<F> en = buf[e - base];
        const unsigned j = en.j & ~QF_ROW_START;
        F g = xs[(size_t)j * SP + s];
        if (SHIFT) g -= mus[j];
        yi = (en.j & QF_ROW_START) ? g : yi;
        acc[0] += (en.v * g) * yi;
      }
      __syncwarp(slot_mask);   // the buffer is refilled two steps from now
    }
    cp_async_wait0();
    const double q = (double)acc[0] + (double)acc[1];
    red[(size_t)slot * S + s] = q;
    if (dots) {
      F d = F(0);
      for (int i = slot; i < n; i += nslots) d += us[i] * (xs[(size_t)i * SP + s] - (SHIFT ? mus[i] : F(0)));
      red[(size_t)(nslots + slot) * S + s] = (double)d;
    }
    __syncthreads();
    if (tid < S) {
      double tot = 0.0, dot = 0.0;
      for (int w = 0; w < nslots; ++w) tot += red[(size_t)w * S + tid];
      if (dots)
        for (int w = 0; w < nslots; ++w) dot += red[(size_t)(nslots + w) * S + tid];
      const int64_t gs = s0 + tid;
      if (gs < a.m) {
        static_cast<F *>(a.Q)[gs * K + k] = (F)tot;
        if (dots) static_cast<F *>(a.D)[gs * K + k] = (F)dot;
      }
      const double cmp = (double)(F)tot;   // the argmin sees the stored value
      if (k == 0 || cmp < best) {
        best = cmp;
        best_k = k;
      }
    }
  }
  if (tid < S && a.labels != nullptr && s0 + tid < a.m) a.labels[s0 + tid] = best_k;
}

template <typename F, int S> static size_t qf_smem(int n) {
  constexpr int R = 32 / S;
  return sizeof(F) * ((size_t)n * (S + 1) + 2 * (size_t)n + (n & 1)) + sizeof(double) * 2 * QF_WARPS * R * S +
         (size_t)QF_WARPS * R * 2 * S * 16 + 16;   // + the entry rings (aligned up)
}

template <typename F, int S> static int qf_launch(const QfArgs &a, cudaStream_t st) {
  const size_t smem = qf_smem<F, S>(a.n);
  const int64_t grid = (a.m + S - 1) / S;
  if (a.shift != nullptr) {
    GL_CUDA(cudaFuncSetAttribute(quadform_kernel<F, S, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    quadform_kernel<F, S, true><<<(unsigned)grid, QF_THREADS, smem, st>>>(a);
  } else {
    GL_CUDA(cudaFuncSetAttribute(quadform_kernel<F, S, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    quadform_kernel<F, S, false><<<(unsigned)grid, QF_THREADS, smem, st>>>(a);
  }
  GL_LAUNCH_CHECK();
  return GL_OK;
}

template <typename F> static int qf_apply_t(const QfArgs &a, cudaStream_t st) {
  // widest block of samples that fits (S = 32: one slot per warp, conflict-free gathers)
  if (qf_smem<F, 32>(a.n) <= QF_SMEM_CAP) return qf_launch<F, 32>(a, st);
  if (qf_smem<F, 16>(a.n) <= QF_SMEM_CAP) return qf_launch<F, 16>(a, st);
  if (qf_smem<F, 8>(a.n) <= QF_SMEM_CAP) return qf_launch<F, 8>(a, st);
  if (qf_smem<F, 4>(a.n) <= QF_SMEM_CAP) return qf_launch<F, 4>(a, st);
  return fail(GL_E_SHAPE, "n=%d: a block of 4 samples does not fit in shared memory", a.n);
}

template <typename F> static int qf_create_t(gl_quadform *q, const void *A, double diag_add, cudaStream_t st) {
  const int rows = q->K * q->n;
  int32_t *cnt = nullptr;
  void *tmp = nullptr;
  size_t tmp_bytes = 0;
  GL_CUDA(cudaMalloc(&q->rowoff, sizeof(int32_t) * ((size_t)rows + 1)));
  GL_CUDA(cudaMalloc(&cnt, sizeof(int32_t) * ((size_t)rows + 1)));
  GL_CUDA(cudaMemsetAsync(cnt, 0, sizeof(int32_t) * ((size_t)rows + 1), st));
  const int wpb = 8, grid = (rows + wpb - 1) / wpb;
  qf_count_kernel<F><<<grid, wpb * 32, 0, st>>>(static_cast<const F *>(A), q->K, q->n, cnt);
  ::gl::g_launches.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, cnt, q->rowoff, rows + 1, st);
  if (e == cudaSuccess) e = cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 1);
  if (e == cudaSuccess) e = cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, cnt, q->rowoff, rows + 1, st);
  int32_t total = 0;
  if (e == cudaSuccess) e = cudaMemcpyAsync(&total, q->rowoff + rows, sizeof(int32_t), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(cnt);
  cudaFree(tmp);
  if (e != cudaSuccess) return fail(GL_E_CUDA, "quadform compaction failed: %s", cudaGetErrorString(e));
  q->nnz = total;
  GL_CUDA(cudaMalloc(&q->ent, sizeof(QfEntry<F>) * ((size_t)total + 2)));   // + slack for the last 16-byte piece
  qf_fill_kernel<F><<<grid, wpb * 32, 0, st>>>(static_cast<const F *>(A), q->K, q->n, (F)diag_add, q->rowoff,
                                              static_cast<QfEntry<F> *>(q->ent));
  GL_LAUNCH_CHECK();
  const int nb = q->K * (QF_MAX_SLOTS + 1);
  GL_CUDA(cudaMalloc(&q->bounds, sizeof(int32_t) * (size_t)nb));
  qf_bounds_kernel<<<(nb + 255) / 256, 256, 0, st>>>(q->rowoff, q->K, q->n, q->bounds);
  GL_LAUNCH_CHECK();
  return GL_OK;
}

}  // namespace gl
using namespace gl;

extern "C" int gl_quadform_create(gl_quadform **out, const void *A_dev, int K, int n, int dtype, double diag_add,
                                  void *stream) {
  if (out == nullptr) return fail(GL_E_ARG, "out is null");
  *out = nullptr;
  if (K < 1 || n < 1) return fail(GL_E_SHAPE, "need K >= 1 matrices of n >= 1 nodes (K=%d, n=%d)", K, n);
  if ((int64_t)K * n * n > (int64_t)INT32_MAX) return fail(GL_E_SHAPE, "K*n*n = %lld entries exceed int32 offsets",
                                                           (long long)K * n * n);
  if (dtype != GL_F32 && dtype != GL_F64) return fail(GL_E_ARG, "bad dtype %d", dtype);
  gl_quadform *q = new gl_quadform{K, n, dtype, 0, nullptr, nullptr, nullptr};
  const int rc = dtype == GL_F32 ? qf_create_t<float>(q, A_dev, diag_add, as_stream(stream))
                                 : qf_create_t<double>(q, A_dev, diag_add, as_stream(stream));
  if (rc != GL_OK) {
    gl_quadform_destroy(q);
    return rc;
  }
  *out = q;
  return GL_OK;
}

extern "C" int gl_quadform_destroy(gl_quadform *q) {
  if (q == nullptr) return GL_OK;
  cudaFree(q->rowoff);
  cudaFree(q->ent);
  cudaFree(q->bounds);
  delete q;
  return GL_OK;
}

extern "C" int64_t gl_quadform_nnz(const gl_quadform *q) { return q ? q->nnz : -1; }

extern "C" int gl_quadform_apply(gl_quadform *q, const void *X_dev, int64_t m, int64_t ldx, const void *shift_dev,
                                 const void *U_dev, void *Q_dev, void *D_dev, int32_t *labels_dev, void *stream) {
  if (q == nullptr) return fail(GL_E_ARG, "quadform handle is null");
  if (m < 0 || ldx < q->n) return fail(GL_E_SHAPE, "bad signal matrix shape (m=%lld, ldx=%lld, n=%d)", (long long)m,
                                       (long long)ldx, q->n);
  if (m == 0) return GL_OK;
  if (Q_dev == nullptr) return fail(GL_E_ARG, "Q_out is null");
  if ((U_dev == nullptr) != (D_dev == nullptr)) return fail(GL_E_ARG, "U and D_out go together");
  QfArgs a{X_dev, m, ldx, q->n, q->K, q->bounds, q->ent, shift_dev, U_dev, Q_dev, D_dev, labels_dev};
  return q->dtype == GL_F32 ? qf_apply_t<float>(a, as_stream(stream)) : qf_apply_t<double>(a, as_stream(stream));
}
