// gram.cu -- K1 (tensor-core form): pairwise squared distances between the COLUMNS of X as a
// tcgen05 / TMA Gram GEMM with the norm epilogue fused, written straight into tiled or packed
// edge order (BASELINE north_star (1)).
//
//   y[t,i] = row_scale[t] * (X[t,i] - col_shift[i])         (GLMM form, clustering/glmm.py:61-68)
//   Z_ij   = |y_i|^2 + |y_j|^2 - 2 <y_i, y_j>               (= pdist(x.T)**2, smooth_learning.py:278)
//
// Precision (SURVEY 7.2 "Z precision"): the Gram form cancels on smooth signals (neighbours have
// Z_ij << |y_i|^2), and the tensor core adds into its fp32 accumulator with truncation, which biases
// same-sign sums by ~1 ulp per MMA (measured on B200: theta off by 3e-4 with a plain accumulator).
// Four measures, all measured in tools/gram_weights_probe.py:
//   * error-compensated operands: y = y0 + y1 (+ y2), each slice bf16, products expanded into the
//     cross terms of weight >= 2^-16 (2^-24): 3 (6) bf16 MMAs per k-step;
//   * exact leading slice (3-term form): y0 is a multiple of a per-node power-of-two quantum with
//     |y0 / quantum| <= 256, so the products y0_i * y0_j are integers <= 2^16 and up to 256 samples
//     accumulate EXACTLY in fp32 whatever the rounding mode (an Ozaki-style slice);
//   * separate TMEM accumulators for the leading term and for the small cross terms, so the small
//     terms are truncated at the ulp of a 2^-8 smaller sum;
//   * chunked accumulation: every `chunk` (<= 256) samples the epilogue warps fold both TMEM tiles
//     into float64 registers while the MMA pipe fills the other TMEM buffer pair.
// Result on B200 (n=5,000, m=20,000 smooth signals): |dZ| <= 5e-7 |y|^2 (the exact fp32 kernel:
// 1.4e-6), theta within 3e-8, learned weights within 2.6e-6 of the float64 solve.
// Norms |y_i|^2 are summed in float64 by the split pre-pass from the same fp32 y values.
//
// Stage 1  gram_split_kernel: X (m,n) fp32 -> TERMS bf16 planes, TRANSPOSED to (n_pad, m_pad) so
//          both GEMM operands are K-major (sample index contiguous), zero padded; partial norms.
// Stage 2  gram_kernel: persistent, one CTA per SM, warp-specialised:
//            warp 0     TMA producer   (cp.async.bulk.tensor.3d: one box = all planes of a k-block)
//            warp 1     MMA issuer     (one thread; tcgen05.mma.cta_group::1.kind::f16, M=128, N=BN)
//            warp 2     TMEM allocator
//            warps 4..  epilogue       (tcgen05.ld chunk folds; final norm epilogue + stores)
//          Variants: split accumulators + float64 folds on 128-wide units (default, precise), or one
//          fp32 accumulator on 128/256-wide units (faster, ~1e-5 |y|^2 bias).
//          Work unit = tile row I x BN/128 adjacent column tiles; units are ordered so that the
//          ~148 units in flight cover a compact block of the triangle (operand panels are shared
//          through the 126 MB L2).
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <type_traits>
#include <vector>

#include "solver.cuh"

namespace gl {

constexpr int GR_BM = 128;          // rows of a unit = one solver tile row
constexpr int GR_UMMA_K = 16;       // bf16 MMA depth
constexpr int GR_CTRL_WARPS = 4;    // producer, MMA, TMEM allocator, (spare: keeps epilogue warp%4 aligned)
constexpr int GR_SPLIT_TILE = 64;   // split pre-pass: 64 samples x 64 nodes per step
constexpr int GR_GROUP_ROWS = 8;    // tile rows per L2 rasterisation group

// ------------------------------------------------------------------------------------------------
// PTX wrappers (sm_100a)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
// Bounded wait: a protocol bug must trap (and fail the launch) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  const long long t0 = clock64();
  for (;;) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
    if (done) return;
    if (clock64() - t0 > 8000000000LL) {  // ~4 s: far beyond any legitimate wait
      printf("gram_kernel: mbarrier timeout (block %d thread %d)\n", blockIdx.x, threadIdx.x);
      __trap();
    }
  }
}

__device__ __forceinline__ void tma_load_3d(void *smem_dst, const CUtensorMap *map, uint64_t *bar, int c0,
                                            int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t *slot_in_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot_in_smem)),
               "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem], bf16 x bf16 -> fp32, issued by ONE thread for the CTA.
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// mbarrier arrives once all MMAs issued so far by this thread have completed.
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// 32 lanes x 16 consecutive fp32 columns: thread t of the warp reads lane (lane_base + t).
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float *v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// Same for 8 columns (keeps the staging registers of the two-tile fold small).
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float *v) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}

// K-major shared-memory matrix descriptor (cute::UMMA::SmemDescriptor, mma_sm100_desc.hpp): rows of
// ROW_BYTES (= swizzle span) bytes, 8-row groups SBO = 8*ROW_BYTES apart, sm_100 version bit.
template <int ROW_BYTES>
__device__ __forceinline__ uint64_t smem_desc_kmajor(uint32_t saddr) {
  static_assert(ROW_BYTES == 64 || ROW_BYTES == 128, "swizzle span");
  constexpr uint64_t layout = ROW_BYTES == 128 ? 2 /*SWIZZLE_128B*/ : 4 /*SWIZZLE_64B*/;
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);                 // start address, bits [0,14)
  d |= (uint64_t)1 << 16;                                  // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)((8 * ROW_BYTES) >> 4) << 32;             // stride byte offset, bits [32,46)
  d |= (uint64_t)1 << 46;                                  // descriptor version (Blackwell)
  d |= layout << 61;                                       // layout type, bits [61,64)
  return d;
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): D=F32, A=B=BF16, both K-major, M=128, N.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int n_dim) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n_dim >> 3) << 17) | ((uint32_t)(GR_BM >> 4) << 24);
}

// Same with A = B = F16 (format code 0): the two-slice normalised form (gram_split16_kernel).
__host__ __device__ constexpr uint32_t umma_idesc_f16(int n_dim) {
  return (1u << 4) | ((uint32_t)(n_dim >> 3) << 17) | ((uint32_t)(GR_BM >> 4) << 24);
}
// Power-of-two scale exponent of a node from its column maximum (float bits): 2^e > max_t |y_ti|.
__host__ __device__ inline int gram_node_exponent(int colmax_bits) {
  float mx;
#ifdef __CUDA_ARCH__
  mx = __int_as_float(colmax_bits);
#else
  memcpy(&mx, &colmax_bits, sizeof(mx));
#endif
  int e = 0;
  if (mx > 0.f) frexpf(mx, &e);   // mx = f * 2^e, f in [0.5, 1)
  return e < -100 ? -100 : (e > 127 ? 127 : e);
}

// ------------------------------------------------------------------------------------------------
// Stage 1: split + transpose + norms
// ------------------------------------------------------------------------------------------------
// Column maxima max_t |y[t,i]| (as float bit patterns: non-negative floats order like ints), for the
// exact leading slice.  grid (n_pad/64, S); block 256; colmax zero-initialised by the caller.
__global__ void __launch_bounds__(256)
gram_colmax_kernel(const float *__restrict__ X, int64_t m, int n, int64_t ldx, const float *__restrict__ row_scale,
                   const float *__restrict__ col_shift, int *__restrict__ colmax, int64_t seg_len) {
  __shared__ float s_max[4][GR_SPLIT_TILE];
  const int tid = threadIdx.x, lx = tid & 63, ly = tid >> 6;
  const int gnode = blockIdx.x * GR_SPLIT_TILE + lx;
  const int64_t t_begin = (int64_t)blockIdx.y * seg_len;
  const int64_t t_end = min(t_begin + seg_len, m);
  float mx = 0.f;
  if (gnode < n) {
    const float sh = col_shift ? col_shift[gnode] : 0.f;
    for (int64_t tt = t_begin + ly; tt < t_end; tt += 4) {
      const float sc = row_scale ? row_scale[tt] : 1.f;
      mx = fmaxf(mx, fabsf(sc * (X[tt * ldx + gnode] - sh)));
    }
  }
  s_max[ly][lx] = mx;
  __syncthreads();
  if (tid < GR_SPLIT_TILE && gnode < n) {
    mx = fmaxf(fmaxf(s_max[0][tid], s_max[1][tid]), fmaxf(s_max[2][tid], s_max[3][tid]));
    if (mx == mx && mx > 0.f) atomicMax(colmax + gnode, __float_as_int(fminf(mx, 3.0e38f)));
  }
}

// grid (n_pad/64, S); block 256.  planes[p][node][sample] bf16, leading dimension m_pad.
// EXACT0: the leading slice y0 is rounded to a multiple of the per-node quantum 2^(e_i - 8)
// (2^e_i > max_t |y_ti|), so y0 = k * quantum with |k| <= 256: products of leading slices are
// integers <= 2^16 in units of quantum_i * quantum_j and up to 256 of them sum EXACTLY in the fp32
// TMEM accumulator, whatever its rounding mode (Ozaki-style slice).  The residual goes to the
// ordinary bf16 slices y1 (, y2), whose accumulator is 2^-8 smaller.
template <int TERMS, bool EXACT0>
__global__ void __launch_bounds__(256)
gram_split_kernel(const float *__restrict__ X, int64_t m, int n, int64_t ldx, const float *__restrict__ row_scale,
                  const float *__restrict__ col_shift, __nv_bfloat16 *__restrict__ planes, int64_t m_pad, int n_pad,
                  double *__restrict__ norm_part, int64_t seg_len, const int *__restrict__ colmax) {
  __shared__ float tile[GR_SPLIT_TILE][GR_SPLIT_TILE + 1];
  __shared__ double s_norm[4][GR_SPLIT_TILE];
  const int tid = threadIdx.x;
  const int node0 = blockIdx.x * GR_SPLIT_TILE;
  const int64_t t_begin = (int64_t)blockIdx.y * seg_len;
  const int64_t t_end = min(t_begin + seg_len, m_pad);
  // loader mapping: column lx (node), rows ly, ly+4, ...
  const int lx = tid & 63, ly = tid >> 6;
  const int gnode = node0 + lx;
  const float sh = (col_shift && gnode < n) ? col_shift[gnode] : 0.f;
  // writer mapping: node wx, 16 consecutive samples starting at wy*16
  const int wx = tid >> 2, wy = tid & 3;
  const int64_t plane_stride = (int64_t)n_pad * m_pad;
  double nsum = 0.0;
  float q_up = 1.f, q_down = 1.f;   // 1/quantum, quantum (powers of two) of the writer's node
  if (EXACT0) {
    const float mx = __int_as_float(colmax[node0 + wx]);
    int e = 0;
    if (mx > 0.f) frexpf(mx, &e);   // mx = f * 2^e, f in [0.5, 1)
    e = max(-100, min(e, 127));
    q_up = ldexpf(1.f, 8 - e);
    q_down = ldexpf(1.f, e - 8);
  }

  for (int64_t t0 = t_begin; t0 < t_end; t0 += GR_SPLIT_TILE) {
#pragma unroll
    for (int r = 0; r < GR_SPLIT_TILE / 4; ++r) {
      const int row = ly + 4 * r;
      const int64_t tt = t0 + row;
      float y = 0.f;
      if (tt < m && gnode < n) {
        const float sc = row_scale ? row_scale[tt] : 1.f;
        y = sc * (X[tt * ldx + gnode] - sh);
      }
      tile[row][lx] = y;
      nsum += (double)y * (double)y;
    }
    __syncthreads();
    float v[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) v[e] = tile[wy * 16 + e][wx];
#pragma unroll
    for (int p = 0; p < TERMS; ++p) {
      uint32_t packed[8];
#pragma unroll
      for (int e = 0; e < 16; e += 2) {
        const bool quant = EXACT0 && p == 0;
        const __nv_bfloat16 b0 = __float2bfloat16_rn(quant ? rintf(v[e] * q_up) * q_down : v[e]);
        const __nv_bfloat16 b1 = __float2bfloat16_rn(quant ? rintf(v[e + 1] * q_up) * q_down : v[e + 1]);
        v[e] -= __bfloat162float(b0);        // exact: the residual fits fp32
        v[e + 1] -= __bfloat162float(b1);
        packed[e >> 1] = (uint32_t)__bfloat16_as_ushort(b0) | ((uint32_t)__bfloat16_as_ushort(b1) << 16);
      }
      uint4 *dst = reinterpret_cast<uint4 *>(planes + p * plane_stride + (int64_t)(node0 + wx) * m_pad + t0 + wy * 16);
      dst[0] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
      dst[1] = make_uint4(packed[4], packed[5], packed[6], packed[7]);
    }
    __syncthreads();
  }
  s_norm[ly][lx] = nsum;
  __syncthreads();
  if (tid < GR_SPLIT_TILE)
    norm_part[(int64_t)blockIdx.y * n_pad + node0 + tid] =
        ((s_norm[0][tid] + s_norm[1][tid]) + s_norm[2][tid]) + s_norm[3][tid];
}

// Two-slice fp16 form (variant 7).  Every node is normalised by its own power of two, yh = y * 2^-e_i with
// |yh| < 1, and cut into  yh0 = k / 256 (|k| <= 256: exact in fp16; products of leading slices are integers
// <= 2^16 in units of 2^-16, so 256 of them sum EXACTLY in the fp32 TMEM accumulator, as in the bf16 form)
// and yh1 = fp16(yh - yh0), |yh1| <= 2^-9, whose 11 significant bits leave an absolute error <= 2^-21 --
// fp16 carries 3 more bits per slice than bf16, which is what lets two slices replace three.  The GEMM then
// needs the four products (0,0) | (0,1) (1,0) (1,1) instead of six, and the epilogue scales the normalised
// Gram entry back by 2^(e_i + e_j) (exact).  fp16's narrow exponent range is why the planes are
// normalised; the residual plane may use fp16 subnormals (spacing 2^-24: finer than needed).
__global__ void __launch_bounds__(256)
gram_split16_kernel(const float *__restrict__ X, int64_t m, int n, int64_t ldx, const float *__restrict__ row_scale,
                    const float *__restrict__ col_shift, __half *__restrict__ planes, int64_t m_pad, int n_pad,
                    double *__restrict__ norm_part, int64_t seg_len, const int *__restrict__ colmax) {
  __shared__ float tile[GR_SPLIT_TILE][GR_SPLIT_TILE + 1];
  __shared__ double s_norm[4][GR_SPLIT_TILE];
  const int tid = threadIdx.x;
  const int node0 = blockIdx.x * GR_SPLIT_TILE;
  const int64_t t_begin = (int64_t)blockIdx.y * seg_len;
  const int64_t t_end = min(t_begin + seg_len, m_pad);
  const int lx = tid & 63, ly = tid >> 6;
  const int gnode = node0 + lx;
  const float sh = (col_shift && gnode < n) ? col_shift[gnode] : 0.f;
  const int wx = tid >> 2, wy = tid & 3;
  const int64_t plane_stride = (int64_t)n_pad * m_pad;
  double nsum = 0.0;
  const int e = gram_node_exponent(colmax[node0 + wx]);
  const float q_up = ldexpf(1.f, 8 - e), q_down = ldexpf(1.f, e - 8), inv_scale = ldexpf(1.f, -e);

  for (int64_t t0 = t_begin; t0 < t_end; t0 += GR_SPLIT_TILE) {
#pragma unroll
    for (int r = 0; r < GR_SPLIT_TILE / 4; ++r) {
      const int row = ly + 4 * r;
      const int64_t tt = t0 + row;
      float y = 0.f;
      if (tt < m && gnode < n) {
        const float sc = row_scale ? row_scale[tt] : 1.f;
        y = sc * (X[tt * ldx + gnode] - sh);
      }
      tile[row][lx] = y;
      nsum += (double)y * (double)y;
    }
    __syncthreads();
    uint32_t p0[8], p1[8];
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
      __half h0[2], h1[2];
#pragma unroll
      for (int x = 0; x < 2; ++x) {
        const float v = tile[wy * 16 + i + x][wx];
        const float k = rintf(v * q_up);                    // |k| <= 256
        h0[x] = __float2half_rn(k * (1.f / 256.f));         // exact
        h1[x] = __float2half_rn((v - k * q_down) * inv_scale);   // residual (exact in fp32) in normalised units
      }
      p0[i >> 1] = (uint32_t)__half_as_ushort(h0[0]) | ((uint32_t)__half_as_ushort(h0[1]) << 16);
      p1[i >> 1] = (uint32_t)__half_as_ushort(h1[0]) | ((uint32_t)__half_as_ushort(h1[1]) << 16);
    }
    uint4 *d0 = reinterpret_cast<uint4 *>(planes + (int64_t)(node0 + wx) * m_pad + t0 + wy * 16);
    uint4 *d1 = reinterpret_cast<uint4 *>(planes + plane_stride + (int64_t)(node0 + wx) * m_pad + t0 + wy * 16);
    d0[0] = make_uint4(p0[0], p0[1], p0[2], p0[3]);
    d0[1] = make_uint4(p0[4], p0[5], p0[6], p0[7]);
    d1[0] = make_uint4(p1[0], p1[1], p1[2], p1[3]);
    d1[1] = make_uint4(p1[4], p1[5], p1[6], p1[7]);
    __syncthreads();
  }
  s_norm[ly][lx] = nsum;
  __syncthreads();
  if (tid < GR_SPLIT_TILE)
    norm_part[(int64_t)blockIdx.y * n_pad + node0 + tid] =
        ((s_norm[0][tid] + s_norm[1][tid]) + s_norm[2][tid]) + s_norm[3][tid];
}

// norms[i] = sum over segments, fixed order (deterministic)
__global__ void gram_norm_reduce_kernel(const double *__restrict__ norm_part, int n_pad, int segs,
                                        double *__restrict__ norms) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_pad) return;
  double s = 0.0;
  for (int g = 0; g < segs; ++g) s += norm_part[(int64_t)g * n_pad + i];
  norms[i] = s;
}

// ------------------------------------------------------------------------------------------------
// Stage 2: the GEMM
// ------------------------------------------------------------------------------------------------
// FMT 0: bf16 slices in true units (3 or 6 products); FMT 1: fp16 normalised slices, TERMS = 2, 4 products
template <int TERMS, int BN, int BK, bool SPLIT, int FMT = 0>
struct GramCfg {
  static constexpr int ROW_BYTES = BK * 2;                             // bf16
  static constexpr int A_PLANE = GR_BM * ROW_BYTES;                    // bytes
  static constexpr int B_PLANE = BN * ROW_BYTES;
  static constexpr int STAGE_BYTES = TERMS * (A_PLANE + B_PLANE);
  static constexpr int SMEM_BUDGET = 227 * 1024 - 2048;                // barriers + alignment slack
  static constexpr int STAGES_RAW = SMEM_BUDGET / STAGE_BYTES;
  static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
  // SPLIT: the leading term y0*y0 and the small cross terms accumulate in separate TMEM tiles and
  // are folded into float64 registers, 32 columns per epilogue warp; otherwise one TMEM tile,
  // fp32 registers, 64 columns per warp.
  static constexpr int COLS_PER_WARP = SPLIT ? 32 : 64;
  static constexpr int EPI_WARPS = 4 * (BN / COLS_PER_WARP);
  static constexpr int THREADS = 32 * (GR_CTRL_WARPS + EPI_WARPS);
  static constexpr int TMEM_COLS = (SPLIT ? 4 : 2) * BN;               // double-buffered accumulator(s)
  static constexpr int SMALL_OFF = 2 * BN;                             // first column of the small-term tiles
  static constexpr int PAIRS = TERMS == 2 ? (FMT == 1 ? 4 : 3) : 6;
  static_assert(FMT == 0 || (TERMS == 2 && SPLIT), "the fp16 form is the 2-slice split-accumulator scheme");
  static constexpr size_t SMEM_BYTES = (size_t)STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
  static_assert(STAGES >= 2, "need at least a double-buffered operand pipeline");
  static_assert(TMEM_COLS <= 512, "TMEM has 512 columns");
  static_assert(THREADS <= 1024, "block size");
};

struct GramArgs {
  const int2 *units;      // (I, Jp) work list
  int n_units;
  int num_kb;             // k-blocks (m_pad / BK)
  int kb_per_chunk;       // k-blocks accumulated in TMEM before the fold
  const double *norms;    // (n_pad,)
  float *Z;
  int layout, n, nt, tile_begin, tile_end;
  int debug;              // diagnostics only (GL_GRAM_DEBUG): 1 = skip the TMEM folds, 2 = skip the MMAs, 4 = 1 pair
  const int *colmax;      // FMT 1: column maxima (float bits) -> per-node scale exponents for the epilogue
};

template <int TERMS, int BN, int BK, bool SPLIT, int FMT = 0>
__global__ void __launch_bounds__(GramCfg<TERMS, BN, BK, SPLIT, FMT>::THREADS, 1)
gram_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
            const GramArgs args) {
  using C = GramCfg<TERMS, BN, BK, SPLIT, FMT>;
  extern __shared__ unsigned char smem_raw[];
  unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + (size_t)C::STAGES * C::STAGE_BYTES);
  uint64_t *full = bars, *empty = bars + C::STAGES, *acc_full = bars + 2 * C::STAGES,
           *acc_empty = bars + 2 * C::STAGES + 2;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * C::STAGES + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&acc_full[b], 1);
      mbar_init(&acc_empty[b], C::EPI_WARPS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 2) tmem_alloc(tmem_slot, C::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int num_kb = args.num_kb, kb_per_chunk = args.kb_per_chunk;
  const int num_chunks = (num_kb + kb_per_chunk - 1) / kb_per_chunk;

  if (warp == 0) {
    // ===== TMA producer ===========================================================================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      for (int u = blockIdx.x; u < args.n_units; u += gridDim.x) {
        const int2 unit = args.units[u];
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty[s], ph ^ 1);
          unsigned char *sa = smem + (size_t)s * C::STAGE_BYTES;
          unsigned char *sb = sa + TERMS * C::A_PLANE;
          mbar_expect_tx(&full[s], C::STAGE_BYTES);
          tma_load_3d(sa, &map_a, &full[s], kb * BK, unit.x * GR_BM, 0);
          tma_load_3d(sb, &map_b, &full[s], kb * BK, unit.y * BN, 0);
          if (++s == C::STAGES) {
            s = 0;
            ph ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) ==================================================================
    if (lane == 0) {
      constexpr uint32_t idesc = FMT == 1 ? umma_idesc_f16(BN) : umma_idesc_bf16(BN);
      // cross terms in decreasing weight: (a-plane, b-plane)
      constexpr int pa[6] = {0, 0, 1, 1, 0, 2};
      constexpr int pb[6] = {0, 1, 0, 1, 2, 0};
      int s = 0;
      uint32_t ph = 0;
      int ab = 0;
      uint32_t aph = 0;
      for (int u = blockIdx.x; u < args.n_units; u += gridDim.x) {
        for (int c = 0; c < num_chunks; ++c) {
          mbar_wait(&acc_empty[ab], aph ^ 1);
          tc_fence_after();
          const uint32_t tmem_d = tmem_base + (uint32_t)(ab * BN);
          const int kb_end = min(num_kb, (c + 1) * kb_per_chunk);
          for (int kb = c * kb_per_chunk; kb < kb_end; ++kb) {
            mbar_wait(&full[s], ph);
            tc_fence_after();
            const uint32_t sa = smem_u32(smem + (size_t)s * C::STAGE_BYTES);
            const uint32_t sb = sa + TERMS * C::A_PLANE;
            // accumulate flags: the first MMA into each TMEM tile of a chunk overwrites it
            uint32_t acc_main = kb > c * kb_per_chunk, acc_small = acc_main;
#pragma unroll
            for (int p = 0; p < C::PAIRS; ++p) {
              if ((args.debug & 2) || ((args.debug & 4) && p > 0)) continue;   // timing diagnostics (wrong results)
              const uint64_t da = smem_desc_kmajor<C::ROW_BYTES>(sa + pa[p] * C::A_PLANE);
              const uint64_t db = smem_desc_kmajor<C::ROW_BYTES>(sb + pb[p] * C::B_PLANE);
              const bool to_small = SPLIT && p > 0;
#pragma unroll
              for (int k = 0; k < BK / GR_UMMA_K; ++k) {
                // advance along K inside the swizzle span: +32 B per k-step (encoded >> 4)
                const uint64_t adv = (uint64_t)(k * GR_UMMA_K * 2 / 16);
                if (to_small) {
                  umma_bf16(tmem_d + C::SMALL_OFF, da + adv, db + adv, idesc, acc_small);
                  acc_small = 1;
                } else {
                  umma_bf16(tmem_d, da + adv, db + adv, idesc, acc_main);
                  acc_main = 1;
                }
              }
            }
            umma_commit(&empty[s]);   // frees the operand stage once these MMAs have read it
            if (++s == C::STAGES) {
              s = 0;
              ph ^= 1;
            }
          }
          umma_commit(&acc_full[ab]);  // chunk accumulated: hand the TMEM buffer to the epilogue
          if ((ab ^= 1) == 0) aph ^= 1;
        }
      }
    }
  } else if (warp >= GR_CTRL_WARPS) {
    // ===== epilogue warps ===========================================================================
    using AccT = typename std::conditional<SPLIT, double, float>::type;
    constexpr int CPW = C::COLS_PER_WARP;
    const int e = warp - GR_CTRL_WARPS;
    const int q = warp & 3;                  // TMEM lane quarter this warp may access
    const int cb = e >> 2;                   // column block of the unit
    const uint32_t lane_sel = (uint32_t)(q * 32) << 16;
    int ab = 0;
    uint32_t aph = 0;
    for (int u = blockIdx.x; u < args.n_units; u += gridDim.x) {
      const int2 unit = args.units[u];
      AccT acc[CPW];
#pragma unroll
      for (int i = 0; i < CPW; ++i) acc[i] = AccT(0);
      for (int c = 0; c < num_chunks; ++c) {
        mbar_wait(&acc_full[ab], aph);
        tc_fence_after();
        const uint32_t taddr = tmem_base + lane_sel + (uint32_t)(ab * BN + cb * CPW);
        if (args.debug & 1) {
          // timing diagnostic: hand the buffer back without reading it (wrong results)
        } else if (SPLIT) {
#pragma unroll
          for (int g = 0; g < CPW / 8; ++g) {
            float v[8], vs[8];
            tmem_ld8(taddr + g * 8, v);
            tmem_ld8(taddr + C::SMALL_OFF + g * 8, vs);
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[g * 8 + i] += (AccT)(v[i] + vs[i]);
          }
        } else {
#pragma unroll
          for (int g = 0; g < CPW / 16; ++g) {
            float v[16];
            tmem_ld16(taddr + g * 16, v);
#pragma unroll
            for (int i = 0; i < 16; ++i) acc[g * 16 + i] += (AccT)v[i];
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&acc_empty[ab]);
        if ((ab ^= 1) == 0) aph ^= 1;
      }
      // ---- norm epilogue: Z = n_i + n_j - 2 G_ij, clamped at 0, masked to the strict upper triangle
      const int I = unit.x;
      const int col0 = unit.y * BN + cb * CPW;         // first global column of this warp's block
      const int J = col0 / GR_BM;                       // the solver tile this block falls into
      if (J < I || J >= args.nt) continue;
      const int tile = tile_index(I, J, args.nt);
      if (tile < args.tile_begin || tile >= args.tile_end) continue;
      const int r = q * 32 + lane;                      // row inside the tile
      const int gi = I * GR_BM + r;
      const int c0 = col0 - J * GR_BM;
      const double ni = args.norms[gi];
      if (FMT == 1) {   // normalised Gram entries back to true units: x 2^(e_i + e_j), exact
        const int ei = gram_node_exponent(args.colmax[gi]);
#pragma unroll
        for (int x = 0; x < CPW; ++x) {
          const int ex = ei + gram_node_exponent(args.colmax[col0 + x]);    // in [-200, 254]: a normal double
          acc[x] = (AccT)((double)acc[x] * __longlong_as_double((long long)(ex + 1023) << 52));
        }
      }
      if (args.layout == GL_LAYOUT_TILED) {
        float *dst = args.Z + (int64_t)(tile - args.tile_begin) * TT + r * T + c0;
#pragma unroll
        for (int g = 0; g < CPW / 4; ++g) {
          float o[4];
#pragma unroll
          for (int x = 0; x < 4; ++x) {
            const int gj = col0 + g * 4 + x;
            const double z = (ni + args.norms[gj]) - 2.0 * (double)acc[g * 4 + x];
            o[x] = (gj > gi && gj < args.n) ? fmaxf((float)z, 0.f) : 0.f;
          }
          *reinterpret_cast<float4 *>(dst + g * 4) = make_float4(o[0], o[1], o[2], o[3]);
        }
      } else if (gi < args.n) {
#pragma unroll
        for (int x = 0; x < CPW; ++x) {
          const int gj = col0 + x;
          if (gj > gi && gj < args.n) {
            const double z = (ni + args.norms[gj]) - 2.0 * (double)acc[x];
            args.Z[packed_index(gi, gj, args.n)] = fmaxf((float)z, 0.f);
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
}

}  // namespace gl
#include "gram_2cta.cuh"
namespace gl {

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int get_encode_fn(EncodeTiledFn *out) {
  static EncodeTiledFn cached = nullptr;
  if (!cached) {
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    GL_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (qres != cudaDriverEntryPointSuccess || !fn) return fail(GL_E_CUDA, "cuTensorMapEncodeTiled unavailable");
    cached = reinterpret_cast<EncodeTiledFn>(fn);
  }
  *out = cached;
  return GL_OK;
}

struct GramLayout {
  int64_t m_pad;
  int n_pad, segs;
  int64_t seg_len;
  size_t off_planes, off_norm_part, off_norms, off_colmax, off_units, total;
  int max_units;
};

static GramLayout gram_layout(int64_t m, int n, int terms) {
  GramLayout L;
  L.m_pad = ((m + 63) / 64) * 64;
  if (L.m_pad == 0) L.m_pad = 64;
  L.n_pad = tiles_per_side(n) * T;
  const int node_blocks = L.n_pad / GR_SPLIT_TILE;
  const int64_t steps = L.m_pad / GR_SPLIT_TILE;
  int64_t segs = (1184 + node_blocks - 1) / node_blocks;   // ~8 blocks per SM overall
  if (segs > steps) segs = steps;
  if (segs < 1) segs = 1;
  L.seg_len = ((steps + segs - 1) / segs) * GR_SPLIT_TILE;
  L.segs = (int)((L.m_pad + L.seg_len - 1) / L.seg_len);
  auto align = [](size_t x) { return (x + 1023) & ~(size_t)1023; };
  size_t off = 0;
  L.off_planes = off;
  off = align(off + (size_t)terms * L.n_pad * L.m_pad * sizeof(__nv_bfloat16));
  L.off_norm_part = off;
  off = align(off + (size_t)L.segs * L.n_pad * sizeof(double));
  L.off_norms = off;
  off = align(off + (size_t)(L.n_pad + 256) * sizeof(double));   // + slack: B box may overhang by one tile
  L.off_colmax = off;
  off = align(off + (size_t)L.n_pad * sizeof(int));
  L.off_units = off;
  const int nt = L.n_pad / T;
  L.max_units = nt * (nt + 1) / 2 + nt;
  off = align(off + (size_t)L.max_units * sizeof(int2));
  L.total = off;
  return L;
}

// Units (I, Jp) covering tiles [tile_begin, tile_end), rasterised in groups of GR_GROUP_ROWS tile
// rows, column-major inside a group, so that concurrently running CTAs share operand panels in L2.
static void build_units(int nt, int nb /*tiles per unit*/, int tile_begin, int tile_end, std::vector<int2> &units) {
  units.clear();
  const int njp = (nt + nb - 1) / nb;
  for (int I0 = 0; I0 < nt; I0 += GR_GROUP_ROWS) {
    const int I1 = std::min(nt, I0 + GR_GROUP_ROWS);
    for (int jp = I0 / nb; jp < njp; ++jp)
      for (int I = I0; I < I1; ++I) {
        bool any = false;
        for (int b = 0; b < nb; ++b) {
          const int J = jp * nb + b;
          if (J < I || J >= nt) continue;
          const int t = tile_index(I, J, nt);
          any |= (t >= tile_begin && t < tile_end);
        }
        if (any) units.push_back(make_int2(I, jp));
      }
  }
}

template <int TERMS, int BN, int BK, bool SPLIT, int FMT = 0>
static int launch_gram(const GramLayout &L, unsigned char *ws, const GramArgs &base, int n_sms, cudaStream_t st) {
  using C = GramCfg<TERMS, BN, BK, SPLIT, FMT>;
  EncodeTiledFn encode;
  GL_TRY(get_encode_fn(&encode));
  CUtensorMap map_a, map_b;
  const cuuint64_t gdim[3] = {(cuuint64_t)L.m_pad, (cuuint64_t)L.n_pad, (cuuint64_t)TERMS};
  const cuuint64_t gstride[2] = {(cuuint64_t)L.m_pad * 2, (cuuint64_t)L.n_pad * (cuuint64_t)L.m_pad * 2};
  const cuuint32_t estride[3] = {1, 1, 1};
  const CUtensorMapSwizzle sw = BK == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
  const cuuint32_t box_a[3] = {BK, GR_BM, TERMS};
  const cuuint32_t box_b[3] = {BK, BN, TERMS};
  const CUtensorMapDataType dtp = FMT == 1 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  CUresult r1 = encode(&map_a, dtp, 3, ws + L.off_planes, gdim, gstride, box_a, estride,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  CUresult r2 = encode(&map_b, dtp, 3, ws + L.off_planes, gdim, gstride, box_b, estride,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r1 != CUDA_SUCCESS || r2 != CUDA_SUCCESS)
    return fail(GL_E_CUDA, "cuTensorMapEncodeTiled failed (%d, %d)", (int)r1, (int)r2);
  GramArgs a = base;
  a.num_kb = (int)(L.m_pad / BK);
  a.kb_per_chunk = std::max(1, a.kb_per_chunk / BK);   // caller passed samples per chunk
  auto kern = gram_kernel<TERMS, BN, BK, SPLIT, FMT>;
  GL_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM_BYTES));
  const int grid = std::min(n_sms, a.n_units);
  kern<<<grid, C::THREADS, C::SMEM_BYTES, st>>>(map_a, map_b, a);
  GL_LAUNCH_CHECK();
  return GL_OK;
}

template <int TERMS_ = 3, int FMT_ = 0>
static int launch_gram_2cta(const GramLayout &L, unsigned char *ws, const GramArgs &base, int n_sms, cudaStream_t st) {
  using C = Gram2CfgT<TERMS_, FMT_>;
  EncodeTiledFn encode;
  GL_TRY(get_encode_fn(&encode));
  CUtensorMap map_a, map_b;
  const cuuint64_t gdim[3] = {(cuuint64_t)L.m_pad, (cuuint64_t)L.n_pad, (cuuint64_t)C::TERMS};
  const cuuint64_t gstride[2] = {(cuuint64_t)L.m_pad * 2, (cuuint64_t)L.n_pad * (cuuint64_t)L.m_pad * 2};
  const cuuint32_t estride[3] = {1, 1, 1};
  const cuuint32_t box_a[3] = {C::BK, GR_BM, C::TERMS};
  const cuuint32_t box_b[3] = {C::BK, C::BN / 2, C::TERMS};
  const CUtensorMapDataType dtp = FMT_ == 1 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  CUresult r1 = encode(&map_a, dtp, 3, ws + L.off_planes, gdim, gstride, box_a, estride,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  CUresult r2 = encode(&map_b, dtp, 3, ws + L.off_planes, gdim, gstride, box_b, estride,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r1 != CUDA_SUCCESS || r2 != CUDA_SUCCESS)
    return fail(GL_E_CUDA, "cuTensorMapEncodeTiled failed (%d, %d)", (int)r1, (int)r2);
  GramArgs a = base;
  a.num_kb = (int)(L.m_pad / C::BK);
  a.kb_per_chunk = std::max(1, a.kb_per_chunk / C::BK);
  auto kern = gram_kernel_2cta<TERMS_, FMT_>;
  GL_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM_BYTES));
  const int clusters = std::max(1, std::min(n_sms / 2, a.n_units));
  kern<<<2 * clusters, C::THREADS, C::SMEM_BYTES, st>>>(map_a, map_b, a);
  GL_LAUNCH_CHECK();
  return GL_OK;
}

}  // namespace gl
using namespace gl;

extern "C" int64_t gl_gram_workspace_bytes(int64_t m, int n, int terms) {
  if (m < 0 || n < 2 || (terms != 2 && terms != 3)) return -1;
  return (int64_t)gram_layout(m, n, terms).total + 1024;   // + slack: the base is aligned up internally
}

extern "C" int gl_pdist_sq_gram(const void *X, int64_t m, int n, int64_t ldx, const void *row_scale,
                                const void *col_shift, void *Z, int layout, int tile_begin, int tile_end,
                                int terms, int chunk, int variant, void *workspace, int64_t workspace_bytes,
                                void *stream) {
  if (n < 2) return fail(GL_E_SHAPE, "need at least 2 nodes, got n=%d", n);
  if (m < 0 || ldx < n) return fail(GL_E_SHAPE, "bad signal matrix shape (m=%lld, ldx=%lld, n=%d)",
                                    (long long)m, (long long)ldx, n);
  if (layout != GL_LAYOUT_PACKED && layout != GL_LAYOUT_TILED) return fail(GL_E_ARG, "bad layout");
  if (terms != 2 && terms != 3) return fail(GL_E_ARG, "terms must be 2 or 3");
  if (variant < 0 || variant > 8) return fail(GL_E_ARG, "variant must be 0..8");
  if (variant == 6 && terms != 3) return fail(GL_E_ARG, "the CTA-pair variant needs terms = 3");
  if ((variant == 7 || variant == 8) && terms != 2) return fail(GL_E_ARG, "the fp16 two-slice variants need terms = 2");
  if (layout == GL_LAYOUT_PACKED) {
    tile_begin = 0;
    tile_end = num_tiles(n);
  }
  if (tile_begin < 0 || tile_end > num_tiles(n) || tile_end < tile_begin)
    return fail(GL_E_ARG, "bad tile range [%d,%d)", tile_begin, tile_end);
  if (tile_end == tile_begin) return GL_OK;
  const GramLayout L = gram_layout(m, n, terms);
  // TMA and the swizzled operand planes want a 1024-byte aligned base: align up inside the buffer
  const uintptr_t ws_raw = reinterpret_cast<uintptr_t>(workspace);
  const uintptr_t ws_al = (ws_raw + 1023) & ~(uintptr_t)1023;
  if (!workspace || workspace_bytes < (int64_t)(L.total + (ws_al - ws_raw)))
    return fail(GL_E_ARG, "workspace too small: need %lld bytes (gl_gram_workspace_bytes)",
                (long long)L.total + 1024);
  if (chunk <= 0) chunk = 256;
  cudaStream_t st = as_stream(stream);
  unsigned char *ws = reinterpret_cast<unsigned char *>(ws_al);
  int dev = 0, n_sms = 0, cc_major = 0;
  GL_CUDA(cudaGetDevice(&dev));
  GL_CUDA(cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev));
  GL_CUDA(cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, dev));
  if (cc_major != 10) return fail(GL_E_CUDA, "the tcgen05 Gram kernel needs an sm_100 device (got sm_%d)", cc_major * 10);

  // ---- stage 1: split / transpose / norms ----------------------------------------------------------
  __nv_bfloat16 *planes = reinterpret_cast<__nv_bfloat16 *>(ws + L.off_planes);
  double *norm_part = reinterpret_cast<double *>(ws + L.off_norm_part);
  double *norms = reinterpret_cast<double *>(ws + L.off_norms);
  GL_CUDA(cudaMemsetAsync(norms, 0, (size_t)(L.n_pad + 256) * sizeof(double), st));
  const dim3 sgrid(L.n_pad / GR_SPLIT_TILE, L.segs);
  // The split-accumulator variants of the 3-term form carry the exact leading slice.  With 2 terms
  // the residual slice alone cannot hold the bits the quantisation pushes into it, so the 2-term
  // form keeps plain round-to-nearest bf16 slices (still with separate accumulators).
  const bool exact0 = ((variant <= 1 || variant == 6) && terms == 3) || variant == 7 || variant == 8;
  int *colmax = reinterpret_cast<int *>(ws + L.off_colmax);
  const float *Xf = (const float *)X, *rsf = (const float *)row_scale, *csf = (const float *)col_shift;
  if (exact0) {
    GL_CUDA(cudaMemsetAsync(colmax, 0, (size_t)L.n_pad * sizeof(int), st));
    gram_colmax_kernel<<<sgrid, 256, 0, st>>>(Xf, m, n, ldx, rsf, csf, colmax, L.seg_len);
    GL_LAUNCH_CHECK();
  }
#define GL_SPLIT(TERMS_, EX_)                                                                                \
  gram_split_kernel<TERMS_, EX_><<<sgrid, 256, 0, st>>>(Xf, m, n, ldx, rsf, csf, planes, L.m_pad, L.n_pad, \
                                                        norm_part, L.seg_len, colmax)
  if (variant == 7 || variant == 8) {
    gram_split16_kernel<<<sgrid, 256, 0, st>>>(Xf, m, n, ldx, rsf, csf, reinterpret_cast<__half *>(planes), L.m_pad, L.n_pad,
                                               norm_part, L.seg_len, colmax);
  } else if (terms == 2) {
    GL_SPLIT(2, false);
  } else {
    if (exact0) GL_SPLIT(3, true); else GL_SPLIT(3, false);
  }
#undef GL_SPLIT
  GL_LAUNCH_CHECK();
  gram_norm_reduce_kernel<<<(L.n_pad + 255) / 256, 256, 0, st>>>(norm_part, L.n_pad, L.segs, norms);
  GL_LAUNCH_CHECK();

  // ---- stage 2: work list + GEMM -------------------------------------------------------------------
  const int bn = (variant == 2 || variant == 4) ? 256 : 128;
  const int nt = tiles_per_side(n);
  std::vector<int2> units;
  if (variant == 6 || variant == 8)
    build_units_2cta(nt, tile_begin, tile_end, units);
  else
    build_units(nt, bn / T, tile_begin, tile_end, units);
  if ((int)units.size() > L.max_units) return fail(GL_E_STATE, "unit list overflow");
  int2 *units_dev = reinterpret_cast<int2 *>(ws + L.off_units);
  // pageable source: the runtime stages it before returning, so the vector may die afterwards
  GL_CUDA(cudaMemcpyAsync(units_dev, units.data(), units.size() * sizeof(int2), cudaMemcpyHostToDevice, st));
  GramArgs a;
  a.units = units_dev;
  a.n_units = (int)units.size();
  a.num_kb = 0;
  a.kb_per_chunk = chunk;
  a.norms = norms;
  a.Z = static_cast<float *>(Z);
  a.layout = layout;
  a.n = n;
  a.nt = nt;
  a.tile_begin = tile_begin;
  a.tile_end = tile_end;
  a.colmax = colmax;
  {
    const char *dbg = getenv("GL_GRAM_DEBUG");   // timing diagnostics of the pipeline stages; results are wrong
    a.debug = dbg ? atoi(dbg) : 0;
    if (a.debug) fprintf(stderr, "gl_b200: GL_GRAM_DEBUG=%d -- timing diagnostics, the distances are WRONG\n", a.debug);
  }
  // variant: 0 = split accumulators (precise, default), 128-wide units, 32-sample k-blocks (SWIZZLE_64B)
  //          1 = same with 64-sample k-blocks (SWIZZLE_128B)
  //          2 / 3 = single accumulator, 256- / 128-wide units, 32-sample k-blocks
  //          4 / 5 = single accumulator, 256- / 128-wide units, 64-sample k-blocks
  if (variant == 6) return launch_gram_2cta<3, 0>(L, ws, a, n_sms, st);
  if (variant == 8) return launch_gram_2cta<2, 1>(L, ws, a, n_sms, st);
  if (variant == 7) return launch_gram<2, 128, 32, true, 1>(L, ws, a, n_sms, st);
  if (terms == 3) {
    switch (variant) {
      case 0: return launch_gram<3, 128, 32, true>(L, ws, a, n_sms, st);
      case 1: return launch_gram<3, 128, 64, true>(L, ws, a, n_sms, st);
      case 2: return launch_gram<3, 256, 32, false>(L, ws, a, n_sms, st);
      case 3: return launch_gram<3, 128, 32, false>(L, ws, a, n_sms, st);
      case 4: return fail(GL_E_ARG, "terms=3 with 256-wide units and 64-sample k-blocks does not fit shared memory");
      default: return launch_gram<3, 128, 64, false>(L, ws, a, n_sms, st);
    }
  }
  switch (variant) {
    case 0: return launch_gram<2, 128, 32, true>(L, ws, a, n_sms, st);
    case 1: return launch_gram<2, 128, 64, true>(L, ws, a, n_sms, st);
    case 2: return launch_gram<2, 256, 32, false>(L, ws, a, n_sms, st);
    case 3: return launch_gram<2, 128, 32, false>(L, ws, a, n_sms, st);
    case 4: return launch_gram<2, 256, 64, false>(L, ws, a, n_sms, st);
    default: return launch_gram<2, 128, 64, false>(L, ws, a, n_sms, st);
  }
}
