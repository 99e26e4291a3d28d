// edge_tile.cuh -- device code of the fused edge sweep shared by the per-iteration kernel
// (edge_pass.cu) and the persistent whole-iteration kernel (persistent.cu).
#pragma once
#include "solver.cuh"

namespace gl {

template <typename F> struct EdgeArgs {
  F *w;
  const F *z;
  F *r;  // batch mode: q - y
  F *w_out = nullptr;  // two-buffer mode (PP): w' goes here and A.w keeps the previous iterate; nullptr = in place
  int64_t gstride;
  const F *a, *b;          // gamma*d, gamma*pb   [K][npad]
  double *xch;             // exchange buffer 0: [S p | S w' | G_r | G_w]; buffer 1 at +xch_stride
  int64_t xch_stride;
  double *cta_partial;     // [ntl][2]
  int *hdr;                // comm header (peer mode)
  int world, rank;
  PeerView peers;
  const int2 *tile_ij;
  int *flags;
  int n, npad, ntl;
  F c, tgz;
};

// one stage of a reduce-scatter butterfly over `CNT` values held per lane
template <typename F, int CNT> __device__ __forceinline__ void rs_stage(F *v, int lane, int mask) {
  const bool upper = (lane & mask) != 0;
#pragma unroll
  for (int i = 0; i < CNT / 2; ++i) {
    const F send = upper ? v[i] : v[i + CNT / 2];
    const F keep = upper ? v[i + CNT / 2] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, mask);
  }
}

template <typename F, bool BATCH, bool MASKED, bool PP = false>
__device__ __forceinline__ void edge_tile(const EdgeArgs<F> &A, F *__restrict__ wt,
                                          const F *__restrict__ zt, F *__restrict__ rt, int I, int J,
                                          const F (*s_ab)[2], const F *aj, const F *bj, F *csp, F *csw,
                                          F *rowfinal, F &res, F &wsq) {
  using V = typename Vec<F>::type;
  constexpr int VEC = Vec<F>::N;
  constexpr int NV = T / (32 * VEC);
  constexpr int CH = 4;                       // rows per chunk
  constexpr int NCH = ROWS_PER_WARP / CH;     // 4 chunks
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  const bool diag = I == J;
  const F c = A.c, tgz = A.tgz;
  const F zero = F(0);

  V wv[2][CH][NV], zv[2][CH][NV];
  F hold[NCH];

  auto active = [&](int ch, int v) -> bool {
    // in a diagonal tile a vector is live only if its last column lies right of the chunk's first row
    const int r0 = wp * ROWS_PER_WARP + ch * CH;
    return !MASKED || !diag || (v * 32 * VEC + lane * VEC + VEC - 1 > r0);
  };
  auto load_chunk = [&](int buf, int ch) {
    const int r0 = wp * ROWS_PER_WARP + ch * CH;
#pragma unroll
    for (int rr = 0; rr < CH; ++rr)
#pragma unroll
      for (int v = 0; v < NV; ++v) {
        const int off = (r0 + rr) * T + v * 32 * VEC + lane * VEC;
        if (active(ch, v)) {
          wv[buf][rr][v] = ld_stream(reinterpret_cast<const V *>(wt + off));
          zv[buf][rr][v] = ld_stream(reinterpret_cast<const V *>(zt + off));
        } else {
          F zz[VEC];
#pragma unroll
          for (int e = 0; e < VEC; ++e) zz[e] = zero;
          wv[buf][rr][v] = vec_pack<F>(zz);
          zv[buf][rr][v] = vec_pack<F>(zz);
        }
      }
  };

  load_chunk(0, 0);
#pragma unroll
  for (int ch = 0; ch < NCH; ++ch) {
    if (ch + 1 < NCH) load_chunk((ch + 1) & 1, ch + 1);
    const int buf = ch & 1;
    const int r0 = wp * ROWS_PER_WARP + ch * CH;
    F rowv[2 * CH];
#pragma unroll
    for (int rr = 0; rr < CH; ++rr) {
      const int r = r0 + rr;
      const F ai = s_ab[r][0], bi = s_ab[r][1];
      F rsp = zero, rsw = zero;
#pragma unroll
      for (int v = 0; v < NV; ++v) {
        F wq[VEC], zq[VEC], wn[VEC], rq[VEC];
        vec_unpack<F>(wv[buf][rr][v], wq);
        vec_unpack<F>(zv[buf][rr][v], zq);
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
          const int ci = v * VEC + e;
          const F wo = wq[e];
          const F y = c * wo - (ai + aj[ci]);                 // :147
          F p = y - tgz * zq[e];                              // :150
          p = p > zero ? p : zero;
          const F q = c * p - (bi + bj[ci]);                  // :153
          F rr_ = q - y;
          F wnew = (wo - y) + q;                              // :163
          if (MASKED) {
            const int gi = I * T + r, gj = J * T + v * 32 * VEC + lane * VEC + e;
            const bool ok = (gj > gi) && (gj < A.n);
            p = ok ? p : zero;
            rr_ = ok ? rr_ : zero;
            wnew = ok ? wnew : zero;
          }
          res += rr_ * rr_;
          wsq += wnew * wnew;
          rsp += p;
          rsw += wnew;
          csp[ci] += p;
          csw[ci] += wnew;
          wn[e] = wnew;
          rq[e] = rr_;
        }
        if (active(ch, v)) {
          const int off = r * T + v * 32 * VEC + lane * VEC;
          st_stream(reinterpret_cast<V *>((PP ? rt : wt) + off), vec_pack<F>(wn));   // PP: rt is the second w buffer
          if (BATCH) st_stream(reinterpret_cast<V *>(rt + off), vec_pack<F>(rq));
        }
      }
      rowv[2 * rr] = rsp;
      rowv[2 * rr + 1] = rsw;
    }
    // reduce-scatter the chunk's 8 row values over lane bits 4,3,2
    rs_stage<F, 8>(rowv, lane, 16);
    rs_stage<F, 4>(rowv, lane, 8);
    rs_stage<F, 2>(rowv, lane, 4);
    hold[ch] = rowv[0];
  }
  // finish over lane bits 1,0: lane ends with the total of
  //   row = wp*16 + (lane&3)*4 + ((lane>>3)&3),  quantity = (lane>>2)&1  (0: S p, 1: S w')
  rs_stage<F, 4>(hold, lane, 2);
  rs_stage<F, 2>(hold, lane, 1);
  *rowfinal = hold[0];
}

// Shared memory of one edge CTA.
template <typename F> struct EdgeSmem {
  F ab[T][2];                          // a_i, b_i of the tile's rows
  F col[EDGE_THREADS / 32][2][T];      // column partials per warp
  double red[EDGE_THREADS / 32][2];
};

// One CTA sweeps tile `t` of graph `k`: loads, update, stores, and the tile's 512 float64 atomics
// into the degree accumulators g_sp / g_sw.  COHERENT_AB: read a/b through L2 (persistent kernel:
// they were rewritten by other SMs during this launch).  On return thread 0 holds the tile's
// residual sums in (r2, w2).  Ends with a __syncthreads-safe state for reuse.
template <typename F, bool BATCH, bool COHERENT_AB, bool PP = false>
__device__ __forceinline__ void edge_cta_tile(const EdgeArgs<F> &A, EdgeSmem<F> &sm, int t, int k,
                                              double *g_sp, double *g_sw, double &r2, double &w2) {
  using V = typename Vec<F>::type;
  constexpr int VEC = Vec<F>::N;
  constexpr int NV = T / (32 * VEC);
  constexpr int NC = NV * VEC;
  const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
  const int2 ij = A.tile_ij[t];
  const int I = ij.x, J = ij.y;
  const bool masked = (I == J) || ((J + 1) * T > A.n);
  const F *a = A.a + (int64_t)k * A.npad, *b = A.b + (int64_t)k * A.npad;
  if (tid < T) {
    sm.ab[tid][0] = COHERENT_AB ? __ldcg(a + I * T + tid) : a[I * T + tid];
    sm.ab[tid][1] = COHERENT_AB ? __ldcg(b + I * T + tid) : b[I * T + tid];
  }
  F aj[NC], bj[NC], csp[NC], csw[NC];
#pragma unroll
  for (int v = 0; v < NV; ++v) {
    const int col = J * T + v * 32 * VEC + lane * VEC;
    const V va = COHERENT_AB ? __ldcg(reinterpret_cast<const V *>(a + col)) : *reinterpret_cast<const V *>(a + col);
    const V vb = COHERENT_AB ? __ldcg(reinterpret_cast<const V *>(b + col)) : *reinterpret_cast<const V *>(b + col);
    vec_unpack<F>(va, aj + v * VEC);
    vec_unpack<F>(vb, bj + v * VEC);
  }
#pragma unroll
  for (int i = 0; i < NC; ++i) csp[i] = csw[i] = F(0);
  __syncthreads();

  const int64_t base = (int64_t)k * A.gstride + (int64_t)t * TT;
  F *wt = A.w + base;
  const F *zt = A.z + base;
  static_assert(!(BATCH && PP), "the two-buffer mode keeps the previous iterate instead of q - y");
  F *rt = BATCH ? A.r + base : (PP ? A.w_out + base : nullptr);
  F rowfinal, res = F(0), wsq = F(0);
  if (masked)
    edge_tile<F, BATCH, true, PP>(A, wt, zt, rt, I, J, sm.ab, aj, bj, csp, csw, &rowfinal, res, wsq);
  else
    edge_tile<F, BATCH, false, PP>(A, wt, zt, rt, I, J, sm.ab, aj, bj, csp, csw, &rowfinal, res, wsq);

  // ---- row totals: one float64 atomic per (row, quantity) -----------------------------------
  {
    const int r = wp * ROWS_PER_WARP + (lane & 3) * 4 + ((lane >> 3) & 3);
    const int qty = (lane >> 2) & 1;
    const int gi = I * T + r;
    if (gi < A.n) atomicAdd((qty ? g_sw : g_sp) + (int64_t)k * A.npad + gi, (double)rowfinal);
  }
  // ---- column totals: 8 warps -> shared -> one atomic per (column, quantity) ------------------
#pragma unroll
  for (int v = 0; v < NV; ++v) {
    const int col = v * 32 * VEC + lane * VEC;
    *reinterpret_cast<V *>(&sm.col[wp][0][col]) = vec_pack<F>(csp + v * VEC);
    *reinterpret_cast<V *>(&sm.col[wp][1][col]) = vec_pack<F>(csw + v * VEC);
  }
  {
    const double rr = warp_sum((double)res), ww = warp_sum((double)wsq);
    if (lane == 0) {
      sm.red[wp][0] = rr;
      sm.red[wp][1] = ww;
    }
  }
  __syncthreads();
  {
    const int qty = tid >> 7, col = tid & (T - 1);
    double acc = 0.0;
#pragma unroll
    for (int q = 0; q < EDGE_THREADS / 32; ++q) acc += (double)sm.col[q][qty][col];
    const int gj = J * T + col;
    if (gj < A.n) atomicAdd((qty ? g_sw : g_sp) + (int64_t)k * A.npad + gj, acc);
  }
  if (tid == 0) {
    r2 = w2 = 0.0;
#pragma unroll
    for (int q = 0; q < EDGE_THREADS / 32; ++q) {
      r2 += sm.red[q][0];
      w2 += sm.red[q][1];
    }
  }
}

}  // namespace gl
