// gram_2cta.cuh -- CTA-pair form of the Gram GEMM (included by gram.cu after its PTX helpers).
//
// Two CTAs of a cluster (one TPC) compute a 256 x 128 block of the Gram matrix with
// tcgen05.mma.cta_group::2 (M = 256, N = 128): CTA r owns tile row 2*Ip + r (its 128 operand rows of A
// and its 128 accumulator lanes) and loads only HALF of the shared B operand (64 of the 128 columns'
// rows); the tensor core reads the other half out of the peer's shared memory.  Operand traffic per
// CTA and k-block drops from 256 to 192 rows (the single-CTA kernel is L2-bandwidth bound: 71 % tensor
// pipe, profiles/r1_gram_kernel_full.txt) and the smaller stage allows a 6-deep TMA ring.
//
// Protocol (all barriers exist in both CTAs at the same shared-memory offsets):
//   full[s]      leader's copy only: expect_tx of BOTH CTAs' bytes; the peer's TMA signals it through
//                the cta_group::2 form with the peer bit of the barrier address cleared
//   empty[s]     tcgen05.commit ... multicast::cluster 0b11: each CTA's producer waits on its own copy
//   acc_full[b]  likewise multicast to both epilogues
//   acc_empty[b] leader's copy, count 2 x epilogue warps: the peer's warps arrive remotely (mapa)
// Same precision scheme as the default variant: exact leading slice, leading / small-term TMEM
// accumulators, float64 folds every chunk.
#pragma once

namespace gl {

constexpr uint32_t GR_PEER_BIT_MASK = 0xFEFFFFFFu;   // clears the CTA-pair rank bit of a shared::cluster address

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_load_3d_2sm(void *smem_dst, const CUtensorMap *map, uint32_t leader_bar, int c0,
                                                int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(leader_bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t *slot_in_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot_in_smem)),
               "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_bf16_2sm(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                              uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at this shared-memory offset in BOTH CTAs once the MMAs issued so far are done
__device__ __forceinline__ void umma_commit_2sm(uint64_t *bar) {
  const uint16_t mask = 3;
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(mask)
               : "memory");
}
// arrive on the barrier at this offset in CTA `rank` of the cluster
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t *bar, uint32_t rank) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}"
      ::"r"(smem_u32(bar)), "r"(rank)
      : "memory");
}
__host__ __device__ constexpr uint32_t umma_idesc_bf16_m256(int n_dim) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n_dim >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
}

__host__ __device__ constexpr uint32_t umma_idesc_f16_m256(int n_dim) {
  return (1u << 4) | ((uint32_t)(n_dim >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
}

// TERMS_ / FMT_ as in GramCfg: (3, 0) = three bf16 slices, six products; (2, 1) = two normalised fp16 slices,
// four products (gram_split16_kernel), scaled back by 2^(e_i + e_j) in the epilogue.
template <int TERMS_ = 3, int FMT_ = 0>
struct Gram2CfgT {
  static constexpr int TERMS = TERMS_, FMT = FMT_, BN = 128, BK = 32;
  static constexpr int ROW_BYTES = BK * 2;
  static constexpr int A_PLANE = GR_BM * ROW_BYTES;            // 8 KB: this CTA's 128 rows
  static constexpr int B_PLANE = (BN / 2) * ROW_BYTES;         // 4 KB: this CTA's half of the columns' rows
  static constexpr int STAGE_BYTES = TERMS * (A_PLANE + B_PLANE);   // 36 KB (24 KB with two slices) per CTA
  static constexpr int STAGES = TERMS == 3 ? 6 : 8;
  static constexpr int COLS_PER_WARP = 32;
  static constexpr int EPI_WARPS = 4 * (BN / COLS_PER_WARP);
  static constexpr int THREADS = 32 * (GR_CTRL_WARPS + EPI_WARPS);
  static constexpr int TMEM_COLS = 4 * BN;
  static constexpr int SMALL_OFF = 2 * BN;
  static constexpr int PAIRS = TERMS == 3 ? 6 : (FMT == 1 ? 4 : 3);
  static constexpr size_t SMEM_BYTES = (size_t)STAGES * STAGE_BYTES + 1024 + 256;
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory");
};
using Gram2Cfg = Gram2CfgT<3, 0>;

// units: (Ip, J): tile rows 2*Ip and 2*Ip + 1, tile column J
template <int TERMS_ = 3, int FMT_ = 0>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__((Gram2CfgT<TERMS_, FMT_>::THREADS), 1)
gram_kernel_2cta(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                 const GramArgs args) {
  using C = Gram2CfgT<TERMS_, FMT_>;
  constexpr int BN = C::BN, BK = C::BK;
  extern __shared__ unsigned char smem_raw[];
  unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + (size_t)C::STAGES * C::STAGE_BYTES);
  uint64_t *full = bars, *empty = bars + C::STAGES, *acc_full = bars + 2 * C::STAGES,
           *acc_empty = bars + 2 * C::STAGES + 2;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * C::STAGES + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

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
      mbar_init(&acc_empty[b], 2 * C::EPI_WARPS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 2) tmem_alloc_2sm(tmem_slot, C::TMEM_COLS);
  tc_fence_before();
  cluster_sync_all();   // both CTAs' barriers are initialised before any remote arrive / TMA signal
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int num_kb = args.num_kb, kb_per_chunk = args.kb_per_chunk;
  const int num_chunks = (num_kb + kb_per_chunk - 1) / kb_per_chunk;

  if (warp == 0) {
    // ===== TMA producer (both CTAs): own A rows, own half of B, signalling the LEADER's full barrier ====
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      for (int u = cluster_id; u < args.n_units; u += n_clusters) {
        const int2 unit = args.units[u];
        const int row_a = (2 * unit.x + (int)rank) * GR_BM;
        const int row_b = unit.y * BN + (int)rank * (BN / 2);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty[s], ph ^ 1);
          unsigned char *sa = smem + (size_t)s * C::STAGE_BYTES;
          unsigned char *sb = sa + C::TERMS * C::A_PLANE;
          if (leader) mbar_expect_tx(&full[s], 2 * C::STAGE_BYTES);
          const uint32_t lbar = smem_u32(&full[s]) & GR_PEER_BIT_MASK;
          tma_load_3d_2sm(sa, &map_a, lbar, kb * BK, row_a, 0);
          tma_load_3d_2sm(sb, &map_b, lbar, kb * BK, row_b, 0);
          if (++s == C::STAGES) {
            s = 0;
            ph ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: one thread of the LEADER CTA drives both tensor cores ==========================
    if (lane == 0 && leader) {
      constexpr uint32_t idesc = C::FMT == 1 ? umma_idesc_f16_m256(BN) : umma_idesc_bf16_m256(BN);
      constexpr int pa[6] = {0, 0, 1, 1, 0, 2};
      constexpr int pb[6] = {0, 1, 0, 1, 2, 0};
      int s = 0;
      uint32_t ph = 0;
      int ab = 0;
      uint32_t aph = 0;
      for (int u = cluster_id; u < args.n_units; u += n_clusters) {
        for (int c = 0; c < num_chunks; ++c) {
          mbar_wait(&acc_empty[ab], aph ^ 1);
          tc_fence_after();
          const uint32_t tmem_d = tmem_base + (uint32_t)(ab * BN);
          const int kb_end = min(num_kb, (c + 1) * kb_per_chunk);
          for (int kb = c * kb_per_chunk; kb < kb_end; ++kb) {
            mbar_wait(&full[s], ph);
            tc_fence_after();
            const uint32_t sa = smem_u32(smem + (size_t)s * C::STAGE_BYTES);
            const uint32_t sb = sa + C::TERMS * C::A_PLANE;
            uint32_t acc_main = kb > c * kb_per_chunk, acc_small = acc_main;
#pragma unroll
            for (int p = 0; p < C::PAIRS; ++p) {
              const uint64_t da = smem_desc_kmajor<C::ROW_BYTES>(sa + pa[p] * C::A_PLANE);
              const uint64_t db = smem_desc_kmajor<C::ROW_BYTES>(sb + pb[p] * C::B_PLANE);
#pragma unroll
              for (int k = 0; k < BK / GR_UMMA_K; ++k) {
                const uint64_t adv = (uint64_t)(k * GR_UMMA_K * 2 / 16);
                if (p > 0) {
                  umma_bf16_2sm(tmem_d + C::SMALL_OFF, da + adv, db + adv, idesc, acc_small);
                  acc_small = 1;
                } else {
                  umma_bf16_2sm(tmem_d, da + adv, db + adv, idesc, acc_main);
                  acc_main = 1;
                }
              }
            }
            umma_commit_2sm(&empty[s]);
            if (++s == C::STAGES) {
              s = 0;
              ph ^= 1;
            }
          }
          umma_commit_2sm(&acc_full[ab]);
          if ((ab ^= 1) == 0) aph ^= 1;
        }
      }
    }
  } else if (warp >= GR_CTRL_WARPS) {
    // ===== epilogue warps (both CTAs): fold the own 128 accumulator lanes, norm epilogue, store =======
    constexpr int CPW = C::COLS_PER_WARP;
    const int e = warp - GR_CTRL_WARPS;
    const int q = warp & 3;
    const int cb = e >> 2;
    const uint32_t lane_sel = (uint32_t)(q * 32) << 16;
    int ab = 0;
    uint32_t aph = 0;
    for (int u = cluster_id; u < args.n_units; u += n_clusters) {
      const int2 unit = args.units[u];
      double acc[CPW];
#pragma unroll
      for (int i = 0; i < CPW; ++i) acc[i] = 0.0;
      for (int c = 0; c < num_chunks; ++c) {
        mbar_wait(&acc_full[ab], aph);
        tc_fence_after();
        const uint32_t taddr = tmem_base + lane_sel + (uint32_t)(ab * BN + cb * CPW);
#pragma unroll
        for (int g = 0; g < CPW / 8; ++g) {
          float v[8], vs[8];
          tmem_ld8(taddr + g * 8, v);
          tmem_ld8(taddr + C::SMALL_OFF + g * 8, vs);
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[g * 8 + i] += (double)(v[i] + vs[i]);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(&acc_empty[ab], 0);   // the leader's MMA thread waits for both CTAs
        if ((ab ^= 1) == 0) aph ^= 1;
      }
      const int I = 2 * unit.x + (int)rank, J = unit.y;
      if (I >= args.nt || J < I) continue;
      const int tile = tile_index(I, J, args.nt);
      if (tile < args.tile_begin || tile >= args.tile_end) continue;
      const int col0 = J * BN + cb * CPW;
      const int r = q * 32 + lane;
      const int gi = I * GR_BM + r;
      const int c0 = cb * CPW;
      const double ni = args.norms[gi];
      if (C::FMT == 1) {   // normalised Gram entries back to true units: x 2^(e_i + e_j), exact
        const int ei = gram_node_exponent(args.colmax[gi]);
#pragma unroll
        for (int x = 0; x < CPW; ++x) {
          const int ex = ei + gram_node_exponent(args.colmax[col0 + x]);
          acc[x] *= __longlong_as_double((long long)(ex + 1023) << 52);
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
            const double z = (ni + args.norms[gj]) - 2.0 * acc[g * 4 + x];
            o[x] = (gj > gi && gj < args.n) ? fmaxf((float)z, 0.f) : 0.f;
          }
          *reinterpret_cast<float4 *>(dst + g * 4) = make_float4(o[0], o[1], o[2], o[3]);
        }
      } else if (gi < args.n) {
#pragma unroll
        for (int x = 0; x < CPW; ++x) {
          const int gj = col0 + x;
          if (gj > gi && gj < args.n) {
            const double z = (ni + args.norms[gj]) - 2.0 * acc[x];
            args.Z[packed_index(gi, gj, args.n)] = fmaxf((float)z, 0.f);
          }
        }
      }
    }
  }

  tc_fence_before();
  cluster_sync_all();   // no CTA leaves (or frees TMEM) while its peer may still signal it
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_2sm(tmem_base, C::TMEM_COLS);
  }
}

// Units (Ip, J) of the pair kernel, rasterised like build_units: groups of GR_GROUP_ROWS tile rows
// (= GR_GROUP_ROWS / 2 pair rows), column-major inside a group.
static void build_units_2cta(int nt, int tile_begin, int tile_end, std::vector<int2> &units) {
  units.clear();
  const int npairs = (nt + 1) / 2, gp = GR_GROUP_ROWS / 2;
  for (int P0 = 0; P0 < npairs; P0 += gp) {
    const int P1 = std::min(npairs, P0 + gp);
    for (int J = 2 * P0; J < nt; ++J)
      for (int P = P0; P < P1; ++P) {
        bool any = false;
        for (int r = 0; r < 2; ++r) {
          const int I = 2 * P + r;
          if (I >= nt || J < I) continue;
          const int t = tile_index(I, J, nt);
          any |= (t >= tile_begin && t < tile_end);
        }
        if (any) units.push_back(make_int2(P, J));
      }
  }
}

}  // namespace gl
