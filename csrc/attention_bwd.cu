// K4 backward: flash attention backward on tcgen05 tensor cores.
//
// Three launches:
//   1. prep   : delta[b,h,i] = sum_d dO*O (fp32) and zero the fp32 dQ accumulator           (bandwidth-bound)
//   2. main   : one CTA per (batch, head, 128-row K/V tile), looping over the query tiles that see it.
//               Per query tile the single MMA-issuer thread runs five tcgen05 GEMMs, all accumulators in TMEM:
//                 S  = Q K^T          dP = dO V^T                        (both operands K-major)
//                 dV += P^T dO        dK += dS^T Q                       (A MN-major from the P/dS smem tiles)
//                 dQ  = dS K                                             (B MN-major = the resident K tile)
//               Eight compute warps (thread == query row x column half) turn S,dP into P and dS
//               (exp2 from the saved LSE, causal mask, regenerated dropout mask) and write them once, as bf16,
//               into 128B-swizzled smem tiles that serve as BOTH the K-major and the MN-major operand above.
//               dQ tiles are drained TMEM -> smem -> TMA reduce-add (fp32) into the accumulator, so no atomics
//               instructions are issued by threads and no second pass over K/V is needed.
//   3. finish : dQ accumulator fp32 -> bf16 in the layout of q.
// dK / dV stay in TMEM for the whole CTA lifetime and are written once with TMA stores.
//
// Head dims 129..192 (the reference grid's 4-head rows: 768 / 4): S | dP | dV | dK would need 640 TMEM columns and all
// seven tiles 256 KB of shared memory, so the main step runs as TWO specialisations of the same kernel (PART):
//   PART 1: S -> P, dV += P^T dO                 (no V, no dP, no dS: 176 KB, 448 columns)
//   PART 2: S, dP -> dS, dK += dS^T Q, dQ = dS K (no P tile; dQ re-uses the S | dP columns and leaves through the dS
//           tile in three 64-column passes: 224 KB, 448 columns)
// Both recompute S (six tile GEMMs instead of five); dV / dK leave through the Q stage, which is free by then.
//
// Replaces the autograd backward of aten.bmm x2 + _softmax_backward_data + dropout backward
//  for the reference's attention blocks (/root/reference/LLMsDistributedTrainingHelper.py:40-44).
#include "common.cuh"
#include "host_common.h"
#include "ppb200.h"

namespace ppb {

bool make_head_tmap(CUtensorMap* tm, const void* ptr, int dh, int H, int S, int B, long long ld, int box_rows);
int attn_tc_max_head_dim();

struct AttnBwdParams {
  int B, H, Sq, Skv, dh;
  int causal;
  float scale;
  float sl2;
  unsigned long long seed;
  const unsigned long long* seed_off;
  unsigned int drop_thresh;
  float drop_scale;
  const float* lse;    // [B, H, Sq]
  const float* delta;  // [B, H, Sq]
};

template <int DQK, int DV, int PART>
struct AttnBwdCfg {
  static constexpr int NQK = DQK / 64, NV = DV / 64;
  static constexpr int CHUNK = 128 * 128;  // bytes of a [128 rows x 128 B] swizzled chunk
  static constexpr bool WIDE = (DQK > 128) || (DV > 128);
  static constexpr bool HAS_P = PART != 2;    // P tile (dV) wanted
  static constexpr bool HAS_DS = PART != 1;   // dP, dS tile (dK, dQ) wanted; V is only read for dP
  static_assert(WIDE == (PART != 0), "wide heads run as PART 1 + PART 2, the others as one kernel");
  static constexpr int K_BYTES = NQK * CHUNK, V_BYTES = HAS_DS ? NV * CHUNK : 0;
  static constexpr int Q_BYTES = NQK * CHUNK, DO_BYTES = NV * CHUNK;
  static constexpr int PD_BYTES = HAS_P ? 2 * CHUNK : 0, DS_BYTES = HAS_DS ? 2 * CHUNK : 0;
  static constexpr int STG_BYTES = (DQK / 32) * CHUNK;  // fp32 dQ staging: DQK/32 chunks of [128 x 32 fp32]
  static constexpr bool BIG = (DQK > 64) || (DV > 64);
  static constexpr int STAGES = BIG ? 1 : 2;
  static constexpr bool ALIAS_STG = BIG;  // big heads: the dQ staging re-uses the P/dS tiles
  static constexpr int SMEM_BYTES =
      K_BYTES + V_BYTES + STAGES * (Q_BYTES + DO_BYTES) + PD_BYTES + DS_BYTES + (ALIAS_STG ? 0 : STG_BYTES) + 512 + 1024;
  static_assert(SMEM_BYTES <= 232448, "shared memory budget of one CTA");
  // TMEM columns: S | dP | dV | dK | dQ  (dQ aliases dP -- or, 192 wide, S | dP -- when the budget would exceed 512)
  static constexpr int S_COL = 0, DP_COL = 128, DV_COL = 256, DK_COL = (PART == 2) ? 256 : 256 + DV;
  static constexpr bool ALIAS_DQ = WIDE || (256 + DV + DQK + DQK) > 512;
  static constexpr int DQ_COL = ALIAS_DQ ? (DQK > 128 ? S_COL : DP_COL) : (256 + DV + DQK);
  static_assert((!HAS_DS || DK_COL + DQK <= 512) && (!HAS_P || DV_COL + DV <= 512) && (!ALIAS_DQ || DQ_COL + DQK <= 256), "TMEM budget");
  static constexpr int TMEM_COLS = 512;
  static constexpr int THREADS = 10 * 32;
};

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int DQK, int DV, int PART>
__global__ void __launch_bounds__(AttnBwdCfg<DQK, DV, PART>::THREADS, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                const __grid_constant__ CUtensorMap tmDQacc, const __grid_constant__ CUtensorMap tmDK,
                const __grid_constant__ CUtensorMap tmDV, const __grid_constant__ CUtensorMap tmDQ32,
                const AttnBwdParams p) {
  using Cfg = AttnBwdCfg<DQK, DV, PART>;
  constexpr int STAGES = Cfg::STAGES;
  constexpr bool HAS_P = Cfg::HAS_P, HAS_DS = Cfg::HAS_DS;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t k_smem = smem_base;
  const uint32_t v_smem = k_smem + Cfg::K_BYTES;
  const uint32_t q_smem = v_smem + Cfg::V_BYTES;                 // [STAGES][Q_BYTES]
  const uint32_t do_smem = q_smem + STAGES * Cfg::Q_BYTES;       // [STAGES][DO_BYTES]
  const uint32_t pd_smem = do_smem + STAGES * Cfg::DO_BYTES;
  const uint32_t ds_smem = pd_smem + Cfg::PD_BYTES;
  const uint32_t stg_smem = Cfg::ALIAS_STG ? pd_smem : ds_smem + Cfg::DS_BYTES;
  const uint32_t bar_base = (Cfg::ALIAS_STG ? ds_smem + Cfg::DS_BYTES : stg_smem + Cfg::STG_BYTES);
  const uint32_t kv_full = bar_base;
  auto qdo_full = [&](int s) { return bar_base + 8u * (1 + s); };
  auto qdo_empty = [&](int s) { return bar_base + 8u * (3 + s); };
  const uint32_t sdp_full = bar_base + 8u * 5;    // S and dP ready in TMEM            (tcgen05.commit)
  const uint32_t sdp_empty = bar_base + 8u * 6;   // compute warps have read S and dP   (8 warp-elected arrivals)
  const uint32_t pds_full = bar_base + 8u * 7;    // P and dS tiles written to smem     (8 warp-elected arrivals)
  const uint32_t dq_full = bar_base + 8u * 8;     // dV/dK/dQ GEMMs of this tile done   (tcgen05.commit)
  const uint32_t dq_empty = bar_base + 8u * 9;    // dQ drained from TMEM               (8 warp-elected arrivals)
  const uint32_t fin_full = bar_base + 8u * 10;   // all GEMMs of the CTA done
  const uint32_t tmem_ptr_smem = bar_base + 8u * 11;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int TMA_WARP = 8, MMA_WARP = 9;

  const int kv_tiles = (p.Skv + 127) / 128;
  const int jt = blockIdx.x % kv_tiles;
  const int bh = blockIdx.x / kv_tiles;
  const int h = bh % p.H, b = bh / p.H;
  const int kv0 = jt * 128;
  const int off = p.Skv - p.Sq;
  const int n_q = (p.Sq + 127) / 128;
  int i_start = 0;
  if (p.causal) i_start = max(0, (kv0 - off) / 128);
  const int n_it = n_q - i_start;

  if (warp == TMA_WARP && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmDO);
    tma_prefetch_desc(&tmDQacc);
    mbar_init(kv_full, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(qdo_full(s), 1);
      mbar_init(qdo_empty(s), 1);
    }
    mbar_init(sdp_full, 1);
    mbar_init(sdp_empty, 8);
    mbar_init(pds_full, 8);
    mbar_init(dq_full, 1);
    mbar_init(dq_empty, 8);
    mbar_init(fin_full, 1);
    fence_mbar_init();
  }
  if (warp == MMA_WARP) tmem_alloc(tmem_ptr_smem, Cfg::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_ptr_smem - smem_base));

  if (warp == TMA_WARP) {
    if (lane == 0) {
      mbar_expect_tx(kv_full, Cfg::K_BYTES + Cfg::V_BYTES);
      for (int c = 0; c < Cfg::NQK; ++c) tma_load_4d(k_smem + c * Cfg::CHUNK, &tmK, kv_full, c * 64, h, kv0, b);
      if constexpr (HAS_DS)
        for (int c = 0; c < Cfg::NV; ++c) tma_load_4d(v_smem + c * Cfg::CHUNK, &tmV, kv_full, c * 64, h, kv0, b);
      for (int it = 0; it < n_it; ++it) {
        const int st = it % STAGES;
        const int qrow = (i_start + it) * 128;
        mbar_wait(qdo_empty(st), ((it / STAGES) & 1) ^ 1);
        mbar_expect_tx(qdo_full(st), Cfg::Q_BYTES + Cfg::DO_BYTES);
        for (int c = 0; c < Cfg::NQK; ++c)
          tma_load_4d(q_smem + st * Cfg::Q_BYTES + c * Cfg::CHUNK, &tmQ, qdo_full(st), c * 64, h, qrow, b);
        for (int c = 0; c < Cfg::NV; ++c)
          tma_load_4d(do_smem + st * Cfg::DO_BYTES + c * Cfg::CHUNK, &tmDO, qdo_full(st), c * 64, h, qrow, b);
      }
    }
  } else if (warp == MMA_WARP) {
    if (lane == 0) {
      constexpr uint32_t id_s = make_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t id_dv = make_idesc_bf16(128, DV, 1, 1);
      constexpr uint32_t id_dk = make_idesc_bf16(128, DQK, 1, 1);
      constexpr uint32_t id_dq = make_idesc_bf16(128, DQK, 0, 1);
      auto issue_sdp = [&](int it) {   // S = Q K^T and dP = dO V^T of query tile `it`
        const int st = it % STAGES;
        const uint32_t qa = q_smem + st * Cfg::Q_BYTES;
        const uint32_t da = do_smem + st * Cfg::DO_BYTES;
#pragma unroll
        for (int kk = 0; kk < DQK / 16; ++kk) {
          const uint32_t o = (kk >> 2) * Cfg::CHUNK + (kk & 3) * 32;
          umma_f16_ss(tmem_base + Cfg::S_COL, make_smem_desc_sw128(qa + o, 16, 1024),
                      make_smem_desc_sw128(k_smem + o, 16, 1024), id_s, kk > 0 ? 1u : 0u);
        }
        if constexpr (HAS_DS) {
#pragma unroll
          for (int kk = 0; kk < DV / 16; ++kk) {
            const uint32_t o = (kk >> 2) * Cfg::CHUNK + (kk & 3) * 32;
            umma_f16_ss(tmem_base + Cfg::DP_COL, make_smem_desc_sw128(da + o, 16, 1024),
                        make_smem_desc_sw128(v_smem + o, 16, 1024), id_s, kk > 0 ? 1u : 0u);
          }
        }
        umma_commit(sdp_full);
      };
      auto issue_grads = [&](int it) {  // dV += P^T dO, dK += dS^T Q, dQ = dS K of query tile `it`
        const int st = it % STAGES;
        const uint32_t qa = q_smem + st * Cfg::Q_BYTES;
        const uint32_t da = do_smem + st * Cfg::DO_BYTES;
        if constexpr (HAS_P) {
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {  // K = 128 query rows; A: P tile MN-major (M = kv), B: dO tile MN-major (N = dv)
            umma_f16_ss(tmem_base + Cfg::DV_COL, make_smem_desc_sw128(pd_smem + kk * 2048, Cfg::CHUNK, 1024),
                        make_smem_desc_sw128(da + kk * 2048, Cfg::CHUNK, 1024), id_dv, (it > 0 || kk > 0) ? 1u : 0u);
          }
        }
        if constexpr (HAS_DS) {
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {  // A: dS tile MN-major, B: Q tile MN-major (N = dqk)
            umma_f16_ss(tmem_base + Cfg::DK_COL, make_smem_desc_sw128(ds_smem + kk * 2048, Cfg::CHUNK, 1024),
                        make_smem_desc_sw128(qa + kk * 2048, Cfg::CHUNK, 1024), id_dk, (it > 0 || kk > 0) ? 1u : 0u);
          }
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {  // K = 128 kv rows; A: dS tile K-major (M = q), B: K tile MN-major (N = dqk)
            umma_f16_ss(tmem_base + Cfg::DQ_COL,
                        make_smem_desc_sw128(ds_smem + (kk >> 2) * Cfg::CHUNK + (kk & 3) * 32, 16, 1024),
                        make_smem_desc_sw128(k_smem + kk * 2048, Cfg::CHUNK, 1024), id_dq, kk > 0 ? 1u : 0u);
          }
        }
        umma_commit(dq_full);
        umma_commit(qdo_empty(st));
      };
      mbar_wait(kv_full, 0);
      if constexpr (Cfg::ALIAS_DQ || STAGES < 2) {
        // dQ shares the dP columns (wide heads): strictly alternate  [S,dP](it) -> compute -> [dV,dK,dQ](it)
        for (int it = 0; it < n_it; ++it) {
          mbar_wait(qdo_full(it % STAGES), (it / STAGES) & 1);
          mbar_wait(sdp_empty, (it & 1) ^ 1);   // compute warps have consumed S/dP of tile it-1
          mbar_wait(dq_empty, (it & 1) ^ 1);    // dQ(it-1) has left the columns it shares with dP
          tc_fence_after();
          issue_sdp(it);
          mbar_wait(pds_full, it & 1);
          tc_fence_after();
          issue_grads(it);
        }
      } else {
        // Software pipeline: the compute warps release S/dP as soon as they hold them in registers, so
        // [S,dP](it+1) is issued while tile `it` is still in its exp/dropout phase, and [dV,dK,dQ](it) runs while
        // the warps already work on tile it+1.  The tensor pipe never waits for the softmax math and vice versa.
        mbar_wait(qdo_full(0), 0);
        tc_fence_after();
        issue_sdp(0);
        for (int it = 0; it < n_it; ++it) {
          if (it + 1 < n_it) {
            mbar_wait(qdo_full((it + 1) % STAGES), ((it + 1) / STAGES) & 1);
            mbar_wait(sdp_empty, it & 1);       // S/dP of tile `it` are in the compute warps' registers
            tc_fence_after();
            issue_sdp(it + 1);
          }
          mbar_wait(pds_full, it & 1);          // P / dS tiles of tile `it` are in shared memory
          mbar_wait(dq_empty, (it & 1) ^ 1);    // dQ(it-1) drained from TMEM
          tc_fence_after();
          issue_grads(it);
        }
      }
      umma_commit(fin_full);
    }
    __syncwarp();
  } else {
    // ======================= compute warps: thread = (query row, column half) =======================
    const int hf = warp >> 2;
    const int row = (warp & 3) * 32 + lane;
    const uint32_t lane_off = uint32_t((warp & 3) * 32) << 16;
    const uint32_t t_s = tmem_base + lane_off + Cfg::S_COL + hf * 64;
    const uint32_t t_dp = tmem_base + lane_off + Cfg::DP_COL + hf * 64;
    const uint32_t t_dq = tmem_base + lane_off + Cfg::DQ_COL + hf * (DQK / 2);
    const float LOG2E = 1.4426950408889634f;
    const unsigned long long seed_eff = p.seed + ((p.drop_thresh && p.seed_off) ? *p.seed_off : 0ull);

    auto drain_dq = [&](int it_prev) {
      // dQ tile of iteration it_prev: TMEM -> fp32 staging -> TMA reduce-add into the accumulator
      mbar_wait(dq_full, it_prev & 1);
      tc_fence_after();
      if constexpr (!HAS_DS) {   // PART 1 has no dQ: the barrier pair only says "the dV GEMM has read the P tile"
        __syncwarp();
        if (lane == 0) mbar_arrive(dq_empty);
        return;
      }
      if constexpr (DQK > 128) {
        // 192 columns of dQ through the 32 KB dS tile (free: its readers, the GEMMs of this tile, are done), 64 columns
        // per pass: thread (row, hf) stages columns [64 * pass + 32 * hf, + 32) into chunk hf
        const int qrow = (i_start + it_prev) * 128;
#pragma unroll 1
        for (int pass = 0; pass < DQK / 64; ++pass) {
          if (warp == 0 && lane == 0) tma_wait_group_read<0>();   // the previous reduce has read the staging
          named_bar_sync(1, 256);
          uint32_t r[32];
          tmem_ld_x32(tmem_base + lane_off + Cfg::DQ_COL + pass * 64 + hf * 32, r);
          tmem_ld_wait();
          const uint32_t dst_row = stg_smem + hf * Cfg::CHUNK + row * 128;
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst_row + ((q ^ (row & 7)) << 4)),
                         "r"(__float_as_uint(__uint_as_float(r[4 * q]) * p.scale)),
                         "r"(__float_as_uint(__uint_as_float(r[4 * q + 1]) * p.scale)),
                         "r"(__float_as_uint(__uint_as_float(r[4 * q + 2]) * p.scale)),
                         "r"(__float_as_uint(__uint_as_float(r[4 * q + 3]) * p.scale))
                         : "memory");
          }
          if (pass == DQK / 64 - 1) {   // last read of the dQ columns: the next tile's S / dP may overwrite them
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(dq_empty);
          }
          fence_proxy_async_smem();
          named_bar_sync(1, 256);
          if (warp == 0 && lane == 0) {
            for (int c = 0; c < 2; ++c)
              if (pass * 64 + c * 32 < p.dh) tma_reduce_add_4d(&tmDQacc, stg_smem + c * Cfg::CHUNK, pass * 64 + c * 32, h, qrow, b);
            tma_commit_group();
          }
        }
        if (warp == 0 && lane == 0) tma_wait_group_read<0>();
        named_bar_sync(1, 256);   // the staging is the dS tile that is written next
        return;
      }
      if constexpr (!Cfg::ALIAS_STG) {
        // Warp-local drain (dedicated staging): every warp owns the 32 rows x (DQK/2) columns it reads from TMEM, stages
        // them in its own 4 KB slices and issues its own TMA reduce-add (32-row boxes), so no CTA-wide barrier sits between
        // the eight compute warps in the steady state.
        if (lane == 0) tma_wait_group_read<0>();   // this warp's previous reduce has read its slices
        __syncwarp();
#pragma unroll
        for (int c = 0; c < DQK / 64; ++c) {
          uint32_t r[32];
          tmem_ld_x32(t_dq + c * 32, r);
          tmem_ld_wait();
          const int chunk = hf * (DQK / 64) + c;
          const uint32_t dst_row = stg_smem + chunk * Cfg::CHUNK + row * 128;
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst_row + ((q ^ (row & 7)) << 4)),
                         "r"(__float_as_uint(__uint_as_float(r[4 * q]) * p.scale)),
                         "r"(__float_as_uint(__uint_as_float(r[4 * q + 1]) * p.scale)),
                         "r"(__float_as_uint(__uint_as_float(r[4 * q + 2]) * p.scale)),
                         "r"(__float_as_uint(__uint_as_float(r[4 * q + 3]) * p.scale))
                         : "memory");
          }
        }
        tc_fence_before();
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(dq_empty);
          const int qrow = (i_start + it_prev) * 128 + (warp & 3) * 32;
          if (qrow < p.Sq) {
#pragma unroll
            for (int c = 0; c < DQK / 64; ++c) {
              const int chunk = hf * (DQK / 64) + c;
              if (chunk * 32 < p.dh)
                tma_reduce_add_4d(&tmDQ32, stg_smem + chunk * Cfg::CHUNK + (warp & 3) * 4096, chunk * 32, h, qrow, b);
            }
          }
          tma_commit_group();
        }
        return;
      }
      if (warp == 0 && lane == 0) tma_wait_group_read<0>();  // previous reduce has finished reading the staging
      named_bar_sync(1, 256);
#pragma unroll
      for (int c = 0; c < DQK / 64; ++c) {
        uint32_t r[32];
        tmem_ld_x32(t_dq + c * 32, r);
        tmem_ld_wait();
        const int chunk = hf * (DQK / 64) + c;  // 32-column fp32 chunk index
        const uint32_t dst_row = stg_smem + chunk * Cfg::CHUNK + row * 128;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst_row + ((q ^ (row & 7)) << 4)),
                       "r"(__float_as_uint(__uint_as_float(r[4 * q]) * p.scale)),
                       "r"(__float_as_uint(__uint_as_float(r[4 * q + 1]) * p.scale)),
                       "r"(__float_as_uint(__uint_as_float(r[4 * q + 2]) * p.scale)),
                       "r"(__float_as_uint(__uint_as_float(r[4 * q + 3]) * p.scale))
                       : "memory");
        }
      }
      tc_fence_before();
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(dq_empty);
      named_bar_sync(1, 256);
      if (warp == 0 && lane == 0) {
        const int qrow = (i_start + it_prev) * 128;
        for (int c = 0; c < DQK / 32; ++c)
          if (c * 32 < p.dh) tma_reduce_add_4d(&tmDQacc, stg_smem + c * Cfg::CHUNK, c * 32, h, qrow, b);
        tma_commit_group();
        if (Cfg::ALIAS_STG) tma_wait_group_read<0>();
      }
      if (Cfg::ALIAS_STG) named_bar_sync(1, 256);  // staging aliases the P/dS tiles that are written next
    };

    // lse / delta of the NEXT query tile are fetched while the current one is processed (two dependent global loads per
    // tile would otherwise sit at the head of every iteration)
    const long long lbase = ((long long)(b * p.H + h)) * p.Sq;
    float lse_nxt = p.lse[lbase + min(i_start * 128 + row, p.Sq - 1)];
    float dlt_nxt = p.delta[lbase + min(i_start * 128 + row, p.Sq - 1)];
    for (int it = 0; it < n_it; ++it) {
      const int q_idx = (i_start + it) * 128 + row;
      const bool row_ok = q_idx < p.Sq;
      const float lse2 = lse_nxt * LOG2E;
      const float dlt = dlt_nxt;
      if (it + 1 < n_it) {
        const long long nidx = lbase + min(q_idx + 128, p.Sq - 1);
        lse_nxt = p.lse[nidx];
        dlt_nxt = p.delta[nidx];
      }
      const int causal_lim = q_idx + off;
      constexpr bool PIPELINED = !(Cfg::ALIAS_DQ || STAGES < 2);   // must match the MMA issuer's ordering
      if (!PIPELINED && it > 0) drain_dq(it - 1);
      mbar_wait(sdp_full, it & 1);
      tc_fence_after();
      const unsigned long long drop_row = ((unsigned long long)(b * p.H + h) * p.Sq + (unsigned long long)min(q_idx, p.Sq - 1)) *
                                          (unsigned long long)p.Skv;
      // this thread's 64 columns of S and dP in two 32-column halves: the second half is in flight (TMEM reads are
      // bandwidth-bound, all compute warps ask at the same moment) while the arithmetic of the first one runs
      uint32_t rs[64], rd[64];
      tmem_ld_x32(t_s, rs);
      if constexpr (HAS_DS) tmem_ld_x32(t_dp, rd);
      tmem_ld_wait();
      tmem_ld_x32(t_s + 32, rs + 32);
      if constexpr (HAS_DS) tmem_ld_x32(t_dp + 32, rd + 32);
      float* pv = reinterpret_cast<float*>(rs);    // P (then dropped-out P) in place of S
      float* dsv = reinterpret_cast<float*>(rd);   // dS in place of dP
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        if (c == 1) {
          tmem_ld_wait();
          if (PIPELINED) {   // S/dP may be overwritten by the next tile's GEMMs while we do the rest of the math
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(sdp_empty);
          }
        }
        const int kvb = kv0 + hf * 64 + c * 32;
        const bool need_mask = !row_ok || (kvb + 32 > p.Skv) || (p.causal && (kvb + 31 > causal_lim));
#pragma unroll
        for (int i = 0; i < 32; ++i) pv[c * 32 + i] = ex2f(fmaf(pv[c * 32 + i], p.sl2, -lse2));
        if (need_mask) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int kv = kvb + i;
            const bool ok = row_ok && (kv < p.Skv) && (!p.causal || kv <= causal_lim);
            pv[c * 32 + i] = ok ? pv[c * 32 + i] : 0.f;
          }
        }
        if (p.drop_thresh) {
          const DropoutStream dstr(seed_eff, (drop_row + (unsigned long long)kvb) >> 1);
          const bool nw = dstr.nowrap(16);
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const DropPair hsh = nw ? dstr.bits_nowrap(i >> 1) : dstr.bits(i >> 1);
            const float k0 = (hsh.a >= p.drop_thresh) ? p.drop_scale : 0.f;
            const float k1 = (hsh.b >= p.drop_thresh) ? p.drop_scale : 0.f;
            const int e = c * 32 + i;
            if constexpr (HAS_DS) {
              dsv[e] = pv[e] * fmaf(dsv[e], k0, -dlt);        // softmax scale is applied when dQ / dK leave TMEM
              dsv[e + 1] = pv[e + 1] * fmaf(dsv[e + 1], k1, -dlt);
            }
            pv[e] *= k0;
            pv[e + 1] *= k1;
          }
        } else if constexpr (HAS_DS) {
#pragma unroll
          for (int i = 0; i < 32; ++i) dsv[c * 32 + i] = pv[c * 32 + i] * (dsv[c * 32 + i] - dlt);
        }
      }
      // the previous tile's [dV,dK,dQ] GEMMs must have finished reading the P/dS tiles before they are rewritten;
      // waiting for them (and draining dQ) only now keeps that latency off the critical path
      if (PIPELINED && it > 0) drain_dq(it - 1);
      const uint32_t prow = pd_smem + hf * Cfg::CHUNK + row * 128;
      const uint32_t drow = ds_smem + hf * Cfg::CHUNK + row * 128;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const uint32_t sw = (((uint32_t)q) ^ (row & 7)) << 4;
        if constexpr (HAS_P)
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(prow + sw), "r"(pack_bf16(pv[8 * q], pv[8 * q + 1])),
                       "r"(pack_bf16(pv[8 * q + 2], pv[8 * q + 3])), "r"(pack_bf16(pv[8 * q + 4], pv[8 * q + 5])),
                       "r"(pack_bf16(pv[8 * q + 6], pv[8 * q + 7]))
                       : "memory");
        if constexpr (HAS_DS)
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(drow + sw), "r"(pack_bf16(dsv[8 * q], dsv[8 * q + 1])),
                       "r"(pack_bf16(dsv[8 * q + 2], dsv[8 * q + 3])), "r"(pack_bf16(dsv[8 * q + 4], dsv[8 * q + 5])),
                       "r"(pack_bf16(dsv[8 * q + 6], dsv[8 * q + 7]))
                       : "memory");
      }
      tc_fence_before();
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        if (!PIPELINED) mbar_arrive(sdp_empty);
        mbar_arrive(pds_full);
      }
    }
    drain_dq(n_it - 1);
    // ---- dV, dK: TMEM -> bf16 -> swizzled staging (the P / dS tiles are free now) -> TMA store ----
    mbar_wait(fin_full, 0);
    tc_fence_after();
    if (lane == 0) tma_wait_group_read<0>();   // (every warp issued its own dQ reduces)
    named_bar_sync(1, 256);
    auto stage_out = [&](uint32_t tcol, int ncols, uint32_t dst_base, float sc) {
      // this thread owns columns [hf*ncols/2, (hf+1)*ncols/2) of its kv row
      const int per = ncols / 2;
#pragma unroll
      for (int c = 0; c < per / 32; ++c) {
        uint32_t r[32];
        tmem_ld_x32(tmem_base + lane_off + tcol + hf * per + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int col = hf * per + c * 32 + g * 8;
          const uint32_t dst = dst_base + (col >> 6) * Cfg::CHUNK + row * 128 + (((uint32_t)((col & 63) >> 3) ^ (row & 7)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst),
                       "r"(pack_bf16(__uint_as_float(r[8 * g]) * sc, __uint_as_float(r[8 * g + 1]) * sc)),
                       "r"(pack_bf16(__uint_as_float(r[8 * g + 2]) * sc, __uint_as_float(r[8 * g + 3]) * sc)),
                       "r"(pack_bf16(__uint_as_float(r[8 * g + 4]) * sc, __uint_as_float(r[8 * g + 5]) * sc)),
                       "r"(pack_bf16(__uint_as_float(r[8 * g + 6]) * sc, __uint_as_float(r[8 * g + 7]) * sc))
                       : "memory");
        }
      }
    };
    // wide heads (three 16 KB chunks per output) leave through the Q stage: every GEMM that read it has completed
    const uint32_t dv_out = Cfg::WIDE ? q_smem : pd_smem, dk_out = Cfg::WIDE ? q_smem : ds_smem;
    if constexpr (HAS_P) stage_out(Cfg::DV_COL, DV, dv_out, 1.f);
    if constexpr (HAS_DS) stage_out(Cfg::DK_COL, DQK, dk_out, p.scale);
    fence_proxy_async_smem();
    named_bar_sync(1, 256);
    if (warp == 0 && lane == 0) {
      if constexpr (HAS_P)
        for (int c = 0; c < Cfg::NV; ++c)
          if (c * 64 < p.dh) tma_store_4d(&tmDV, dv_out + c * Cfg::CHUNK, c * 64, h, kv0, b);
      if constexpr (HAS_DS)
        for (int c = 0; c < Cfg::NQK; ++c)
          if (c * 64 < p.dh) tma_store_4d(&tmDK, dk_out + c * Cfg::CHUNK, c * 64, h, kv0, b);
      tma_commit_group();
      tma_wait_group<0>();
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

// delta[b,h,s] = sum_d dO[b,s,h,d] * O[b,s,h,d];  one warp per (token, head).  Also zeroes the dQ accumulator.
__global__ void __launch_bounds__(256) attn_bwd_prep_kernel(const bf16* __restrict__ o, const bf16* __restrict__ dO,
                                                            long long ldo, float* __restrict__ delta,
                                                            float* __restrict__ dq_acc, int B, int H, int S, int dh) {
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  const long long w0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const long long total = (long long)B * S * H;
  for (long long w = w0; w < total; w += nwarps) {
    const int h = (int)(w % H);
    const long long tok = w / H;
    const bf16* po = o + tok * ldo + (long long)h * dh;
    const bf16* pd = dO + tok * ldo + (long long)h * dh;
    float s = 0.f;
    for (int c = lane; c < dh / 8; c += 32) {
      const uint4 a = *reinterpret_cast<const uint4*>(po + c * 8);
      const uint4 g = *reinterpret_cast<const uint4*>(pd + c * 8);
      const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 x = unpack_bf16(aw[q]), y = unpack_bf16(gw[q]);
        s += x.x * y.x + x.y * y.y;
      }
    }
    s = warp_sum(s);
    const int bb = (int)(tok / S), ss = (int)(tok % S);
    if (lane == 0) delta[((long long)bb * H + h) * S + ss] = s;
    float* acc = dq_acc + (tok * H + h) * (long long)dh;
    for (int c = lane; c < dh / 4; c += 32) reinterpret_cast<float4*>(acc)[c] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// dq[tok, h*dh + d] (bf16, row stride ldq) = dq_acc[tok, h, d]
__global__ void __launch_bounds__(256) attn_bwd_finish_kernel(const float* __restrict__ dq_acc, bf16* __restrict__ dq,
                                                              long long ldq, long long tokens, int D) {
  const long long n8 = tokens * (D / 8);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    const long long tok = i / (D / 8);
    const int c = (int)(i % (D / 8));
    const float4 a = reinterpret_cast<const float4*>(dq_acc + tok * D + c * 8)[0];
    const float4 bq = reinterpret_cast<const float4*>(dq_acc + tok * D + c * 8)[1];
    uint4 o;
    o.x = pack_bf16(a.x, a.y);
    o.y = pack_bf16(a.z, a.w);
    o.z = pack_bf16(bq.x, bq.y);
    o.w = pack_bf16(bq.z, bq.w);
    *reinterpret_cast<uint4*>(dq + tok * ldq + c * 8) = o;
  }
}

template <int DQK, int DV, int PART>
static int launch_attn_bwd(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const CUtensorMap& tdo,
                           const CUtensorMap& tdq, const CUtensorMap& tdk, const CUtensorMap& tdv, const CUtensorMap& tdq32,
                           const AttnBwdParams& p, cudaStream_t stream) {
  using Cfg = AttnBwdCfg<DQK, DV, PART>;
  auto kfn = attn_bwd_kernel<DQK, DV, PART>;
  PPB_CUDA_OK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
  const int kv_tiles = (p.Skv + 127) / 128;
  kfn<<<p.B * p.H * kv_tiles, Cfg::THREADS, Cfg::SMEM_BYTES, stream>>>(tq, tk, tv, tdo, tdq, tdk, tdv, tdq32, p);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

int attn_bwd_generic(const void* q, const void* k, const void* v, const void* o, const void* dO, const float* lse,
                     void* dq, void* dk, void* dv, void* scratch, int B, int H, int Sq, int Skv, int dh, long long ldq,
                     long long ldk, long long ldv, long long ldo, int causal, float scale, float dropout_p,
                     unsigned long long seed, cudaStream_t stream);

}  // namespace ppb

using namespace ppb;

// `S` = max(Sq, Skv).  tcgen05 path: delta + fp32 dQ accumulator; generic path: fp32 dK and dV accumulators (sized for
// either path when dh > 128, so the switch below never changes a caller's buffer).
extern "C" long long ppb_attn_bwd_scratch_bytes(int B, int H, int S, int dh) {
  const long long delta = ((long long)B * H * S * 4 + 255) / 256 * 256;
  return delta + (long long)B * S * H * dh * 4 * (dh > 128 ? 2 : 1) + 256;
}

extern "C" int ppb_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* dO,
                            const float* lse, void* dq, void* dk, void* dv, void* scratch, void* reserved_ptr, int B,
                            int H, int Sq, int Skv, int dh, long long ldq, long long ldk, long long ldv, long long ldo,
                            int causal, int reserved0, int reserved1, float scale, float dropout_p,
                            unsigned long long seed, void* stream_) {
  (void)reserved_ptr;
  (void)reserved0;
  (void)reserved1;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(B > 0 && H > 0 && Sq > 0 && Skv > 0, "ppb_attn_bwd: empty problem");
  PPB_REQUIRE(dh % 8 == 0 && dh >= 8 && dh <= 256, "ppb_attn_bwd: head dim must be a multiple of 8 in [8, 256] (got %d)", dh);
  PPB_REQUIRE((ldq % 8) == 0 && (ldk % 8) == 0 && (ldv % 8) == 0 && (ldo % 8) == 0, "ppb_attn_bwd: row strides must be multiples of 8");
  PPB_REQUIRE(!causal || Skv >= Sq, "ppb_attn_bwd: causal attention needs Skv >= Sq");
  PPB_REQUIRE(dropout_p <= 0.f || (Skv % 2) == 0, "ppb_attn_bwd: dropout needs an even Skv");
  PPB_REQUIRE(scratch != nullptr && (reinterpret_cast<uintptr_t>(scratch) % 256) == 0, "ppb_attn_bwd: scratch must be 256-byte aligned");
  if (dh > attn_tc_max_head_dim())   // CUDA-core path (attention_generic.cu)
    return attn_bwd_generic(q, k, v, o, dO, lse, dq, dk, dv, scratch, B, H, Sq, Skv, dh, ldq, ldk, ldv, ldo, causal, scale,
                            dropout_p, seed, stream);
  float* delta = reinterpret_cast<float*>(scratch);
  const long long delta_bytes = ((long long)B * H * Sq * 4 + 255) / 256 * 256;
  float* dq_acc = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(scratch) + delta_bytes);
  const int D = H * dh;

  {
    const long long warps = (long long)B * Sq * H;
    long long blocks = (warps + 7) / 8;
    const long long cap = (long long)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    attn_bwd_prep_kernel<<<(int)blocks, 256, 0, stream>>>((const bf16*)o, (const bf16*)dO, ldo, delta, dq_acc, B, H, Sq, dh);
    PPB_CUDA_OK(cudaGetLastError());
  }

  CUtensorMap tq, tk, tv, tdo, tdq, tdk, tdv, tdq32;
  if (!make_head_tmap(&tq, q, dh, H, Sq, B, ldq, 128) || !make_head_tmap(&tk, k, dh, H, Skv, B, ldk, 128) ||
      !make_head_tmap(&tv, v, dh, H, Skv, B, ldv, 128) || !make_head_tmap(&tdo, dO, dh, H, Sq, B, ldo, 128) ||
      !make_head_tmap(&tdk, dk, dh, H, Skv, B, ldk, 128) || !make_head_tmap(&tdv, dv, dh, H, Skv, B, ldv, 128))
    return -3;
  {
    uint64_t dims[4] = {(uint64_t)dh, (uint64_t)H, (uint64_t)Sq, (uint64_t)B};
    uint64_t strides[3] = {(uint64_t)dh * 4, (uint64_t)D * 4, (uint64_t)Sq * (uint64_t)D * 4};
    uint32_t box[4] = {32, 1, 128, 1};
    if (!make_tmap(&tdq, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, dq_acc, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B)) return -3;
    uint32_t box32[4] = {32, 1, 32, 1};   // one warp's 32 query rows
    if (!make_tmap(&tdq32, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, dq_acc, dims, strides, box32, CU_TENSOR_MAP_SWIZZLE_128B)) return -3;
  }
  AttnBwdParams p;
  p.B = B;
  p.H = H;
  p.Sq = Sq;
  p.Skv = Skv;
  p.dh = dh;
  p.causal = causal;
  p.scale = scale;
  p.sl2 = scale * 1.4426950408889634f;
  p.seed = seed;
  p.seed_off = seed_offset_ptr();
  p.drop_thresh = dropout_p > 0.f ? dropout_thr32(dropout_p) : 0u;
  p.drop_scale = dropout_p < 1.f ? 1.f / (1.f - dropout_p) : 0.f;
  p.lse = lse;
  p.delta = delta;
  int rc;
  if (dh <= 64) {
    rc = launch_attn_bwd<64, 64, 0>(tq, tk, tv, tdo, tdq, tdk, tdv, tdq32, p, stream);
  } else if (dh <= 128) {
    rc = launch_attn_bwd<128, 128, 0>(tq, tk, tv, tdo, tdq, tdk, tdv, tdq32, p, stream);
  } else {   // 129..192: dV, then dK + dQ (see the header comment)
    rc = launch_attn_bwd<192, 192, 1>(tq, tk, tv, tdo, tdq, tdk, tdv, tdq32, p, stream);
    if (rc == 0) rc = launch_attn_bwd<192, 192, 2>(tq, tk, tv, tdo, tdq, tdk, tdv, tdq32, p, stream);
  }
  if (rc) return rc;
  {
    const long long n8 = (long long)B * Sq * (D / 8);
    long long blocks = (n8 + 255) / 256;
    const long long cap = (long long)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    attn_bwd_finish_kernel<<<(int)blocks, 256, 0, stream>>>(dq_acc, (bf16*)dq, ldq, (long long)B * Sq, D);
    PPB_CUDA_OK(cudaGetLastError());
  }
  return 0;
}
