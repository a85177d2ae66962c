// K4 forward: flash attention on tcgen05 tensor cores with TMEM accumulators.
//
// One CTA handles MT (1 or 2) query tiles of 128 rows for one (batch, head) and streams 128-row K/V tiles:
//   * warp 2*MT*... layout: warps [0, 4*MT) = softmax warpgroups (one per query tile; thread == query row,
//     which is exactly the TMEM lane it may read with tcgen05.ld 32x32b, so row max / row sum need no shuffles),
//     warp 4*MT = TMA producer, warp 4*MT+1 = MMA issuer (+ TMEM allocation);
//   * S = Q K^T   : tcgen05.mma, both operands K-major from 128B-swizzled smem, accumulator S_t in TMEM;
//   * softmax     : online (running max / sum in the exp2 domain), optional causal mask and probability
//                   dropout (counter-based mask, regenerated in backward), P written as bf16 into swizzled smem;
//   * O_tile = P V: tcgen05.mma with V as an MN-major B operand (no transpose of V anywhere), result in TMEM,
//                   drained into fp32 registers and rescaled there (O never round-trips through HBM);
//   * with MT = 2 the two softmax warpgroups ping-pong against the single MMA issuer, so the tensor pipe
//     works on one query tile while the other one is in its exp2 phase.
// Inputs are column-sliced views of the packed QKV activation ([tokens, 3*d]); heads are addressed through 4-D
// TMA tensor maps (dh, head, seq, batch), so no split/transposes/.contiguous() copies exist.  Head dims that are
// not multiples of 64 (96 in the reference sweep) are zero-padded by TMA out-of-bounds fill.  Head dims 129..192 (the
// sweep's 4-head rows, 768 / 4) run one query tile per CTA with TWO threads per query row (96 output columns each,
// so the fp32 accumulator still lives in registers), single-stage K / V tiles of three 64-column chunks, 320 TMEM
// columns; O leaves through the Q tile (three chunks), which is dead after the last Q K^T.
//
// Replaces aten.bmm x2 + mul.Scalar + _safe_softmax + dropout of the reference's attention blocks
// (self- and cross-attention of nn.TransformerDecoderLayer,
// /root/reference/LLMsDistributedTrainingHelper.py:40-44, :52).
#include <stdlib.h>

#include "common.cuh"
#include "host_common.h"
#include "ppb200.h"

namespace ppb {

struct AttnParams {
  int B, H, Sq, Skv, dh;
  int causal;
  float sl2;  // softmax scale * log2(e)
  unsigned long long seed;
  const unsigned long long* seed_off;
  unsigned int drop_thresh;
  float drop_scale;
  float* lse;  // [B, H, Sq] natural-log LSE
};

template <int DQK, int DV, int MT, int SPLIT>
struct AttnFwdCfg {
  static_assert(DV <= 128 || (MT == 1 && SPLIT == 2 && DQK >= DV), "wide heads: one query tile, two threads per row, O staged in Q");
  static constexpr int NQK = DQK / 64, NV = DV / 64;
  static constexpr int CHUNK_BYTES = 128 * 128;  // 128 rows x 64 bf16 (one 128B-swizzled column chunk)
  static constexpr int Q_BYTES = NQK * CHUNK_BYTES;
  static constexpr int K_BYTES = NQK * CHUNK_BYTES;
  static constexpr int V_BYTES = NV * CHUNK_BYTES;
  static constexpr int P_BYTES = 2 * CHUNK_BYTES;
  static constexpr int FIXED = MT * (Q_BYTES + P_BYTES);
  static constexpr int STAGES = (FIXED + 2 * (K_BYTES + V_BYTES) <= 220 * 1024) ? 2 : 1;
  static constexpr int SMEM_BYTES = FIXED + STAGES * (K_BYTES + V_BYTES) + 512 + 1024;
  static constexpr int TMEM_NEED = MT * (128 + DV);
  static constexpr int TMEM_COLS = TMEM_NEED <= 128 ? 128 : (TMEM_NEED <= 256 ? 256 : 512);
  static constexpr int SWARPS = 4 * MT * SPLIT;   // softmax warps: SPLIT threads share one query row (column halves)
  static constexpr int THREADS = (SWARPS + 2) * 32;
  // RESIDENT (head dims up to 64, one thread per row): O stays in TMEM for the whole kv loop (every P V accumulates onto
  // it) and the row's 128 scores are read from TMEM ONCE and kept in registers for both the max and the exp2 pass.  TMEM
  // reads (64 B/clk per SM) are this kernel's scarcest resource: scores twice + the P V read-back were 160 KB per tile,
  // now 64 KB.  The exponent reference of a row only moves when the running maximum grew by more than 8 (factor 256, safe
  // in bf16 / fp32), so the in-TMEM rescale (tcgen05.ld, multiply, tcgen05.st) runs on the first tiles only.
  static constexpr bool RESIDENT = (SPLIT == 1 && DQK == 64 && DV == 64);
};

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int DQK, int DV, int MT, int SPLIT>
__global__ void __launch_bounds__(AttnFwdCfg<DQK, DV, MT, SPLIT>::THREADS, 1)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO, const AttnParams p) {
  using Cfg = AttnFwdCfg<DQK, DV, MT, SPLIT>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  __shared__ float xch[2][2][2][128];  // [parity][q tile][column half][row]: row-max / row-sum exchange (SPLIT == 2)
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t q_smem = smem_base;                          // [MT][Q_BYTES]
  const uint32_t p_smem = q_smem + MT * Cfg::Q_BYTES;         // [MT][P_BYTES]
  const uint32_t k_smem = p_smem + MT * Cfg::P_BYTES;         // [STAGES][K_BYTES]
  const uint32_t v_smem = k_smem + STAGES * Cfg::K_BYTES;     // [STAGES][V_BYTES]
  const uint32_t bar_base = v_smem + STAGES * Cfg::V_BYTES;
  // barriers
  const uint32_t q_full = bar_base;
  auto k_full = [&](int s) { return bar_base + 8u * (1 + s); };
  auto k_empty = [&](int s) { return bar_base + 8u * (3 + s); };
  auto v_full = [&](int s) { return bar_base + 8u * (5 + s); };
  auto v_empty = [&](int s) { return bar_base + 8u * (7 + s); };
  auto s_full = [&](int t) { return bar_base + 8u * (9 + t); };
  auto s_empty = [&](int t) { return bar_base + 8u * (11 + t); };
  auto p_full = [&](int t) { return bar_base + 8u * (13 + t); };
  auto pv_full = [&](int t) { return bar_base + 8u * (15 + t); };
  auto pv_empty = [&](int t) { return bar_base + 8u * (17 + t); };
  const uint32_t tmem_ptr_smem = bar_base + 8u * 19;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int TMA_WARP = Cfg::SWARPS, MMA_WARP = Cfg::SWARPS + 1;

  const int q_tiles = (p.Sq + 128 * MT - 1) / (128 * MT);
  const int qt = blockIdx.x % q_tiles;
  const int bh = blockIdx.x / q_tiles;
  const int h = bh % p.H, b = bh / p.H;
  const int q0 = qt * 128 * MT;
  int kv_len = p.Skv;
  if (p.causal) kv_len = min(p.Skv, q0 + 128 * MT + (p.Skv - p.Sq));
  const int n_kv = (kv_len + 127) / 128;

  if (warp == TMA_WARP && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmO);
    mbar_init(q_full, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(k_full(s), 1);
      mbar_init(k_empty(s), 1);
      mbar_init(v_full(s), 1);
      mbar_init(v_empty(s), 1);
    }
    for (int t = 0; t < 2; ++t) {
      mbar_init(s_full(t), 1);
      mbar_init(s_empty(t), 4 * SPLIT);
      mbar_init(p_full(t), 4 * SPLIT);
      mbar_init(pv_full(t), 1);
      mbar_init(pv_empty(t), 4 * SPLIT);
    }
    fence_mbar_init();
  }
  if (warp == MMA_WARP) tmem_alloc(tmem_ptr_smem, Cfg::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_ptr_smem - smem_base));
  auto s_tmem = [&](int t) { return tmem_base + t * (128 + DV); };
  auto pv_tmem = [&](int t) { return tmem_base + t * (128 + DV) + 128; };

  if (warp == TMA_WARP) {
    if (lane == 0) {
      mbar_expect_tx(q_full, MT * Cfg::Q_BYTES);
      for (int t = 0; t < MT; ++t)
        for (int c = 0; c < Cfg::NQK; ++c)
          tma_load_4d(q_smem + t * Cfg::Q_BYTES + c * Cfg::CHUNK_BYTES, &tmQ, q_full, c * 64, h, q0 + t * 128, b);
      for (int j = 0; j < n_kv; ++j) {
        const int st = j % STAGES;
        const uint32_t par = ((j / STAGES) & 1) ^ 1;
        mbar_wait(k_empty(st), par);
        mbar_expect_tx(k_full(st), Cfg::K_BYTES);
        for (int c = 0; c < Cfg::NQK; ++c)
          tma_load_4d(k_smem + st * Cfg::K_BYTES + c * Cfg::CHUNK_BYTES, &tmK, k_full(st), c * 64, h, j * 128, b);
        mbar_wait(v_empty(st), par);
        mbar_expect_tx(v_full(st), Cfg::V_BYTES);
        for (int c = 0; c < Cfg::NV; ++c)
          tma_load_4d(v_smem + st * Cfg::V_BYTES + c * Cfg::CHUNK_BYTES, &tmV, v_full(st), c * 64, h, j * 128, b);
      }
    }
  } else if (warp == MMA_WARP) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t idesc_pv = make_idesc_bf16(128, DV, 0, 1);
      auto issue_s = [&](int t, int j) {
        const int st = j % STAGES;
        const uint32_t qa = q_smem + t * Cfg::Q_BYTES;
        const uint32_t kb = k_smem + st * Cfg::K_BYTES;
#pragma unroll
        for (int kk = 0; kk < DQK / 16; ++kk) {
          const uint32_t off = (kk >> 2) * Cfg::CHUNK_BYTES + (kk & 3) * 32;
          umma_f16_ss(s_tmem(t), make_smem_desc_sw128(qa + off, 16, 1024), make_smem_desc_sw128(kb + off, 16, 1024),
                      idesc_s, kk > 0 ? 1u : 0u);
        }
        umma_commit(s_full(t));
      };
      mbar_wait(q_full, 0);
      mbar_wait(k_full(0), 0);
      tc_fence_after();
      for (int t = 0; t < MT; ++t) issue_s(t, 0);
      umma_commit(k_empty(0));
      for (int j = 0; j < n_kv; ++j) {
        const int st = j % STAGES;
        const uint32_t par = (j / STAGES) & 1;
        mbar_wait(v_full(st), par);
        if (j + 1 < n_kv) mbar_wait(k_full((j + 1) % STAGES), ((j + 1) / STAGES) & 1);
        for (int t = 0; t < MT; ++t) {
          mbar_wait(p_full(t), j & 1);
          if constexpr (!Cfg::RESIDENT) mbar_wait(pv_empty(t), (j & 1) ^ 1);
          tc_fence_after();
          const uint32_t pa = p_smem + t * Cfg::P_BYTES;
          const uint32_t vb = v_smem + st * Cfg::V_BYTES;
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const uint32_t acc = (kk > 0 || (Cfg::RESIDENT && j > 0)) ? 1u : 0u;   // RESIDENT: O accumulates over all kv tiles
            umma_f16_ss(pv_tmem(t), make_smem_desc_sw128(pa + (kk >> 2) * Cfg::CHUNK_BYTES + (kk & 3) * 32, 16, 1024),
                        make_smem_desc_sw128(vb + kk * 2048, Cfg::CHUNK_BYTES, 1024), idesc_pv, acc);
          }
          // RESIDENT: pv_full is a one-shot "O of this query tile is final" barrier
          if (!Cfg::RESIDENT || j == n_kv - 1) umma_commit(pv_full(t));
          if (t == MT - 1) umma_commit(v_empty(st));
          if (j + 1 < n_kv) {
            mbar_wait(s_empty(t), j & 1);
            tc_fence_after();
            issue_s(t, j + 1);
            if (t == MT - 1) umma_commit(k_empty((j + 1) % STAGES));
          }
        }
      }
    }
    __syncwarp();
  } else {
    // ============ softmax warps of query tile t: SPLIT threads per query row (column halves) ============
    constexpr int COLS = 128 / SPLIT;   // kv columns of S handled by this thread
    constexpr int OD = DV / SPLIT;      // output columns accumulated by this thread
    const int t = warp / (4 * SPLIT);
    const int wl = warp % (4 * SPLIT);
    const int half = wl >> 2;
    const int row = (wl & 3) * 32 + lane;
    const int q_idx = q0 + t * 128 + row;
    const uint32_t lane_off = uint32_t((wl & 3) * 32) << 16;
    const int causal_lim = q_idx + (p.Skv - p.Sq);  // last visible kv index for this row
    const unsigned long long drop_row = ((unsigned long long)(b * p.H + h) * p.Sq + (unsigned long long)min(q_idx, p.Sq - 1)) *
                                        (unsigned long long)p.Skv;
    const unsigned long long seed_eff = p.seed + ((p.drop_thresh && p.seed_off) ? *p.seed_off : 0ull);
    float o[OD];
#pragma unroll
    for (int i = 0; i < OD; ++i) o[i] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;
    const uint32_t p_row = p_smem + t * Cfg::P_BYTES + row * 128;
    const uint32_t s_cols = s_tmem(t) + lane_off + half * COLS;
    const uint32_t pv_cols = pv_tmem(t) + lane_off + half * OD;

    if constexpr (Cfg::RESIDENT) {
      float m_ref = -INFINITY;   // exponent reference of this row (log2 domain): a lagging upper-ish bound of the running maximum
      for (int j = 0; j < n_kv; ++j) {
        mbar_wait(s_full(t), j & 1);
        tc_fence_after();
        uint32_t sc[128];
#pragma unroll
        for (int c = 0; c < 4; ++c) tmem_ld_x32(s_cols + c * 32, sc + c * 32);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(s_empty(t));   // S_t may be overwritten by the next Q K^T
        const int kv0 = j * 128;
        if ((kv0 + 128 > p.Skv) || (p.causal && (kv0 + 127 > causal_lim))) {   // sequence tail / causal diagonal
#pragma unroll
          for (int i = 0; i < 128; ++i) {
            const int kv = kv0 + i;
            if (!((kv < p.Skv) && (!p.causal || kv <= causal_lim))) sc[i] = 0xff800000u;   // exp2(-inf) == 0
          }
        }
        float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
        for (int i = 0; i < 128; i += 4) {
          m0 = fmaxf(m0, __uint_as_float(sc[i]));
          m1 = fmaxf(m1, __uint_as_float(sc[i + 1]));
          m2 = fmaxf(m2, __uint_as_float(sc[i + 2]));
          m3 = fmaxf(m3, __uint_as_float(sc[i + 3]));
        }
        const float mx = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3)) * p.sl2;
        const bool grow = mx > m_ref + 8.f;   // (first tile: m_ref == -inf)
        const float m_new = grow ? mx : m_ref;
        if (j > 0 && __any_sync(0xffffffffu, grow)) {
          // move the reference: O row *= 2^(m_ref - m_new) in TMEM.  P V of tile j-1 has completed: S of tile j, which this
          // thread has just read, was issued after it and tcgen05 operations of one thread complete in order.
          const float alpha = grow ? fast_exp2(m_ref - m_new) : 1.f;
          tc_fence_after();
#pragma unroll
          for (int c = 0; c < OD / 16; ++c) {
            uint32_t r[16];
            tmem_ld_x16(pv_cols + c * 16, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
            tmem_st_x16(pv_cols + c * 16, r);
          }
          tmem_st_wait();
          l_run *= alpha;
        }
        m_ref = m_new;
        const float m_use = (m_ref == -INFINITY) ? 0.f : m_ref;
        // ---- probabilities (in place of the scores) -> bf16 P tile in swizzled smem ----
        float ls0 = 0.f, ls1 = 0.f, ls2 = 0.f, ls3 = 0.f;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          float* pr = reinterpret_cast<float*>(sc + c * 32);
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            pr[i] = fast_exp2(fmaf(pr[i], p.sl2, -m_use));
            pr[i + 1] = fast_exp2(fmaf(pr[i + 1], p.sl2, -m_use));
            pr[i + 2] = fast_exp2(fmaf(pr[i + 2], p.sl2, -m_use));
            pr[i + 3] = fast_exp2(fmaf(pr[i + 3], p.sl2, -m_use));
            ls0 += pr[i];
            ls1 += pr[i + 1];
            ls2 += pr[i + 2];
            ls3 += pr[i + 3];
          }
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) pk[i] = pack_bf16(pr[2 * i], pr[2 * i + 1]);
          if (p.drop_thresh) {   // mask convention: see the two-pass loop below
            const DropoutStream ds(seed_eff, (drop_row + (unsigned long long)(kv0 + c * 32)) >> 1);
            if (ds.nowrap(16)) {
#pragma unroll
              for (int i = 0; i < 16; ++i) pk[i] &= drop_pair_mask(ds.word_nowrap(i), p.drop_thresh);
            } else {
#pragma unroll
              for (int i = 0; i < 16; ++i) pk[i] &= drop_pair_mask(ds.word(i), p.drop_thresh);
            }
          }
          const uint32_t atom = p_row + (c >> 1) * Cfg::CHUNK_BYTES;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint32_t cc = (c & 1) * 4 + q;
            const uint32_t dst = atom + ((cc ^ (row & 7)) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(pk[4 * q]), "r"(pk[4 * q + 1]),
                         "r"(pk[4 * q + 2]), "r"(pk[4 * q + 3])
                         : "memory");
          }
        }
        l_run += (ls0 + ls1) + (ls2 + ls3);
        tc_fence_before();
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(p_full(t));   // P_t is visible to the tensor core, the O row carries the current reference
      }
      m_run = m_ref;
    } else
    for (int j = 0; j < n_kv; ++j) {
      mbar_wait(s_full(t), j & 1);
      tc_fence_after();
      const int kv0 = j * 128 + half * COLS;
      const bool need_mask = (kv0 + COLS > p.Skv) || (p.causal && (kv0 + COLS - 1 > causal_lim));
      // ---- pass 1: row max over this thread's columns (+ exchange with the other half) ----
      float mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < COLS / 32; ++c) {
        uint32_t r[32];
        tmem_ld_x32(s_cols + c * 32, r);
        tmem_ld_wait();
        if (need_mask) {  // rare (sequence tail / causal diagonal): overwrite masked scores with -inf
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int kv = kv0 + c * 32 + i;
            if (!((kv < p.Skv) && (!p.causal || kv <= causal_lim))) r[i] = 0xff800000u;
          }
        }
        float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          m0 = fmaxf(m0, __uint_as_float(r[i]));
          m1 = fmaxf(m1, __uint_as_float(r[i + 1]));
        }
        mx = fmaxf(mx, fmaxf(m0, m1));
      }
      if constexpr (SPLIT == 2) {
        xch[j & 1][t][half][row] = mx;
        named_bar_sync(1 + t, 128 * SPLIT);
        mx = fmaxf(mx, xch[j & 1][t][half ^ 1][row]);
      }
      const float m_new = fmaxf(m_run, mx * p.sl2);
      const float alpha = (m_new == -INFINITY) ? 1.f : fast_exp2(m_run - m_new);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      // ---- fold the previous tile's P*V into the register accumulator, then rescale ----
      if (j > 0) {
        mbar_wait(pv_full(t), (j - 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < OD / 32; ++c) {
          uint32_t r[32];
          tmem_ld_x32(pv_cols + c * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[c * 32 + i] += __uint_as_float(r[i]);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(pv_empty(t));
      }
      if (alpha != 1.f) {
#pragma unroll
        for (int i = 0; i < OD; ++i) o[i] *= alpha;
      }
      l_run *= alpha;
      m_run = m_new;
      // ---- pass 2: probabilities -> bf16 P tile in swizzled smem ----
      float ls0 = 0.f, ls1 = 0.f, ls2 = 0.f, ls3 = 0.f;
#pragma unroll 1
      for (int c = 0; c < COLS / 32; ++c) {
        uint32_t r[32];
        tmem_ld_x32(s_cols + c * 32, r);
        tmem_ld_wait();
        if (need_mask) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int kv = kv0 + c * 32 + i;
            if (!((kv < p.Skv) && (!p.causal || kv <= causal_lim))) r[i] = 0xff800000u;  // exp2(-inf) == 0
          }
        }
        float pr[32];
#pragma unroll
        for (int i = 0; i < 32; i += 4) {   // four independent partial sums instead of one 32-long FADD chain
          pr[i] = fast_exp2(fmaf(__uint_as_float(r[i]), p.sl2, -m_use));
          pr[i + 1] = fast_exp2(fmaf(__uint_as_float(r[i + 1]), p.sl2, -m_use));
          pr[i + 2] = fast_exp2(fmaf(__uint_as_float(r[i + 2]), p.sl2, -m_use));
          pr[i + 3] = fast_exp2(fmaf(__uint_as_float(r[i + 3]), p.sl2, -m_use));
          ls0 += pr[i];
          ls1 += pr[i + 1];
          ls2 += pr[i + 2];
          ls3 += pr[i + 3];
        }
        // bf16 pairs first, then probability dropout as ONE half2 compare + ONE AND per pair (drop_pair_mask); kept
        // probabilities are NOT rescaled here: the constant 1/(1-p) is folded into the final 1/l
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) pk[i] = pack_bf16(pr[2 * i], pr[2 * i + 1]);
        if (p.drop_thresh) {
          const DropoutStream ds(seed_eff, (drop_row + (unsigned long long)(kv0 + c * 32)) >> 1);
          if (ds.nowrap(16)) {
#pragma unroll
            for (int i = 0; i < 16; ++i) pk[i] &= drop_pair_mask(ds.word_nowrap(i), p.drop_thresh);
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) pk[i] &= drop_pair_mask(ds.word(i), p.drop_thresh);
          }
        }
        const int col = half * COLS + c * 32;   // first kv column (within the 128-wide tile) of this chunk
        const uint32_t atom = p_row + (col >> 6) * Cfg::CHUNK_BYTES;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint32_t cc = ((col & 63) >> 3) + q;
          const uint32_t dst = atom + ((cc ^ (row & 7)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(pk[4 * q]), "r"(pk[4 * q + 1]),
                       "r"(pk[4 * q + 2]), "r"(pk[4 * q + 3])
                       : "memory");
        }
      }
      l_run += (ls0 + ls1) + (ls2 + ls3);
      tc_fence_before();
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(s_empty(t));  // S_t may be overwritten by the next QK^T
        mbar_arrive(p_full(t));   // P_t is visible to the tensor core
      }
    }
    // ---- last P*V (RESIDENT: the one-shot "O is final" phase), normalise, store ----
    mbar_wait(pv_full(t), Cfg::RESIDENT ? 0 : ((n_kv - 1) & 1));
    tc_fence_after();
#pragma unroll
    for (int c = 0; c < OD / 32; ++c) {
      uint32_t r[32];
      tmem_ld_x32(pv_cols + c * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) o[c * 32 + i] += __uint_as_float(r[i]);
    }
    if constexpr (SPLIT == 2) {   // the two halves of a row hold partial sums of the same running maximum
      xch[n_kv & 1][t][half][row] = l_run;
      named_bar_sync(1 + t, 128 * SPLIT);
      l_run += xch[n_kv & 1][t][half ^ 1][row];
    }
    const float inv_l = (l_run > 0.f ? 1.f / l_run : 0.f) * (p.drop_thresh ? p.drop_scale : 1.f);
    if (half == 0 && q_idx < p.Sq)
      p.lse[((long long)(b * p.H + h)) * p.Sq + q_idx] = (m_run + log2f(l_run)) * 0.6931471805599453f;
    // stage O (bf16) through the (now free) P region of this query tile: NV chunks of [128 rows x 128 B]; three chunks
    // do not fit there, they go through the Q tile (every Q K^T of this CTA has completed: its last S was consumed)
    const uint32_t o_stage = (Cfg::NV > 2) ? q_smem + t * Cfg::Q_BYTES : p_smem + t * Cfg::P_BYTES;
#pragma unroll
    for (int g = 0; g < OD / 8; ++g) {
      const int col = half * OD + g * 8;
      const uint32_t dst = o_stage + (col >> 6) * Cfg::CHUNK_BYTES + row * 128 +
                           (((uint32_t)((col & 63) >> 3) ^ (row & 7)) << 4);
      const float* sp = &o[g * 8];
      asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(pack_bf16(sp[0] * inv_l, sp[1] * inv_l)),
                   "r"(pack_bf16(sp[2] * inv_l, sp[3] * inv_l)), "r"(pack_bf16(sp[4] * inv_l, sp[5] * inv_l)),
                   "r"(pack_bf16(sp[6] * inv_l, sp[7] * inv_l))
                   : "memory");
    }
    fence_proxy_async_smem();
    named_bar_sync(1 + t, 128 * SPLIT);
    if (wl == 0 && lane == 0 && q0 + t * 128 < p.Sq) {
      for (int c = 0; c < DV / 64; ++c)
        if (c * 64 < p.dh) tma_store_4d(&tmO, o_stage + c * Cfg::CHUNK_BYTES, c * 64, h, q0 + t * 128, b);
      tma_commit_group();
      tma_wait_group<0>();
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

bool make_head_tmap(CUtensorMap* tm, const void* ptr, int dh, int H, int S, int B, long long ld, int box_rows) {
  uint64_t dims[4] = {(uint64_t)dh, (uint64_t)H, (uint64_t)S, (uint64_t)B};
  uint64_t strides[3] = {(uint64_t)dh * 2, (uint64_t)ld * 2, (uint64_t)S * (uint64_t)ld * 2};
  uint32_t box[4] = {64, 1, (uint32_t)box_rows, 1};
  return make_tmap(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, ptr, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

template <int DQK, int DV, int MT, int SPLIT>
static int launch_attn_fwd(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const CUtensorMap& to,
                           const AttnParams& p, cudaStream_t stream) {
  using Cfg = AttnFwdCfg<DQK, DV, MT, SPLIT>;
  auto kfn = attn_fwd_kernel<DQK, DV, MT, SPLIT>;
  PPB_CUDA_OK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
  const int q_tiles = (p.Sq + 128 * MT - 1) / (128 * MT);
  kfn<<<p.B * p.H * q_tiles, Cfg::THREADS, Cfg::SMEM_BYTES, stream>>>(tq, tk, tv, to, p);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

// Largest head dim that takes the tcgen05 kernels (forward and backward); larger ones (up to 256) run on CUDA cores.
// PPB_ATTN_TC_MAX_DH=128 sends 129..192 back to the CUDA-core path (A/B measurements).
int attn_tc_max_head_dim() {
  static const int v = [] {
    const char* e = getenv("PPB_ATTN_TC_MAX_DH");
    const int x = e ? atoi(e) : 192;
    return x >= 192 ? 192 : 128;
  }();
  return v;
}

int attn_fwd_generic(const void* q, const void* k, const void* v, void* o, float* lse, int B, int H, int Sq, int Skv,
                     int dh, long long ldq, long long ldk, long long ldv, long long ldo, int causal, float scale,
                     float dropout_p, unsigned long long seed, cudaStream_t stream);

}  // namespace ppb

using namespace ppb;

extern "C" int ppb_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B, int H, int Sq,
                            int Skv, int dh, long long ldq, long long ldk, long long ldv, long long ldo, int causal,
                            int reserved0, int reserved1, float scale, float dropout_p, unsigned long long seed,
                            void* stream_) {
  (void)reserved0;
  (void)reserved1;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(B > 0 && H > 0 && Sq > 0 && Skv > 0, "ppb_attn_fwd: empty problem");
  PPB_REQUIRE(dh % 8 == 0 && dh >= 8 && dh <= 256, "ppb_attn_fwd: head dim must be a multiple of 8 in [8, 256] (got %d)", dh);
  PPB_REQUIRE((ldq % 8) == 0 && (ldk % 8) == 0 && (ldv % 8) == 0 && (ldo % 8) == 0, "ppb_attn_fwd: row strides must be multiples of 8");
  PPB_REQUIRE(!causal || Skv >= Sq, "ppb_attn_fwd: causal attention needs Skv >= Sq");
  PPB_REQUIRE(dropout_p <= 0.f || (Skv % 2) == 0, "ppb_attn_fwd: dropout needs an even Skv");
  if (dh > attn_tc_max_head_dim())  // CUDA-core fallback (attention_generic.cu); the tcgen05 path covers head dims up to 192
    return attn_fwd_generic(q, k, v, o, lse, B, H, Sq, Skv, dh, ldq, ldk, ldv, ldo, causal, scale, dropout_p, seed, stream);
  CUtensorMap tq, tk, tv, to;
  if (!make_head_tmap(&tq, q, dh, H, Sq, B, ldq, 128) || !make_head_tmap(&tk, k, dh, H, Skv, B, ldk, 128) ||
      !make_head_tmap(&tv, v, dh, H, Skv, B, ldv, 128) || !make_head_tmap(&to, o, dh, H, Sq, B, ldo, 128))
    return -3;
  AttnParams p;
  p.B = B;
  p.H = H;
  p.Sq = Sq;
  p.Skv = Skv;
  p.dh = dh;
  p.causal = causal;
  p.sl2 = scale * 1.4426950408889634f;
  p.seed = seed;
  p.seed_off = seed_offset_ptr();
  p.drop_thresh = dropout_p > 0.f ? dropout_thr32(dropout_p) : 0u;
  p.drop_scale = dropout_p < 1.f ? 1.f / (1.f - dropout_p) : 0.f;
  p.lse = lse;
  const bool two = Sq > 128;
  // SPLIT (threads per query row) = 1: measured on B200 (B=4,H=16,S=1024,dh=64) two threads per row are SLOWER
  // (66.6 vs 59.4 us without dropout, 81.9 vs 73.7 us with): the softmax is bound by FMA/ALU pipe throughput, not
  // by latency, so extra warps only add the row-max exchange barrier and register spills.
  if (dh > 128) return launch_attn_fwd<192, 192, 1, 2>(tq, tk, tv, to, p, stream);
  if (dh <= 64) return two ? launch_attn_fwd<64, 64, 2, 1>(tq, tk, tv, to, p, stream) : launch_attn_fwd<64, 64, 1, 1>(tq, tk, tv, to, p, stream);
  return two ? launch_attn_fwd<128, 128, 2, 1>(tq, tk, tv, to, p, stream) : launch_attn_fwd<128, 128, 1, 1>(tq, tk, tv, to, p, stream);
}
