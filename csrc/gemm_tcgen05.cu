// K1/K2/K3: persistent, warp-specialised bf16 GEMM on the 5th-gen tensor cores.
//
//   C[M,N] (+)= epilogue( op(A)[M,K] * op(B)[N,K]^T )          fp32 accumulation in TMEM
//
// * operands are staged by TMA (128B swizzle) into a multi-stage smem ring; both K-major ("row-major
//   [rows,K]") and MN-major ("stored [K,rows]") operands are supported through the UMMA smem/instruction
//   descriptors, which covers forward (x*W^T), dgrad (dy*W) and wgrad (dy^T*x) without any transposes;
// * one elected thread issues tcgen05.mma (M=128, N=BN, K=16) into a double-buffered TMEM accumulator
//   (2*BN columns), so the epilogue of tile i overlaps the main loop of tile i+1;
// * four epilogue warps read the accumulator with tcgen05.ld (one row per thread), apply the fused
//   epilogue (bias / ReLU / GELU / dropout / activation-gradient masks), stage through swizzled smem and
//   write with TMA stores -- or TMA reduce-add in fp32, which is what makes split-K wgrad accumulation
//   across micro-batches free of read-modify-write traffic.
//
// Replaces aten.mm/addmm (+add, +relu, +sum) of the reference's op stream; reference call
// sites: nn.TransformerDecoderLayer linears and the LM head, /root/reference/LLMsDistributedTrainingHelper.py:40-46.
#include "common.cuh"
#include "host_common.h"
#include "ppb200.h"

namespace ppb {

#ifndef PPB_GEMM_PAIR_AUTO
#define PPB_GEMM_PAIR_AUTO 1   // measured (profiles/gemm_bench_r2.json): pairs win or tie whenever every SM pair gets a tile
#endif
constexpr int BM = 128;
constexpr int BK = 64;
constexpr int GEMM_THREADS = 192;  // warps 0-3 epilogue, warp 4 TMA producer, warp 5 MMA issuer + TMEM owner
constexpr int EPI_BUF_BYTES = 4096;  // 32 rows x 128 B per warp per buffer

struct GemmParams {
  int M, N, K;
  int tiles_m, tiles_n, splits, kb_per_split;
  int epi;
  float alpha;
  const bf16* bias;
  const bf16* aux;
  long long ld_aux;
  unsigned long long seed;
  const unsigned long long* seed_off;
  unsigned int drop_thresh;
  float drop_scale;
  // OUT == 3 (fused all-reduce): per-(tile, epilogue-warp) arrival counters + number of replicas to push to
  unsigned int* tile_counters;
  int n_dst;
  // tensor-parallel fusions (0 = off): rows of A gathered from / rows of C scattered to `rows_per_rank`-row shards
  int ag_rows_per_rank;   // K12: A tile rows [m0, m0+128) live in shard m0 / ag_rows_per_rank (maps in a_maps[])
  int ag_rank;            //      own rank: its shard's tiles are scheduled first
  int rs_rows_per_rank;   // K11: C slab rows go to owner (row / rs_rows_per_rank)'s buffer (maps in dst_maps[])
};

// K12 (fused all-gather): what the copier warps move and how the GEMM tiles learn that a panel has landed.
struct AgCopy {
  const unsigned char* src[8];  // every rank's [rows_per_rank, K] bf16 shard (own buffer or a peer's mapped symmetric buffer)
  unsigned char* full;          // local [M, K] gathered activations (an OUTPUT of the kernel; the A tiles of remote panels are read from here)
  unsigned int* flags;          // one counter per 128-row panel of `full`, monotonically increasing across calls
  unsigned int target;          // value every panel counter reaches once this call's chunks have landed
  int chunks_per_panel;         // 16 KB chunks per 128-row panel (= K / 64)
  int chunks_per_shard;
  long long shard_bytes;
  int tp, rank;
};
constexpr int AG_CHUNK = 16384;

struct DstMaps {
  CUtensorMap m[8];  // fp32 accumulators of every data-parallel replica (own + peers), peer-mapped over NVLink
};

// CG = CTAs per tile (cta_group): 1 = one CTA computes a 128 x BN tile; 2 = a CTA pair computes a 256 x BN tile, each
// CTA staging its own 128 rows of A and HALF of the B tile (BN/2 rows), so the operand bytes per FLOP that each SM
// pulls through L2 drop by a third at BN = 256.
template <int BN, int CG = 1, int EXTRA = 0>
struct GemmCfg {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = (BN / CG) * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int EPI_BYTES = 4 * 2 * EPI_BUF_BYTES;
  static constexpr int BAR_BYTES = 512;
  static constexpr int FIT = (232448 - 1024 - BAR_BYTES - EPI_BYTES - EXTRA) / STAGE_BYTES;   // 227 KB per CTA
  static constexpr int STAGES = FIT > 8 ? 8 : FIT;   // 4 / 6 / 8 at BN = 256 / 128 / 64 (pairs: 6 / 8)
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + BAR_BYTES + EXTRA + 1024;  // +1024 align slack
  static constexpr int TMEM_COLS = 2 * BN;
};

template <int BN, bool A_MN, bool B_MN, int OUT, int CG = 1, bool AG2 = false>
__device__ __forceinline__ void gemm_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                                          const GemmParams& p, const CUtensorMap* dst_maps,
                                          const CUtensorMap* a_maps = nullptr, const AgCopy* agc = nullptr) {
  using Cfg = GemmCfg<BN, CG, AG2 ? 2 * AG_CHUNK : 0>;
  // (AG2 with CTA pairs is fine: the tiles read A from THIS GPU's gathered copy; only the copier warps touch peer memory,
  //  with plain non-multicast bulk loads)
  static_assert(CG == 1 || (OUT != 3 && BN >= 128), "CTA pairs: plain / accumulate outputs, BN >= 128");
  constexpr int TM = BM * CG;          // rows of the (pair) tile
  constexpr int BNH = BN / CG;         // B rows staged by one CTA
  const int cta_rank = (CG == 2) ? (int)cluster_ctarank() : 0;
  const bool leader = cta_rank == 0;
  const int tile0 = blockIdx.x / CG, tile_stride = gridDim.x / CG;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t epi_base = smem_base + STAGES * Cfg::STAGE_BYTES;
  const uint32_t bar_base = epi_base + Cfg::EPI_BYTES;
  // barrier layout: full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2], tmem ptr
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 2 + a); };
  const uint32_t tmem_ptr_smem = bar_base + 8u * (2 * STAGES + 4);
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));  // generic pointer to the aligned base

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 4 * CG);   // epilogue warps of BOTH CTAs release the accumulator to the leader's issuer
    }
    for (int w = 0; w < 8; ++w) mbar_init(bar_base + 8u * (2 * STAGES + 5 + w), 1);  // fused all-reduce slab loads
    for (int w = 0; w < 2; ++w) mbar_init(bar_base + 8u * (2 * STAGES + 13 + w), 1);  // fused all-gather chunk loads
    fence_mbar_init();
  }
  if (warp == 5) {
    if constexpr (CG == 2) tmem_alloc_cg2(tmem_ptr_smem, Cfg::TMEM_COLS);
    else tmem_alloc(tmem_ptr_smem, Cfg::TMEM_COLS);
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (CG == 2) cluster_sync_all();   // the peer's barriers exist before anything signals them
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_ptr_smem - smem_base));

  const int total_tiles = p.tiles_m * p.tiles_n * p.splits;
  const int num_kb_total = (p.K + BK - 1) / BK;
  // all-gather fusion: rotate the row-block order so that every rank starts on its OWN shard (already local)
  // and touches the peers' shards afterwards, spreading the NVLink reads over time
  const int m_rot = p.ag_rows_per_rank ? (p.ag_rank * (p.ag_rows_per_rank / TM)) : 0;
  auto m_block_of = [&](int mn) { return (mn % p.tiles_m + m_rot) % p.tiles_m; };

  if (warp == 4) {
    // ======================= TMA producer =======================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = tile0; t < total_tiles; t += tile_stride) {
        const int mn = t % (p.tiles_m * p.tiles_n);
        const int sp = t / (p.tiles_m * p.tiles_n);
        const int m0 = m_block_of(mn) * TM + cta_rank * BM;       // this CTA's 128 rows of the (pair) tile
        const int n0 = (mn / p.tiles_m) * BN + cta_rank * BNH;    // this CTA's share of the B rows
        const int kb0 = sp * p.kb_per_split;
        const int kb1 = min(num_kb_total, kb0 + p.kb_per_split);
        // K12: the A rows of this tile come from the shard (possibly on a peer GPU) that owns them
        const CUtensorMap* ta = &tmA;
        int a_row = m0;
        if constexpr (AG2) {
          // own-shard panels come straight from the own buffer; a remote panel is read from the local gathered copy
          // once the copier warps (of whichever CTAs moved its chunks) have reported all of them
          const int owner = m0 / p.ag_rows_per_rank;
          if (owner == p.ag_rank) {
            a_row = m0 - owner * p.ag_rows_per_rank;
          } else {
            ta = a_maps;
            const unsigned int* flag = agc->flags + m0 / BM;
            while (ld_acquire_gpu(flag) < agc->target) {
            }
            fence_proxy_async_all();
          }
        } else if (p.ag_rows_per_rank) {
          const int owner = m0 / p.ag_rows_per_rank;
          ta = a_maps + owner;
          a_row = m0 - owner * p.ag_rows_per_rank;
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          // CTA pair: both CTAs' bytes are counted by the LEADER's barrier (the only one the MMA issuer waits on)
          if (leader) mbar_expect_tx(full_bar(stage), CG * Cfg::STAGE_BYTES);
          const uint32_t sa = smem_base + stage * Cfg::STAGE_BYTES;
          const uint32_t sb = sa + Cfg::A_BYTES;
          auto load = [&](uint32_t dst, const CUtensorMap* tm, int c0, int c1) {
            if constexpr (CG == 2) tma_load_2d_cg2(dst, tm, full_bar(stage), c0, c1);
            else tma_load_2d(dst, tm, full_bar(stage), c0, c1);
          };
          if (!A_MN) {
            load(sa, ta, kb * BK, a_row);  // box {64 k, 128 rows}
          } else {
#pragma unroll
            for (int i = 0; i < BM / 64; ++i)  // box {64 m, 64 k}
              load(sa + i * (BK * 128), &tmA, m0 + i * 64, kb * BK);
          }
          if (!B_MN) {
            load(sb, &tmB, kb * BK, n0);  // box {64 k, BN/CG rows}
          } else {
#pragma unroll
            for (int i = 0; i < BNH / 64; ++i)
              load(sb + i * (BK * 128), &tmB, n0 + i * 64, kb * BK);
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 5) {
    // ======================= MMA issuer =======================
    if (lane == 0 && leader) {
      constexpr uint32_t idesc = make_idesc_bf16(TM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int t = tile0; t < total_tiles; t += tile_stride, ++it) {
        const int sp = t / (p.tiles_m * p.tiles_n);
        const int kb0 = sp * p.kb_per_split;
        const int kb1 = min(num_kb_total, kb0 + p.kb_per_split);
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        mbar_wait(tempty_bar(acc), acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = smem_base + stage * Cfg::STAGE_BYTES;
          const uint32_t sb = sa + Cfg::A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // K-major: 16 k-elements = 32 B inside the 128 B swizzled row; SBO = 8 rows * 128 B.
            // MN-major: 16 k-rows = 2 atoms of 1024 B; LBO = stride between 64-wide MN chunks (one TMA box).
            const uint64_t da = A_MN ? make_smem_desc_sw128(sa + k * 2048, BK * 128, 1024)
                                     : make_smem_desc_sw128(sa + k * 32, 16, 1024);
            const uint64_t db = B_MN ? make_smem_desc_sw128(sb + k * 2048, BK * 128, 1024)
                                     : make_smem_desc_sw128(sb + k * 32, 16, 1024);
            if constexpr (CG == 2) umma_f16_ss_cg2(d_tmem, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            else umma_f16_ss(d_tmem, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          // smem slot is free once these MMAs have read it (pair: the slot of BOTH CTAs)
          if constexpr (CG == 2) umma_commit_cg2(empty_bar(stage));
          else umma_commit(empty_bar(stage));
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        // accumulator complete -> epilogue (of both CTAs: each drains its own 128 TMEM lanes)
        if constexpr (CG == 2) umma_commit_cg2(tfull_bar(acc));
        else umma_commit(tfull_bar(acc));
      }
    }
    __syncwarp();
  } else if (warp == 6) {
    // ======================= all-gather copier (K12) =======================
    // Every CTA moves its share of the shards' 16 KB chunks: bulk copy (peer or own buffer) -> smem -> bulk copy into
    // the local gathered buffer, then one release-increment of the panel's counter.  Remote shards go first, in the
    // order the tiles will ask for them; the own shard (nobody waits for it) last.  Plain, non-multicast bulk loads.
    if constexpr (AG2) {
      if (lane == 0) {
        const uint32_t cbar0 = bar_base + 8u * (2 * STAGES + 13);
        const uint32_t cbuf0 = bar_base + Cfg::BAR_BYTES;
        const int total = agc->tp * agc->chunks_per_shard;
        const int panels_per_shard = agc->chunks_per_shard / agc->chunks_per_panel;
        int nb = 0, prev_panel = -1;
        uint32_t ph[2] = {0, 0};
        for (int g = blockIdx.x; g < total; g += gridDim.x, ++nb) {
          const int s_idx = g / agc->chunks_per_shard, c = g % agc->chunks_per_shard;
          const int shard = (agc->rank + 1 + s_idx) % agc->tp;
          const int b = nb & 1;
          if (nb >= 2) tma_wait_group_read<1>();   // the store that used this buffer two chunks ago has read it out
          mbar_expect_tx(cbar0 + 8u * b, AG_CHUNK);
          bulk_load_1d(cbuf0 + b * AG_CHUNK, agc->src[shard] + (size_t)c * AG_CHUNK, AG_CHUNK, cbar0 + 8u * b);
          mbar_wait(cbar0 + 8u * b, ph[b]);
          ph[b] ^= 1;
          bulk_store_1d(agc->full + (size_t)shard * agc->shard_bytes + (size_t)c * AG_CHUNK, cbuf0 + b * AG_CHUNK, AG_CHUNK);
          tma_commit_group();
          if (prev_panel >= 0) {                  // the PREVIOUS chunk's store is complete: publish it
            tma_wait_group<1>();
            fence_proxy_async_all();
            red_release_gpu_add(agc->flags + prev_panel, 1u);
          }
          prev_panel = shard * panels_per_shard + c / agc->chunks_per_panel;
        }
        if (prev_panel >= 0) {
          tma_wait_group<0>();
          fence_proxy_async_all();
          red_release_gpu_add(agc->flags + prev_panel, 1u);
        }
      }
      __syncwarp();
    }
  } else {
    // ======================= epilogue warps 0..3 =======================
    constexpr int CHUNK = (OUT == 0) ? 64 : 32;  // columns per staged 128-byte row (bf16: 64, fp32: 32)
    const uint32_t my_epi = epi_base + warp * 2 * EPI_BUF_BYTES;
    int it = 0;
    int buf = 0;
    uint32_t push_phase = 0;
    for (int t = tile0; t < total_tiles; t += tile_stride, ++it) {
      const int mn = t % (p.tiles_m * p.tiles_n);
      const int m0 = m_block_of(mn) * TM + cta_rank * BM;
      const int n0 = (mn / p.tiles_m) * BN;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + acc * BN + (uint32_t(warp * 32) << 16);
      const int row = m0 + warp * 32 + lane;
#pragma unroll 1
      for (int c = 0; c < BN / CHUNK; ++c) {
        if (n0 + c * CHUNK >= p.N) break;  // fully out-of-range column chunk (uniform across the CTA)
        uint32_t r[CHUNK];
        tmem_ld_x32(t_row + c * CHUNK, r);
        if constexpr (CHUNK == 64) tmem_ld_x32(t_row + c * CHUNK + 32, r + 32);
        tmem_ld_wait();
        float v[CHUNK];
#pragma unroll
        for (int j = 0; j < CHUNK; ++j) v[j] = __uint_as_float(r[j]) * p.alpha;
        const int nc = n0 + c * CHUNK;
        if (p.epi == PPB_EPI_BIAS || p.epi == PPB_EPI_BIAS_RELU || p.epi == PPB_EPI_BIAS_GELU ||
            p.epi == PPB_EPI_BIAS_RELU_DROPOUT) {
#pragma unroll
          for (int j = 0; j < CHUNK; ++j) {
            const int n = nc + j;
            const float b = (n < p.N) ? __bfloat162float(__ldg(p.bias + n)) : 0.f;
            v[j] += b;
          }
          if (p.epi == PPB_EPI_BIAS_RELU || p.epi == PPB_EPI_BIAS_RELU_DROPOUT) {
#pragma unroll
            for (int j = 0; j < CHUNK; ++j) v[j] = fmaxf(v[j], 0.f);
          } else if (p.epi == PPB_EPI_BIAS_GELU) {
#pragma unroll
            for (int j = 0; j < CHUNK; ++j) v[j] = gelu_erf(v[j]);
          }
          if (p.epi == PPB_EPI_BIAS_RELU_DROPOUT) {
            const unsigned long long seed_eff = p.seed + (p.seed_off ? *p.seed_off : 0ull);
            // N is even for every dropout caller, so (row*N + nc + j) is even for even j and shares one pair hash
            const DropoutStream ds(seed_eff, ((unsigned long long)row * (unsigned long long)p.N + nc) >> 1);
#pragma unroll
            for (int j = 0; j < CHUNK; j += 2) {
              const DropPair h = ds.bits(j >> 1);
              v[j] = (h.a >= p.drop_thresh) ? v[j] * p.drop_scale : 0.f;
              v[j + 1] = (h.b >= p.drop_thresh) ? v[j + 1] * p.drop_scale : 0.f;
            }
          }
        } else if (p.epi == PPB_EPI_MASK_POS || p.epi == PPB_EPI_GELU_BWD) {
          // activation backward: aux holds the stashed (post-dropout) ReLU output, or the GELU pre-activation
          if (row < p.M) {
            const bf16* ap = p.aux + (long long)row * p.ld_aux + nc;
#pragma unroll
            for (int j = 0; j < CHUNK; j += 8) {
              if (nc + j + 8 <= p.N) {
                const uint4 a4 = *reinterpret_cast<const uint4*>(ap + j);
                const uint32_t aw[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  const float2 a2 = unpack_bf16(aw[q]);
                  if (p.epi == PPB_EPI_MASK_POS) {
                    v[j + 2 * q] = a2.x > 0.f ? v[j + 2 * q] : 0.f;
                    v[j + 2 * q + 1] = a2.y > 0.f ? v[j + 2 * q + 1] : 0.f;
                  } else {
                    v[j + 2 * q] *= gelu_erf_grad(a2.x);
                    v[j + 2 * q + 1] *= gelu_erf_grad(a2.y);
                  }
                }
              } else {
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                  const int n = nc + j + q;
                  if (n < p.N) {
                    const float a = __bfloat162float(ap[j + q]);
                    v[j + q] = (p.epi == PPB_EPI_MASK_POS) ? (a > 0.f ? v[j + q] : 0.f) : v[j + q] * gelu_erf_grad(a);
                  }
                }
              }
            }
          }
        }
        // ---- stage into swizzled smem (row = lane, 8 x 16B chunks, chunk position XOR (row & 7)) ----
        if (lane == 0) tma_wait_group_read<1>();  // the store issued from this buffer two chunks ago is done
        __syncwarp();
        const uint32_t srow = my_epi + buf * EPI_BUF_BYTES + lane * 128;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          uint32_t w0, w1, w2, w3;
          if constexpr (OUT == 0) {
            w0 = pack_bf16(v[8 * q + 0], v[8 * q + 1]);
            w1 = pack_bf16(v[8 * q + 2], v[8 * q + 3]);
            w2 = pack_bf16(v[8 * q + 4], v[8 * q + 5]);
            w3 = pack_bf16(v[8 * q + 6], v[8 * q + 7]);
          } else {
            w0 = __float_as_uint(v[4 * q + 0]);
            w1 = __float_as_uint(v[4 * q + 1]);
            w2 = __float_as_uint(v[4 * q + 2]);
            w3 = __float_as_uint(v[4 * q + 3]);
          }
          const uint32_t dst = srow + ((q ^ (lane & 7)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(w0), "r"(w1), "r"(w2), "r"(w3)
                       : "memory");
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          if constexpr (OUT == 2 || OUT == 3) {
            if (p.rs_rows_per_rank) {
              // K11: this 32-row slab belongs to one tensor-parallel rank; reduce-add it into THAT rank's buffer
              const int srow = m0 + warp * 32;
              const int owner = srow / p.rs_rows_per_rank;
              if (srow < p.M) tma_reduce_add_2d(dst_maps + owner, my_epi + buf * EPI_BUF_BYTES, nc, srow - owner * p.rs_rows_per_rank);
            } else {
              tma_reduce_add_2d(&tmC, my_epi + buf * EPI_BUF_BYTES, nc, m0 + warp * 32);
            }
          } else {
            tma_store_2d(&tmC, my_epi + buf * EPI_BUF_BYTES, nc, m0 + warp * 32);
          }
          tma_commit_group();
        }
        buf ^= 1;
      }
      // all TMEM reads of this accumulator are complete -> hand it back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (CG == 2) mbar_arrive_leader(tempty_bar(acc));
        else mbar_arrive(tempty_bar(acc));
      }
      if constexpr (OUT == 3) {
        // ---- fused all-reduce: the LAST K-split to finish this 32-row slab streams the completed gradient slab
        //      to every replica (own reduced buffer through L2, the peers' over NVLink) with TMA only ----
        uint32_t last = 0;
        if (lane == 0) {
          tma_wait_group<0>();   // my reduce-adds into the local accumulator are complete
          __threadfence();
          const unsigned int old = atomicAdd(p.tile_counters + (size_t)mn * 4 + warp, 1u);
          last = (old == (unsigned int)p.splits - 1u) ? 1u : 0u;
          if (last) {
            p.tile_counters[(size_t)mn * 4 + warp] = 0u;   // self-reset for the next launch
            __threadfence();
          }
        }
        last = __shfl_sync(0xffffffffu, last, 0);
        if (last && lane == 0) {
          // two-deep pipeline: the TMA load of slab chunk c+1 is in flight while chunk c is pushed
          const uint32_t lbar0 = bar_base + 8u * (2 * STAGES + 5 + 2 * warp);
          int nchunk = 0;
          for (int c = 0; c < BN / 32 && n0 + c * 32 < p.N; ++c) ++nchunk;
          tma_wait_group_read<0>();
          mbar_expect_tx(lbar0 + 8u * (buf & 1), 32 * 32 * 4);
          tma_load_2d(my_epi + buf * EPI_BUF_BYTES, &tmC, lbar0 + 8u * (buf & 1), n0, m0 + warp * 32);
          for (int c = 0; c < nchunk; ++c) {
            const int nc = n0 + c * 32;
            if (c + 1 < nchunk) {
              tma_wait_group_read<0>();  // the push that last used the other buffer has been read out
              mbar_expect_tx(lbar0 + 8u * ((buf ^ 1) & 1), 32 * 32 * 4);
              tma_load_2d(my_epi + (buf ^ 1) * EPI_BUF_BYTES, &tmC, lbar0 + 8u * ((buf ^ 1) & 1), nc + 32, m0 + warp * 32);
            }
            mbar_wait(lbar0 + 8u * (buf & 1), (push_phase >> (buf & 1)) & 1u);
            push_phase ^= (1u << (buf & 1));
            for (int r = 0; r < p.n_dst; ++r)
              tma_reduce_add_2d(dst_maps + r, my_epi + buf * EPI_BUF_BYTES, nc, m0 + warp * 32);
            tma_commit_group();
            buf ^= 1;
          }
        }
        __syncwarp();
      }
    }
    if (lane == 0) tma_wait_group_read<0>();  // smem may be released; global writes complete before grid end
  }

  tc_fence_before();
  __syncthreads();
  if constexpr (CG == 2) cluster_sync_all();   // no CTA of the pair leaves while the other may still signal it
  if (warp == 5) {
    if constexpr (CG == 2) tmem_dealloc_cg2(tmem_base, Cfg::TMEM_COLS);
    else tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

// CTA-pair variant: 2-CTA clusters, cta_group::2 UMMA on 256 x BN tiles (see GemmCfg).  Operands are this GPU's
// memory only (the tensor-parallel kernels that read peer memory keep one CTA per tile and plain TMA loads).
template <int BN, bool A_MN, bool B_MN, int OUT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const __grid_constant__ CUtensorMap tmC, const GemmParams p) {
  gemm_body<BN, A_MN, B_MN, OUT, 2>(tmA, tmB, tmC, p, nullptr);
}

template <int BN, bool A_MN, bool B_MN, int OUT>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmC, const GemmParams p) {
  gemm_body<BN, A_MN, B_MN, OUT>(tmA, tmB, tmC, p, nullptr);
}

// K10: the LAST micro-batch's weight-gradient GEMM of a stage, fused with the data-parallel
// all-reduce.  The GEMM accumulates into the local fp32 gradient like every other micro-batch (split-K, TMA
// reduce-add).  Every (tile, 32-row slab) has an arrival counter; the K-split that completes a slab last reads the
// finished slab back with TMA (it is still in L2) and reduce-adds it into the reduced-gradient buffer of EVERY
// replica -- its own through L2, the peers' through NVLink (peer-mapped symmetric memory,
// cp.reduce.async.bulk.tensor) -- while the other CTAs keep computing their tiles, so the collective overlaps the
// math tile by tile and each gradient byte crosses the link exactly once.  No in-kernel waits on other GPUs:
// addition commutes, every replica pushes into every replica, and one cross-replica barrier before the optimizer
// makes the sums complete.
template <int BN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_allreduce_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                              const __grid_constant__ CUtensorMap tmC, const __grid_constant__ DstMaps dst,
                              const GemmParams p) {
  gemm_body<BN, true, true, 3>(tmA, tmB, tmC, p, dst.m);
}

// K12: all-gather -> GEMM.  C[M,N] = epi(A_gathered[M,K] * op(B)^T) where the rows of A are sharded
// over the tensor-parallel ranks (sequence-parallel activations in symmetric memory): the TMA producer reads every
// A tile straight from the shard that owns it -- the local one first, then the peers' over NVLink -- so no gathered
// copy of the activations is ever materialised and the transfer overlaps the tensor-core work tile by tile.
template <int BN, bool B_MN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_ag_kernel(const __grid_constant__ DstMaps amaps, const __grid_constant__ CUtensorMap tmB,
                       const __grid_constant__ CUtensorMap tmC, const GemmParams p) {
  gemm_body<BN, false, B_MN, 0>(amaps.m[0], tmB, tmC, p, nullptr, amaps.m);
}

// K12, fused form: ONE kernel gathers and multiplies.  A seventh warp per CTA streams the shards (the peers' over
// NVLink, each byte exactly once) into the local gathered buffer and publishes 128-row panels through release
// counters; the TMA producer multiplies the own-shard panels first and picks up each remote panel as soon as its
// counter says it has landed, so the transfer hides under the tensor-core work.  amaps.m[0] = own shard, m[1] = the
// local gathered buffer.  The gathered activations stay behind for the weight-gradient GEMM of the backward pass.
constexpr int GEMM_AG_THREADS = GEMM_THREADS + 32;
template <int BN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_AG_THREADS, 1)
gemm_tcgen05_agfused_pair_kernel(const __grid_constant__ DstMaps amaps, const __grid_constant__ CUtensorMap tmB,
                                 const __grid_constant__ CUtensorMap tmC, const __grid_constant__ AgCopy agc,
                                 const GemmParams p) {
  gemm_body<BN, false, B_MN, 0, 2, true>(amaps.m[0], tmB, tmC, p, nullptr, amaps.m + 1, &agc);
}
// K11 with CTA pairs (256-row tiles): the epilogue's TMA reduce-adds into the owners' buffers are unchanged.
template <int BN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_rs_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                            const __grid_constant__ DstMaps dst, const GemmParams p) {
  gemm_body<BN, false, B_MN, 2, 2>(tmA, tmB, dst.m[0], p, dst.m);
}
template <int BN, bool B_MN>
__global__ void __launch_bounds__(GEMM_AG_THREADS, 1)
gemm_tcgen05_agfused_kernel(const __grid_constant__ DstMaps amaps, const __grid_constant__ CUtensorMap tmB,
                            const __grid_constant__ CUtensorMap tmC, const __grid_constant__ AgCopy agc,
                            const GemmParams p) {
  gemm_body<BN, false, B_MN, 0, 1, true>(amaps.m[0], tmB, tmC, p, nullptr, amaps.m + 1, &agc);
}

// K11: GEMM -> reduce-scatter.  Every rank computes its partial product (its K-slice of a
// row-parallel layer) and the epilogue reduce-adds each 32-row slab directly into the fp32 buffer of the rank that
// owns those rows (own slabs through L2, the others over NVLink).  After a barrier each rank holds the reduced
// [M/tp, N] block; no full-size partial output and no separate reduce-scatter kernel exist.
template <int BN, bool B_MN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_rs_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                       const __grid_constant__ DstMaps dst, const GemmParams p) {
  gemm_body<BN, false, B_MN, 2>(tmA, tmB, dst.m[0], p, dst.m);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
template <int BN, bool A_MN, bool B_MN, int OUT>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const GemmParams& p,
                       int grid, cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  auto kfn = gemm_tcgen05_kernel<BN, A_MN, B_MN, OUT>;
  static bool attr_done[64] = {false};
  int dev = 0;
  PPB_CUDA_OK(cudaGetDevice(&dev));
  if (dev >= 64 || !attr_done[dev]) {
    PPB_CUDA_OK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    if (dev < 64) attr_done[dev] = true;
  }
  kfn<<<grid, GEMM_THREADS, Cfg::SMEM_BYTES, stream>>>(tmA, tmB, tmC, p);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

template <int BN, bool A_MN, bool B_MN, int OUT>
static int launch_gemm_pair(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const GemmParams& p,
                            int grid, cudaStream_t stream) {
  using Cfg = GemmCfg<BN, 2>;
  auto kfn = gemm_tcgen05_pair_kernel<BN, A_MN, B_MN, OUT>;
  static bool attr_done[64] = {false};
  int dev = 0;
  PPB_CUDA_OK(cudaGetDevice(&dev));
  if (dev >= 64 || !attr_done[dev]) {
    PPB_CUDA_OK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    if (dev < 64) attr_done[dev] = true;
  }
  kfn<<<grid, GEMM_THREADS, Cfg::SMEM_BYTES, stream>>>(tmA, tmB, tmC, p);   // grid is even: __cluster_dims__(2,1,1)
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

template <int BN, int OUT>
static int dispatch_major_pair(int a_mn, int b_mn, const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                               const GemmParams& p, int grid, cudaStream_t s) {
  if (!a_mn && !b_mn) return launch_gemm_pair<BN, false, false, OUT>(tmA, tmB, tmC, p, grid, s);
  if (!a_mn && b_mn) return launch_gemm_pair<BN, false, true, OUT>(tmA, tmB, tmC, p, grid, s);
  if (a_mn && b_mn) return launch_gemm_pair<BN, true, true, OUT>(tmA, tmB, tmC, p, grid, s);
  return launch_gemm_pair<BN, true, false, OUT>(tmA, tmB, tmC, p, grid, s);
}

// CTAs per tile: 0 = automatic, 1 / 2 = forced (ppb_set_gemm_cta_group / $PPB_GEMM_CG)
static int g_gemm_cg = -1;
static int gemm_cg_setting() {
  if (g_gemm_cg < 0) {
    const char* e = getenv("PPB_GEMM_CG");
    g_gemm_cg = e ? atoi(e) : 0;
    if (g_gemm_cg < 0 || g_gemm_cg > 2) g_gemm_cg = 0;
  }
  return g_gemm_cg;
}
// What "automatic" means: pairs whenever the problem has at least two 128-row blocks and the tile is >= 128 wide.
constexpr bool kPairAuto = PPB_GEMM_PAIR_AUTO;

template <int BN, int OUT>
static int dispatch_major(int a_mn, int b_mn, const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                          const GemmParams& p, int grid, cudaStream_t s) {
  if (!a_mn && !b_mn) return launch_gemm<BN, false, false, OUT>(tmA, tmB, tmC, p, grid, s);
  if (!a_mn && b_mn) return launch_gemm<BN, false, true, OUT>(tmA, tmB, tmC, p, grid, s);
  if (a_mn && b_mn) return launch_gemm<BN, true, true, OUT>(tmA, tmB, tmC, p, grid, s);
  return launch_gemm<BN, true, false, OUT>(tmA, tmB, tmC, p, grid, s);
}

// Pick (BN, splits).  Calibrated on a measured sweep (benchmarks/gemm_sweep.py, profiles/gemm_sweep_r1.json):
// the widest tile that N can fill always wins (BN = 256 halves the smem operand traffic per FLOP; BN = 64 is smem-
// bandwidth bound and 2x slower), and for the accumulate mode the split count minimises
// waves * (k-blocks per split * cycles per k-block + fixed tile overhead).
static void choose_config(int M, int N, int K, int out_mode, int force_bn, int force_splits, int* bn_out,
                          int* splits_out, int* cg_out = nullptr) {
  const int num_kb = (K + BK - 1) / BK;
  int bn = force_bn ? force_bn : (N > 128 ? 256 : (N > 64 ? 128 : 64));
  int cg = gemm_cg_setting();
  if (cg == 0) {
    // Pairs only for the big problems (the vocabulary-sized LM-head GEMMs: many tiles or a very long K): measured
    // +11..14 % there and +3 % at 8192^3 (profiles/gemm_bench_r2.json).  The layer-sized GEMMs tie in isolation and
    // lose ~3 % inside the captured stage program (a cluster launch needs both SMs of a pair free at once, which
    // delays the hand-over between back-to-back kernels), so they keep one CTA per tile.
    const long long pair_tiles = (long long)((M + 2 * BM - 1) / (2 * BM)) * ((N + bn - 1) / bn);
    cg = (kPairAuto && bn >= 128 && pair_tiles >= 32 && pair_tiles * num_kb >= 16384) ? 2 : 1;
  }
  if (bn < 128 || cg_out == nullptr) cg = 1;
  if (cg_out) *cg_out = cg;
  const int sms = sm_count() / cg;   // tiles that run concurrently
  const long long tiles = (long long)((M + BM * cg - 1) / (BM * cg)) * ((N + bn - 1) / bn);
  const int max_sp = (out_mode == PPB_OUT_F32_ACCUM) ? 16 : 1;
  double best = 1e30;
  int best_sp = 1;
  for (int sp = 1; sp <= max_sp; ++sp) {
    if (force_splits && sp != force_splits) continue;
    const int kbs = (num_kb + sp - 1) / sp;
    if (sp > 1 && kbs < 4) break;
    if ((long long)(sp - 1) * kbs >= num_kb) continue;  // an empty split
    const long long waves = (tiles * sp + sms - 1) / sms;
    const double cost = waves * (kbs * 2.0 * bn + 3000.0);
    if (cost < best) {
      best = cost;
      best_sp = sp;
    }
  }
  *bn_out = bn;
  *splits_out = best_sp;
}

}  // namespace ppb

using namespace ppb;

extern "C" int ppb_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, long long lda, long long ldb,
                             long long ldc, int a_mn_major, int b_mn_major, int out_mode, int epi, float alpha,
                             const void* bias, const void* aux, long long ld_aux, unsigned long long seed,
                             float dropout_p, int force_bn, int force_splits, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(M > 0 && N > 0 && K > 0, "ppb_gemm_bf16: empty problem M=%d N=%d K=%d", M, N, K);
  PPB_REQUIRE((lda % 8) == 0 && (ldb % 8) == 0, "ppb_gemm_bf16: lda/ldb must be multiples of 8 elements (TMA 16B stride)");
  PPB_REQUIRE(out_mode == PPB_OUT_BF16 ? (ldc % 8) == 0 : (ldc % 4) == 0, "ppb_gemm_bf16: ldc must give a 16B-aligned row stride");
  PPB_REQUIRE((reinterpret_cast<uintptr_t>(A) % 16) == 0 && (reinterpret_cast<uintptr_t>(B) % 16) == 0 &&
                  (reinterpret_cast<uintptr_t>(C) % 16) == 0,
              "ppb_gemm_bf16: operands must be 16-byte aligned");
  PPB_REQUIRE(force_bn == 0 || force_bn == 64 || force_bn == 128 || force_bn == 256, "ppb_gemm_bf16: force_bn must be 0/64/128/256");
  if (epi == PPB_EPI_BIAS || epi == PPB_EPI_BIAS_RELU || epi == PPB_EPI_BIAS_GELU || epi == PPB_EPI_BIAS_RELU_DROPOUT)
    PPB_REQUIRE(bias != nullptr, "ppb_gemm_bf16: epilogue %d needs a bias vector", epi);
  if (epi == PPB_EPI_MASK_POS || epi == PPB_EPI_GELU_BWD)
    PPB_REQUIRE(aux != nullptr && (ld_aux % 8) == 0, "ppb_gemm_bf16: epilogue %d needs an aux matrix with ld %% 8 == 0", epi);
  if (epi == PPB_EPI_BIAS_RELU_DROPOUT) PPB_REQUIRE((N % 2) == 0, "ppb_gemm_bf16: dropout epilogue needs even N");
  if (out_mode != PPB_OUT_F32_ACCUM) PPB_REQUIRE(force_splits <= 1, "ppb_gemm_bf16: split-K needs the fp32 accumulate output mode");

  int bn = 128, splits = 1, cg = 1;
  choose_config(M, N, K, out_mode, force_bn, force_splits, &bn, &splits, &cg);

  CUtensorMap tmA, tmB, tmC;
  bool ok = true;
  // operand maps: K-major -> dims {K, rows}, box {64, rows_per_tile}; MN-major -> dims {rows, K}, box {64, 64}
  if (!a_mn_major)
    ok = ok && make_tmap_2d(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 2, BK, BM);
  else
    ok = ok && make_tmap_2d(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda * 2, 64, BK);
  if (!b_mn_major)
    ok = ok && make_tmap_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb * 2, BK, bn / cg);
  else
    ok = ok && make_tmap_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb * 2, 64, BK);
  if (out_mode == PPB_OUT_BF16)
    ok = ok && make_tmap_2d(&tmC, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, C, (uint64_t)N, (uint64_t)M, (uint64_t)ldc * 2, 64, 32);
  else
    ok = ok && make_tmap_2d(&tmC, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, C, (uint64_t)N, (uint64_t)M, (uint64_t)ldc * 4, 32, 32);
  if (!ok) return -3;

  GemmParams p;
  p.M = M;
  p.N = N;
  p.K = K;
  p.tiles_m = (M + BM * cg - 1) / (BM * cg);
  p.tiles_n = (N + bn - 1) / bn;
  p.splits = splits;
  p.kb_per_split = (((K + BK - 1) / BK) + splits - 1) / splits;
  p.epi = epi;
  p.alpha = alpha;
  p.bias = reinterpret_cast<const bf16*>(bias);
  p.aux = reinterpret_cast<const bf16*>(aux);
  p.ld_aux = ld_aux;
  p.seed = seed;
  p.seed_off = seed_offset_ptr();
  p.tile_counters = nullptr;
  p.n_dst = 0;
  p.ag_rows_per_rank = 0;
  p.ag_rank = 0;
  p.rs_rows_per_rank = 0;
  p.drop_thresh = dropout_thr32(dropout_p);
  p.drop_scale = dropout_p < 1.f ? 1.f / (1.f - dropout_p) : 0.f;
  const long long total = (long long)p.tiles_m * p.tiles_n * splits;
  const int slots = sm_count() / cg;
  const int grid = (int)(total < slots ? total : slots) * cg;

  if (cg == 2) {
#define PPB_DISPATCH_PAIR(BN_)                                                                                   \
  switch (out_mode) {                                                                                            \
    case PPB_OUT_BF16: return dispatch_major_pair<BN_, 0>(a_mn_major, b_mn_major, tmA, tmB, tmC, p, grid, stream); \
    case PPB_OUT_F32: return dispatch_major_pair<BN_, 1>(a_mn_major, b_mn_major, tmA, tmB, tmC, p, grid, stream);  \
    case PPB_OUT_F32_ACCUM: return dispatch_major_pair<BN_, 2>(a_mn_major, b_mn_major, tmA, tmB, tmC, p, grid, stream); \
    default: break;                                                                                              \
  }
    if (bn == 256) {
      PPB_DISPATCH_PAIR(256)
    } else {
      PPB_DISPATCH_PAIR(128)
    }
#undef PPB_DISPATCH_PAIR
    set_error("ppb_gemm_bf16: bad out_mode %d", out_mode);
    return -2;
  }

#define PPB_DISPATCH_OUT(BN_)                                                                               \
  switch (out_mode) {                                                                                       \
    case PPB_OUT_BF16: return dispatch_major<BN_, 0>(a_mn_major, b_mn_major, tmA, tmB, tmC, p, grid, stream); \
    case PPB_OUT_F32: return dispatch_major<BN_, 1>(a_mn_major, b_mn_major, tmA, tmB, tmC, p, grid, stream);  \
    case PPB_OUT_F32_ACCUM: return dispatch_major<BN_, 2>(a_mn_major, b_mn_major, tmA, tmB, tmC, p, grid, stream); \
    default: break;                                                                                         \
  }
  if (bn == 256) {
    PPB_DISPATCH_OUT(256)
  } else if (bn == 128) {
    PPB_DISPATCH_OUT(128)
  } else {
    PPB_DISPATCH_OUT(64)
  }
#undef PPB_DISPATCH_OUT
  set_error("ppb_gemm_bf16: bad out_mode %d", out_mode);
  return -2;
}


extern "C" int ppb_set_gemm_cta_group(int cg) {
  PPB_REQUIRE(cg >= 0 && cg <= 2, "ppb_set_gemm_cta_group: 0 (automatic), 1 or 2");
  ppb::g_gemm_cg = cg;
  return 0;
}

extern "C" long long ppb_gemm_allreduce_scratch_bytes(int M, int N) {
  return (long long)((M + BM - 1) / BM) * ((N + 63) / 64) * 4 * sizeof(unsigned int) + 256;
}

extern "C" int ppb_gemm_allreduce_bf16(const void* A, const void* B, float* accum, void* const* dst_ptrs, int n_dst,
                                       int M, int N, int K, long long lda, long long ldb, long long ldc, void* scratch,
                                       long long scratch_bytes, int force_bn, int force_splits, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(M > 0 && N > 0 && K > 0, "ppb_gemm_allreduce_bf16: empty problem");
  PPB_REQUIRE(n_dst >= 1 && n_dst <= 8, "ppb_gemm_allreduce_bf16: 1..8 destination replicas supported (got %d)", n_dst);
  PPB_REQUIRE((lda % 8) == 0 && (ldb % 8) == 0 && (ldc % 4) == 0, "ppb_gemm_allreduce_bf16: misaligned leading dimension");
  PPB_REQUIRE(accum != nullptr && (reinterpret_cast<uintptr_t>(accum) % 16) == 0, "ppb_gemm_allreduce_bf16: accum must be 16-byte aligned");
  PPB_REQUIRE(scratch != nullptr && scratch_bytes >= ppb_gemm_allreduce_scratch_bytes(M, N),
              "ppb_gemm_allreduce_bf16: scratch too small (%lld bytes, need %lld)", scratch_bytes,
              ppb_gemm_allreduce_scratch_bytes(M, N));
  PPB_REQUIRE(force_bn == 0 || force_bn == 64 || force_bn == 128 || force_bn == 256, "ppb_gemm_allreduce_bf16: bad force_bn");
  int bn = 128, splits = 1;
  choose_config(M, N, K, PPB_OUT_F32_ACCUM, force_bn, force_splits, &bn, &splits);
  CUtensorMap tmA, tmB, tmC;
  DstMaps dst;
  bool ok = make_tmap_2d(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda * 2, 64, BK) &&
            make_tmap_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb * 2, 64, BK) &&
            make_tmap_2d(&tmC, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, accum, (uint64_t)N, (uint64_t)M, (uint64_t)ldc * 4, 32, 32);
  for (int r = 0; ok && r < n_dst; ++r)
    ok = make_tmap_2d(&dst.m[r], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, dst_ptrs[r], (uint64_t)N, (uint64_t)M, (uint64_t)ldc * 4, 32, 32);
  for (int r = n_dst; ok && r < 8; ++r) dst.m[r] = dst.m[0];
  if (!ok) return -3;
  GemmParams p;
  p.M = M;
  p.N = N;
  p.K = K;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (N + bn - 1) / bn;
  p.splits = splits;
  p.kb_per_split = (((K + BK - 1) / BK) + splits - 1) / splits;
  p.epi = PPB_EPI_NONE;
  p.alpha = 1.f;
  p.bias = nullptr;
  p.aux = nullptr;
  p.ld_aux = 0;
  p.seed = 0;
  p.seed_off = nullptr;
  p.drop_thresh = 0;
  p.drop_scale = 1.f;
  p.tile_counters = reinterpret_cast<unsigned int*>(scratch);
  p.n_dst = n_dst;
  p.ag_rows_per_rank = 0;
  p.ag_rank = 0;
  p.rs_rows_per_rank = 0;
  // the arrival counters must be zero on entry: the caller zero-fills the scratch once, the kernel resets every
  // counter it used (no per-call memset launch)
  const long long total = (long long)p.tiles_m * p.tiles_n * splits;
  const int grid = (int)(total < sm_count() ? total : sm_count());
#define PPB_LAUNCH_AR(BN_)                                                                                       \
  {                                                                                                              \
    auto kfn = gemm_tcgen05_allreduce_kernel<BN_>;                                                               \
    PPB_CUDA_OK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, GemmCfg<BN_>::SMEM_BYTES)); \
    kfn<<<grid, GEMM_THREADS, GemmCfg<BN_>::SMEM_BYTES, stream>>>(tmA, tmB, tmC, dst, p);                         \
  }
  if (bn == 256) PPB_LAUNCH_AR(256) else if (bn == 128) PPB_LAUNCH_AR(128) else PPB_LAUNCH_AR(64)
#undef PPB_LAUNCH_AR
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}


namespace ppb {
static void fill_plain_params(GemmParams& p, int M, int N, int K, int bn, int splits) {
  p.M = M;
  p.N = N;
  p.K = K;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (N + bn - 1) / bn;
  p.splits = splits;
  p.kb_per_split = (((K + BK - 1) / BK) + splits - 1) / splits;
  p.epi = PPB_EPI_NONE;
  p.alpha = 1.f;
  p.bias = nullptr;
  p.aux = nullptr;
  p.ld_aux = 0;
  p.seed = 0;
  p.seed_off = nullptr;
  p.drop_thresh = 0;
  p.drop_scale = 1.f;
  p.tile_counters = nullptr;
  p.n_dst = 0;
  p.ag_rows_per_rank = 0;
  p.ag_rank = 0;
  p.rs_rows_per_rank = 0;
}
}  // namespace ppb

extern "C" int ppb_gemm_allgather_bf16(void* const* a_shards, const void* B, void* C, int M, int N, int K, long long lda,
                                       long long ldb, long long ldc, int tp, int rank, int b_mn_major, int epi,
                                       const void* bias, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(tp >= 1 && tp <= 8 && rank >= 0 && rank < tp, "ppb_gemm_allgather_bf16: bad tp/rank %d/%d", tp, rank);
  PPB_REQUIRE(M > 0 && N > 0 && K > 0 && (M % tp) == 0 && ((M / tp) % BM) == 0,
              "ppb_gemm_allgather_bf16: rows per rank (M/tp = %d) must be a multiple of %d", M / (tp ? tp : 1), BM);
  PPB_REQUIRE((lda % 8) == 0 && (ldb % 8) == 0 && (ldc % 8) == 0, "ppb_gemm_allgather_bf16: misaligned leading dimension");
  PPB_REQUIRE(epi == PPB_EPI_NONE || epi == PPB_EPI_BIAS || epi == PPB_EPI_BIAS_RELU || epi == PPB_EPI_BIAS_GELU,
              "ppb_gemm_allgather_bf16: unsupported epilogue %d", epi);
  PPB_REQUIRE(epi == PPB_EPI_NONE || bias != nullptr, "ppb_gemm_allgather_bf16: epilogue needs a bias");
  int bn, splits;
  choose_config(M, N, K, PPB_OUT_BF16, 0, 0, &bn, &splits);
  const int rows = M / tp;
  DstMaps am;
  CUtensorMap tmB, tmC;
  bool ok = true;
  for (int r = 0; ok && r < tp; ++r)
    ok = make_tmap_2d(&am.m[r], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, a_shards[r], (uint64_t)K, (uint64_t)rows, (uint64_t)lda * 2, BK, BM);
  for (int r = tp; ok && r < 8; ++r) am.m[r] = am.m[0];
  ok = ok && (b_mn_major ? make_tmap_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb * 2, 64, BK)
                         : make_tmap_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb * 2, BK, bn));
  ok = ok && make_tmap_2d(&tmC, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, C, (uint64_t)N, (uint64_t)M, (uint64_t)ldc * 2, 64, 32);
  if (!ok) return -3;
  GemmParams p;
  fill_plain_params(p, M, N, K, bn, 1);
  p.epi = epi;
  p.bias = reinterpret_cast<const bf16*>(bias);
  p.ag_rows_per_rank = rows;
  p.ag_rank = rank;
  const long long total = (long long)p.tiles_m * p.tiles_n;
  const int grid = (int)(total < sm_count() ? total : sm_count());
#define PPB_LAUNCH_AG(BN_, BMN_)                                                                                 \
  {                                                                                                              \
    auto kfn = gemm_tcgen05_ag_kernel<BN_, BMN_>;                                                                \
    PPB_CUDA_OK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, GemmCfg<BN_>::SMEM_BYTES)); \
    kfn<<<grid, GEMM_THREADS, GemmCfg<BN_>::SMEM_BYTES, stream>>>(am, tmB, tmC, p);                               \
  }
  if (bn == 256) { if (b_mn_major) PPB_LAUNCH_AG(256, true) else PPB_LAUNCH_AG(256, false) }
  else if (bn == 128) { if (b_mn_major) PPB_LAUNCH_AG(128, true) else PPB_LAUNCH_AG(128, false) }
  else { if (b_mn_major) PPB_LAUNCH_AG(64, true) else PPB_LAUNCH_AG(64, false) }
#undef PPB_LAUNCH_AG
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

// K12, fused form (see gemm_tcgen05_agfused_kernel).  `a_shards[r]` = rank r's [M/tp, K] bf16 shard, contiguous rows
// (own buffer or peer-mapped symmetric memory); `full` = local [M, K] buffer that receives the gathered activations;
// `flags` = M/128 counters (zero-filled once, then owned by this entry point); `target` = chunks_per_panel * (number
// of calls made with these flags, this one included) -- ppb_gemm_allgather_fused_chunks_per_panel(K) gives the factor.
extern "C" int ppb_gemm_allgather_fused_chunks_per_panel(int K) { return (K * 2 * BM) / AG_CHUNK; }

extern "C" int ppb_gemm_allgather_fused_bf16(void* const* a_shards, void* full, unsigned int* flags, unsigned int target,
                                             const void* B, void* C, int M, int N, int K, long long ldb, long long ldc,
                                             int tp, int rank, int b_mn_major, int epi, const void* bias, const void* aux,
                                             long long ld_aux, float alpha, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(tp >= 1 && tp <= 8 && rank >= 0 && rank < tp, "ppb_gemm_allgather_fused_bf16: bad tp/rank %d/%d", tp, rank);
  PPB_REQUIRE(M > 0 && N > 0 && K > 0 && (M % tp) == 0 && ((M / tp) % BM) == 0,
              "ppb_gemm_allgather_fused_bf16: rows per rank (M/tp = %d) must be a multiple of %d", M / (tp ? tp : 1), BM);
  PPB_REQUIRE((K % 64) == 0, "ppb_gemm_allgather_fused_bf16: K must be a multiple of 64 (16 KB chunks of whole rows), got %d", K);
  PPB_REQUIRE((ldb % 8) == 0 && (ldc % 8) == 0, "ppb_gemm_allgather_fused_bf16: misaligned leading dimension");
  PPB_REQUIRE(full != nullptr && flags != nullptr && (reinterpret_cast<uintptr_t>(full) % 128) == 0,
              "ppb_gemm_allgather_fused_bf16: need a 128-byte aligned gathered buffer and the panel counters");
  PPB_REQUIRE(epi == PPB_EPI_NONE || epi == PPB_EPI_BIAS || epi == PPB_EPI_BIAS_RELU || epi == PPB_EPI_BIAS_GELU ||
                  epi == PPB_EPI_MASK_POS || epi == PPB_EPI_GELU_BWD,
              "ppb_gemm_allgather_fused_bf16: unsupported epilogue %d", epi);
  if (epi == PPB_EPI_MASK_POS || epi == PPB_EPI_GELU_BWD)
    PPB_REQUIRE(aux != nullptr && (ld_aux % 8) == 0, "ppb_gemm_allgather_fused_bf16: epilogue %d needs an aux matrix with ld %% 8 == 0", epi);
  else
    PPB_REQUIRE(epi == PPB_EPI_NONE || bias != nullptr, "ppb_gemm_allgather_fused_bf16: epilogue needs a bias");
  int bn, splits, cg = 1;
  choose_config(M, N, K, PPB_OUT_BF16, 0, 0, &bn, &splits, &cg);
  const int rows = M / tp;
  if (cg == 2 && ((rows % (2 * BM)) != 0 || bn != 256)) cg = 1;   // a pair tile must not straddle two shards
  DstMaps am;
  CUtensorMap tmB, tmC;
  bool ok = make_tmap_2d(&am.m[0], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, a_shards[rank], (uint64_t)K, (uint64_t)rows, (uint64_t)K * 2, BK, BM) &&
            make_tmap_2d(&am.m[1], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, full, (uint64_t)K, (uint64_t)M, (uint64_t)K * 2, BK, BM);
  for (int r = 2; ok && r < 8; ++r) am.m[r] = am.m[0];
  ok = ok && (b_mn_major ? make_tmap_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb * 2, 64, BK)
                         : make_tmap_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb * 2, BK, bn / cg));
  ok = ok && make_tmap_2d(&tmC, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, C, (uint64_t)N, (uint64_t)M, (uint64_t)ldc * 2, 64, 32);
  if (!ok) return -3;
  GemmParams p;
  fill_plain_params(p, M, N, K, bn, 1);
  p.tiles_m = (M + BM * cg - 1) / (BM * cg);
  p.epi = epi;
  p.bias = reinterpret_cast<const bf16*>(bias);
  p.aux = reinterpret_cast<const bf16*>(aux);
  p.ld_aux = ld_aux;
  p.alpha = alpha;
  p.ag_rows_per_rank = rows;
  p.ag_rank = rank;
  AgCopy agc;
  for (int r = 0; r < 8; ++r) agc.src[r] = reinterpret_cast<const unsigned char*>(a_shards[r < tp ? r : 0]);
  agc.full = reinterpret_cast<unsigned char*>(full);
  agc.flags = flags;
  agc.target = target;
  agc.chunks_per_panel = ppb_gemm_allgather_fused_chunks_per_panel(K);
  agc.shard_bytes = (long long)rows * K * 2;
  agc.chunks_per_shard = (int)(agc.shard_bytes / AG_CHUNK);
  agc.tp = tp;
  agc.rank = rank;
  // every CTA must be resident at once (tiles wait for chunks that other CTAs move): at most one CTA per SM
  const long long total = (long long)p.tiles_m * p.tiles_n;
  const int slots = sm_count() / cg;
  const int grid = (int)(total < slots ? total : slots) * cg;
  if (cg == 2) {
    using CfgP = GemmCfg<256, 2, 2 * AG_CHUNK>;
    if (b_mn_major) {
      auto kfn = gemm_tcgen05_agfused_pair_kernel<256, true>;
      PPB_CUDA_OK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, CfgP::SMEM_BYTES));
      kfn<<<grid, GEMM_AG_THREADS, CfgP::SMEM_BYTES, stream>>>(am, tmB, tmC, agc, p);
    } else {
      auto kfn = gemm_tcgen05_agfused_pair_kernel<256, false>;
      PPB_CUDA_OK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, CfgP::SMEM_BYTES));
      kfn<<<grid, GEMM_AG_THREADS, CfgP::SMEM_BYTES, stream>>>(am, tmB, tmC, agc, p);
    }
    PPB_CUDA_OK(cudaGetLastError());
    return 0;
  }
#define PPB_LAUNCH_AGF(BN_, BMN_)                                                                                   \
  {                                                                                                                 \
    using CfgF = GemmCfg<BN_, 1, 2 * AG_CHUNK>;                                                                     \
    auto kfn = gemm_tcgen05_agfused_kernel<BN_, BMN_>;                                                              \
    PPB_CUDA_OK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, CfgF::SMEM_BYTES));          \
    kfn<<<grid, GEMM_AG_THREADS, CfgF::SMEM_BYTES, stream>>>(am, tmB, tmC, agc, p);                                 \
  }
  if (bn == 256) { if (b_mn_major) PPB_LAUNCH_AGF(256, true) else PPB_LAUNCH_AGF(256, false) }
  else if (bn == 128) { if (b_mn_major) PPB_LAUNCH_AGF(128, true) else PPB_LAUNCH_AGF(128, false) }
  else { if (b_mn_major) PPB_LAUNCH_AGF(64, true) else PPB_LAUNCH_AGF(64, false) }
#undef PPB_LAUNCH_AGF
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int ppb_gemm_reducescatter_bf16(const void* A, const void* B, void* const* dst_shards, int M, int N, int K,
                                           long long lda, long long ldb, long long ldc, int tp, int b_mn_major,
                                           void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(tp >= 1 && tp <= 8, "ppb_gemm_reducescatter_bf16: bad tp %d", tp);
  PPB_REQUIRE(M > 0 && N > 0 && K > 0 && (M % tp) == 0 && ((M / tp) % 32) == 0,
              "ppb_gemm_reducescatter_bf16: rows per rank (M/tp) must be a multiple of 32");
  PPB_REQUIRE((lda % 8) == 0 && (ldb % 8) == 0 && (ldc % 4) == 0, "ppb_gemm_reducescatter_bf16: misaligned leading dimension");
  int bn, splits, cg = 1;
  choose_config(M, N, K, PPB_OUT_F32, 0, 0, &bn, &splits, &cg);
  if (bn != 256) cg = 1;
  const int rows = M / tp;
  CUtensorMap tmA, tmB;
  DstMaps dm;
  bool ok = make_tmap_2d(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 2, BK, BM);
  ok = ok && (b_mn_major ? make_tmap_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb * 2, 64, BK)
                         : make_tmap_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb * 2, BK, bn / cg));
  for (int r = 0; ok && r < tp; ++r)
    ok = make_tmap_2d(&dm.m[r], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, dst_shards[r], (uint64_t)N, (uint64_t)rows, (uint64_t)ldc * 4, 32, 32);
  for (int r = tp; ok && r < 8; ++r) dm.m[r] = dm.m[0];
  if (!ok) return -3;
  GemmParams p;
  fill_plain_params(p, M, N, K, bn, 1);
  p.tiles_m = (M + BM * cg - 1) / (BM * cg);
  p.rs_rows_per_rank = rows;
  const long long total = (long long)p.tiles_m * p.tiles_n;
  const int slots = sm_count() / cg;
  const int grid = (int)(total < slots ? total : slots) * cg;
  if (cg == 2) {
    using CfgP = GemmCfg<256, 2>;
    if (b_mn_major) {
      auto kfn = gemm_tcgen05_rs_pair_kernel<256, true>;
      PPB_CUDA_OK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, CfgP::SMEM_BYTES));
      kfn<<<grid, GEMM_THREADS, CfgP::SMEM_BYTES, stream>>>(tmA, tmB, dm, p);
    } else {
      auto kfn = gemm_tcgen05_rs_pair_kernel<256, false>;
      PPB_CUDA_OK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, CfgP::SMEM_BYTES));
      kfn<<<grid, GEMM_THREADS, CfgP::SMEM_BYTES, stream>>>(tmA, tmB, dm, p);
    }
    PPB_CUDA_OK(cudaGetLastError());
    return 0;
  }
#define PPB_LAUNCH_RS(BN_, BMN_)                                                                                 \
  {                                                                                                              \
    auto kfn = gemm_tcgen05_rs_kernel<BN_, BMN_>;                                                                \
    PPB_CUDA_OK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, GemmCfg<BN_>::SMEM_BYTES)); \
    kfn<<<grid, GEMM_THREADS, GemmCfg<BN_>::SMEM_BYTES, stream>>>(tmA, tmB, dm, p);                               \
  }
  if (bn == 256) { if (b_mn_major) PPB_LAUNCH_RS(256, true) else PPB_LAUNCH_RS(256, false) }
  else if (bn == 128) { if (b_mn_major) PPB_LAUNCH_RS(128, true) else PPB_LAUNCH_RS(128, false) }
  else { if (b_mn_major) PPB_LAUNCH_RS(64, true) else PPB_LAUNCH_RS(64, false) }
#undef PPB_LAUNCH_RS
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}
