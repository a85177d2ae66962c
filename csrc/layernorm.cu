// K5: fused residual-add + dropout + LayerNorm / RMSNorm, forward and backward.
//
//   forward :  z = resid + dropout(a) ;  y = norm(z) * gamma (+ beta)          one pass over HBM
//   backward:  dy = dy0 (+ dy1 + dy2)   (the fan-in of the post-LN residual graph is summed on load)
//              dz = norm'(dy) ;  da = dropout'(dz) ;  dgamma += .., dbeta += ..
//
// A row lives entirely in registers (d <= 2048), so every tensor is read exactly once and written exactly once
// (forward: one warp per row; backward: a group of d/8 threads per row, each thread owning 8 columns of every row
// its group visits).  The dropout mask is regenerated from (seed, element index) and is never stored.
// dgamma/dbeta partials are reduced thread -> CTA (smem) -> scratch, and a second small kernel folds the scratch,
// column slab by column slab, into the fp32 gradient accumulators (which persist across micro-batches).
//
// Replaces aten.add + bernoulli_/div_/mul + native_layer_norm(_backward) of the reference's op stream
// (torch nn.TransformerDecoderLayer post-LN blocks used at
// /root/reference/LLMsDistributedTrainingHelper.py:40-45 and the final norm at :45,:53).
#include "common.cuh"
#include "host_common.h"
#include "ppb200.h"

namespace ppb {

constexpr int LN_WARPS = 8;
constexpr int LN_THREADS = LN_WARPS * 32;

struct LnFwdParams {
  const bf16* resid;  // [T, d] or null (then z = dropout(a))
  const bf16* a;      // [T, d] or null (then z = resid)
  const bf16* gamma;
  const bf16* beta;  // null for RMSNorm
  bf16* z;           // [T, d] optional output (pre-norm sum, needed by backward)
  bf16* y;           // [T, d]
  float* mean;       // [T] (unused for RMSNorm)
  float* rstd;       // [T]
  int T, d;
  float eps;
  unsigned long long seed;
  const unsigned long long* seed_off;
  unsigned int thresh;
  float scale;
  int rms;
};

template <int MAXV>
__global__ void __launch_bounds__(LN_THREADS, (MAXV <= 4) ? 4 : 1) ln_fwd_kernel(const LnFwdParams p) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nchunks = p.d >> 3;
  const float inv_d = 1.f / float(p.d);
  const unsigned long long seed_eff = p.seed + ((p.thresh && p.seed_off) ? *p.seed_off : 0ull);
  for (int row = blockIdx.x * LN_WARPS + warp; row < p.T; row += gridDim.x * LN_WARPS) {
    float v[MAXV][8];
    const long long rbase = (long long)row * p.d;
    const DropoutStream dstr(seed_eff, (unsigned long long)rbase >> 1);  // d is even: a row starts on a pair boundary
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int c = lane + 32 * i;
      if (c < nchunks) {
        uint4 r4 = make_uint4(0, 0, 0, 0), a4 = make_uint4(0, 0, 0, 0);
        if (p.resid) r4 = *reinterpret_cast<const uint4*>(p.resid + rbase + c * 8);
        if (p.a) a4 = *reinterpret_cast<const uint4*>(p.a + rbase + c * 8);
        const uint32_t rw[4] = {r4.x, r4.y, r4.z, r4.w};
        const uint32_t aw[4] = {a4.x, a4.y, a4.z, a4.w};
        uint32_t zw[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float2 r2 = unpack_bf16(rw[q]);
          float2 a2 = unpack_bf16(aw[q]);
          if (p.thresh) {
            const DropPair h = dstr.bits(c * 4 + q);
            a2.x = (h.a >= p.thresh) ? a2.x * p.scale : 0.f;
            a2.y = (h.b >= p.thresh) ? a2.y * p.scale : 0.f;
          }
          zw[q] = pack_bf16(r2.x + a2.x, r2.y + a2.y);
          const float2 z2 = unpack_bf16(zw[q]);  // statistics are taken on the bf16-rounded z
          v[i][2 * q] = z2.x;
          v[i][2 * q + 1] = z2.y;
          sum += z2.x + z2.y;
        }
        if (p.z) *reinterpret_cast<uint4*>(p.z + rbase + c * 8) = make_uint4(zw[0], zw[1], zw[2], zw[3]);
      }
    }
    float mean = 0.f;
    if (!p.rms) mean = warp_sum(sum) * inv_d;
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int c = lane + 32 * i;
      if (c < nchunks) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float dlt = v[i][j] - mean;
          sq += dlt * dlt;
        }
      }
    }
    const float var = warp_sum(sq) * inv_d;
    const float rstd = rsqrtf(var + p.eps);
    if (lane == 0) {
      if (p.mean) p.mean[row] = mean;
      p.rstd[row] = rstd;
    }
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int c = lane + 32 * i;
      if (c < nchunks) {
        const uint4 g4 = *reinterpret_cast<const uint4*>(p.gamma + c * 8);
        uint4 b4 = make_uint4(0, 0, 0, 0);
        if (p.beta) b4 = *reinterpret_cast<const uint4*>(p.beta + c * 8);
        const uint32_t gw[4] = {g4.x, g4.y, g4.z, g4.w};
        const uint32_t bw[4] = {b4.x, b4.y, b4.z, b4.w};
        uint32_t yw[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float2 g2 = unpack_bf16(gw[q]);
          const float2 b2 = unpack_bf16(bw[q]);
          yw[q] = pack_bf16((v[i][2 * q] - mean) * rstd * g2.x + b2.x, (v[i][2 * q + 1] - mean) * rstd * g2.y + b2.y);
        }
        *reinterpret_cast<uint4*>(p.y + rbase + c * 8) = make_uint4(yw[0], yw[1], yw[2], yw[3]);
      }
    }
  }
}

struct LnBwdParams {
  const bf16* dy0;
  const bf16* dy1;  // nullable
  const bf16* dy2;  // nullable
  const bf16* z;
  const float* mean;  // null for RMSNorm
  const float* rstd;
  const bf16* gamma;
  bf16* dz;  // [T, d] gradient w.r.t. z (residual branch)
  bf16* da;  // [T, d] nullable: dropout'(dz) for the sub-layer branch (null => caller aliases dz)
  float* dgamma;  // [d] fp32 accumulators (+=)
  float* dbeta;   // nullable
  float* scratch;  // 256-byte reserved header + [grid, 2, d] partials
  int T, d;
  unsigned long long seed;
  const unsigned long long* seed_off;
  unsigned int thresh;
  float scale;
  int rms;
};

// Backward, "column-owner" layout: a group of TPR threads works on one row at a time and thread t of the group owns
// the 8 columns [8t, 8t+8) of EVERY row its group visits, so the dgamma / dbeta partial sums live in 16 registers per
// thread for the whole kernel (no per-warp copies of the full row width).  Rows are processed in batches of RB: all
// 16-byte loads of a batch (up to 4 tensors x RB rows per thread) are in flight together, then ONE group-wide
// reduction delivers the two row statistics of all RB rows.  256 threads x 2 CTAs per SM keep > 100 KB of loads in
// flight per SM, which is what a bandwidth-bound kernel needs on HBM3e.
template <int TPR>
__global__ void __launch_bounds__(LN_THREADS, 2) ln_bwd_kernel(const LnBwdParams p) {
  constexpr int G = LN_THREADS / TPR;   // row groups per CTA
  constexpr int WPR = TPR / 32;         // warps per row group
  constexpr int RB = 4;                 // rows per batch
  extern __shared__ float red_smem[];   // [G][2][d] CTA-level partials of dgamma / dbeta
  __shared__ float stat[2][G][WPR][2 * RB];
  const int tid = threadIdx.x, lane = tid & 31;
  const int grp = tid / TPR, t = tid % TPR, wig = t >> 5;
  const int nchunks = p.d >> 3;
  const bool active = t < nchunks;
  const float inv_d = 1.f / float(p.d);
  const unsigned long long seed_eff = p.seed + ((p.thresh && p.seed_off) ? *p.seed_off : 0ull);
  const int rows_per_cta = (p.T + gridDim.x - 1) / gridDim.x;
  const int r_begin = blockIdx.x * rows_per_cta;
  const int r_end = min(p.T, r_begin + rows_per_cta);
  const int n_iter = (rows_per_cta + G * RB - 1) / (G * RB);   // uniform over the CTA (the groups meet at named barriers)

  float gam[8], dg[8], db[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) gam[j] = dg[j] = db[j] = 0.f;
  if (active) {
    const uint4 g4 = *reinterpret_cast<const uint4*>(p.gamma + t * 8);
    const uint32_t gw[4] = {g4.x, g4.y, g4.z, g4.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 g2 = unpack_bf16(gw[q]);
      gam[2 * q] = g2.x;
      gam[2 * q + 1] = g2.y;
    }
  }

  for (int it = 0; it < n_iter; ++it) {
    const int row0 = r_begin + (it * G + grp) * RB;
    float xh[RB][8], gdy[RB][8], cs[2 * RB], rs[RB];
    // ---- issue every load of the batch before using any of them ----
    uint4 d4[RB], e4[RB], f4[RB], z4[RB];
    float mu[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      const int row = row0 + r;
      const bool ok = active && row < r_end;
      d4[r] = e4[r] = f4[r] = z4[r] = make_uint4(0, 0, 0, 0);
      mu[r] = 0.f;
      rs[r] = 0.f;
      if (ok) {
        const long long off = (long long)row * p.d + t * 8;
        d4[r] = *reinterpret_cast<const uint4*>(p.dy0 + off);
        if (p.dy1) e4[r] = *reinterpret_cast<const uint4*>(p.dy1 + off);
        if (p.dy2) f4[r] = *reinterpret_cast<const uint4*>(p.dy2 + off);
        z4[r] = *reinterpret_cast<const uint4*>(p.z + off);
        if (!p.rms) mu[r] = __ldg(p.mean + row);
        rs[r] = __ldg(p.rstd + row);
      }
    }
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      const uint32_t dw[4] = {d4[r].x, d4[r].y, d4[r].z, d4[r].w};
      const uint32_t ew[4] = {e4[r].x, e4[r].y, e4[r].z, e4[r].w};
      const uint32_t fw[4] = {f4[r].x, f4[r].y, f4[r].z, f4[r].w};
      const uint32_t zw[4] = {z4[r].x, z4[r].y, z4[r].z, z4[r].w};
      float c1 = 0.f, c2 = 0.f;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 d2 = unpack_bf16(dw[q]), e2 = unpack_bf16(ew[q]), f2 = unpack_bf16(fw[q]);
        const float2 z2 = unpack_bf16(zw[q]);
        const float dyx = d2.x + e2.x + f2.x, dyy = d2.y + e2.y + f2.y;
        // rs == 0 on rows / columns this thread does not own, which zeroes every contribution below
        const float xx = (z2.x - mu[r]) * rs[r], xy = (z2.y - mu[r]) * rs[r];
        xh[r][2 * q] = xx;
        xh[r][2 * q + 1] = xy;
        gdy[r][2 * q] = dyx * gam[2 * q];
        gdy[r][2 * q + 1] = dyy * gam[2 * q + 1];
        dg[2 * q] += dyx * xx;
        dg[2 * q + 1] += dyy * xy;
        db[2 * q] += dyx;
        db[2 * q + 1] += dyy;
        c1 += gdy[r][2 * q] + gdy[r][2 * q + 1];
        c2 += gdy[r][2 * q] * xx + gdy[r][2 * q + 1] * xy;
      }
      cs[2 * r] = c1;
      cs[2 * r + 1] = c2;
    }
    // ---- the two statistics of all RB rows: warp reduce, then across the group's warps ----
#pragma unroll
    for (int i = 0; i < 2 * RB; ++i) cs[i] = warp_sum(cs[i]);
    if constexpr (WPR > 1) {
      float* my = stat[it & 1][grp][wig];
      if (lane == 0) {
#pragma unroll
        for (int i = 0; i < 2 * RB; ++i) my[i] = cs[i];
      }
      named_bar_sync(1 + grp, TPR);
#pragma unroll
      for (int i = 0; i < 2 * RB; ++i) {
        float a = 0.f;
#pragma unroll
        for (int w = 0; w < WPR; ++w) a += stat[it & 1][grp][w][i];
        cs[i] = a;
      }
    }
    // ---- dz (and da) ----
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      const int row = row0 + r;
      if (!(active && row < r_end)) continue;
      const float c1 = p.rms ? 0.f : cs[2 * r] * inv_d;
      const float c2 = cs[2 * r + 1] * inv_d;
      const long long off = (long long)row * p.d + t * 8;
      const DropoutStream dstr(seed_eff, (unsigned long long)off >> 1);  // d and 8t are even: a pair boundary
      uint32_t zo[4], ao[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float gx = rs[r] * (gdy[r][2 * q] - c1 - xh[r][2 * q] * c2);
        const float gy = rs[r] * (gdy[r][2 * q + 1] - c1 - xh[r][2 * q + 1] * c2);
        zo[q] = pack_bf16(gx, gy);
        float ax = gx, ay = gy;
        if (p.da && p.thresh) {
          const DropPair h = dstr.bits(q);
          ax = (h.a >= p.thresh) ? gx * p.scale : 0.f;
          ay = (h.b >= p.thresh) ? gy * p.scale : 0.f;
        }
        ao[q] = pack_bf16(ax, ay);
      }
      *reinterpret_cast<uint4*>(p.dz + off) = make_uint4(zo[0], zo[1], zo[2], zo[3]);
      if (p.da) *reinterpret_cast<uint4*>(p.da + off) = make_uint4(ao[0], ao[1], ao[2], ao[3]);
    }
  }

  // group partials -> smem [G][2][d] -> CTA partial (plain adds, no atomics) -> scratch[cta][2][d]
  if (active) {
    float* gsm = red_smem + (size_t)grp * 2 * p.d + t * 8;
    *reinterpret_cast<float4*>(gsm) = make_float4(dg[0], dg[1], dg[2], dg[3]);
    *reinterpret_cast<float4*>(gsm + 4) = make_float4(dg[4], dg[5], dg[6], dg[7]);
    *reinterpret_cast<float4*>(gsm + p.d) = make_float4(db[0], db[1], db[2], db[3]);
    *reinterpret_cast<float4*>(gsm + p.d + 4) = make_float4(db[4], db[5], db[6], db[7]);
  }
  __syncthreads();
  float* parts = p.scratch + 64;  // the first 256 bytes of the scratch are a reserved header
  float* part = parts + (size_t)blockIdx.x * 2 * p.d;
  for (int i = tid * 4; i < 2 * p.d; i += LN_THREADS * 4) {
    float4 acc = *reinterpret_cast<const float4*>(red_smem + i);
#pragma unroll
    for (int g = 1; g < G; ++g) {
      const float4 v = *reinterpret_cast<const float4*>(red_smem + (size_t)g * 2 * p.d + i);
      acc.x += v.x;
      acc.y += v.y;
      acc.z += v.z;
      acc.w += v.w;
    }
    *reinterpret_cast<float4*>(part + i) = acc;
  }
}

// Second phase: column slabs in parallel, each CTA folds `nparts` partial rows into the fp32 accumulators.
__global__ void __launch_bounds__(256) ln_bwd_finalize_kernel(const float* __restrict__ parts, int nparts, int d,
                                                              float* __restrict__ dgamma, float* __restrict__ dbeta) {
  __shared__ float red[8][32 * 2 + 2];
  // block = 32 columns x 8 row-lanes; blockIdx.x covers 2*d columns (dgamma then dbeta)
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + cx;
  float acc = 0.f;
  if (col < 2 * d) {
    for (int b = ry; b < nparts; b += 8) acc += __ldcg(parts + (size_t)b * 2 * d + col);
  }
  red[ry][cx] = acc;
  __syncthreads();
  if (ry == 0 && col < 2 * d) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += red[i][cx];
    if (col < d)
      dgamma[col] += s;
    else if (dbeta)
      dbeta[col - d] += s;
  }
}

// ---- wide rows (2048 < d <= 8192, e.g. Llama-width models): nothing is cached in registers --------------------
// Kernel A: one warp per row, two sweeps over the row (the second one hits L2): dz / da only.
__global__ void __launch_bounds__(LN_THREADS) ln_bwd_wide_rows_kernel(const LnBwdParams p) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nchunks = p.d >> 3;
  const float inv_d = 1.f / float(p.d);
  const unsigned long long seed_eff = p.seed + ((p.thresh && p.seed_off) ? *p.seed_off : 0ull);
  for (int row = blockIdx.x * LN_WARPS + warp; row < p.T; row += gridDim.x * LN_WARPS) {
    const long long rbase = (long long)row * p.d;
    const float mean = p.rms ? 0.f : p.mean[row];
    const float rstd = p.rstd[row];
    float c1 = 0.f, c2 = 0.f;
    for (int c = lane; c < nchunks; c += 32) {
      const bf16* z8 = p.z + rbase + c * 8;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float dy = __bfloat162float(p.dy0[rbase + c * 8 + j]);
        if (p.dy1) dy += __bfloat162float(p.dy1[rbase + c * 8 + j]);
        if (p.dy2) dy += __bfloat162float(p.dy2[rbase + c * 8 + j]);
        const float g = dy * __bfloat162float(p.gamma[c * 8 + j]);
        const float xh = (__bfloat162float(z8[j]) - mean) * rstd;
        c1 += g;
        c2 += g * xh;
      }
    }
    c1 = p.rms ? 0.f : warp_sum(c1) * inv_d;
    c2 = warp_sum(c2) * inv_d;
    const DropoutStream dstr(seed_eff, (unsigned long long)rbase >> 1);
    for (int c = lane; c < nchunks; c += 32) {
      uint32_t zo[4], ao[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float gx[2];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const long long i = rbase + c * 8 + 2 * q + e;
          float dy = __bfloat162float(p.dy0[i]);
          if (p.dy1) dy += __bfloat162float(p.dy1[i]);
          if (p.dy2) dy += __bfloat162float(p.dy2[i]);
          const float g = dy * __bfloat162float(p.gamma[c * 8 + 2 * q + e]);
          const float xh = (__bfloat162float(p.z[i]) - mean) * rstd;
          gx[e] = rstd * (g - c1 - xh * c2);
        }
        zo[q] = pack_bf16(gx[0], gx[1]);
        float ax = gx[0], ay = gx[1];
        if (p.da && p.thresh) {
          const DropPair h = dstr.bits(c * 4 + q);
          ax = (h.a >= p.thresh) ? gx[0] * p.scale : 0.f;
          ay = (h.b >= p.thresh) ? gx[1] * p.scale : 0.f;
        }
        ao[q] = pack_bf16(ax, ay);
      }
      *reinterpret_cast<uint4*>(p.dz + rbase + c * 8) = make_uint4(zo[0], zo[1], zo[2], zo[3]);
      if (p.da) *reinterpret_cast<uint4*>(p.da + rbase + c * 8) = make_uint4(ao[0], ao[1], ao[2], ao[3]);
    }
  }
}

// Kernel B: dgamma[c] += sum_t dy*xhat, dbeta[c] += sum_t dy.  Thread = (column pair, row lane); fp32 atomics.
__global__ void __launch_bounds__(256) ln_bwd_wide_cols_kernel(const LnBwdParams p, int rows_per_cta) {
  __shared__ float2 rg[8][32], rb[8][32];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int col = (blockIdx.x * 32 + cx) * 2;
  const int r0 = blockIdx.y * rows_per_cta, r1 = min(p.T, r0 + rows_per_cta);
  float2 ag = make_float2(0.f, 0.f), ab = make_float2(0.f, 0.f);
  if (col < p.d) {
    for (int r = r0 + ry; r < r1; r += 8) {
      const long long i = (long long)r * p.d + col;
      float2 dy = unpack_bf16(*reinterpret_cast<const uint32_t*>(p.dy0 + i));
      if (p.dy1) {
        const float2 e = unpack_bf16(*reinterpret_cast<const uint32_t*>(p.dy1 + i));
        dy.x += e.x;
        dy.y += e.y;
      }
      if (p.dy2) {
        const float2 e = unpack_bf16(*reinterpret_cast<const uint32_t*>(p.dy2 + i));
        dy.x += e.x;
        dy.y += e.y;
      }
      const float2 z2 = unpack_bf16(*reinterpret_cast<const uint32_t*>(p.z + i));
      const float mean = p.rms ? 0.f : p.mean[r], rstd = p.rstd[r];
      ag.x += dy.x * (z2.x - mean) * rstd;
      ag.y += dy.y * (z2.y - mean) * rstd;
      ab.x += dy.x;
      ab.y += dy.y;
    }
  }
  rg[ry][cx] = ag;
  rb[ry][cx] = ab;
  __syncthreads();
  if (ry == 0 && col < p.d) {
    float2 sg = rg[0][cx], sb = rb[0][cx];
#pragma unroll
    for (int i = 1; i < 8; ++i) {
      sg.x += rg[i][cx].x;
      sg.y += rg[i][cx].y;
      sb.x += rb[i][cx].x;
      sb.y += rb[i][cx].y;
    }
    atomicAdd(p.dgamma + col, sg.x);
    atomicAdd(p.dgamma + col + 1, sg.y);
    if (p.dbeta) {
      atomicAdd(p.dbeta + col, sb.x);
      atomicAdd(p.dbeta + col + 1, sb.y);
    }
  }
}

static int ln_grid(int T) {
  const int want = (T + LN_WARPS - 1) / LN_WARPS;
  const int cap = sm_count() * 2;
  return want < cap ? want : cap;
}

}  // namespace ppb

using namespace ppb;

extern "C" int ppb_add_dropout_norm_fwd(const void* resid, const void* a, const void* gamma, const void* beta, void* z,
                                        void* y, float* mean, float* rstd, int T, int d, float eps, float dropout_p,
                                        unsigned long long seed, int rms, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(T > 0 && d > 0 && (d % 8) == 0 && d <= 2048 * 4, "ppb_add_dropout_norm_fwd: need d %% 8 == 0 and d <= 8192 (got %d)", d);
  PPB_REQUIRE(resid || a, "ppb_add_dropout_norm_fwd: at least one of resid / a must be given");
  LnFwdParams p;
  p.resid = (const bf16*)resid;
  p.a = (const bf16*)a;
  p.gamma = (const bf16*)gamma;
  p.beta = (const bf16*)beta;
  p.z = (bf16*)z;
  p.y = (bf16*)y;
  p.mean = mean;
  p.rstd = rstd;
  p.T = T;
  p.d = d;
  p.eps = eps;
  p.seed = seed;
  p.seed_off = seed_offset_ptr();
  p.thresh = (a && dropout_p > 0.f) ? dropout_thr32(dropout_p) : 0u;
  p.scale = dropout_p < 1.f ? 1.f / (1.f - dropout_p) : 0.f;
  p.rms = rms;
  // one row per warp with every row resident at once (4 CTAs x 8 warps per SM at d <= 1024): the kernel is one wave of
  // 16-byte loads, bandwidth- rather than latency-bound; very large T falls back to a grid-stride loop
  const int want = (T + LN_WARPS - 1) / LN_WARPS, cap = sm_count() * 16;
  const int grid = want < cap ? want : cap;
  const int nchunks = d / 8;
  if (nchunks <= 32 * 4)
    ln_fwd_kernel<4><<<grid, LN_THREADS, 0, stream>>>(p);
  else if (nchunks <= 32 * 8)
    ln_fwd_kernel<8><<<grid, LN_THREADS, 0, stream>>>(p);
  else if (nchunks <= 32 * 16)
    ln_fwd_kernel<16><<<grid, LN_THREADS, 0, stream>>>(p);
  else
    ln_fwd_kernel<32><<<grid, LN_THREADS, 0, stream>>>(p);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" long long ppb_norm_bwd_scratch_bytes(int T, int d) {
  (void)T;
  // sized for the largest grid the launcher may pick, so one allocation serves every T
  return (long long)sm_count() * 2 * 2 * d * sizeof(float) + 256;
}

namespace ppb {
static int launch_ln_bwd(const LnBwdParams& p, int grid, cudaStream_t stream) {
  const int nchunks = p.d / 8;   // <= 256 (d <= 2048)
#define PPB_LN_BWD(TPR_)                                                                        \
  {                                                                                             \
    const size_t smem = (size_t)(LN_THREADS / TPR_) * 2 * p.d * sizeof(float); /* <= 32 KB */   \
    ln_bwd_kernel<TPR_><<<grid, LN_THREADS, smem, stream>>>(p);                                 \
  }
  if (nchunks <= 32) PPB_LN_BWD(32)
  else if (nchunks <= 64) PPB_LN_BWD(64)
  else if (nchunks <= 128) PPB_LN_BWD(128)
  else PPB_LN_BWD(256)
#undef PPB_LN_BWD
  PPB_CUDA_OK(cudaGetLastError());
  ln_bwd_finalize_kernel<<<(2 * p.d + 31) / 32, 256, 0, stream>>>(p.scratch + 64, grid, p.d, p.dgamma, p.dbeta);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}
static void fill_ln_bwd_params(LnBwdParams& p, const void* dy0, const void* dy1, const void* dy2, const void* z,
                               const float* mean, const float* rstd, const void* gamma, void* dz, void* da, float* dgamma,
                               float* dbeta, float* scratch, int T, int d, float dropout_p, unsigned long long seed,
                               int rms) {
  p.dy0 = (const bf16*)dy0;
  p.dy1 = (const bf16*)dy1;
  p.dy2 = (const bf16*)dy2;
  p.z = (const bf16*)z;
  p.mean = mean;
  p.rstd = rstd;
  p.gamma = (const bf16*)gamma;
  p.dz = (bf16*)dz;
  p.da = (bf16*)da;
  p.dgamma = dgamma;
  p.dbeta = dbeta;
  p.scratch = scratch;
  p.T = T;
  p.d = d;
  p.seed = seed;
  p.seed_off = seed_offset_ptr();
  p.thresh = (da && dropout_p > 0.f) ? dropout_thr32(dropout_p) : 0u;
  p.scale = dropout_p < 1.f ? 1.f / (1.f - dropout_p) : 0.f;
  p.rms = rms;
}
}  // namespace ppb

extern "C" int ppb_norm_bwd(const void* dy0, const void* dy1, const void* dy2, const void* z, const float* mean,
                            const float* rstd, const void* gamma, void* dz, void* da, float* dgamma, float* dbeta,
                            void* scratch, long long scratch_bytes, int T, int d, float dropout_p,
                            unsigned long long seed, int rms, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(T > 0 && d > 0 && (d % 8) == 0 && d <= 8192, "ppb_norm_bwd: need d %% 8 == 0 and d <= 8192 (got %d)", d);
  if (d > 2048) {  // wide rows: row pass + column pass, no scratch needed
    LnBwdParams pw;
    fill_ln_bwd_params(pw, dy0, dy1, dy2, z, mean, rstd, gamma, dz, da, dgamma, dbeta, nullptr, T, d, dropout_p, seed, rms);
    ln_bwd_wide_rows_kernel<<<ln_grid(T), LN_THREADS, 0, stream>>>(pw);
    const int col_blocks = (d + 63) / 64;
    int slabs = (sm_count() * 4 + col_blocks - 1) / col_blocks;
    if (slabs > (T + 63) / 64) slabs = (T + 63) / 64;
    if (slabs < 1) slabs = 1;
    const int rows_per = (T + slabs - 1) / slabs;
    ln_bwd_wide_cols_kernel<<<dim3(col_blocks, (T + rows_per - 1) / rows_per), 256, 0, stream>>>(pw, rows_per);
    PPB_CUDA_OK(cudaGetLastError());
    return 0;
  }
  PPB_REQUIRE(scratch != nullptr && scratch_bytes >= ppb_norm_bwd_scratch_bytes(T, d),
              "ppb_norm_bwd: scratch too small (%lld bytes, need %lld); it must also be zero-initialised once",
              scratch_bytes, ppb_norm_bwd_scratch_bytes(T, d));
  LnBwdParams p;
  fill_ln_bwd_params(p, dy0, dy1, dy2, z, mean, rstd, gamma, dz, da, dgamma, dbeta, (float*)scratch, T, d, dropout_p,
                     seed, rms);
  return launch_ln_bwd(p, ln_grid(T), stream);
}

// ------------------------------------------------------------------------------------------------------------
// Allocation-free variant of ppb_norm_bwd.
//
// Some hosts of this library are not allowed to create device memory at all: a serving or training runtime that
// hands every byte of HBM to its own arena at start-up, a plug-in entry called from inside another framework's
// custom-op hook, or a thread that must never trigger cudaMalloc's implicit device-wide synchronisation.  Those
// callers cannot produce the scratch buffer that ppb_norm_bwd wants.  For them the dgamma/dbeta partial sums go
// into a fixed array that lives inside the library image instead: every GPU that loads libppb200.so keeps these
// PPB_NORM_NOALLOC_SCRATCH_BYTES (about 2.4 MiB) reserved until the process exits.
//
// Limits that follow from the fixed size: d <= PPB_NORM_NOALLOC_MAX_D, and the launch uses at most
// PPB_NORM_NOALLOC_MAX_CTAS CTAs.  Larger d is refused with an error (use ppb_norm_bwd and bring a buffer).
// ------------------------------------------------------------------------------------------------------------
#define PPB_NORM_NOALLOC_MAX_D 1024
#define PPB_NORM_NOALLOC_MAX_CTAS 304
#define PPB_NORM_NOALLOC_SCRATCH_BYTES (256 + PPB_NORM_NOALLOC_MAX_CTAS * 2 * PPB_NORM_NOALLOC_MAX_D * 4)

// Host code must never pass this name to a kernel: on the host it is only a stand-in for the symbol, not a device
// address.  The address is looked up once per device below.
static __device__ __align__(256) unsigned char g_norm_noalloc_scratch[PPB_NORM_NOALLOC_SCRATCH_BYTES];

namespace ppb {
static float* g_noalloc_ptr[64] = {nullptr};

// One-time, per-device set-up: resolve the device address of the static array with the target device current.
static int norm_noalloc_setup(float** out) {
  int dev = 0;
  PPB_CUDA_OK(cudaGetDevice(&dev));
  PPB_REQUIRE(dev >= 0 && dev < 64, "ppb_norm_bwd_noalloc: device index %d out of range", dev);
  if (g_noalloc_ptr[dev] == nullptr) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_norm_noalloc_scratch);
    if (e != cudaSuccess || p == nullptr) {
      set_error("ppb_norm_bwd_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return -1;
    }
    g_noalloc_ptr[dev] = reinterpret_cast<float*>(p);
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}
}  // namespace ppb

extern "C" long long ppb_norm_bwd_noalloc_max_d(void) { return PPB_NORM_NOALLOC_MAX_D; }

// Same contract as ppb_norm_bwd, minus the scratch arguments.  Makes no cudaMalloc/cudaFree call and calls no
// allocating library routine; inputs and outputs are still the caller's.
//  * The FIRST call on each device resolves the address of the built-in scratch array (cudaGetSymbolAddress);
//    make that first call outside of any stream capture.
//  * All calls on a device share the one array.  It still holds whatever the previous call left behind, so its
//    256-byte header is cleared on `stream` before every launch; the partial-sum area is fully rewritten by each
//    launch before it is read.
//  * Because the array is shared, two calls whose kernels can overlap on the GPU (different streams, no ordering
//    between them) race with each other.  Order them, or use ppb_norm_bwd with separate buffers.
extern "C" int ppb_norm_bwd_noalloc(const void* dy0, const void* dy1, const void* dy2, const void* z, const float* mean,
                                    const float* rstd, const void* gamma, void* dz, void* da, float* dgamma,
                                    float* dbeta, int T, int d, float dropout_p, unsigned long long seed, int rms,
                                    void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(T > 0 && d > 0 && (d % 8) == 0, "ppb_norm_bwd_noalloc: need T > 0 and d %% 8 == 0 (got T=%d d=%d)", T, d);
  PPB_REQUIRE(d <= PPB_NORM_NOALLOC_MAX_D,
              "ppb_norm_bwd_noalloc: d=%d exceeds the built-in scratch (max d = %d); call ppb_norm_bwd with your own buffer",
              d, PPB_NORM_NOALLOC_MAX_D);
  float* scratch = nullptr;
  int rc = norm_noalloc_setup(&scratch);
  if (rc) return rc;
  int grid = ln_grid(T);
  if (grid > PPB_NORM_NOALLOC_MAX_CTAS) grid = PPB_NORM_NOALLOC_MAX_CTAS;
  PPB_CUDA_OK(cudaMemsetAsync(scratch, 0, 256, stream));  // the header must start at zero
  LnBwdParams p;
  fill_ln_bwd_params(p, dy0, dy1, dy2, z, mean, rstd, gamma, dz, da, dgamma, dbeta, scratch, T, d, dropout_p, seed, rms);
  return launch_ln_bwd(p, grid, stream);
}
