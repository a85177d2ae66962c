// K4 fallback: attention for head dims the tcgen05 path does not cover (128 < dh <= 256, e.g. dh = 192 = the
// reference sweep's 4-head / dim-768 configuration, /root/reference/LLMsDistributedTraining.ipynb nb:346-349).
// Plain CUDA-core kernels, one warp per query row, fp32 math, same masking / dropout-index / LSE conventions as
// attention_fwd.cu / attention_bwd.cu so the two paths are interchangeable.  These shapes only occur with short
// sequences (the reference uses S = 128), so simplicity wins over speed here.
#include "common.cuh"
#include "host_common.h"
#include "ppb200.h"

namespace ppb {

struct GenAttnParams {
  const bf16 *q, *k, *v, *o, *dO;
  bf16 *out, *dq;
  float *lse, *dk_acc, *dv_acc;
  int B, H, Sq, Skv, dh;
  long long ldq, ldk, ldv, ldo;
  int causal;
  float scale;
  unsigned long long seed;
  const unsigned long long* seed_off;
  unsigned int thr32;
  float drop_scale;
};

constexpr int GA_WARPS = 4;
constexpr int GA_MAXD = 8;  // dh <= 256: each lane owns up to 8 output dims

// forward: warp per query row, lanes sweep 32 kv positions at a time
__global__ void __launch_bounds__(GA_WARPS * 32) attn_fwd_generic_kernel(const GenAttnParams p) {
  extern __shared__ float gsm[];  // [GA_WARPS][dh] query rows
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qi = blockIdx.x * GA_WARPS + warp;
  const int h = blockIdx.y, b = blockIdx.z;
  if (qi >= p.Sq) return;
  float* qs = gsm + warp * p.dh;
  const bf16* qrow = p.q + ((long long)b * p.Sq + qi) * p.ldq + (long long)h * p.dh;
  for (int d = lane; d < p.dh; d += 32) qs[d] = __bfloat162float(qrow[d]);
  __syncwarp();
  const unsigned long long seed_eff = p.seed + ((p.thr32 && p.seed_off) ? *p.seed_off : 0ull);
  const unsigned long long drow = ((unsigned long long)(b * p.H + h) * p.Sq + qi) * (unsigned long long)p.Skv;
  const int lim = p.causal ? min(p.Skv - 1, qi + (p.Skv - p.Sq)) : p.Skv - 1;
  float m = -INFINITY, l = 0.f;
  float acc[GA_MAXD];
#pragma unroll
  for (int i = 0; i < GA_MAXD; ++i) acc[i] = 0.f;
  for (int j0 = 0; j0 <= lim; j0 += 32) {
    const int j = j0 + lane;
    float s = -INFINITY;
    if (j <= lim) {
      const bf16* krow = p.k + ((long long)b * p.Skv + j) * p.ldk + (long long)h * p.dh;
      float dot = 0.f;
      for (int d = 0; d < p.dh; d += 8) {
        const uint4 u = *reinterpret_cast<const uint4*>(krow + d);
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const float2 f = unpack_bf16(w[t]);
          dot += f.x * qs[d + 2 * t] + f.y * qs[d + 2 * t + 1];
        }
      }
      s = dot * p.scale;
    }
    const float m_new = fmaxf(m, warp_max(s));
    const float alpha = __expf(m - m_new);
    float pj = (j <= lim) ? __expf(s - m_new) : 0.f;
    l = l * alpha + warp_sum(pj);
    m = m_new;
    if (p.thr32 && j <= lim) pj = dropout_keep(seed_eff, drow + j, p.thr32) ? pj * p.drop_scale : 0.f;
#pragma unroll
    for (int i = 0; i < GA_MAXD; ++i) acc[i] *= alpha;
    for (int jj = 0; jj < 32 && j0 + jj <= lim; ++jj) {
      const float pb = __shfl_sync(0xffffffffu, pj, jj);
      const bf16* vrow = p.v + ((long long)b * p.Skv + j0 + jj) * p.ldv + (long long)h * p.dh;
#pragma unroll
      for (int i = 0; i < GA_MAXD; ++i) {
        const int d = lane + 32 * i;
        if (d < p.dh) acc[i] += pb * __bfloat162float(vrow[d]);
      }
    }
  }
  const float inv = l > 0.f ? 1.f / l : 0.f;
  bf16* orow = p.out + ((long long)b * p.Sq + qi) * p.ldo + (long long)h * p.dh;
#pragma unroll
  for (int i = 0; i < GA_MAXD; ++i) {
    const int d = lane + 32 * i;
    if (d < p.dh) orow[d] = __float2bfloat16(acc[i] * inv);
  }
  if (lane == 0) p.lse[((long long)(b * p.H + h)) * p.Sq + qi] = m + __logf(l);
}

// backward: warp per query row; dq written directly, dk/dv accumulated with fp32 atomics into scratch
__global__ void __launch_bounds__(GA_WARPS * 32) attn_bwd_generic_kernel(const GenAttnParams p) {
  extern __shared__ float gsm[];  // [GA_WARPS][2][dh] : q row, dO row
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qi = blockIdx.x * GA_WARPS + warp;
  const int h = blockIdx.y, b = blockIdx.z;
  if (qi >= p.Sq) return;
  float* qs = gsm + warp * 2 * p.dh;
  float* ds_ = qs + p.dh;
  const long long qoff = ((long long)b * p.Sq + qi);
  const bf16* qrow = p.q + qoff * p.ldq + (long long)h * p.dh;
  const bf16* dorow = p.dO + qoff * p.ldo + (long long)h * p.dh;
  const bf16* orow = p.o + qoff * p.ldo + (long long)h * p.dh;
  float dpart = 0.f;
  for (int d = lane; d < p.dh; d += 32) {
    qs[d] = __bfloat162float(qrow[d]);
    ds_[d] = __bfloat162float(dorow[d]);
    dpart += ds_[d] * __bfloat162float(orow[d]);
  }
  const float delta = warp_sum(dpart);
  __syncwarp();
  const float lse = p.lse[((long long)(b * p.H + h)) * p.Sq + qi];
  const unsigned long long seed_eff = p.seed + ((p.thr32 && p.seed_off) ? *p.seed_off : 0ull);
  const unsigned long long drow = ((unsigned long long)(b * p.H + h) * p.Sq + qi) * (unsigned long long)p.Skv;
  const int lim = p.causal ? min(p.Skv - 1, qi + (p.Skv - p.Sq)) : p.Skv - 1;
  float dq[GA_MAXD];
#pragma unroll
  for (int i = 0; i < GA_MAXD; ++i) dq[i] = 0.f;
  for (int j0 = 0; j0 <= lim; j0 += 32) {
    const int j = j0 + lane;
    float pj = 0.f, dsj = 0.f, pdj = 0.f;
    if (j <= lim) {
      const bf16* krow = p.k + ((long long)b * p.Skv + j) * p.ldk + (long long)h * p.dh;
      const bf16* vrow = p.v + ((long long)b * p.Skv + j) * p.ldv + (long long)h * p.dh;
      float dot = 0.f, dp = 0.f;
      for (int d = 0; d < p.dh; d += 8) {
        const uint4 u = *reinterpret_cast<const uint4*>(krow + d);
        const uint4 w4 = *reinterpret_cast<const uint4*>(vrow + d);
        const uint32_t kw[4] = {u.x, u.y, u.z, u.w}, vw[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const float2 kf = unpack_bf16(kw[t]), vf = unpack_bf16(vw[t]);
          dot += kf.x * qs[d + 2 * t] + kf.y * qs[d + 2 * t + 1];
          dp += vf.x * ds_[d + 2 * t] + vf.y * ds_[d + 2 * t + 1];
        }
      }
      pj = __expf(dot * p.scale - lse);
      float keep = 1.f;
      if (p.thr32) keep = dropout_keep(seed_eff, drow + j, p.thr32) ? p.drop_scale : 0.f;
      pdj = pj * keep;
      dsj = pj * (dp * keep - delta) * p.scale;
    }
    for (int jj = 0; jj < 32 && j0 + jj <= lim; ++jj) {
      const float dsb = __shfl_sync(0xffffffffu, dsj, jj);
      const float pdb = __shfl_sync(0xffffffffu, pdj, jj);
      const long long kvoff = ((long long)b * p.Skv + j0 + jj);
      const bf16* krow = p.k + kvoff * p.ldk + (long long)h * p.dh;
      float* dkr = p.dk_acc + (kvoff * p.H + h) * (long long)p.dh;
      float* dvr = p.dv_acc + (kvoff * p.H + h) * (long long)p.dh;
#pragma unroll
      for (int i = 0; i < GA_MAXD; ++i) {
        const int d = lane + 32 * i;
        if (d < p.dh) {
          dq[i] += dsb * __bfloat162float(krow[d]);
          atomicAdd(dkr + d, dsb * qs[d]);
          atomicAdd(dvr + d, pdb * ds_[d]);
        }
      }
    }
  }
  bf16* dqrow = p.dq + qoff * p.ldq + (long long)h * p.dh;
#pragma unroll
  for (int i = 0; i < GA_MAXD; ++i) {
    const int d = lane + 32 * i;
    if (d < p.dh) dqrow[d] = __float2bfloat16(dq[i]);
  }
}

// acc[tok, h, d] fp32 -> out[tok, h*dh + d] bf16 (row stride ld)
__global__ void __launch_bounds__(256) gen_acc_to_bf16_kernel(const float* __restrict__ acc, bf16* __restrict__ out,
                                                              long long ld, long long tokens, int D) {
  const long long n = tokens * D;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[(i / D) * ld + (i % D)] = __float2bfloat16(acc[i]);
}

static void fill_gen(GenAttnParams& p, int B, int H, int Sq, int Skv, int dh, long long ldq, long long ldk, long long ldv,
                     long long ldo, int causal, float scale, float dropout_p, unsigned long long seed) {
  p.B = B;
  p.H = H;
  p.Sq = Sq;
  p.Skv = Skv;
  p.dh = dh;
  p.ldq = ldq;
  p.ldk = ldk;
  p.ldv = ldv;
  p.ldo = ldo;
  p.causal = causal;
  p.scale = scale;
  p.seed = seed;
  p.seed_off = seed_offset_ptr();
  p.thr32 = dropout_p > 0.f ? dropout_thr32(dropout_p) : 0u;
  p.drop_scale = dropout_p < 1.f ? 1.f / (1.f - dropout_p) : 0.f;
}

int attn_fwd_generic(const void* q, const void* k, const void* v, void* o, float* lse, int B, int H, int Sq, int Skv,
                     int dh, long long ldq, long long ldk, long long ldv, long long ldo, int causal, float scale,
                     float dropout_p, unsigned long long seed, cudaStream_t stream) {
  GenAttnParams p = {};
  fill_gen(p, B, H, Sq, Skv, dh, ldq, ldk, ldv, ldo, causal, scale, dropout_p, seed);
  p.q = (const bf16*)q;
  p.k = (const bf16*)k;
  p.v = (const bf16*)v;
  p.out = (bf16*)o;
  p.lse = lse;
  dim3 grid((Sq + GA_WARPS - 1) / GA_WARPS, H, B);
  attn_fwd_generic_kernel<<<grid, GA_WARPS * 32, GA_WARPS * dh * sizeof(float), stream>>>(p);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

int attn_bwd_generic(const void* q, const void* k, const void* v, const void* o, const void* dO, const float* lse,
                     void* dq, void* dk, void* dv, void* scratch, int B, int H, int Sq, int Skv, int dh, long long ldq,
                     long long ldk, long long ldv, long long ldo, int causal, float scale, float dropout_p,
                     unsigned long long seed, cudaStream_t stream) {
  GenAttnParams p = {};
  fill_gen(p, B, H, Sq, Skv, dh, ldq, ldk, ldv, ldo, causal, scale, dropout_p, seed);
  p.q = (const bf16*)q;
  p.k = (const bf16*)k;
  p.v = (const bf16*)v;
  p.o = (const bf16*)o;
  p.dO = (const bf16*)dO;
  p.dq = (bf16*)dq;
  p.lse = const_cast<float*>(lse);
  const long long kvn = (long long)B * Skv * H * dh;
  p.dk_acc = reinterpret_cast<float*>(scratch);
  p.dv_acc = p.dk_acc + kvn;
  PPB_CUDA_OK(cudaMemsetAsync(scratch, 0, (size_t)kvn * 2 * sizeof(float), stream));
  dim3 grid((Sq + GA_WARPS - 1) / GA_WARPS, H, B);
  attn_bwd_generic_kernel<<<grid, GA_WARPS * 32, GA_WARPS * 2 * dh * sizeof(float), stream>>>(p);
  PPB_CUDA_OK(cudaGetLastError());
  const int D = H * dh;
  long long blocks = (kvn + 255) / 256;
  const long long cap = (long long)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  gen_acc_to_bf16_kernel<<<(int)blocks, 256, 0, stream>>>(p.dk_acc, (bf16*)dk, ldk, (long long)B * Skv, D);
  gen_acc_to_bf16_kernel<<<(int)blocks, 256, 0, stream>>>(p.dv_acc, (bf16*)dv, ldv, (long long)B * Skv, D);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace ppb
