// K7: fused softmax-cross-entropy forward + backward for the last pipeline stage.
//
// One CTA per token row.  The bf16 logits row (up to ~100K vocabulary entries) is pulled into shared memory
// once; max / sum-exp / target logit are reduced from there and the gradient
//      dlogits = (softmax(logits) - onehot(target)) * grad_scale
// is written back IN PLACE over the logits (bf16), so the (tokens, vocab) matrix crosses HBM exactly once in
// each direction and is never materialised in fp32.  Per-row losses go to a small fp32 vector.
//
// Replaces aten._log_softmax + nll_loss_forward + both backwards + the logits .clone() of the reference
// (/root/reference/LLMsDistributedTrainingHelper.py:54 and :197-201).
#include "common.cuh"
#include "host_common.h"
#include "ppb200.h"

namespace ppb {

constexpr int XE_THREADS = 512;

__device__ __forceinline__ float block_reduce_max(float v, float* red) {
  v = warp_max(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = (threadIdx.x < XE_THREADS / 32) ? red[threadIdx.x] : -INFINITY;
  if (threadIdx.x < 32) {
    r = warp_max(r);
    if (threadIdx.x == 0) red[0] = r;
  }
  __syncthreads();
  r = red[0];
  __syncthreads();
  return r;
}
__device__ __forceinline__ float block_reduce_sum(float v, float* red) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = (threadIdx.x < XE_THREADS / 32) ? red[threadIdx.x] : 0.f;
  if (threadIdx.x < 32) {
    r = warp_sum(r);
    if (threadIdx.x == 0) red[0] = r;
  }
  __syncthreads();
  r = red[0];
  __syncthreads();
  return r;
}

// USE_SMEM: the row is cached in shared memory; otherwise it is re-read from global/L2.
// Two sweeps over the row instead of three: the first one loads it, keeps a per-thread ONLINE (max, sum-exp) pair --
// one rescale per 8-element chunk, no bounds checks except in the last chunk -- and the block combines the pairs;
// the second one turns the cached row into the gradient.  exp2 on log2(e)-prescaled logits: one FFMA + one MUFU.
template <bool USE_SMEM>
__global__ void __launch_bounds__(XE_THREADS) xent_fwd_bwd_kernel(bf16* __restrict__ logits, long long ld,
                                                                   const long long* __restrict__ targets,
                                                                   float* __restrict__ row_loss, int T, int V,
                                                                   float grad_scale, int write_grad) {
  extern __shared__ __align__(16) uint8_t xe_smem[];
  __shared__ float red[XE_THREADS / 32];
  bf16* srow = reinterpret_cast<bf16*>(xe_smem);
  const int row = blockIdx.x;
  if (row >= T) return;
  bf16* grow = logits + (long long)row * ld;
  const int nchunks = (V + 7) >> 3;  // ld is a multiple of 8, so reading the padded tail chunk is in-bounds
  const int tail_valid = V - (nchunks - 1) * 8;   // valid elements of the last chunk (1..8)
  const float LOG2E = 1.4426950408889634f;

  float m_t = -INFINITY, s_t = 0.f;   // running max (log2 domain) and sum of exp2(y - m_t) of this thread's chunks
  for (int c = threadIdx.x; c < nchunks; c += XE_THREADS) {
    const uint4 u = *reinterpret_cast<const uint4*>(grow + c * 8);
    if (USE_SMEM) *reinterpret_cast<uint4*>(srow + c * 8) = u;
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
    float y[8];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 f = unpack_bf16(w[q]);
      y[2 * q] = f.x * LOG2E;
      y[2 * q + 1] = f.y * LOG2E;
    }
    if (c == nchunks - 1) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (j >= tail_valid) y[j] = -INFINITY;
    }
    const float cm = fmaxf(fmaxf(fmaxf(y[0], y[1]), fmaxf(y[2], y[3])), fmaxf(fmaxf(y[4], y[5]), fmaxf(y[6], y[7])));
    if (cm > m_t) {                    // rare after the first few chunks
      s_t *= exp2f(m_t - cm);          // m_t = -inf on the first chunk: exp2(-inf) = 0, s_t stays 0
      m_t = cm;
    }
    float a0 = 0.f, a1 = 0.f;
#pragma unroll
    for (int j = 0; j < 8; j += 2) {
      a0 += exp2f(y[j] - m_t);
      a1 += exp2f(y[j + 1] - m_t);
    }
    s_t += a0 + a1;
  }
  const float mx2 = block_reduce_max(m_t, red);  // (contains the __syncthreads that publishes srow)
  const float se = block_reduce_sum(m_t == -INFINITY ? 0.f : s_t * exp2f(m_t - mx2), red);
  const bf16* src = USE_SMEM ? srow : grow;
  const long long tgt = targets[row];
  const bool has_tgt = tgt >= 0 && tgt < V;
  if (threadIdx.x == 0) {
    const float lt2 = has_tgt ? __bfloat162float(src[tgt]) * LOG2E : mx2;
    row_loss[row] = has_tgt ? (log2f(se) + mx2 - lt2) * 0.6931471805599453f : 0.f;
  }
  if (!write_grad) return;
  if (!USE_SMEM) __syncthreads();  // thread 0 has read the target logit before the row is overwritten in place
  const float inv = has_tgt ? grad_scale / se : 0.f;  // ignored rows get zero gradient
  const float gs = has_tgt ? grad_scale : 0.f;
  const int tgt_chunk = has_tgt ? (int)(tgt >> 3) : -1;
  for (int c = threadIdx.x; c < nchunks; c += XE_THREADS) {
    const uint4 u = *reinterpret_cast<const uint4*>(src + c * 8);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
    float g[8];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 f = unpack_bf16(w[q]);
      g[2 * q] = exp2f(fmaf(f.x, LOG2E, -mx2)) * inv;
      g[2 * q + 1] = exp2f(fmaf(f.y, LOG2E, -mx2)) * inv;
    }
    if (c == tgt_chunk) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (j == (int)(tgt & 7)) g[j] -= gs;
    }
    if (c == nchunks - 1) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (j >= tail_valid) g[j] = 0.f;
    }
    *reinterpret_cast<uint4*>(grow + c * 8) =
        make_uint4(pack_bf16(g[0], g[1]), pack_bf16(g[2], g[3]), pack_bf16(g[4], g[5]), pack_bf16(g[6], g[7]));
  }
}

// deterministic single-CTA reduction: out (=|+=) scale * sum(in[0..n))
__global__ void __launch_bounds__(1024) reduce_sum_kernel(const float* __restrict__ in, long long n,
                                                          float* __restrict__ out, float scale, int accumulate) {
  __shared__ float red[32];
  float s = 0.f;
  for (long long i = threadIdx.x; i < n; i += 1024) s += in[i];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    float r = red[threadIdx.x];
    r = warp_sum(r);
    if (threadIdx.x == 0) out[0] = (accumulate ? out[0] : 0.f) + r * scale;
  }
}

}  // namespace ppb

using namespace ppb;

extern "C" int ppb_xent_fwd_bwd(void* logits, long long ld, const long long* targets, float* row_loss, int T, int V,
                                float grad_scale, int write_grad, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(T > 0 && V > 0 && (ld % 8) == 0 && ld >= ((V + 7) / 8) * 8,
              "ppb_xent_fwd_bwd: ld (%lld) must be a multiple of 8 and cover V=%d rounded up to 8", ld, V);
  const size_t smem = (size_t)((V + 7) / 8) * 16;
  if (smem <= 200 * 1024) {
    auto k = xent_fwd_bwd_kernel<true>;
    PPB_CUDA_OK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    k<<<T, XE_THREADS, smem, stream>>>((bf16*)logits, ld, targets, row_loss, T, V, grad_scale, write_grad);
  } else {
    xent_fwd_bwd_kernel<false><<<T, XE_THREADS, 0, stream>>>((bf16*)logits, ld, targets, row_loss, T, V, grad_scale,
                                                            write_grad);
  }
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int ppb_reduce_sum_f32(const float* in, long long n, float* out, float scale, int accumulate, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(n >= 0, "ppb_reduce_sum_f32: negative length");
  reduce_sum_kernel<<<1, 1024, 0, stream>>>(in, n, out, scale, accumulate);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}
