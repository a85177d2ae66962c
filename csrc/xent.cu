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
  const float LOG2E = 1.4426950408889634f;

  float mx = -INFINITY;
  for (int c = threadIdx.x; c < nchunks; c += XE_THREADS) {
    const uint4 u = *reinterpret_cast<const uint4*>(grow + c * 8);
    if (USE_SMEM) *reinterpret_cast<uint4*>(srow + c * 8) = u;
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 f = unpack_bf16(w[q]);
      const int e = c * 8 + 2 * q;
      if (e < V) mx = fmaxf(mx, f.x);
      if (e + 1 < V) mx = fmaxf(mx, f.y);
    }
  }
  mx = block_reduce_max(mx, red);  // (contains the __syncthreads that publishes srow)
  const bf16* src = USE_SMEM ? srow : grow;
  float se = 0.f;
  for (int c = threadIdx.x; c < nchunks; c += XE_THREADS) {
    const uint4 u = *reinterpret_cast<const uint4*>(src + c * 8);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 f = unpack_bf16(w[q]);
      const int e = c * 8 + 2 * q;
      if (e < V) se += exp2f((f.x - mx) * LOG2E);
      if (e + 1 < V) se += exp2f((f.y - mx) * LOG2E);
    }
  }
  se = block_reduce_sum(se, red);
  const long long tgt = targets[row];
  if (threadIdx.x == 0) {
    const float lt = (tgt >= 0 && tgt < V) ? __bfloat162float(src[tgt]) : mx;
    row_loss[row] = (tgt >= 0 && tgt < V) ? (logf(se) + mx - lt) : 0.f;
  }
  if (!write_grad) return;
  if (!USE_SMEM) __syncthreads();  // thread 0 has read the target logit before the row is overwritten in place
  const float inv = (tgt >= 0 && tgt < V) ? grad_scale / se : 0.f;  // ignored rows get zero gradient
  const float gs = (tgt >= 0 && tgt < V) ? grad_scale : 0.f;
  for (int c = threadIdx.x; c < nchunks; c += XE_THREADS) {
    const uint4 u = *reinterpret_cast<const uint4*>(src + c * 8);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
    uint32_t o[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 f = unpack_bf16(w[q]);
      const int e = c * 8 + 2 * q;
      float gx = (e < V) ? exp2f((f.x - mx) * LOG2E) * inv : 0.f;
      float gy = (e + 1 < V) ? exp2f((f.y - mx) * LOG2E) * inv : 0.f;
      if (e == tgt) gx -= gs;
      if (e + 1 == tgt) gy -= gs;
      o[q] = pack_bf16(gx, gy);
    }
    *reinterpret_cast<uint4*>(grow + c * 8) = make_uint4(o[0], o[1], o[2], o[3]);
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
