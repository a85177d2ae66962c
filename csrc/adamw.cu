// K9: fused AdamW over a flat parameter arena.
//
// A pipeline stage keeps all of its parameters in one contiguous bf16 arena with matching fp32 arenas for the
// master copy, both moments and the accumulated gradient, so the whole optimizer step is ONE launch:
//   g   = grad * grad_scale * (*clip_coef)         (1/(m*dp) micro-batch averaging folded in; the reference does
//                                                   this as a separate p.grad.div_(m), torch stage.py:589-591)
//   m,v = moments update;  master -= lr * (m_hat / (sqrt(v_hat) + eps) + wd * master)
//   param(bf16) = master ;  grad = 0                (the next step's wgrad reduce-adds start from zero)
// The reference has no optimizer at all (/root/reference/LLMsDistributedTrainingHelper.py:113-131); BASELINE.json
// requires one inside the timed step.
#include "common.cuh"
#include "host_common.h"
#include "ppb200.h"

namespace ppb {

__global__ void __launch_bounds__(256) adamw_kernel(bf16* __restrict__ param, float* __restrict__ master,
                                                    float* __restrict__ m, float* __restrict__ v,
                                                    float* __restrict__ grad, long long n, float lr, float beta1,
                                                    float beta2, float eps, float wd, float bc1, float bc2,
                                                    float grad_scale, const float* __restrict__ clip_coef,
                                                    int zero_grad) {
  const float gs = grad_scale * (clip_coef ? *clip_coef : 1.f);
  const long long n4 = n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 g = reinterpret_cast<float4*>(grad)[i];
    float4 w = reinterpret_cast<float4*>(master)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float* gp = &g.x;
    float* wp = &w.x;
    float* mp = &mm.x;
    float* vp = &vv.x;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float gj = gp[j] * gs;
      mp[j] = beta1 * mp[j] + (1.f - beta1) * gj;
      vp[j] = beta2 * vp[j] + (1.f - beta2) * gj * gj;
      const float mh = mp[j] * bc1, vh = vp[j] * bc2;
      wp[j] = wp[j] - lr * (mh / (sqrtf(vh) + eps) + wd * wp[j]);
    }
    reinterpret_cast<float4*>(master)[i] = w;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
    if (zero_grad) reinterpret_cast<float4*>(grad)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (param) {
      uint2 o;
      o.x = pack_bf16(w.x, w.y);
      o.y = pack_bf16(w.z, w.w);
      reinterpret_cast<uint2*>(param)[i] = o;
    }
  }
  // tail (n % 4 elements)
  const long long tail0 = n4 << 2;
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid < n - tail0) {
    const long long i = tail0 + gid;
    const float gj = grad[i] * gs;
    const float mj = beta1 * m[i] + (1.f - beta1) * gj;
    const float vj = beta2 * v[i] + (1.f - beta2) * gj * gj;
    const float wj = master[i] - lr * ((mj * bc1) / (sqrtf(vj * bc2) + eps) + wd * master[i]);
    m[i] = mj;
    v[i] = vj;
    master[i] = wj;
    if (zero_grad) grad[i] = 0.f;
    if (param) param[i] = __float2bfloat16(wj);
  }
}

// sum of squares of a flat fp32 buffer -> out[0] (+=), used for global-norm clipping
__global__ void __launch_bounds__(256) sumsq_kernel(const float* __restrict__ x, long long n, float* __restrict__ out) {
  __shared__ float red[8];
  float s = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float t = x[i];
    s += t * t;
  }
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    float r = threadIdx.x < 8 ? red[threadIdx.x] : 0.f;
    r = warp_sum(r);
    if (threadIdx.x == 0) atomicAdd(out, r);
  }
}

// coef[0] = min(1, max_norm / (sqrt(sumsq[0]) * grad_scale + 1e-6))
__global__ void clip_coef_kernel(const float* sumsq, float grad_scale, float max_norm, float* coef) {
  const float nrm = sqrtf(sumsq[0]) * grad_scale;
  coef[0] = fminf(1.f, max_norm / (nrm + 1e-6f));
}

}  // namespace ppb

using namespace ppb;

extern "C" int ppb_adamw_step(void* param_bf16, float* master, float* m, float* v, float* grad, long long n, float lr,
                              float beta1, float beta2, float eps, float weight_decay, int step, float grad_scale,
                              const float* clip_coef, int zero_grad, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(n >= 0 && step >= 1, "ppb_adamw_step: bad n=%lld / step=%d", n, step);
  if (n == 0) return 0;
  PPB_REQUIRE((reinterpret_cast<uintptr_t>(master) % 16) == 0 && (reinterpret_cast<uintptr_t>(m) % 16) == 0 &&
                  (reinterpret_cast<uintptr_t>(v) % 16) == 0 && (reinterpret_cast<uintptr_t>(grad) % 16) == 0 &&
                  (reinterpret_cast<uintptr_t>(param_bf16) % 8) == 0,
              "ppb_adamw_step: arenas must be 16-byte aligned");
  const float bc1 = 1.f / (1.f - powf(beta1, (float)step));
  const float bc2 = 1.f / (1.f - powf(beta2, (float)step));
  long long blocks = ((n >> 2) + 255) / 256;
  const long long cap = (long long)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  adamw_kernel<<<(int)blocks, 256, 0, stream>>>((bf16*)param_bf16, master, m, v, grad, n, lr, beta1, beta2, eps,
                                                weight_decay, bc1, bc2, grad_scale, clip_coef, zero_grad);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int ppb_sumsq_f32(const float* x, long long n, float* out, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if (n <= 0) return 0;
  long long blocks = (n + 255) / 256;
  const long long cap = (long long)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  sumsq_kernel<<<(int)blocks, 256, 0, stream>>>(x, n, out);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int ppb_clip_coef(const float* sumsq, float grad_scale, float max_norm, float* coef, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  clip_coef_kernel<<<1, 1, 0, stream>>>(sumsq, grad_scale, max_norm, coef);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}
