// Llama-style block pieces that are not a GEMM epilogue (no counterpart in the reference, whose block is
// nn.TransformerDecoderLayer with ReLU and no positional encoding, /root/reference/LLMsDistributedTrainingHelper.py:38-46;
// needed for BASELINE.json config 4 "Llama-style 7B"):
//   * SwiGLU:  h = silu(g) * u  on the packed [T, 2f] output of linear1 (gate columns 0..f-1, up columns f..2f-1),
//              and its backward (d g, d u from d h), both pure bandwidth: one pass, 16-byte vectors;
//   * RoPE:    rotary position embedding applied IN PLACE to the q and k column blocks of the packed qkv activation
//              (rotate-half convention: pairs (i, i + dh/2) of each head), forward and inverse (= backward), with
//              the cos/sin table [S, dh/2] precomputed once on the host side of the framework.
#include "common.cuh"
#include "host_common.h"
#include "ppb200.h"

namespace ppb {

__device__ __forceinline__ float sigmoid_f(float x) { return 1.f / (1.f + __expf(-x)); }

// gu: [T, 2f] with row stride ld_gu; h: [T, f] with row stride ld_h.  One thread = 8 columns.
__global__ void __launch_bounds__(256) swiglu_fwd_kernel(const bf16* __restrict__ gu, long long ld_gu, bf16* __restrict__ h,
                                                         long long ld_h, long long T, int f8) {
  const long long total = T * f8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / f8;
    const int c = (int)(i - row * f8);
    const uint4 g4 = *reinterpret_cast<const uint4*>(gu + row * ld_gu + (long long)c * 8);
    const uint4 u4 = *reinterpret_cast<const uint4*>(gu + row * ld_gu + (long long)(f8 + c) * 8);
    const uint32_t gw[4] = {g4.x, g4.y, g4.z, g4.w}, uw[4] = {u4.x, u4.y, u4.z, u4.w};
    uint32_t o[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 g = unpack_bf16(gw[q]), u = unpack_bf16(uw[q]);
      o[q] = pack_bf16(g.x * sigmoid_f(g.x) * u.x, g.y * sigmoid_f(g.y) * u.y);
    }
    *reinterpret_cast<uint4*>(h + row * ld_h + (long long)c * 8) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// dgu[:, :f] = dh * u * silu'(g),  dgu[:, f:] = dh * silu(g);  silu'(g) = s + g s (1 - s), s = sigmoid(g)
__global__ void __launch_bounds__(256) swiglu_bwd_kernel(const bf16* __restrict__ dh, long long ld_dh,
                                                         const bf16* __restrict__ gu, long long ld_gu,
                                                         bf16* __restrict__ dgu, long long ld_dgu, long long T, int f8) {
  const long long total = T * f8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / f8;
    const int c = (int)(i - row * f8);
    const uint4 d4 = *reinterpret_cast<const uint4*>(dh + row * ld_dh + (long long)c * 8);
    const uint4 g4 = *reinterpret_cast<const uint4*>(gu + row * ld_gu + (long long)c * 8);
    const uint4 u4 = *reinterpret_cast<const uint4*>(gu + row * ld_gu + (long long)(f8 + c) * 8);
    const uint32_t dw[4] = {d4.x, d4.y, d4.z, d4.w}, gw[4] = {g4.x, g4.y, g4.z, g4.w}, uw[4] = {u4.x, u4.y, u4.z, u4.w};
    uint32_t og[4], ou[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 d = unpack_bf16(dw[q]), g = unpack_bf16(gw[q]), u = unpack_bf16(uw[q]);
      const float sx = sigmoid_f(g.x), sy = sigmoid_f(g.y);
      og[q] = pack_bf16(d.x * u.x * (sx + g.x * sx * (1.f - sx)), d.y * u.y * (sy + g.y * sy * (1.f - sy)));
      ou[q] = pack_bf16(d.x * g.x * sx, d.y * g.y * sy);
    }
    *reinterpret_cast<uint4*>(dgu + row * ld_dgu + (long long)c * 8) = make_uint4(og[0], og[1], og[2], og[3]);
    *reinterpret_cast<uint4*>(dgu + row * ld_dgu + (long long)(f8 + c) * 8) = make_uint4(ou[0], ou[1], ou[2], ou[3]);
  }
}

// x: [T, ld] rows; the rotated region is n_heads consecutive heads of dh columns starting at column col0.
// Position of a row = row % S.  cs: [S, dh/2] float2 (cos, sin).  sign = +1 forward, -1 inverse.
// One thread = 8 pairs: columns [8j, 8j+8) and [dh/2 + 8j, dh/2 + 8j + 8) of one head of one row.
__global__ void __launch_bounds__(256) rope_kernel(bf16* __restrict__ x, long long ld, int col0, long long T, int S,
                                                   int n_heads, int dh, const float2* __restrict__ cs, float sign) {
  const int half = dh >> 1, j8 = half >> 3;
  const long long total = T * n_heads * j8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(i % j8);
    const long long rh = i / j8;
    const int head = (int)(rh % n_heads);
    const long long row = rh / n_heads;
    const int pos = (int)(row % S);
    bf16* p = x + row * ld + col0 + (long long)head * dh + j * 8;
    const uint4 a4 = *reinterpret_cast<const uint4*>(p);
    const uint4 b4 = *reinterpret_cast<const uint4*>(p + half);
    const uint32_t aw[4] = {a4.x, a4.y, a4.z, a4.w}, bw[4] = {b4.x, b4.y, b4.z, b4.w};
    const float2* t = cs + (long long)pos * half + j * 8;
    uint32_t oa[4], ob[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 a = unpack_bf16(aw[q]), b = unpack_bf16(bw[q]);
      const float2 c0 = __ldg(t + 2 * q), c1 = __ldg(t + 2 * q + 1);
      const float s0 = sign * c0.y, s1 = sign * c1.y;
      oa[q] = pack_bf16(a.x * c0.x - b.x * s0, a.y * c1.x - b.y * s1);
      ob[q] = pack_bf16(b.x * c0.x + a.x * s0, b.y * c1.x + a.y * s1);
    }
    *reinterpret_cast<uint4*>(p) = make_uint4(oa[0], oa[1], oa[2], oa[3]);
    *reinterpret_cast<uint4*>(p + half) = make_uint4(ob[0], ob[1], ob[2], ob[3]);
  }
}

static int llama_grid(long long work_items) {
  long long b = (work_items + 255) / 256;
  const long long cap = (long long)sm_count() * 8;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace ppb

using namespace ppb;

extern "C" int ppb_swiglu_fwd(const void* gu, long long ld_gu, void* h, long long ld_h, long long T, int f, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(T > 0 && f > 0 && (f % 8) == 0 && (ld_gu % 8) == 0 && (ld_h % 8) == 0,
              "ppb_swiglu_fwd: f and the row strides must be multiples of 8 (f=%d)", f);
  swiglu_fwd_kernel<<<llama_grid(T * (f / 8)), 256, 0, stream>>>((const bf16*)gu, ld_gu, (bf16*)h, ld_h, T, f / 8);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int ppb_swiglu_bwd(const void* dh, long long ld_dh, const void* gu, long long ld_gu, void* dgu, long long ld_dgu,
                              long long T, int f, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(T > 0 && f > 0 && (f % 8) == 0 && (ld_gu % 8) == 0 && (ld_dh % 8) == 0 && (ld_dgu % 8) == 0,
              "ppb_swiglu_bwd: f and the row strides must be multiples of 8 (f=%d)", f);
  swiglu_bwd_kernel<<<llama_grid(T * (f / 8)), 256, 0, stream>>>((const bf16*)dh, ld_dh, (const bf16*)gu, ld_gu, (bf16*)dgu,
                                                                 ld_dgu, T, f / 8);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int ppb_rope_inplace(void* x, long long ld, int col0, long long T, int S, int n_heads, int dh,
                                const void* cos_sin, int inverse, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(T > 0 && S > 0 && n_heads > 0 && dh > 0 && (dh % 16) == 0 && (ld % 8) == 0 && (col0 % 8) == 0,
              "ppb_rope_inplace: head_dim must be a multiple of 16 and offsets 16-byte aligned (dh=%d, ld=%lld, col0=%d)",
              dh, ld, col0);
  PPB_REQUIRE(cos_sin != nullptr, "ppb_rope_inplace: missing cos/sin table");
  rope_kernel<<<llama_grid(T * n_heads * (dh / 16)), 256, 0, stream>>>((bf16*)x, ld, col0, T, S, n_heads, dh,
                                                                      (const float2*)cos_sin, inverse ? -1.f : 1.f);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}
