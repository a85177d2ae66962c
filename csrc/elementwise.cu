// K6 / K8: the bandwidth-bound helpers that are not absorbed by a GEMM epilogue.
//   * embedding gather forward and scatter-add backward (aten.embedding / embedding_dense_backward;
//     /root/reference/LLMsDistributedTrainingHelper.py:36,50)
//   * bias gradient (column sums of dy, fp32 accumulate; aten.sum.dim_IntList of the reference op stream)
//   * fan-in sum of up to three bf16 gradients (aten.add.Tensor in autograd's accumulation)
//   * GELU forward (only for the GPT presets, where the pre-activation has to be kept for backward)
//   * fp32 -> bf16 cast, fill
#include "common.cuh"
#include "host_common.h"
#include "ppb200.h"

namespace ppb {

// one warp per token row; 16-byte vectors
__global__ void __launch_bounds__(256) embedding_fwd_kernel(const long long* __restrict__ idx,
                                                            const bf16* __restrict__ W, bf16* __restrict__ out, int T,
                                                            int d, int V) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  const int nch = d >> 3;
  for (int t = warp; t < T; t += nwarps) {
    long long id = idx[t];
    id = id < 0 ? 0 : (id >= V ? V - 1 : id);
    const uint4* src = reinterpret_cast<const uint4*>(W + id * (long long)d);
    uint4* dst = reinterpret_cast<uint4*>(out + (long long)t * d);
    for (int c = lane; c < nch; c += 32) dst[c] = __ldg(src + c);
  }
}

// dW[idx[t], :] += dy0[t,:] (+ dy1 + dy2)   fp32 vector atomics (red.global.add.v4.f32)
__global__ void __launch_bounds__(256) embedding_bwd_kernel(const long long* __restrict__ idx,
                                                            const bf16* __restrict__ dy0, const bf16* __restrict__ dy1,
                                                            const bf16* __restrict__ dy2, float* __restrict__ dW, int T,
                                                            int d, int V) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  const int nch = d >> 3;
  for (int t = warp; t < T; t += nwarps) {
    long long id = idx[t];
    if (id < 0 || id >= V) continue;
    float* dst = dW + id * (long long)d;
    for (int c = lane; c < nch; c += 32) {
      const long long off = (long long)t * d + c * 8;
      const uint4 a = *reinterpret_cast<const uint4*>(dy0 + off);
      uint4 b = make_uint4(0, 0, 0, 0), e = make_uint4(0, 0, 0, 0);
      if (dy1) b = *reinterpret_cast<const uint4*>(dy1 + off);
      if (dy2) e = *reinterpret_cast<const uint4*>(dy2 + off);
      const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w}, ew[4] = {e.x, e.y, e.z, e.w};
      float f[8];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 x = unpack_bf16(aw[q]), y = unpack_bf16(bw[q]), z = unpack_bf16(ew[q]);
        f[2 * q] = x.x + y.x + z.x;
        f[2 * q + 1] = x.y + y.y + z.y;
      }
      atomicAdd(reinterpret_cast<float4*>(dst + c * 8), make_float4(f[0], f[1], f[2], f[3]));
      atomicAdd(reinterpret_cast<float4*>(dst + c * 8 + 4), make_float4(f[4], f[5], f[6], f[7]));
    }
  }
}

// db[n] += sum_t dy[t, n].  grid = (column blocks of 256, row slabs); each thread owns 2 columns... one bf162
__global__ void __launch_bounds__(256) colsum_kernel(const bf16* __restrict__ dy, long long ld, float* __restrict__ db,
                                                     int T, int N, int rows_per_cta) {
  __shared__ float2 red[8][32];
  // thread layout: 32 column-pairs (64 columns) x 8 row lanes
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int col = (blockIdx.x * 32 + cx) * 2;
  const int r0 = blockIdx.y * rows_per_cta;
  const int r1 = min(T, r0 + rows_per_cta);
  float2 acc = make_float2(0.f, 0.f);
  if (col < N) {
    for (int r = r0 + ry; r < r1; r += 8) {
      const uint32_t u = *reinterpret_cast<const uint32_t*>(dy + (long long)r * ld + col);
      const float2 f = unpack_bf16(u);
      acc.x += f.x;
      acc.y += f.y;
    }
  }
  red[ry][cx] = acc;
  __syncthreads();
  if (ry == 0 && col < N) {
    float2 s = red[0][cx];
#pragma unroll
    for (int i = 1; i < 8; ++i) {
      s.x += red[i][cx].x;
      s.y += red[i][cx].y;
    }
    atomicAdd(db + col, s.x);
    if (col + 1 < N) atomicAdd(db + col + 1, s.y);
  }
}

__global__ void __launch_bounds__(256) add_n_kernel(const bf16* __restrict__ a, const bf16* __restrict__ b,
                                                    const bf16* __restrict__ c, bf16* __restrict__ out, long long n8) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    const uint4 x = reinterpret_cast<const uint4*>(a)[i];
    uint4 y = make_uint4(0, 0, 0, 0), z = make_uint4(0, 0, 0, 0);
    if (b) y = reinterpret_cast<const uint4*>(b)[i];
    if (c) z = reinterpret_cast<const uint4*>(c)[i];
    const uint32_t xw[4] = {x.x, x.y, x.z, x.w}, yw[4] = {y.x, y.y, y.z, y.w}, zw[4] = {z.x, z.y, z.z, z.w};
    uint32_t o[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 f = unpack_bf16(xw[q]), g = unpack_bf16(yw[q]), h = unpack_bf16(zw[q]);
      o[q] = pack_bf16(f.x + g.x + h.x, f.y + g.y + h.y);
    }
    reinterpret_cast<uint4*>(out)[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

__global__ void __launch_bounds__(256) gelu_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, long long n8) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    const uint4 u = reinterpret_cast<const uint4*>(x)[i];
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
    uint32_t o[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 f = unpack_bf16(w[q]);
      o[q] = pack_bf16(gelu_erf(f.x), gelu_erf(f.y));
    }
    reinterpret_cast<uint4*>(y)[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

__global__ void __launch_bounds__(256) cast_f32_bf16_kernel(const float* __restrict__ x, bf16* __restrict__ y,
                                                            long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    y[i] = __float2bfloat16(x[i]);
}

// Residual branch without a norm (pre-LN blocks): z = resid + dropout(a); one pass, mask from (seed, flat index).
__global__ void __launch_bounds__(256) add_dropout_fwd_kernel(const bf16* __restrict__ resid, const bf16* __restrict__ a,
                                                              bf16* __restrict__ z, long long n8, unsigned long long seed,
                                                              const unsigned long long* __restrict__ seed_off,
                                                              unsigned int thresh, float scale) {
  const unsigned long long seed_eff = seed + ((thresh && seed_off) ? *seed_off : 0ull);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    const uint4 r4 = reinterpret_cast<const uint4*>(resid)[i];
    const uint4 a4 = reinterpret_cast<const uint4*>(a)[i];
    const uint32_t rw[4] = {r4.x, r4.y, r4.z, r4.w}, aw[4] = {a4.x, a4.y, a4.z, a4.w};
    uint32_t o[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 r2 = unpack_bf16(rw[q]);
      float2 a2 = unpack_bf16(aw[q]);
      if (thresh) {
        const DropPair h = dropout_pair_bits(seed_eff, (unsigned long long)i * 4 + q);
        a2.x = (h.a >= thresh) ? a2.x * scale : 0.f;
        a2.y = (h.b >= thresh) ? a2.y * scale : 0.f;
      }
      o[q] = pack_bf16(r2.x + a2.x, r2.y + a2.y);
    }
    reinterpret_cast<uint4*>(z)[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// Backward of the above with gradient fan-in: dz = dy0 (+ dy1 + dy2); da = dropout'(dz).
__global__ void __launch_bounds__(256) sum_dropout_bwd_kernel(const bf16* __restrict__ dy0, const bf16* __restrict__ dy1,
                                                              const bf16* __restrict__ dy2, bf16* __restrict__ dz,
                                                              bf16* __restrict__ da, long long n8, unsigned long long seed,
                                                              const unsigned long long* __restrict__ seed_off,
                                                              unsigned int thresh, float scale) {
  const unsigned long long seed_eff = seed + ((thresh && seed_off) ? *seed_off : 0ull);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    const uint4 x = reinterpret_cast<const uint4*>(dy0)[i];
    uint4 y = make_uint4(0, 0, 0, 0), w = make_uint4(0, 0, 0, 0);
    if (dy1) y = reinterpret_cast<const uint4*>(dy1)[i];
    if (dy2) w = reinterpret_cast<const uint4*>(dy2)[i];
    const uint32_t xw[4] = {x.x, x.y, x.z, x.w}, yw[4] = {y.x, y.y, y.z, y.w}, ww[4] = {w.x, w.y, w.z, w.w};
    uint32_t oz[4], oa[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 f = unpack_bf16(xw[q]), g = unpack_bf16(yw[q]), h2 = unpack_bf16(ww[q]);
      const float sx = f.x + g.x + h2.x, sy = f.y + g.y + h2.y;
      oz[q] = pack_bf16(sx, sy);
      float ax = sx, ay = sy;
      if (thresh) {
        const DropPair h = dropout_pair_bits(seed_eff, (unsigned long long)i * 4 + q);
        ax = (h.a >= thresh) ? sx * scale : 0.f;
        ay = (h.b >= thresh) ? sy * scale : 0.f;
      }
      oa[q] = pack_bf16(ax, ay);
    }
    reinterpret_cast<uint4*>(dz)[i] = make_uint4(oz[0], oz[1], oz[2], oz[3]);
    if (da) reinterpret_cast<uint4*>(da)[i] = make_uint4(oa[0], oa[1], oa[2], oa[3]);
  }
}

// dst_r[i] += src[i] * scale for every replica r (own + peers over NVLink): the all-reduce of the small, non-GEMM
// gradients (biases, norm gains, embedding rows) in the same push style as the fused wgrad kernel.
struct PushPtrs {
  float* p[8];
};
__global__ void __launch_bounds__(256) push_add_kernel(const float* __restrict__ src, PushPtrs dst, int n_dst,
                                                       long long n4, float scale) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 v = reinterpret_cast<const float4*>(src)[i];
    v.x *= scale;
    v.y *= scale;
    v.z *= scale;
    v.w *= scale;
    for (int r = 0; r < n_dst; ++r) atomicAdd(reinterpret_cast<float4*>(dst.p[r]) + i, v);
  }
}

// NVLS variant: ONE multimem.red per 16 bytes adds the value into every replica's buffer inside the NVSwitch (the
// destination is the multicast address of the replicas' symmetric buffers), so each GPU sends its gradient once
// instead of once per peer.  Used for data-parallel degrees >= 4, where the per-peer pushes would multiply the bytes.
__global__ void __launch_bounds__(256) push_add_multimem_kernel(const float* __restrict__ src, float* mc_dst, long long n4,
                                                                float scale) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = reinterpret_cast<const float4*>(src)[i];
    asm volatile("multimem.red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_dst + 4 * i),
                 "f"(v.x * scale), "f"(v.y * scale), "f"(v.z * scale), "f"(v.w * scale)
                 : "memory");
  }
}

static int ew_grid(long long work_items) {
  long long b = (work_items + 255) / 256;
  const long long cap = (long long)sm_count() * 8;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace ppb

using namespace ppb;

extern "C" int ppb_embedding_fwd(const long long* idx, const void* W, void* out, int T, int d, int V, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(T > 0 && d > 0 && (d % 8) == 0, "ppb_embedding_fwd: d must be a multiple of 8 (got %d)", d);
  embedding_fwd_kernel<<<ew_grid((long long)T * 32), 256, 0, stream>>>(idx, (const bf16*)W, (bf16*)out, T, d, V);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int ppb_embedding_bwd(const long long* idx, const void* dy0, const void* dy1, const void* dy2, float* dW,
                                 int T, int d, int V, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(T > 0 && d > 0 && (d % 8) == 0, "ppb_embedding_bwd: d must be a multiple of 8 (got %d)", d);
  embedding_bwd_kernel<<<ew_grid((long long)T * 32), 256, 0, stream>>>(idx, (const bf16*)dy0, (const bf16*)dy1,
                                                                       (const bf16*)dy2, dW, T, d, V);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int ppb_colsum_accum(const void* dy, long long ld, float* db, int T, int N, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(T > 0 && N > 0 && (ld % 2) == 0, "ppb_colsum_accum: ld must be even");
  const int col_blocks = (N + 63) / 64;
  int slabs = (sm_count() * 4 + col_blocks - 1) / col_blocks;
  if (slabs > (T + 63) / 64) slabs = (T + 63) / 64;
  if (slabs < 1) slabs = 1;
  const int rows_per = (T + slabs - 1) / slabs;
  dim3 grid(col_blocks, (T + rows_per - 1) / rows_per);
  colsum_kernel<<<grid, 256, 0, stream>>>((const bf16*)dy, ld, db, T, N, rows_per);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int ppb_add_n_bf16(const void* a, const void* b, const void* c, void* out, long long n, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(n > 0 && (n % 8) == 0, "ppb_add_n_bf16: element count must be a multiple of 8");
  add_n_kernel<<<ew_grid(n / 8), 256, 0, stream>>>((const bf16*)a, (const bf16*)b, (const bf16*)c, (bf16*)out, n / 8);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int ppb_gelu_fwd(const void* x, void* y, long long n, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(n > 0 && (n % 8) == 0, "ppb_gelu_fwd: element count must be a multiple of 8");
  gelu_fwd_kernel<<<ew_grid(n / 8), 256, 0, stream>>>((const bf16*)x, (bf16*)y, n / 8);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

// y[t, c] = bf16(x[t, c] + bias[c]): finishes a reduce-scatter (fp32 partial sums -> activation dtype, + the replicated
// bias of a row-parallel layer) in one pass; bias may be null.  4 columns per thread.
__global__ void __launch_bounds__(256) bias_cast_f32_bf16_kernel(const float* __restrict__ x, const bf16* __restrict__ bias,
                                                                 bf16* __restrict__ y, long long n4, int d4) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 v = reinterpret_cast<const float4*>(x)[i];
    if (bias) {
      const uint2 b = reinterpret_cast<const uint2*>(bias)[i % d4];
      const float2 b0 = unpack_bf16(b.x), b1 = unpack_bf16(b.y);
      v.x += b0.x;
      v.y += b0.y;
      v.z += b1.x;
      v.w += b1.y;
    }
    reinterpret_cast<uint2*>(y)[i] = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
  }
}

extern "C" int ppb_bias_cast_f32_to_bf16(const float* x, const void* bias, void* y, long long T, int d, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(T > 0 && d > 0 && (d % 4) == 0, "ppb_bias_cast_f32_to_bf16: need d %% 4 == 0 (got %d)", d);
  bias_cast_f32_bf16_kernel<<<ew_grid(T * d / 4), 256, 0, stream>>>(x, (const bf16*)bias, (bf16*)y, T * d / 4, d / 4);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int ppb_cast_f32_to_bf16(const float* x, void* y, long long n, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if (n <= 0) return 0;
  cast_f32_bf16_kernel<<<ew_grid(n), 256, 0, stream>>>(x, (bf16*)y, n);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}


extern "C" int ppb_push_add_f32(const float* src, void* const* dst_ptrs, int n_dst, long long n, float scale,
                                void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(n_dst >= 1 && n_dst <= 8, "ppb_push_add_f32: 1..8 destinations supported");
  PPB_REQUIRE(n >= 0 && (n % 4) == 0, "ppb_push_add_f32: element count must be a multiple of 4");
  if (n == 0) return 0;
  PushPtrs d;
  for (int r = 0; r < 8; ++r) d.p[r] = reinterpret_cast<float*>(dst_ptrs[r < n_dst ? r : 0]);
  push_add_kernel<<<ew_grid(n / 4), 256, 0, stream>>>(src, d, n_dst, n / 4, scale);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}


extern "C" int ppb_push_add_multimem_f32(const float* src, void* mc_dst, long long n, float scale, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(mc_dst != nullptr && (reinterpret_cast<uintptr_t>(mc_dst) % 16) == 0, "ppb_push_add_multimem_f32: need a 16-byte aligned multicast address");
  PPB_REQUIRE(n >= 0 && (n % 4) == 0, "ppb_push_add_multimem_f32: element count must be a multiple of 4");
  if (n == 0) return 0;
  push_add_multimem_kernel<<<ew_grid(n / 4), 256, 0, stream>>>(src, reinterpret_cast<float*>(mc_dst), n / 4, scale);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int ppb_add_dropout_fwd(const void* resid, const void* a, void* z, long long n, float dropout_p,
                                   unsigned long long seed, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(n > 0 && (n % 8) == 0, "ppb_add_dropout_fwd: element count must be a multiple of 8");
  const unsigned int th = dropout_p > 0.f ? dropout_thr32(dropout_p) : 0u;
  add_dropout_fwd_kernel<<<ew_grid(n / 8), 256, 0, stream>>>((const bf16*)resid, (const bf16*)a, (bf16*)z, n / 8, seed,
                                                             seed_offset_ptr(), th, dropout_p < 1.f ? 1.f / (1.f - dropout_p) : 0.f);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int ppb_sum_dropout_bwd(const void* dy0, const void* dy1, const void* dy2, void* dz, void* da, long long n,
                                   float dropout_p, unsigned long long seed, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  PPB_REQUIRE(n > 0 && (n % 8) == 0, "ppb_sum_dropout_bwd: element count must be a multiple of 8");
  const unsigned int th = (da && dropout_p > 0.f) ? dropout_thr32(dropout_p) : 0u;
  sum_dropout_bwd_kernel<<<ew_grid(n / 8), 256, 0, stream>>>((const bf16*)dy0, (const bf16*)dy1, (const bf16*)dy2, (bf16*)dz,
                                                             (bf16*)da, n / 8, seed, seed_offset_ptr(), th,
                                                             dropout_p < 1.f ? 1.f / (1.f - dropout_p) : 0.f);
  PPB_CUDA_OK(cudaGetLastError());
  return 0;
}
