/* libppb200 -- C ABI of the sm_100a kernel library behind the pipeline-parallel training framework.
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; ppb_last_error() returns the message of the
 *     last failure on the calling thread;
 *   - all pointers are device pointers on the CURRENT device unless stated otherwise; `stream` is a
 *     cudaStream_t passed as void*; nothing synchronises the host;
 *   - bf16 tensors are row-major with a leading dimension (`ld*`, in elements) that keeps rows 16-byte aligned;
 *   - "accum" outputs are fp32 and are added to (they persist across micro-batches);
 *   - no function allocates device memory: work space is passed in by the caller.
 *
 * The reference (aa5490/Distributed-Training-with-Pipeline-Parallelism) has no native code; each entry point
 * cites the op(s) of the reference's ATen op stream it replaces.
 */
#ifndef PPB200_H_
#define PPB200_H_

#ifdef __cplusplus
extern "C" {
#endif

const char* ppb_last_error(void);
int ppb_version(void);
/* Optional device pointer to a 64-bit value that every dropout-using kernel adds to its seed at run time (per
 * device; NULL disables).  Lets CUDA graphs replay with fresh masks. */
int ppb_set_dropout_offset_ptr(const unsigned long long* p);
/* Leave `n` SMs free of persistent-kernel CTAs (for concurrently running NCCL P2P kernels); also $PPB_RESERVED_SMS. */
int ppb_set_reserved_sms(int n);

/* ---- K1/K2/K3: tcgen05 GEMM ------------------------------------------------------------------------------
 * C[M,N] (=|+=) epilogue(alpha * op(A)[M,K] * op(B)[N,K]^T)
 *   a_mn_major = 0: A is stored [M, lda] (K contiguous);  1: A is stored [K, lda] (M contiguous)
 *   b_mn_major = 0: B is stored [N, ldb] (K contiguous);  1: B is stored [K, ldb] (N contiguous)
 * forward x*W^T: (0,0); dgrad dy*W: (0,1); wgrad dy^T*x: (1,1).
 * Replaces aten.mm / aten.addmm (+ bias add, relu, dropout) -- helper:40-46, torch transformer.py:1143-1199. */
enum { PPB_OUT_BF16 = 0, PPB_OUT_F32 = 1, PPB_OUT_F32_ACCUM = 2 };
enum {
  PPB_EPI_NONE = 0,
  PPB_EPI_BIAS = 1,              /* + bias[n]                                      */
  PPB_EPI_BIAS_RELU = 2,         /* relu(. + bias)                                 */
  PPB_EPI_BIAS_GELU = 3,         /* gelu_erf(. + bias)                             */
  PPB_EPI_BIAS_RELU_DROPOUT = 4, /* dropout(relu(. + bias)), mask from (seed, m*N+n) */
  PPB_EPI_MASK_POS = 5,          /* . * (aux[m,n] > 0)   (ReLU(+dropout) backward; pass alpha = 1/(1-p)) */
  PPB_EPI_GELU_BWD = 6           /* . * gelu'(aux[m,n])  (aux = pre-activation)    */
};
int ppb_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, long long lda, long long ldb,
                  long long ldc, int a_mn_major, int b_mn_major, int out_mode, int epi, float alpha, const void* bias,
                  const void* aux, long long ld_aux, unsigned long long seed, float dropout_p, int force_bn,
                  int force_splits, void* stream);

/* CTAs per output tile: 1 = one CTA computes a 128 x BN tile (cta_group::1); 2 = a 2-CTA cluster computes a 256 x BN
 * tile with cta_group::2 UMMAs, each CTA staging its own A rows and half of the B tile; 0 (default) = automatic.
 * Also $PPB_GEMM_CG.  Operands of a pair launch must be this GPU's memory. */
int ppb_set_gemm_cta_group(int cg);

/* ---- K10: last-micro-batch wgrad GEMM fused with the data-parallel gradient all-reduce ---------------------------
 * accum[M,N] += A^T B exactly like the wgrad form of ppb_gemm_bf16 (A stored [K, lda], B stored [K, ldb], split-K,
 * TMA reduce-add); additionally, as soon as a 32-row slab of a tile is complete, the CTA that finished it last
 * reduce-adds the slab into dst_r[M,N] for every replica r in dst_ptrs (host array of n_dst device pointers: the
 * caller's own reduced-gradient buffer and the peer-mapped symmetric buffers of the other replicas, leading
 * dimension ldc).  Every replica calls it on its own data; after a cross-replica barrier every dst holds
 * sum_r accum_r.  `scratch` (ppb_gemm_allreduce_scratch_bytes, zero-filled once by the caller; the kernel re-zeroes what it used) holds the arrival counters.  ppb_push_add_f32 does
 * the same for the small non-GEMM gradients.  The reference has no data parallelism at all (SURVEY.md section
 * 2.5); the baseline this replaces is wgrad GEMM + ncclAllReduce. */
long long ppb_gemm_allreduce_scratch_bytes(int M, int N);
int ppb_gemm_allreduce_bf16(const void* A, const void* B, float* accum, void* const* dst_ptrs, int n_dst, int M, int N,
                            int K, long long lda, long long ldb, long long ldc, void* scratch, long long scratch_bytes,
                            int force_bn, int force_splits, void* stream);
/* ---- K12: all-gather -> GEMM (column-parallel layer under tensor parallelism, sequence-sharded input) ------------
 * C[M,N] = epi(A[M,K] * op(B)^T) where A is the concatenation of `tp` row shards a_shards[r] ([M/tp, lda], local or
 * peer-mapped symmetric memory); the TMA producer reads each tile from the shard that owns it, own shard first.
 * M/tp must be a multiple of 128.  Baseline replaced: ncclAllGather + GEMM (torch _fused_all_gather_matmul). */
int ppb_gemm_allgather_bf16(void* const* a_shards, const void* B, void* C, int M, int N, int K, long long lda,
                            long long ldb, long long ldc, int tp, int rank, int b_mn_major, int epi, const void* bias,
                            void* stream);
/* K12, fused form: ONE kernel gathers and multiplies.  A copier warp per CTA streams every shard (the peers' over
 * NVLink with plain bulk copies, each byte once) into the local gathered buffer `full` [M, K] and publishes 128-row
 * panels through release counters in `flags` (M/128 counters, zero-filled once by the caller, then never touched by
 * anyone else); the GEMM tiles multiply the own-shard panels first and each remote panel as soon as it has landed.
 * Shards must be contiguous ([M/tp, K], K % 64 == 0).  `target` = chunks_per_panel(K) * (calls made with these flags
 * so far, this one included).  `full` keeps the gathered activations (the backward weight-gradient GEMM wants them).
 * Every CTA of the launch must be co-resident (<= one per SM): run it on a stream that has the GPU to itself. */
int ppb_gemm_allgather_fused_chunks_per_panel(int K);
int ppb_gemm_allgather_fused_bf16(void* const* a_shards, void* full, unsigned int* flags, unsigned int target, const void* B,
                                  void* C, int M, int N, int K, long long ldb, long long ldc, int tp, int rank,
                                  int b_mn_major, int epi, const void* bias, const void* aux, long long ld_aux, float alpha,
                                  void* stream);
/* ---- K11: GEMM -> reduce-scatter (row-parallel layer under tensor parallelism) -----------------------------------
 * dst_shards[o][M/tp, N] (fp32, zeroed by the caller) += rows [o*M/tp, (o+1)*M/tp) of A[M,K] * op(B)^T, for every
 * owner o: each 32-row slab of the partial product is TMA-reduce-added into its owner's buffer (local or over
 * NVLink).  After a cross-rank barrier dst_shards[rank] holds the reduced block.  Baseline replaced: GEMM +
 * ncclReduceScatter (torch _fused_matmul_reduce_scatter). */
int ppb_gemm_reducescatter_bf16(const void* A, const void* B, void* const* dst_shards, int M, int N, int K,
                                long long lda, long long ldb, long long ldc, int tp, int b_mn_major, void* stream);
int ppb_push_add_f32(const float* src, void* const* dst_ptrs, int n_dst, long long n, float scale, void* stream);
/* NVLS form of the push: mc_dst is the MULTICAST address of the replicas' symmetric buffers; one multimem.red per 16
 * bytes adds scale * src into every replica inside the NVSwitch, so a GPU sends each gradient byte once whatever the
 * data-parallel degree (the per-peer push sends it dp-1 times).  n % 4 == 0. */
int ppb_push_add_multimem_f32(const float* src, void* mc_dst, long long n, float scale, void* stream);

/* ---- K5: fused residual-add + dropout + LayerNorm / RMSNorm ------------------------------------------------
 * fwd: z = resid + dropout(a); y = norm(z)*gamma(+beta).  Either of resid / a may be NULL; z, mean may be NULL.
 * bwd: dy = dy0(+dy1+dy2); dz = norm'(dy); da = dropout'(dz) (da may be NULL); dgamma/dbeta are fp32 accum.
 * `scratch` needs ppb_norm_bwd_scratch_bytes() bytes, zero-filled once; concurrent launches must not share it.
 * d <= 2048 keeps each row in registers; 2048 < d <= 8192 takes a row-pass + column-pass route (no scratch use).
 * Replaces aten.add + bernoulli_/div_/mul + native_layer_norm(_backward) -- torch transformer.py:1144-1153. */
int ppb_add_dropout_norm_fwd(const void* resid, const void* a, const void* gamma, const void* beta, void* z, void* y,
                             float* mean, float* rstd, int T, int d, float eps, float dropout_p,
                             unsigned long long seed, int rms, void* stream);
long long ppb_norm_bwd_scratch_bytes(int T, int d);
int ppb_norm_bwd(const void* dy0, const void* dy1, const void* dy2, const void* z, const float* mean,
                 const float* rstd, const void* gamma, void* dz, void* da, float* dgamma, float* dbeta, void* scratch,
                 long long scratch_bytes, int T, int d, float dropout_p, unsigned long long seed, int rms,
                 void* stream);

/* Allocation-free variant of ppb_norm_bwd.  No counterpart in the reference (which has no native code and no
 * notion of a caller that cannot allocate).  For hosts that may not create device memory -- a runtime that owns
 * the whole HBM arena, a plug-in hook inside another framework, a thread that must not hit cudaMalloc's implicit
 * synchronisation -- the dgamma/dbeta scratch lives in a fixed ~2.4 MiB array inside the library image (held on
 * every GPU that loads the library, until exit).  Limit: d <= ppb_norm_bwd_noalloc_max_d() (1024); larger d is
 * refused.  The first call on a device looks the array up (do it before any stream capture); calls share the
 * array, so calls that can overlap on the GPU race -- order them. */
long long ppb_norm_bwd_noalloc_max_d(void);
int ppb_norm_bwd_noalloc(const void* dy0, const void* dy1, const void* dy2, const void* z, const float* mean,
                         const float* rstd, const void* gamma, void* dz, void* da, float* dgamma, float* dbeta, int T,
                         int d, float dropout_p, unsigned long long seed, int rms, void* stream);

/* ---- K4: flash attention forward / backward on tcgen05 (csrc/attention_fwd.cu, attention_bwd.cu) ---------------
 * q [B*Sq, ldq], k / v [B*Skv, ldk / ldv], o [B*Sq, ldo]: bf16 rows of H heads x dh columns (column-sliced views of a
 * packed QKV activation are fine; heads are addressed through 4-D TMA maps, nothing is transposed or copied).
 * lse [B, H, Sq] fp32 (natural log).  causal needs Skv >= Sq; dropout (probability dropout, counter-based mask from
 * (seed, element index), regenerated in backward) needs an even Skv.  dh: multiple of 8, 8..256; 8..192 run on the tcgen05
 * kernels (tiles of 64 / 128 / 192 columns, zero-padded by TMA; 129..192: two threads per query row in the forward, the
 * backward as a dV launch and a dK + dQ launch), 193..256 on a CUDA-core path.  PPB_ATTN_TC_MAX_DH=128 in the
 * environment sends 129..192 to the CUDA-core path as well (A/B measurements).  reserved* arguments must be 0 / NULL.
 * Backward writes dq / dk / dv with the row strides of q / k / v; `scratch` = ppb_attn_bwd_scratch_bytes(B, H,
 * max(Sq, Skv), dh) bytes, 256-byte aligned, no initialisation needed (delta rows + the fp32 dQ accumulator).
 * Replaces aten.bmm x2 + mul.Scalar + _safe_softmax + dropout and their backwards -- torch transformer.py:1158-1194,
 * reached from helper:40-44, :52. */
int ppb_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B, int H, int Sq, int Skv, int dh,
                 long long ldq, long long ldk, long long ldv, long long ldo, int causal, int reserved0, int reserved1,
                 float scale, float dropout_p, unsigned long long seed, void* stream);
long long ppb_attn_bwd_scratch_bytes(int B, int H, int S, int dh);
int ppb_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* dO, const float* lse, void* dq,
                 void* dk, void* dv, void* scratch, void* reserved_ptr, int B, int H, int Sq, int Skv, int dh,
                 long long ldq, long long ldk, long long ldv, long long ldo, int causal, int reserved0, int reserved1,
                 float scale, float dropout_p, unsigned long long seed, void* stream);

/* ---- K7: fused softmax-cross-entropy (loss rows + in-place dlogits) ----------------------------------------
 * Replaces aten._log_softmax + nll_loss_forward/backward + logits.clone() -- helper:54, helper:197-201. */
int ppb_xent_fwd_bwd(void* logits, long long ld, const long long* targets, float* row_loss, int T, int V,
                     float grad_scale, int write_grad, void* stream);
int ppb_reduce_sum_f32(const float* in, long long n, float* out, float scale, int accumulate, void* stream);

/* ---- K9: fused AdamW over flat arenas (+ micro-batch grad averaging, + global-norm clip, + grad zeroing) ----
 * The reference has no optimizer (helper:113-131); grad averaging replaces p.grad.div_(m), torch stage.py:589. */
int ppb_adamw_step(void* param_bf16, float* master, float* m, float* v, float* grad, long long n, float lr,
                   float beta1, float beta2, float eps, float weight_decay, int step, float grad_scale,
                   const float* clip_coef, int zero_grad, void* stream);
int ppb_sumsq_f32(const float* x, long long n, float* out, void* stream);
int ppb_clip_coef(const float* sumsq, float grad_scale, float max_norm, float* coef, void* stream);

/* ---- K6/K8: embedding, bias-grad, fan-in add, GELU, cast ---------------------------------------------------
 * Replaces aten.embedding / embedding_dense_backward (helper:36,50), aten.sum.dim_IntList, aten.add.Tensor. */
int ppb_embedding_fwd(const long long* idx, const void* W, void* out, int T, int d, int V, void* stream);
int ppb_embedding_bwd(const long long* idx, const void* dy0, const void* dy1, const void* dy2, float* dW, int T, int d,
                      int V, void* stream);
int ppb_colsum_accum(const void* dy, long long ld, float* db, int T, int N, void* stream);
int ppb_add_n_bf16(const void* a, const void* b, const void* c, void* out, long long n, void* stream);
int ppb_gelu_fwd(const void* x, void* y, long long n, void* stream);
/* residual branch without a norm (pre-LN / GPT blocks): z = resid + dropout(a), and its backward with gradient
 * fan-in: dz = dy0(+dy1+dy2), da = dropout'(dz) (da may be NULL).  Replaces aten.add + dropout of pre-LN layers. */
int ppb_add_dropout_fwd(const void* resid, const void* a, void* z, long long n, float dropout_p, unsigned long long seed,
                        void* stream);
int ppb_sum_dropout_bwd(const void* dy0, const void* dy1, const void* dy2, void* dz, void* da, long long n,
                        float dropout_p, unsigned long long seed, void* stream);
int ppb_cast_f32_to_bf16(const float* x, void* y, long long n, void* stream);
/* y[T, d] (bf16) = x[T, d] (fp32) + bias[d] (bf16, may be NULL): finishes a reduce-scatter in one pass (d % 4 == 0). */
int ppb_bias_cast_f32_to_bf16(const float* x, const void* bias, void* y, long long T, int d, void* stream);

/* ---- Llama-style block pieces (csrc/llama_ops.cu) -------------------------------------------------------------
 * No counterpart in the reference (its block is nn.TransformerDecoderLayer: ReLU FFN, no positional encoding,
 * helper:38-46); needed by BASELINE.json config 4 "Llama-style 7B".
 * SwiGLU on the packed linear1 output gu [T, 2f] (gate columns first): h = silu(g) * u, and its backward.
 * RoPE (rotate-half pairs (i, i + dh/2)) applied in place to n_heads heads starting at column col0 of x [T, ld];
 * position = row % S; cos_sin = float2 [S, dh/2]; inverse != 0 applies the transpose rotation (the backward). */
int ppb_swiglu_fwd(const void* gu, long long ld_gu, void* h, long long ld_h, long long T, int f, void* stream);
int ppb_swiglu_bwd(const void* dh, long long ld_dh, const void* gu, long long ld_gu, void* dgu, long long ld_dgu,
                   long long T, int f, void* stream);
int ppb_rope_inplace(void* x, long long ld, int col0, long long T, int S, int n_heads, int dh, const void* cos_sin,
                     int inverse, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PPB200_H_ */
