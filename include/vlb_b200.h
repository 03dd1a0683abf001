/* libvlb_b200 -- C ABI of the B200-native phantom_vlb training-step hot path.
 *
 * Conventions (every entry point):
 *   - plain C types only; every pointer is a raw CUDA device pointer owned by the caller (16-byte aligned,
 *     contiguous unless a leading dimension is passed); the library never allocates device memory;
 *   - work is enqueued on `stream` (a cudaStream_t passed as void*), the call never synchronises;
 *   - returns 0 on success, non-zero (a cudaError_t or 1 for argument errors) otherwise; the message is
 *     available from vlb_last_error_string(); no exception crosses the boundary;
 *   - sm_100a only: vlb_check_device() fails on anything else; there is no fallback path.
 * "Replaces" cites the reference (courtois-neuromod/phantom_vlb) call site each function stands in for;
 * paths are relative to the reference repository root.
 */
#ifndef VLB_B200_H
#define VLB_B200_H
#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ runtime */
const char* vlb_last_error_string(void);
int vlb_abi_version(void);
int vlb_check_device(void);
void vlb_tmap_cache_clear(void);

/* ------------------------------------------------------------------ tcgen05 GEMM (+ fused LoRA k-block)
 * C[M,N] = epi(alpha * (A·Bᵀ + A2·B2ᵀ)),  epi(v) = bf16(act(v + bias[n])) (+ residual[m,n]).
 * a_mn_major/b_mn_major = 0: operand stored [rows, K] (K contiguous); 1: stored [K, rows].
 * c_is_f32 = 1: C is fp32 [M, ldc] written (accumulate=1: added to) without rounding, bias/act/residual ignored.
 * split_k (c_is_f32=1, accumulate=1 only): 0 = auto, 1 = off, n = contraction split n ways; partial tiles are added with
 * red.global.add.f32, so the summation order (not the value beyond fp32 rounding) varies from run to run.
 * Replaces every nn.Linear / peft lora.Linear on the path: HF MistralDecoderLayer + LoraConfig targets
 * (src/litmodule/videollama2_vlb_litmodule.py:101-120), CLIP encoder linears (VideoLLaMA2/videollama2/model/
 * encoder.py:50), STC 1x1 convs / readout (VideoLLaMA2/videollama2/model/projector.py:153-187). */
int vlb_gemm_bf16(const void* A, int lda, const void* B, int ldb, void* C, int ldc,
                  int M, int N, int K,
                  const void* A2, int lda2, const void* B2, int ldb2, int K2,
                  float alpha, const void* bias, const void* residual, int ldr, int act,
                  int a_mn_major, int b_mn_major, int c_is_f32, int accumulate,
                  int block_n, int max_ctas, int split_k, void* stream);

/* ------------------------------------------------------------------ flash attention (tcgen05, TMEM-resident S/O)
 * q/k/v point into a fused projection buffer with row stride ld_qkv: q [B*S, Hq*D], k/v [B*S, Hkv*D].
 * o [B*S, Hq*D] (row stride ld_o); lse [B, Hq, S] fp32 in the log2 domain (nullable for inference).
 * Replaces flash-attn 2 selected by `_attn_implementation="flash_attention_2"`
 * (src/litmodule/videollama2_vlb_litmodule.py:71-72; VideoLLaMA2/videollama2/model/encoder.py:24). */
int vlb_attn_fwd(const void* q, const void* k, const void* v, int ld_qkv, void* o, int ld_o, void* lse_f32,
                 int B, int S, int Hq, int Hkv, int D, int causal, float scale, void* stream);

/* Debug only (VLB_ATT_TS=1): clock64 stamps of the softmax / MMA-issuer hand-offs of the heaviest forward CTA, 8 per key
 * block (tools/attn_timeline.py). */
int vlb_attn_debug_ts(long long* host_out, int n);
int vlb_attn_bwd_debug_ts(long long* host_out, int n);      /* backward: 16 stamps per 64-query step of CTA 0 */

/* Backward (head_dim 128, S % 128 == 0). delta_f32 [B,Hq,S] and dq_f32 [B*S, Hq*D] are caller-provided scratch/outputs
 * (delta_f32 is left holding scale * rowsum(dO o O): the softmax scale travels inside dS);
 * dq is accumulated in fp32 (zeroed inside) and converted by vlb_rope_bwd_f32; dk/dv are bf16 rows with stride ld_dkv. */
int vlb_attn_bwd(const void* q, const void* k, const void* v, int ld_qkv, const void* o, const void* d_o,
                 int ld_o, const void* lse_f32, void* delta_f32, void* dq_f32, void* dk, void* dv, int ld_dkv,
                 int B, int S, int Hq, int Hkv, int D, int causal, float scale, void* stream);

/* ------------------------------------------------------------------ norms / RoPE / SwiGLU (HBM-bound row kernels)
 * Replace MistralRMSNorm, nn.LayerNorm / timm LayerNorm2d (NHWC rows), apply_rotary_pos_emb and MistralMLP's
 * act_fn(gate)*up reached from VideoLLaMA2/videollama2/model/videollama2_mistral.py:94-104, encoder.py:50,
 * projector.py:153-184. */
int vlb_rmsnorm_fwd(const void* x, const void* w, void* y, void* rstd, long long M, int H, float eps, void* stream);
int vlb_rmsnorm_bwd(const void* dy, const void* x, const void* w, const void* rstd, const void* dres,
                    void* dx, long long M, int H, void* stream);
int vlb_layernorm_fwd(const void* x, const void* w, const void* b, const void* residual, void* y,
                      long long M, int H, float eps, int act, void* stream);
int vlb_rope(void* qkv, int ld, const void* cos_tab, const void* sin_tab, long long M, int S, int nheads,
             int D, int inverse, void* stream);
int vlb_rope_bwd_f32(const void* dq_f32, void* dqkv, int ld, const void* cos_tab, const void* sin_tab,
                     long long M, int S, int nheads, int D, void* stream);
int vlb_swiglu_fwd(const void* gu, void* act, long long M, int I, void* stream);
int vlb_swiglu_bwd(const void* gu, const void* dact, void* dgu, long long M, int I, void* stream);

/* ------------------------------------------------------------------ vision tower / STC connector / splice gathers
 * CLIPVisionEmbeddings + pre_layrnorm (HF, via encoder.py:50), feature_select (encoder.py:31-39), timm RegStage
 * depthwise conv + SE (projector.py:153-184), Conv3d sampler gather (projector.py:164-174), the <video> splice of
 * prepare_inputs_labels_for_multimodal (videollama2_arch.py:161-263), batch["language"].long()
 * (src/litmodule/videollama2_vlb_litmodule.py:269) and make_weight_mask (litmodule.py:178-203). */
int vlb_patch_im2col(const void* pixels_f32, void* cols, long long N, int img, int patch, int kpad, void* stream);
/* uint8 frames [N,3,img,img]: CLIPImageProcessor rescale (1/255) + normalise ((x-mean)/std, HOST pointers to 3 floats
 * each) fused into the gather -- the arithmetic src/preprocessing/videollama2_vlb_extractfeatures.py:344-350 does on the
 * CPU before writing fp32 pixels to HDF5. */
int vlb_patch_im2col_u8(const void* pixels_u8, void* cols, long long N, int img, int patch, int kpad,
                        const float* mean3, const float* std3, void* stream);
int vlb_clip_embed(const void* patch, const void* cls, const void* pos, const void* lnw, const void* lnb,
                   void* out, long long N, int T, int H, float eps, void* stream);
int vlb_drop_cls(const void* in, void* out, long long N, int T, int H, void* stream);
int vlb_dwconv3x3(const void* x, const void* w9c, void* y, long long F, int Hh, int Ww, int C, void* stream);
int vlb_se_pool(const void* x, void* out, long long F, int HW, int C, void* stream);
int vlb_se_scale(const void* x, const void* gate, void* y, long long F, int HW, int C, void* stream);
int vlb_conv3d_gather(const void* x, void* cols, long long B, int T, int Hh, int Ww, int C, void* stream);
int vlb_splice_embed(const void* ids_i64, const void* table, const void* vis, void* out, void* video_pos_i32,
                     long long B, int L, int Nv, int H, int vocab, long long video_id, void* stream);
int vlb_ids_to_long(const void* ids_f32, void* ids_i64, long long n, void* stream);
int vlb_weight_mask(const void* padvals_i64, const void* vis_w_f64, const void* lang_w_f64, void* out_f32,
                    int B, int Tp, int tok_per_frame, int NL, int S, void* stream);

/* ------------------------------------------------------------------ brain head (pooling, readout loss, backward)
 * layer_norm1 -> HRFConvolveLayer einsum('bse,bs->be') -> layer_norm2 -> dropout (src/utils.py:40-56,
 * src/litmodule/videollama2_vlb_litmodule.py:245-251), F.mse_loss + l2_lambda*||W||^2 (src/utils.py:66-71,
 * litmodule.py:302) and their backward. The readout GEMMs themselves run on vlb_gemm_bf16. */
int vlb_hrf_pool_fwd(const void* x, const void* norm_w, const void* ln1_w, const void* ln1_b,
                     const void* wmask_f32, void* partial_f32, int B, int S, int H, int nch, float eps_rms,
                     float eps_ln, void* stream);
int vlb_hrf_pool_finalize(const void* partial_f32, const void* ln2_w, const void* ln2_b, void* pooled,
                          void* hrf, void* keep_scale_f32, int B, int H, int nch, float eps, float drop_p,
                          unsigned long long seed, void* stream);
/* vlb_hrf_pool_fwd followed by vlb_hrf_pool_finalize in ONE call that allocates nothing, for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). NO COUNTERPART IN THE REFERENCE.
 * The partial[B, nch, H] scratch between the two kernels lives in a file-scope
 * `static __device__` array of the library (2 MiB on every GPU that loads it, held until the process exits); its device
 * address is resolved once per device with cudaGetSymbolAddress. Limit: B * nch * H <= vlb_hrf_pool_static_max_floats()
 * (= 524 288 floats; e.g. B=8, H=4096, nch<=16) -- larger requests return non-zero with a message, nothing is launched.
 * Same arguments, meaning and results as the two-call form (eps_ln is used for both LayerNorms, as the engine does).
 * Every call on a device shares the array: calls that overlap on the GPU (different streams without a dependency) RACE;
 * calls on one stream are safe. Nothing needs clearing between calls: the first kernel overwrites all it later reads. */
int vlb_hrf_pool_static_max_floats(void);
int vlb_hrf_pool_fwd_static(const void* x, const void* norm_w, const void* ln1_w, const void* ln1_b,
                            const void* wmask_f32, const void* ln2_w, const void* ln2_b, void* pooled, void* hrf,
                            void* keep_scale_f32, int B, int S, int H, int nch, float eps_rms, float eps_ln,
                            float drop_p, unsigned long long seed, void* stream);
/* HRFConvolveLayer as a stand-alone layer (src/utils.py:40-56, exported by src/__init__.py:3-13):
 * out_f32[b, e] = sum_s w_f32[b, s] * emb[b, s, e]  (emb bf16 [B, S, H], fp32 accumulate; out is zeroed inside).
 * Backward of the same einsum: demb[b, s, :] = w[b, s] * dout[b, :] (bf16, nullable), dw_f32[b, s] = <emb[b, s, :], dout[b, :]>
 * (nullable). Inside the training step the pooling is fused with the final RMSNorm and LN1 (vlb_hrf_pool_fwd). */
int vlb_hrf_convolve_fwd(const void* emb, const void* w_f32, void* out_f32, int B, int S, int H, void* stream);
int vlb_hrf_convolve_bwd(const void* emb, const void* w_f32, const void* dout_f32, void* demb, void* dw_f32,
                         int B, int S, int H, void* stream);

/* n = elements of this rank's [B, P_local] slice; n_total = B * P over all voxel shards (gradient scale 2/n_total). */
int vlb_head_loss(const void* pred, const void* y_f32, void* dpred, void* acc_f32, long long n, long long n_total,
                  void* stream);
int vlb_l2_prefill(const void* W, void* grad_f32, void* acc_f32, long long n, float lambda, int write_grad,
                   void* stream);
int vlb_loss_finalize(const void* acc_f32, void* out_f32, long long n, float lambda, void* stream);
int vlb_colsum(const void* d, void* out_f32, int B, long long P, void* stream);
int vlb_ln2_bwd(const void* dy_f32, const void* keep_scale_f32, const void* pooled, const void* ln2_w,
                void* dpooled_f32, void* dg2_f32, void* db2_f32, int B, int H, float eps, void* stream);
/* dx may be NULL (frozen backbone, reference freeze_backbone=True, litmodule.py:86-90): only dg1/db1 are produced. */
int vlb_hrf_pool_bwd(const void* x, const void* norm_w, const void* ln1_w, const void* wmask_f32,
                     const void* dpooled_f32, void* dx, void* dg1_f32, void* db1_f32, int B, int S, int H,
                     int nch, float eps_rms, float eps_ln, void* stream);

/* ------------------------------------------------------------------ LoRA input dropout on the fused path
 * peft lora.Linear computes y = W x + s B(A(dropout(x))) with one independent Bernoulli mask per adapter (reference
 * lora_dropout 0.1, config/experiment/VLB_vllama2_friends_lora.yaml:30; src/litmodule/videollama2_vlb_litmodule.py:113-120).
 * vlb_dropout_bits draws the keep mask of one adapter over its input [n_tok, n_feat] as packed bits, exact p (32-bit uniform
 * against p * 2^32): word64[fb * n_tok + tok] holds features [64 fb, 64 fb + 64) of token tok. The masks are consumed where the
 * masked value is needed, never materialised as activations:
 *   vlb_lora_down_masked     xa[n_tok, 16 n] = alpha * (keep_a . x) A_a^T for the n <= 3 adapters sharing the input x
 *                            (A [16 n, n_feat] stacked lora_A matrices); x is streamed once.
 *   vlb_lora_grad_a_masked   dA_f32[16 n, n_feat] += alpha * t_a^T (keep_a . x), t [n_tok, 16 n] = s dy B.
 *   vlb_gemm_dx_lora_masked  out[M, N] = dy[M, K] W[K, N] + scale * sum_a keep_a . (t_a A_a)   (W = forward weight
 *                            [out = K, in = N]; the adapter branch of dX is formed in the GEMM epilogue).
 *   vlb_gemm_down_dx_swiglu_bwd_masked   the same for the down projection with the SwiGLU backward in the epilogue. */
int vlb_dropout_bits(void* bits64, long long n_tok, int n_feat, float p, unsigned long long seed, void* stream);
int vlb_lora_down_masked(const void* x, int ldx, const void* A, int lda, void* xa, int ldxa, int n_tok, int n_feat,
                         int n_adapters, const void* bits0, const void* bits1, const void* bits2, float alpha, void* stream);
int vlb_lora_grad_a_masked(const void* x, int ldx, const void* t, int ldt, void* dA_f32, int ldda, int n_tok, int n_feat,
                           int n_adapters, const void* bits0, const void* bits1, const void* bits2, float alpha, void* stream);
int vlb_gemm_dx_lora_masked(const void* dy, int lddy, const void* W, int ldw, void* out, int ldo, int M, int N, int K,
                            const void* t, int ldt, const void* A, int lda, int n_adapters, const void* bits0,
                            const void* bits1, const void* bits2, float scale, void* stream);
int vlb_gemm_down_dx_swiglu_bwd_masked(const void* dy, int lddy, const void* wd, int ldw, const void* gu, int ldgu,
                                       void* dgu, int lddgu, int M, int I, int H, const void* t, int ldt, const void* A,
                                       int lda, const void* bits, float scale, void* stream);

/* ------------------------------------------------------------------ LoRA input dropout (peft lora.Linear: B(A(dropout(x))),
 * reference lora_dropout 0.1, config/experiment/VLB_vllama2_friends_lora.yaml:30). Counter-based mask from (seed, element
 * index), probability quantised to 1/256; vlb_dropout_add applies the same mask to the adapter branch of dX. mask_u8 may be
 * NULL (test hook: 0/1 bytes [M, K] so the oracle can replay the mask). */
int vlb_dropout(const void* x, int ldx, void* out, int ldo, void* mask_u8, long long M, int K, float p,
                unsigned long long seed, void* stream);
int vlb_dropout_add(const void* t, int ldt, void* y, int ldy, long long M, int K, float p, unsigned long long seed,
                    void* stream);

/* ------------------------------------------------------------------ cached pooled features (frozen-backbone baseline)
 * reference: src/preprocessing/videollama2_vlb_extractfeatures.py + VLB_vllama2_friends_baseline.yaml:26-27 recompute
 * the frozen backbone every epoch; here (u, sw) = (sum_s w_s LN1hat(x_s), sum_s w_s) is extracted once per sample
 * (vlb_hrf_pool_fwd with unit LN1 affine -> vlb_pool_reduce) and the head trains from it:
 * pooled = g1*u + b1*sw (vlb_ln1_affine_fwd) -> vlb_hrf_pool_finalize(nch=1) -> readout; vlb_ln1_affine_bwd adds the
 * LN1 affine gradients. */
int vlb_pool_reduce(const void* partial_f32, const void* wmask_f32, void* u_f32, void* sw_f32, int B, int S, int H,
                    int nch, void* stream);
int vlb_ln1_affine_fwd(const void* u_f32, const void* sw_f32, const void* ln1_w, const void* ln1_b, void* out_f32,
                       int B, int H, void* stream);
int vlb_ln1_affine_bwd(const void* dpooled_f32, const void* u_f32, const void* sw_f32, void* dg1_f32, void* db1_f32,
                       int B, int H, void* stream);

/* q/k/v projection of a decoder layer with the rotary embedding fused into the GEMM epilogue (MistralAttention.forward:
 * q_proj / k_proj / v_proj followed by apply_rotary_pos_emb, reached through videollama2_mistral.py:94-104; LoRA wrap
 * src/litmodule/videollama2_vlb_litmodule.py:113-120): qkv[M, N] = x Wqkv^T (+ LoRA k-block A2/B2 as in vlb_gemm_bf16), then
 * RoPE on the first n_rot_heads heads of D columns each (q and k; v passes through), position = row % S, cos/sin bf16 [S, D/2].
 * Same rounding points as vlb_gemm_bf16 followed by vlb_rope (bit-identical). D == 128 with N % 256 == 0 runs fused on the
 * CTA-pair kernel; any other shape runs the two kernels one after the other. */
int vlb_gemm_qkv_rope(const void* x, int ldx, const void* wqkv, int ldw, void* qkv, int ldqkv, int M, int N, int K,
                      const void* A2, int lda2, const void* B2, int ldb2, int K2, float alpha,
                      const void* cos_tab, const void* sin_tab, int S, int n_rot_heads, int D, void* stream);

/* Forward gate/up projection of the decoder MLP with SwiGLU fused into the GEMM epilogue (MistralMLP.forward:
 * act_fn(gate_proj(x)) * up_proj(x), reached through videollama2_mistral.py:94-104): gu[M, 2I] (kept for the backward) and
 * act[M, I] leave the kernel together; wgu is [2I, K] gate rows then up rows; A2/B2 = the fused LoRA k-block as in
 * vlb_gemm_bf16. CTA-pair kernel: the leader stages gate rows, the peer the matching up rows. */
int vlb_gemm_gateup_swiglu(const void* x, int ldx, const void* wgu, int ldw, void* gu, int ldgu, void* act, int ldact,
                           int M, int I, int K, const void* A2, int lda2, const void* B2, int ldb2, int K2, float alpha,
                           void* stream);

/* Backward of the decoder MLP's down projection with the SwiGLU backward fused into the GEMM epilogue (autograd of
 * MistralMLP.forward): d(act) = dy Wd (+ LoRA k-block A2/B2) stays in TMEM and dgu[M, 2I] = [d(act) up silu'(gate) |
 * d(act) silu(gate)] is written directly from the saved gu[M, 2I]. wd is the forward weight [H, I]; I % 256 == 0. */
int vlb_gemm_down_dx_swiglu_bwd(const void* dy, int lddy, const void* wd, int ldw, const void* gu, int ldgu, void* dgu,
                                int lddgu, int M, int I, int H, const void* A2, int lda2, const void* B2, int ldb2, int K2,
                                float alpha, void* stream);

/* ------------------------------------------------------------------ fused head optimiser for very wide readouts
 * (BASELINE configs[3]: P ~ 250 000 voxels). The ridge gradient dW = dpred^T hrf + 2 lambda W has rank <= n_samples, so it is
 * never materialised: vlb_head_gradnorm adds ||dW||^2 to the clip accumulator from two n x n Gram matrices (plain
 * vlb_gemm_bf16 calls) and an elementwise pass over [n, P]; vlb_head_adamw_fused forms dW tile by tile on the tensor cores
 * and applies clip + AdamW (same arithmetic as vlb_adamw_step) in the epilogue. Replaces, for the ridge weight only, the
 * reference's autograd dW + clip_grad_norm_ + torch.optim.AdamW (litmodule.py:345-379). */
int vlb_head_gradnorm(const void* dpred, const void* pred, const void* bias, const void* gram_dpred_f32,
                      const void* gram_hrf_f32, int ld_gram, const void* wsq_f32, void* sumsq_f32, int N, long long P,
                      float l2_lambda, void* stream);
int vlb_head_adamw_fused(const void* dpred, int ld_dpred, const void* hrf, int ld_hrf, int n_samples, long long P, int H,
                         void* w_f32, void* m_f32, void* v_f32, void* w_bf16, const void* sumsq_f32, float lr,
                         float beta1, float beta2, float eps, float weight_decay, int step, float max_norm,
                         float grad_scale, float l2_lambda, void* stream);

/* ------------------------------------------------------------------ optimiser (one flat buffer of trainables)
 * clip_grad_norm_(gradient_clip_val=1) + AdamW (decoupled weight decay on every trainable) with the cosine LR value
 * passed per step; writes the bf16 working copy consumed by the GEMMs in the same sweep. grad_scale = 1/world_size
 * folds the data-parallel mean. (src/litmodule/videollama2_vlb_litmodule.py:345-379; experiment yaml :46) */
int vlb_grad_sumsq(const void* grad_f32, long long n, void* acc_f32, void* stream);
int vlb_adamw_step(void* param_f32, const void* grad_f32, void* m_f32, void* v_f32, void* param_bf16,
                   long long n, const void* sumsq_f32, float lr, float beta1, float beta2, float eps,
                   float weight_decay, int step, float max_norm, float grad_scale, void* stream);
int vlb_f32_to_bf16(const void* src, void* dst, long long n, void* stream);
int vlb_copy2d(const void* src, int lds, void* dst, int ldd, int rows, int cols, int transpose, void* stream);
/* n bf16 transposes dst_i[c, r] = src_i[r, c] in one launch; desc = device int64 [n, 6] rows of (src pointer, dst pointer,
 * lds, ldd, rows, cols). Used once per optimiser step to restage every layer's LoRA B^T [r, N] as the [N, r] column blocks the
 * fused k-block of vlb_gemm_bf16 reads (peft keeps lora_B as nn.Linear weights [N, r]: litmodule.py:113-120). */
int vlb_transpose2d_batch(const void* desc_i64, int n, int blocks_per_item, void* stream);

/* ------------------------------------------------------------------ per-parcel Pearson r (single pass, fp64 moments)
 * stats_f64 [5, P] accumulates (sum x, sum y, sum x^2, sum y^2, sum xy) of nan_to_num'd inputs; the raw sums are
 * additive, so row shards on different GPUs merge with one all-reduce(sum) before finalize.
 * Replaces torchmetrics PearsonCorrCoef in LogValAccuracyCallback (src/utils.py:101-110); vlb_paint_labels is the
 * parcel -> voxel gather of make_acc_brainmaps (src/postprocessing/make_acc_brainmaps.py:50-53). */
int vlb_pearson_update(const void* preds, const void* targets, void* stats_f64, long long N, long long P,
                       int is_bf16, void* stream);
int vlb_pearson_finalize(const void* stats_f64, long long n_rows, void* r_f32, long long P, void* stream);
int vlb_paint_labels(const void* r_f32, const void* labels_i32, void* vol_f32, long long V, int P, int square,
                     void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VLB_B200_H */
