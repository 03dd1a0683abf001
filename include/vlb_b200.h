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
 * Replaces every nn.Linear / peft lora.Linear on the path: HF MistralDecoderLayer + LoraConfig targets
 * (src/litmodule/videollama2_vlb_litmodule.py:101-120), CLIP encoder linears (VideoLLaMA2/videollama2/model/
 * encoder.py:50), STC 1x1 convs / readout (VideoLLaMA2/videollama2/model/projector.py:153-187). */
int vlb_gemm_bf16(const void* A, int lda, const void* B, int ldb, void* C, int ldc,
                  int M, int N, int K,
                  const void* A2, int lda2, const void* B2, int ldb2, int K2,
                  float alpha, const void* bias, const void* residual, int ldr, int act,
                  int a_mn_major, int b_mn_major, int c_is_f32, int accumulate,
                  int block_n, int max_ctas, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VLB_B200_H */
