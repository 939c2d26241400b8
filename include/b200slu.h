/*
 * libb200slu -- C ABI of the B200-native (sm_100a) Conformer CTC/attention train-step kernels.
 *
 * Drop-in boundary for the hot path of BriansIDP/espnet_slurp.  The
 * reference has no FFI of its own (every op is a stock PyTorch call), so each entry point
 * below names the reference Python call site it replaces (paths relative to the reference
 * root).  Conventions:
 *   - plain pointers and sizes only; all tensor pointers are DEVICE pointers unless a
 *     parameter says "host"; nothing here allocates, frees or synchronises;
 *   - `stream` is a cudaStream_t passed as void*;
 *   - return 0 on success, a negative B200SLU_E* code otherwise; the message is available
 *     from b200slu_last_error() (thread-local).  Never throws, never aborts.
 *   - "bf16" = __nv_bfloat16 storage, "f32" = float.
 */
#ifndef B200SLU_H_
#define B200SLU_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200SLU_VERSION 100 /* 0.1.0 */

#define B200SLU_OK 0
#define B200SLU_EINVAL (-1)       /* bad shape / alignment / argument */
#define B200SLU_EUNSUPPORTED (-2) /* configuration outside the hot-path scope */
#define B200SLU_ECUDA (-3)        /* CUDA runtime / driver error */

#define B200SLU_ACT_NONE 0
#define B200SLU_ACT_SWISH 1 /* espnet/nets/pytorch_backend/conformer/swish.py:13-18 */
#define B200SLU_ACT_RELU 2  /* decoder FFN default, positionwise_feed_forward.py:22 */

#define B200SLU_RELPOS_NONE 0   /* MultiHeadedAttention, attention.py:98-114 */
#define B200SLU_RELPOS_LATEST 1 /* RelPositionMultiHeadedAttention.rel_shift, attention.py:240-263 */
#define B200SLU_RELPOS_LEGACY 2 /* LegacyRelPositionMultiHeadedAttention.rel_shift, attention.py:145-165 */

/* ---- runtime ----------------------------------------------------------------------- */
int b200slu_version(void);
const char* b200slu_last_error(void);
/* Per-kernel-family accounting used by bench.py for the roofline line.  When timing is
 * enabled every launch is bracketed by CUDA events on its own stream (adds host syncs at
 * read-out; never enable inside a throughput measurement). */
void b200slu_prof_reset(void);
void b200slu_prof_enable_timing(int on);
int b200slu_prof_num_families(void);
const char* b200slu_prof_family_name(int fam);
/* out[0]=launches out[1]=ms (timing on) out[2]=algorithmic flops out[3]=algorithmic bytes */
int b200slu_prof_get(int fam, double* out4);
/* total kernel launches issued by this library since the last reset */
long long b200slu_prof_total_launches(void);

/* ---- K2: GEMM with fused epilogue ------------------------------------------------------
 * Replaces the nn.Linear / Conv1d(k=1) / im2col-Conv2d GEMMs:
 *   espnet/nets/pytorch_backend/transformer/attention.py:33-36,232 (linear_q/k/v/out/pos)
 *   espnet/nets/pytorch_backend/transformer/positionwise_feed_forward.py:25-26 (w_1, w_2)
 *   espnet/nets/pytorch_backend/conformer/convolution.py:28-53 (pointwise_conv1/2)
 *   espnet/nets/pytorch_backend/transformer/subsampling.py:56-65 (conv.2 as im2col GEMM, out.0)
 *   espnet2/asr/ctc.py:33 (ctc_lo), espnet2/asr/decoder/transformer_decoder.py:85 (output_layer)
 *
 *   y   = acc + bias                      (bias: f32 [N] or NULL)
 *   pre_act_out <- bf16(y)                (optional, ld = ldc)
 *   y   = act(y); y = dropout(y); y *= alpha
 *   y  *= act'(act_grad_input)            (optional bf16 [M,N]; act_grad = SWISH / RELU)
 *   y  += residual                        (optional f32 [M,N])
 *   C   = y  (bf16 or f32), or C += y atomically when accumulate (f32 only; enables split-K)
 */
typedef struct {
  const float* bias;
  int act;
  void* pre_act_out;
  float alpha;
  float dropout_p;
  unsigned long long dropout_seed;
  unsigned int dropout_site;
  const void* act_grad_input;
  long long ld_act_grad_input;
  int act_grad;
  const float* residual;
  long long ld_residual;
  int out_f32;
  int accumulate;
} b200slu_epilogue_t;

/* trans=0: C[M,N] = A[M,K] . B[N,K]^T   (A, B row-major bf16, K contiguous)          fwd / dgrad
 * trans=1: C[M,N] = A[K,M]^T . B[K,N]   (A, B row-major bf16, reduction over rows)   wgrad
 * lda/ldb/ldc in elements.  epilogue may be NULL (plain bf16 store). */
int b200slu_gemm_bf16(int trans, int M, int N, int K, const void* A, long long lda, const void* B, long long ldb, void* C,
                      long long ldc, const b200slu_epilogue_t* epilogue, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200SLU_H_ */
