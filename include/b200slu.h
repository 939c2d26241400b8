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
#define B200SLU_ACT_MUL 3   /* act_grad only: multiply by a saved derivative (see act_grad_out) */

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

/* Device-resident per-step parameters {uint64 dropout seed; float lr, 1-beta1^t, 1-beta2^t, reserved} (24 bytes).
 * When registered, every dropout site XORs the device seed into its seed argument and b200slu_adam_step(step = 0)
 * reads lr / bias corrections from it, so a CUDA graph of the whole step replays with fresh values after a 24-byte
 * copy.  NULL unregisters.  Process-global (one training process per GPU, abs_task.py:1049-1070). */
int b200slu_set_step_params(const void* dev_ptr);

/* Dropout decisions of one site, exported for the parity tests (the reference draws torch's Philox stream, which no
 * other implementation can reproduce; the tests replay THESE decisions in the CPU oracle at every torch.nn.Dropout
 * of encoder_layer.py:97-149, attention.py:91, positionwise_feed_forward.py:32, embedding.py:94,244,
 * decoder_layer.py:88-131).  keep[e] = 1 iff flat element e of site (seed, site) is kept; kept elements are scaled
 * by 1/(1-p).  Element index of a site: row * N + column of the [M,N] tensor the dropout acts on (N = logical width;
 * attention probabilities: ((h*B + b)*Tq + i) * round8(Tk) + j).  Honours the registered device seed like every kernel. */
int b200slu_dropout_mask(float drop_p, unsigned long long seed, unsigned int site, long long n, unsigned char* keep,
                         void* stream);

/* ---- K2: GEMM with fused epilogue ------------------------------------------------------
 * Replaces the nn.Linear / Conv1d(k=1) / im2col-Conv2d GEMMs:
 *   espnet/nets/pytorch_backend/transformer/attention.py:33-36,232 (linear_q/k/v/out/pos)
 *   espnet/nets/pytorch_backend/transformer/positionwise_feed_forward.py:25-26 (w_1, w_2)
 *   espnet/nets/pytorch_backend/conformer/convolution.py:28-53 (pointwise_conv1/2)
 *   espnet/nets/pytorch_backend/transformer/subsampling.py:56-65 (conv.2 as im2col GEMM, out.0)
 *   espnet2/asr/ctc.py:33 (ctc_lo), espnet2/asr/decoder/transformer_decoder.py:85 (output_layer)
 *
 *   y   = acc + bias                      (bias: f32 [N] or NULL)
 *   act_grad_out <- bf16(act'(y) * m)     (optional, ld = ldc; m = dropout multiplier of the element: 0 or 1/(1-p))
 *   y   = act(y); y = dropout(y); y *= alpha
 *   y  *= act'(act_grad_input)            (optional bf16 [M,N]; act_grad = SWISH / RELU: input is the pre-activation
 *                                          (for RELU the post-activation works too); MUL: input is a saved act_grad_out)
 *   y  += residual                        (optional f32 [M,N])
 *   C   = y  (bf16 or f32), or C += y atomically when accumulate (f32 only; enables split-K)
 */
typedef struct {
  const float* bias;
  int act;
  void* act_grad_out;
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
  float* out_colsum_accum; /* optional f32 [N]: += column sums of the (bf16-rounded) output C -- the bias gradient of the
                              Linear whose dY this GEMM produces, taken from the staged tile instead of a pass over C
                              (bf16 output, N % 64 == 0, unbatched) */
} b200slu_epilogue_t;

/* trans=0: C[M,N] = A[M,K] . B[N,K]^T   (A, B row-major bf16, K contiguous)          fwd / dgrad
 * trans=1: C[M,N] = A[K,M]^T . B[K,N]   (A, B row-major bf16, reduction over rows)   wgrad
 * lda/ldb/ldc in elements.  epilogue may be NULL (plain bf16 store). */
int b200slu_gemm_bf16(int trans, int M, int N, int K, const void* A, long long lda, const void* B, long long ldb, void* C,
                      long long ldc, const b200slu_epilogue_t* epilogue, void* stream);

/* General form: per-operand majorness and two batch levels (inner, outer), e.g. (head, utterance)
 * for the attention contractions of attention.py:64-96,281-308 (torch.matmul call sites).
 *   trans_a=0: A[M,K] row-major      trans_a=1: A[K,M] row-major
 *   trans_b=0: B[N,K] row-major      trans_b=1: B[K,N] row-major
 * Batch strides are in elements; 0 broadcasts the operand over that level.  In batched mode
 * only bias / act / alpha / out_f32 / accumulate epilogue fields may be used. */
typedef struct {
  int trans_a, trans_b;
  int M, N, K;
  const void* A;
  long long lda, a_stride_inner, a_stride_outer;
  const void* B;
  long long ldb, b_stride_inner, b_stride_outer;
  void* C;
  long long ldc, c_stride_inner, c_stride_outer;
  int batch_inner, batch_outer;
} b200slu_gemm_t;
int b200slu_gemm_bf16_ex(const b200slu_gemm_t* g, const b200slu_epilogue_t* epilogue, void* stream);

/* ---- K5: LayerNorm ---------------------------------------------------------------------------
 * espnet/nets/pytorch_backend/transformer/layer_norm.py:12-38 at conformer/encoder_layer.py:97-149,
 * espnet2/asr/encoder/conformer_encoder.py:362-363, transformer/decoder_layer.py:63-134.
 * x f32 [M,D]; outputs bf16 and/or f32; stats f32 [M,2] = (mean, rstd). */
int b200slu_layernorm_fwd(const float* x, const float* gamma, const float* beta, int M, int D, float eps, void* y_bf16,
                          float* y_f32, float* stats, void* stream);
/* dx = LN'(dy) (+ dx_add); dgamma/dbeta accumulated; optional g_bf16 = bf16(g_alpha * dropout(dx)),
 * the branch gradient the next dgrad GEMM consumes (encoder_layer.py:101,129,138,144), and with it optionally
 * g_colsum_accum[D] += column sums of g_alpha * dropout(dx): the bias gradient of that branch's last Linear. */
int b200slu_layernorm_bwd(const void* dy, int dy_is_f32, const float* x, const float* gamma, const float* stats,
                          const float* dx_add, int M, int D, float* dx, float* dgamma_accum, float* dbeta_accum, void* g_bf16,
                          float g_alpha, float* g_colsum_accum, float drop_p, unsigned long long seed, unsigned int site,
                          void* stream);

/* ---- small helpers around the GEMMs ---------------------------------------------------------- */
/* y = bf16(alpha * dropout(x)) (weight casts; embedding.py:244 dropout on pos_emb) */
int b200slu_cast_bf16(const float* x, void* y, long long n, float alpha, float drop_p, unsigned long long seed,
                      unsigned int site, void* stream);
/* y = bf16(alpha * dropout(x)) for x f32 [M,N] (N % 8 == 0) and colsum_accum[n] += sum_m of the same values: the backward
   of "x + dropout(branch(x))" (encoder_layer.py:118-175) produces the branch gradient and the bias gradient of the
   branch's last Linear in one pass */
int b200slu_cast_colsum_bf16(const float* x, void* y, int M, int N, float alpha, float drop_p, unsigned long long seed,
                             unsigned int site, float* colsum_accum, void* stream);
/* out[n] += sum_m dy[m,n]: bias gradients of every nn.Linear / Conv1d(k=1) */
int b200slu_colsum_bf16(const void* dy, int M, int N, long long ld, float* out_accum, void* stream);
/* q_u = q + pos_bias_u, q_v = q + pos_bias_v (attention.py:289-291); q = first d columns of rows of stride ldq */
int b200slu_pos_bias_fwd(const void* q, long long ldq, const float* pos_bias_u, const float* pos_bias_v, int M, int d,
                         void* q_u, void* q_v, void* stream);
int b200slu_pos_bias_bwd(const void* dq_u, const void* dq_v, int M, int d, void* dq, long long lddq, float* du_accum,
                         float* dv_accum, void* stream);
/* the same, also accumulating colsum(dq) = colsum(dq_u) + colsum(dq_v) (the q part of linear_q's bias gradient) */
int b200slu_pos_bias_bwd_q(const void* dq_u, const void* dq_v, int M, int d, void* dq, long long lddq, float* du_accum,
                           float* dv_accum, float* dq_colsum_accum, void* stream);
/* decoder input: x = dropout(emb[tok] * scale + pe[pos]) (transformer_decoder.py:72-77, embedding.py:82-94) */
int b200slu_embed_fwd(const float* emb, const int* tokens, const float* pe, int R, int L, int d, float scale, float* x,
                      float drop_p, unsigned long long seed, unsigned int site, void* stream);
int b200slu_embed_bwd(const float* dx, const int* tokens, int R, int d, float scale, float* demb_accum, float drop_p,
                      unsigned long long seed, unsigned int site, void* stream);

/* ---- K1/K10: masked softmax with rel_shift folded into the indexing ---------------------------
 * attention.py:64-96 (forward_attention), :240-263 / :145-165 (rel_shift latest / legacy), :297-306.
 * Score tensors are head-major [H, B, Tq, ld]; ac f32 = (q+u)k^T, bd f32 = un-shifted (q+v)p^T.
 * p = probabilities (bf16), p_drop = dropout(p) (must alias p iff drop_p == 0). */
int b200slu_softmax_fwd(const float* ac, long long ld_ac, const float* bd, long long ld_bd, const int* key_lens, int H, int B,
                        int Tq, int Tk, int relpos_mode, int causal, float scale, void* p, void* p_drop, long long ld_p,
                        float drop_p, unsigned long long seed, unsigned int site, void* stream);
/* d_ac (bf16 [H,B,Tq,ld_p]) and, for rel-pos modes, d_bd (bf16, un-shifted layout [H,B,T,ld_bd]) */
int b200slu_softmax_bwd(const float* dp_drop, const void* p, long long ld_p, int H, int B, int Tq, int Tk, int relpos_mode,
                        float scale, void* d_ac, void* d_bd, long long ld_bd, float drop_p, unsigned long long seed,
                        unsigned int site, void* stream);
/* ---- K1/K10 fused: scores -> rel_shift -> masked softmax -> dropout -> context in ONE kernel (tcgen05 / TMEM / TMA) ----
 * RelPositionMultiHeadedAttention.forward attention.py:281-308 (relpos_mode LATEST: q_u = q + pos_bias_u, q_v = q + pos_bias_v,
 * pos = linear_pos(pos_emb) [2T-1, H*64]) and MultiHeadedAttention.forward :98-114 (relpos_mode NONE: q_u = q; q_v, pos
 * unused; causal = decoder self-attention).  Rows of q_u/q_v (stride ldq) and k/v (stride ldkv) are (utterance, frame),
 * heads are 64-column slices.  Outputs: p / p_drop bf16 head-major [H,B,Tq,ld_p] (saved for the backward; p_drop must
 * alias p iff drop_p == 0; the dropout decisions are those of b200slu_softmax_fwd / _bwd) and the merged-head context
 * bf16 [B*Tq, H*64] (row stride ldctx).  b200slu_relattn_supported tells whether a shape runs here (d_k = 64, keys <= 192,
 * LATEST with Tq == Tk, or NONE); other shapes use the unfused GEMM + softmax entry points above. */
int b200slu_relattn_supported(int Tq, int Tk, int dk, int relpos_mode);
int b200slu_relattn_fwd(const void* q_u, const void* q_v, long long ldq, const void* k, const void* v, long long ldkv,
                        const void* pos, long long ldpos, const int* key_lens, int B, int H, int Tq, int Tk, int relpos_mode,
                        int causal, float scale, void* p, void* p_drop, long long ld_p, void* ctx, long long ldctx, float drop_p,
                        unsigned long long seed, unsigned int site, void* stream);
/* Backward of the same, keys <= 192 (b200slu_relattn_bwd_supported): autograd through attention.py:78-95, the matmuls of
 * :297-302 / :111-113 and rel_shift (:240-263) in ONE kernel -- dP = dO . v^T, dS = p (m dP - <dO, O>) scale, dropout(p)
 * and d(matrix_bd) (dS skewed back into position-window coordinates) stay on the SM (tensor memory / shared memory);
 * dq_u = dS . k, dk = dS^T . q_u, dv = dropout(p)^T . dO and, in rel-pos mode, dq_v = d(bd) . pos are written as bf16
 * rows (utterance, frame) x head slices like their forward counterparts; dpos_accum (fp32 [2T-1, lddpos], gradient of the
 * projected positions, summed over utterances) is ACCUMULATED.  d_ctx = gradient of the merged-head context, p = the
 * probabilities the forward saved, ctx = the forward's context output.  d_ac (optional, bf16 [H,B,Tq,ld_p]) receives dS.
 * q_v / pos / dq_v / dpos_accum: rel-pos mode only (NULL otherwise); dq_v shares dq_u's row stride. */
int b200slu_relattn_bwd_supported(int Tq, int Tk, int dk, int relpos_mode);
int b200slu_relattn_bwd(const void* d_ctx, long long lddo, const void* q_u, long long ldq, const void* k, const void* v,
                        long long ldkv, const void* p, long long ld_p, const void* ctx, long long ldctx, int B, int H, int Tq,
                        int Tk, int relpos_mode, float scale, void* dq_u, long long lddq, void* dk, void* dv, long long lddkv,
                        void* d_ac, float drop_p, unsigned long long seed, unsigned int site, const void* q_v, const void* pos,
                        long long ldpos, void* dq_v, float* dpos_accum, long long lddpos, void* stream);
/* Any number of keys: the key axis is cut into chunks of <= 192 keys (multiples of 16), one launch of the kernels above per
 * chunk, and a combine pass turns the per-chunk softmaxes into the softmax over all keys (attention.py:78-95 over the
 * whole row).  Chunk launch: k, v, p (and dk, dv) point at key / column key_off of tensors with kv_rows keys per utterance;
 * key_off also shifts the rel-pos window (rel_shift source column = T - 1 - i + key_off + j), the valid length and the
 * dropout counter, so every element sees the decisions of the single-launch form.  Forward: p's chunk columns receive the
 * softmax over the chunk, ctx_chunk the chunk's context, lse2 (fp32 [H*B*Tq]) log2 of the chunk's partition sum (-inf when
 * no key of the chunk is valid).  b200slu_relattn_combine: alpha_c = 2^(lse2_c - log2 sum_c 2^lse2_c); p[.., chunk c] *=
 * alpha_c in place and ctx = sum_c alpha_c ctx_c (ctx_chunks [nchunk] slabs chunk_stride elements apart, lse2 [nchunk,
 * H*B*Tq]).  Backward (b200slu_relattn_bwd_chunks: ONE launch over the keys [key_begin, key_end), work unit = (chunk, utterance,
 * head), chunk_keys and key_begin multiples of 64 -- a short last chunk gets its own launch with a smaller chunk_keys so
 * that it does not pay for 192 keys; k, v, p, dk, dv are the whole tensors, the slab pointers those of the launch's first chunk): p = the combined probabilities, ctx = the combined context (for delta =
 * <dO, O>); dk, dv are final, dq_u / dq_v come out as one slab per chunk, dq_slab elements apart (sum them with
 * b200slu_sum_chunks_bf16); dpos_accum accumulates over the chunks like over the utterances.  dk_colsum_accum /
 * dv_colsum_accum (optional, fp32 [H*64], accumulated): the column sums of dk / dv over all rows, i.e. the bias gradients of
 * the K and V projections, taken from the fp32 accumulators on their way out (also valid with ONE chunk of Tk <= 192 keys:
 * chunk_keys = Tk, key range [0, Tk), dq_slab = 0). */
int b200slu_relattn_chunk_supported(int Tq, int Tk_chunk, int dk, int relpos_mode);
int b200slu_relattn_fwd_chunk(const void* qu, const void* qv, long long ldq, const void* k, const void* v, long long ldkv,
                              const void* pos, long long ldpos, const int* key_lens, int B, int H, int Tq, int Tk_chunk,
                              int relpos_mode, int causal, float scale, void* p, long long ld_p, void* ctx_chunk, long long ldctx,
                              float drop_p, unsigned long long seed, unsigned int site, int key_off, int kv_rows, float* lse2,
                              void* stream);
int b200slu_relattn_combine(void* p, long long ld_p, int Tk, int chunk_keys, int nchunk, const float* lse2, const void* ctx_chunks,
                            long long chunk_stride, long long ldctx, void* ctx, long long ldout, int B, int H, int Tq, void* stream);
int b200slu_relattn_bwd_chunks(const void* d_ctx, long long lddo, const void* q_u, long long ldq, const void* k, const void* v,
                               long long ldkv, const void* p, long long ld_p, const void* ctx, long long ldctx, int B, int H, int Tq,
                               int Tk, int chunk_keys, int relpos_mode, float scale, void* dq_u_slabs, long long lddq,
                               long long dq_slab, void* dk, void* dv, long long lddkv, float drop_p, unsigned long long seed,
                               unsigned int site, const void* q_v, const void* pos, long long ldpos, void* dq_v_slabs,
                               float* dpos_accum, long long lddpos, int key_begin, int key_end, float* dk_colsum_accum,
                               float* dv_colsum_accum, void* stream);
int b200slu_sum_chunks_bf16(const void* src, long long chunk_stride, long long ld_src, int nchunk, long long rows, int cols,
                            void* dst, long long ld_dst, void* stream);
/* d(matrix_bd) [H,B,T,ld_bd] (un-shifted) from d(scores) [H,B,T,ld_p] (shifted): the scatter of rel_shift's backward
 * (attention.py:240-263), latest mode. */
int b200slu_rel_unshift(const void* d_scores, long long ld_p, int H, int B, int T, int relpos_mode, void* d_bd, long long ld_bd,
                        void* stream);
/* rel_shift source index (bit-exact index test): returns row * 65536 + col + 1 of the un-shifted matrix that lands at
 * out[i, j], or 0 for the literal zero of the legacy variant. */
long long b200slu_rel_shift_index(int i, int j, int T, int mode);

/* ---- K4: ConvolutionModule middle (convolution.py:67-79), channels-last bf16 ------------------ */
int b200slu_glu_dwconv_fwd(const void* pw1, const float* w, const float* bias, int B, int T, int C, int K, void* out,
                           float* bn_sums_zeroed /* [2C] */, void* stream);
int b200slu_glu_dwconv_bwd(const void* d_out, const void* pw1, const float* w, int B, int T, int C, int K, void* d_pw1,
                           float* dw_accum, float* dbias_accum, void* stream);
/* the same, also accumulating the column sums of d_pw1 (fp32 [2C]) = the bias gradient of pointwise_conv1 */
int b200slu_glu_dwconv_bwd_b(const void* d_out, const void* pw1, const float* w, int B, int T, int C, int K, void* d_pw1,
                             float* dw_accum, float* dbias_accum, float* dpw1_colsum_accum, void* stream);
int b200slu_bn_finalize(const float* sums, int C, long long n, float eps, float momentum, float* mean_rstd /* [2C] */,
                        float* running_mean, float* running_var, long long* num_batches_tracked, void* stream);
int b200slu_bn_swish_fwd(const void* x, const float* mean_rstd, const float* gamma, const float* beta, long long M, int C,
                         void* y, void* stream);
int b200slu_bn_swish_bwd(const void* dy, const void* x, const float* mean_rstd, const float* gamma, const float* beta,
                         long long M, int C, float* sums2_zeroed /* [2C] */, void* dx, float* dgamma_accum, float* dbeta_accum,
                         void* stream);

/* ---- K3: Conv2dSubsampling (subsampling.py:53-87) ---------------------------------------------- */
int b200slu_conv1_fwd(const float* x, const float* w, const float* bias, int B, int Tin, int F, int C, void* h1, void* stream);
/* conv1 weight / bias gradient as a tensor-core reduction over the B*T1*F1 pixels; the 3x3 input patches (bf16) and the
 * [C,16] fp32 accumulator live in a caller-provided workspace (256-byte aligned). */
long long b200slu_conv1_bwd_workspace_size(int B, int Tin, int F, int C);
int b200slu_conv1_bwd(const void* dh1, const float* x, int B, int Tin, int F, int C, float* dw_accum, float* db_accum,
                      void* workspace, long long workspace_bytes, void* stream);
int b200slu_im2col3x3s2(const void* h1, int B, int T1, int F1, int C, void* col, void* stream);
int b200slu_col2im3x3s2_relu(const void* dcol, const void* h1, int B, int T1, int F1, int C, void* dh1, void* stream);
/* The same convolution (subsampling.py:56-65, conv.2: 3x3, stride 2, C -> C) as implicit GEMMs: the [pixels, 9 C] matrix is
 * never written; the tensor-core GEMM's operand loads walk h1 (or dh2) through TMA im2col maps.  h1 bf16 [B,T1,F1,C]
 * channels-last; w bf16 [C, 9 C] with k = (kt*3+kf)*C + ci; h2 / dh2 bf16 [B*T2*F2, C], T2 = (T1-3)/2+1, F2 = (F1-3)/2+1.
 *   fwd:    h2 = relu(conv(h1) + bias)
 *   wgrad:  dw_accum (fp32 [C, 9 C]) += dh2^T * im2col(h1)
 *   dgrad:  dh1 = relu_bits * conv_transpose(dh2, w), one gather GEMM per output parity class (4 / 2 / 2 / 1 taps)
 * Needs C % 256 == 0 (b200slu_conv2_implicit_supported); callers keep the im2col + GEMM path otherwise. */
int b200slu_conv2_implicit_supported(int B, int T1, int F1, int C);
int b200slu_conv2_fwd(const void* h1, int B, int T1, int F1, int C, const void* w, const float* bias, void* h2, void* stream);
int b200slu_conv2_wgrad(const void* dh2, const void* h1, int B, int T1, int F1, int C, float* dw_accum, void* stream);
int b200slu_conv2_dgrad_relu(const void* dh2, const void* w, const unsigned int* relu_bits, int B, int T1, int F1, int C,
                             void* dh1, void* stream);
/* relu_bits: u32 [B*T1*F1, C/32], bit (c % 32) of word c / 32 of a pixel's row = (h1[pixel][c] > 0) -- all dgrad needs of
 * h1, at 1/16 of its bytes.  b200slu_conv1_fwd_bits writes it next to h1 (C % 256 == 0); b200slu_relu_bits derives it from
 * an h1 the caller already holds. */
int b200slu_conv1_fwd_bits(const float* x, const float* w, const float* bias, int B, int Tin, int F, int C, void* h1,
                           unsigned int* relu_bits, void* stream);
int b200slu_relu_bits(const void* h1, long long pixels, int C, unsigned int* relu_bits, void* stream);

/* ---- K6: CTC (espnet2/asr/ctc.py:52-63,82-95; lattice espnet/nets/pytorch_backend/gtn_ctc.py:26-36) -----
 * logits f32 [B,T,ldv] (log_softmax fused); labels int32 [B,Lmax]; all lengths int32 on the device. */
long long b200slu_ctc_workspace_size(int B, int T, int Lmax);
/* The blank-extended lattice exactly as the device code builds it (bit-exact test against gtn_ctc.py:26-36):
 * ext_out int32 [B, 2*Lmax+1] = label of state s (-1 past the utterance's 2L+1 states), skip_out u8 = 1 iff the arc
 * s-2 -> s exists. */
int b200slu_ctc_lattice_states(const int* labels, int Lmax, const int* label_lens, int B, int blank, int* ext_out,
                               unsigned char* skip_out, void* stream);
int b200slu_ctc_fwd(const float* logits, long long ldv, int B, int T, int V, const int* hlens, const int* labels, int Lmax,
                    const int* label_lens, int blank, void* workspace, float* loss_per_utt, float* loss_sum, void* stream);
int b200slu_ctc_bwd(const float* logits, long long ldv, int B, int T, int V, const int* hlens, const int* labels, int Lmax,
                    const int* label_lens, int blank, const void* workspace, float scale, float* grad_f32, void* grad_bf16,
                    long long ldg, void* stream);

/* ---- K9: label-smoothing loss + accuracy (label_smoothing_loss.py:41-63, nets_utils.py:299-320) -------
 * stats3 = {sum of row losses, #correct, #counted rows}; grad = grad_scale * (softmax - true_dist). */
long long b200slu_lsm_workspace_size(long long R);
int b200slu_lsm_loss_fwd_bwd(const float* logits, long long ldv, long long R, int V, const int* targets, int ignore_id,
                             float smoothing, float grad_scale, void* workspace, float* stats3, float* grad_f32, void* grad_bf16,
                             long long ldg, void* stream);

/* Allocation-free variant of b200slu_lsm_loss_fwd_bwd.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * Same arguments minus the workspace: the per-row scratch is a static device array of b200slu_lsm_static_max_rows()
 * (32768) rows = 256 KB per GPU; more rows -> B200SLU_EINVAL with a message.  Allocates nothing.  All calls on a device
 * share the array: calls that overlap on the GPU (different streams) race.  The first call on a device looks the
 * array's address up (cudaGetSymbolAddress): make that call outside a graph capture. */
long long b200slu_lsm_static_max_rows(void);
int b200slu_lsm_loss_fwd_bwd_static(const float* logits, long long ldv, long long R, int V, const int* targets, int ignore_id,
                                    float smoothing, float grad_scale, float* stats3, float* grad_f32, void* grad_bf16,
                                    long long ldg, void* stream);

/* Self-conditioned intermediate CTC (espnet2/asr/encoder/conformer_encoder.py:350-358, espnet2/asr/ctc.py:99-107):
 * probs = softmax(logits[:, :V]) as bf16 [R, ldp] (columns >= V zero), and its backward
 * dlogits = probs * (dprobs - <probs, dprobs>) (+ add), bf16 [R, ldd]. */
int b200slu_vocab_softmax_fwd(const float* logits, long long ldv, long long R, int V, void* probs_bf16, long long ldp,
                              void* stream);
/* N2 (decoding): out = log_softmax(logits[:, :V]) in fp32 -- the scores of TransformerDecoder.forward_one_step
 * (espnet2/asr/decoder/transformer_decoder.py:181) and CTC.log_softmax (espnet2/asr/ctc.py:109-117). */
int b200slu_vocab_log_softmax_fwd(const float* logits, long long ldv, long long R, int V, float* out, long long ldo, void* stream);
int b200slu_vocab_softmax_bwd(const void* probs_bf16, long long ldp, const void* dprobs_bf16, long long lddp, long long R, int V,
                              const void* add_bf16, long long ldadd, void* dlogits_bf16, long long ldd, void* stream);

/* ---- K7/K8: front-end (default.py:82-117, stft.py:83-170, log_mel.py:62-79; specaug.py:89-96,
 * time_warp.py:9-46, mask_along_axis.py:8-68, utterance_mvn.py:45-88) ------------------------------ */
int b200slu_logmel_fwd(const float* wav, long long ldw, int B, int N, const int* lens, const float* melmat, int n_fft, int hop,
                       int n_mels, float* feats, int* feat_lens, void* stream);
/* masks are int32 [B, n, 2] = (pos, len); warp_center < 0 (or NULL) = no time warp for that utterance */
int b200slu_specaug_mvn_fwd(const float* x, int B, int T, int F, const int* lens, const int* warp_center, const int* warp_warped,
                            const int* freq_masks, int n_freq_masks, const int* time_masks, int n_time_masks, int do_mvn,
                            float* y, void* stream);

/* N4 (next row): GlobalMVN, espnet2/layers/global_mvn.py:66-94 -- y = (x - mean[f]) / std[f] on the valid frames, 0 on the
 * padding; mean / inv_std = NULL skip that half (norm_means / norm_vars = False).  F % 4 == 0. */
int b200slu_global_mvn_fwd(const float* x, int B, int T, int F, const int* lens, const float* mean, const float* inv_std,
                           float* y, void* stream);

/* ---- N1: fused optimizer step over the flat buffers (espnet2/train/trainer.py:640-690: clip_grad_norm_ 2-norm,
 * non-finite skip, torch.optim.Adam; lr from WarmupLR espnet2/schedulers/warmup_lr.py:43-50) ----------------
 * state3 (device, 3 doubles): [0] sum of squares, [1] gradient norm, [2] 1 if the update was applied. */
int b200slu_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n, float lr, float beta1,
                      float beta2, float eps, float weight_decay, int step, float grad_inv_scale, float max_norm,
                      double* state3, void* stream);

/* The same with the schedule on the device: state4[3] counts the APPLIED updates (persistent, zeroed once by the caller);
 * lr_t = base_lr * warmup^0.5 * min(t^-0.5, t * warmup^-1.5) (WarmupLR) and Adam's bias corrections use t = state4[3] + 1.
 * A non-finite gradient norm skips the update and leaves t unchanged, as trainer.py:651-690 skips optimizer.step() and
 * scheduler.step() together. */
int b200slu_adam_step_sched(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n, float base_lr,
                            float warmup_steps, float beta1, float beta2, float eps, float weight_decay, float grad_inv_scale,
                            float max_norm, double* state4, void* stream);

/* The same, also writing the bf16 copy of the updated parameters (the GEMM operands of the next step; the reference's
 * autocast re-casts every weight per use, torch/amp autocast_mode) in the same pass. */
int b200slu_adam_step_sched_lp(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n, float base_lr,
                               float warmup_steps, float beta1, float beta2, float eps, float weight_decay, float grad_inv_scale,
                               float max_norm, double* state4, void* params_bf16, void* stream);

/* Local reduction of the peer-memory gradient exchange (8e; reference: DistributedDataParallel's all-reduce,
 * espnet2/train/trainer.py:221-236): dst[i] += sum_k src[k * shard_stride + i], k = 0 .. nshards-1 in that order, fp32.
 * The shards are the other ranks' copies of this rank's slice of a gradient bucket, pulled over NVLink by the copy
 * engines (espnet_slurp_b200/dist.py PeerReducer). */
int b200slu_add_shards_f32(float* dst, const float* src, long long shard_stride, int nshards, long long n, void* stream);

/* Number of SMs the persistent kernels (GEMM, attention, LayerNorm backward) launch their grids for on the current device;
 * 0 = all (default).  Grids are fixed at launch / graph-capture time.  Used by the data-parallel trainer to leave a few SMs
 * to NCCL's kernels while the gradient all-reduce overlaps the backward. */
int b200slu_set_sm_budget(int sms);

#ifdef __cplusplus
}
#endif
#endif /* B200SLU_H_ */
