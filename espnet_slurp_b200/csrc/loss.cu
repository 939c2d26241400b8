// Label-smoothing KL loss + accuracy of the attention decoder, forward and gradient in one pass.
//
// Replaces LabelSmoothingLoss.forward (espnet/nets/pytorch_backend/transformer/label_smoothing_loss.py:41-63,
// normalize_length=False) and th_accuracy (espnet/nets/pytorch_backend/nets_utils.py:299-320):
//   true_dist = eps/(V-1) everywhere, 1-eps at the target; loss = sum KL(true || softmax(x)) over rows whose
//   target != ignore, divided by the batch size by the caller; acc = mean(argmax == target) over those rows.
// The reference makes 7 passes over (B*L, V) (clone, fill, scatter, log_softmax, kl_div, masked_fill, sum);
// here each logit row is read once (it stays in L1/L2 for the second sweep) and its gradient
// scale * (softmax - true_dist) is written once: algorithmic bytes 4*R*V read + (2|4)*R*V written.
#include <math.h>

#include "common.cuh"

namespace {

__global__ void __launch_bounds__(256) lsm_kernel(const float* __restrict__ x, long long ldv, int V, const int* __restrict__ tgt,
                                                  int ignore, float eps, float gscale, float* __restrict__ loss_rows,
                                                  int* __restrict__ correct_rows, float* __restrict__ g32, bf16* __restrict__ g16,
                                                  long long ldg) {
  __shared__ float red[32];
  __shared__ int redi[32];
  const long long r = blockIdx.x;
  const int t = tgt[r];
  const int Vp = (int)ldg;
  if (t == ignore || t < 0 || t >= V) {
    if (threadIdx.x == 0) {
      loss_rows[r] = 0.f;
      correct_rows[r] = -1;  // not counted
    }
    for (int v = threadIdx.x; v < Vp; v += blockDim.x) {
      if (g32 != nullptr) g32[r * ldg + v] = 0.f;
      if (g16 != nullptr) g16[r * ldg + v] = __float2bfloat16(0.f);
    }
    return;
  }
  const float* row = x + r * ldv;
  float mx = -INFINITY, sumx = 0.f;
  int arg = 0x7fffffff;
  for (int v = threadIdx.x; v < V; v += blockDim.x) {
    const float xv = row[v];
    sumx += xv;
    if (xv > mx) {
      mx = xv;
      arg = v;
    }
  }
  // block arg-max (lowest index among equal maxima, like torch.argmax on CPU)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, mx, o);
    const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (om > mx || (om == mx && oa < arg)) {
      mx = om;
      arg = oa;
    }
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if (lane == 0) {
    red[w] = mx;
    redi[w] = arg;
  }
  __syncthreads();
  mx = red[0];
  arg = redi[0];
  for (int k = 1; k < nw; ++k)
    if (red[k] > mx || (red[k] == mx && redi[k] < arg)) {
      mx = red[k];
      arg = redi[k];
    }
  __syncthreads();
  sumx = block_sum(sumx, red);
  float se = 0.f;
  for (int v = threadIdx.x; v < V; v += blockDim.x) se += expf(row[v] - mx);
  se = block_sum(se, red);
  const float lse = mx + logf(se);
  const float conf = 1.f - eps, low = eps / (float)(V - 1);
  if (threadIdx.x == 0) {
    const float lp_t = row[t] - lse;
    const float sum_lp = sumx - (float)V * lse;
    float loss = conf > 0.f ? conf * (logf(conf) - lp_t) : 0.f;
    if (low > 0.f) loss += low * ((float)(V - 1) * logf(low) - (sum_lp - lp_t));
    loss_rows[r] = loss;
    correct_rows[r] = arg == t ? 1 : 0;
  }
  for (int v = threadIdx.x; v < Vp; v += blockDim.x) {
    float g = 0.f;
    if (v < V) g = gscale * (expf(row[v] - lse) - (v == t ? conf : low));
    if (g32 != nullptr) g32[r * ldg + v] = g;
    if (g16 != nullptr) g16[r * ldg + v] = __float2bfloat16(g);
  }
}

// stats[0] = sum loss rows, stats[1] = #correct, stats[2] = #counted rows
__global__ void lsm_finish_kernel(const float* __restrict__ loss_rows, const int* __restrict__ correct_rows, long long R,
                                  float* __restrict__ stats) {
  __shared__ float red[32];
  float s = 0.f, c = 0.f, n = 0.f;
  for (long long r = threadIdx.x; r < R; r += blockDim.x) {
    s += loss_rows[r];
    const int k = correct_rows[r];
    if (k >= 0) {
      n += 1.f;
      c += (float)k;
    }
  }
  s = block_sum(s, red);
  c = block_sum(c, red);
  n = block_sum(n, red);
  if (threadIdx.x == 0) {
    stats[0] = s;
    stats[1] = c;
    stats[2] = n;
  }
}

// Posterior over the vocabulary for self-conditioned intermediate CTC (conformer_encoder.py:350-351, ctc.py:99-107):
// p = softmax(logits[:, :V]) written as bf16 (the A operand of the conditioning Linear), padding columns [V, ldp) = 0.
// One warp per row: V = 600 is 19 elements per lane, the row is read twice (max, then exp + sum) from L1/L2.
__global__ void __launch_bounds__(256) vocab_softmax_fwd_kernel(const float* __restrict__ x, long long ldx, long long R, int V,
                                                                bf16* __restrict__ p, long long ldp) {
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= R) return;
  const float* xr = x + row * ldx;
  float m = -INFINITY;
  for (int j = lane; j < V; j += 32) m = fmaxf(m, xr[j]);
  m = warp_max(m);
  float s = 0.f;
  for (int j = lane; j < V; j += 32) s += __expf(xr[j] - m);
  const float inv = 1.f / warp_sum(s);
  bf16* pr = p + row * ldp;
  for (int j = lane; j < ldp; j += 32) pr[j] = __float2bfloat16(j < V ? __expf(xr[j] - m) * inv : 0.f);
}

// y = log_softmax(logits[:, :V]) in fp32 (decoding: transformer_decoder.py:181, ctc.py:109-117).  One warp per row, exact
// logf / expf (the scores feed a beam search that compares hypotheses).
__global__ void __launch_bounds__(256) vocab_log_softmax_kernel(const float* __restrict__ x, long long ldx, long long R, int V,
                                                                float* __restrict__ y, long long ldy) {
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= R) return;
  const float* xr = x + row * ldx;
  float m = -INFINITY;
  for (int j = lane; j < V; j += 32) m = fmaxf(m, xr[j]);
  m = warp_max(m);
  float s = 0.f;
  for (int j = lane; j < V; j += 32) s += expf(xr[j] - m);
  const float lse = m + logf(warp_sum(s));
  float* yr = y + row * ldy;
  for (int j = lane; j < V; j += 32) yr[j] = xr[j] - lse;
}

// dlogits = p * (dp - sum_j p_j dp_j) (+ add): softmax backward over the vocabulary, bf16 in, bf16 out, padding columns 0.
__global__ void __launch_bounds__(256) vocab_softmax_bwd_kernel(const bf16* __restrict__ p, long long ldp,
                                                                const bf16* __restrict__ dp, long long lddp, long long R, int V,
                                                                const bf16* __restrict__ add, long long ldadd,
                                                                bf16* __restrict__ dx, long long lddx) {
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= R) return;
  const bf16 *pr = p + row * ldp, *dr = dp + row * lddp;
  float dot = 0.f;
  for (int j = lane; j < V; j += 32) dot += __bfloat162float(pr[j]) * __bfloat162float(dr[j]);
  dot = warp_sum(dot);
  bf16* o = dx + row * lddx;
  for (int j = lane; j < lddx; j += 32) {
    float v = 0.f;
    if (j < V) {
      v = __bfloat162float(pr[j]) * (__bfloat162float(dr[j]) - dot);
      if (add != nullptr) v += __bfloat162float(add[row * ldadd + j]);
    }
    o[j] = __float2bfloat16(v);
  }
}

}  // namespace

extern "C" int b200slu_vocab_softmax_fwd(const float* logits, long long ldv, long long R, int V, void* probs_bf16, long long ldp,
                                         void* stream) {
  B2_CHECK_ARG(R > 0 && V > 0 && ldv >= V && ldp >= V, "vocab_softmax_fwd: bad shape R=%lld V=%d", R, V);
  cudaStream_t st = (cudaStream_t)stream;
  B2_SCOPE(FAM_LOSS, st, 0, (double)R * V * 6.0);
  vocab_softmax_fwd_kernel<<<(unsigned)((R + 7) / 8), 256, 0, st>>>(logits, ldv, R, V, (bf16*)probs_bf16, ldp);
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_vocab_log_softmax_fwd(const float* logits, long long ldv, long long R, int V, float* out, long long ldo,
                                             void* stream) {
  B2_CHECK_ARG(R > 0 && V > 0 && ldv >= V && ldo >= V && out != nullptr, "vocab_log_softmax_fwd: bad shape R=%lld V=%d", R, V);
  cudaStream_t st = (cudaStream_t)stream;
  B2_SCOPE(FAM_LOSS, st, 0, (double)R * V * 8.0);
  vocab_log_softmax_kernel<<<(unsigned)((R + 7) / 8), 256, 0, st>>>(logits, ldv, R, V, out, ldo);
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_vocab_softmax_bwd(const void* probs_bf16, long long ldp, const void* dprobs_bf16, long long lddp,
                                         long long R, int V, const void* add_bf16, long long ldadd, void* dlogits_bf16,
                                         long long ldd, void* stream) {
  B2_CHECK_ARG(R > 0 && V > 0 && ldp >= V && lddp >= V && ldd >= V && (add_bf16 == nullptr || ldadd >= V),
               "vocab_softmax_bwd: bad shape R=%lld V=%d", R, V);
  cudaStream_t st = (cudaStream_t)stream;
  B2_SCOPE(FAM_LOSS, st, 0, (double)R * V * (6.0 + (add_bf16 ? 2.0 : 0.0)));
  vocab_softmax_bwd_kernel<<<(unsigned)((R + 7) / 8), 256, 0, st>>>((const bf16*)probs_bf16, ldp, (const bf16*)dprobs_bf16, lddp,
                                                                     R, V, (const bf16*)add_bf16, ldadd, (bf16*)dlogits_bf16, ldd);
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" long long b200slu_lsm_workspace_size(long long R) { return R * 8; }

extern "C" int b200slu_lsm_loss_fwd_bwd(const float* logits, long long ldv, long long R, int V, const int* targets, int ignore_id,
                                        float smoothing, float grad_scale, void* workspace, float* stats3, float* grad_f32,
                                        void* grad_bf16, long long ldg, void* stream) {
  B2_CHECK_ARG(R > 0 && R < (1LL << 31) && V > 1 && ldv >= V, "lsm_loss: bad shape R=%lld V=%d", R, V);
  B2_CHECK_ARG((grad_f32 == nullptr && grad_bf16 == nullptr) || ldg >= V, "lsm_loss: ldg < V");
  cudaStream_t st = (cudaStream_t)stream;
  float* loss_rows = (float*)workspace;
  int* correct_rows = (int*)(loss_rows + R);
  B2_SCOPE(FAM_LOSS, st, 0, (double)R * V * (4.0 + (grad_f32 ? 4.0 : 0.0) + (grad_bf16 ? 2.0 : 0.0)));
  lsm_kernel<<<(unsigned)R, 256, 0, st>>>(logits, ldv, V, targets, ignore_id, smoothing, grad_scale, loss_rows, correct_rows,
                                          grad_f32, (bf16*)grad_bf16, (grad_f32 || grad_bf16) ? ldg : 0);
  B2_LAUNCH_CHECK();
  lsm_finish_kernel<<<1, 256, 0, st>>>(loss_rows, correct_rows, R, stats3);
  B2_LAUNCH_CHECK();
  return 0;
}

// ----------------------------------------------------------------------------------------------------------------
// Allocation-free variant of b200slu_lsm_loss_fwd_bwd for callers that cannot allocate (code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The per-row scratch (loss + "correct" flag, 8 bytes per row) lives in a file-scope device array of
// B200SLU_LSM_STATIC_MAX_ROWS rows = 256 KB, held by every GPU that loads the library until the process exits.
//  * In host code `g_lsm_static_scratch` is only a host-side stand-in for the symbol, not a device address: its device
//    address is looked up with cudaGetSymbolAddress once per device (that device current) and handed to the kernels.
//  * Every call on a device shares the array.  Both kernels touch rows [0, R) only and lsm_kernel writes each of those
//    rows before lsm_finish_kernel reads it, so nothing has to start at zero and stale contents of an earlier call are
//    never read.  Calls that overlap on the GPU (different streams, or graphs replayed concurrently) race on it.
//  * Calls no cudaMalloc / cudaFree / allocating library routine, and leaves the (optional, event-creating) profiling
//    scope out.
// ----------------------------------------------------------------------------------------------------------------
#define B200SLU_LSM_STATIC_MAX_ROWS 32768
static __device__ __align__(256) unsigned char g_lsm_static_scratch[(size_t)B200SLU_LSM_STATIC_MAX_ROWS * 8];

namespace {
constexpr int kMaxDevices = 64;
void* g_lsm_static_ptr[kMaxDevices] = {};  // device address of the array, per device ordinal (0 = not looked up yet)
}  // namespace

extern "C" long long b200slu_lsm_static_max_rows(void) { return B200SLU_LSM_STATIC_MAX_ROWS; }

extern "C" int b200slu_lsm_loss_fwd_bwd_static(const float* logits, long long ldv, long long R, int V, const int* targets,
                                               int ignore_id, float smoothing, float grad_scale, float* stats3, float* grad_f32,
                                               void* grad_bf16, long long ldg, void* stream) {
  B2_CHECK_ARG(R > 0 && V > 1 && ldv >= V, "lsm_loss_static: bad shape R=%lld V=%d", R, V);
  B2_CHECK_ARG(R <= B200SLU_LSM_STATIC_MAX_ROWS,
               "lsm_loss_static: R=%lld rows exceed the static scratch of %d rows; use b200slu_lsm_loss_fwd_bwd with a "
               "caller-provided workspace",
               R, B200SLU_LSM_STATIC_MAX_ROWS);
  B2_CHECK_ARG((grad_f32 == nullptr && grad_bf16 == nullptr) || ldg >= V, "lsm_loss_static: ldg < V");
  int dev = -1;
  B2_CUDA(cudaGetDevice(&dev));
  B2_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "lsm_loss_static: device ordinal %d out of range", dev);
  void* ws = __atomic_load_n(&g_lsm_static_ptr[dev], __ATOMIC_ACQUIRE);
  if (ws == nullptr) {  // set-up, once per device, with that device current (racing threads look up the same address)
    B2_CUDA(cudaGetSymbolAddress(&ws, g_lsm_static_scratch));
    B2_CHECK_ARG(ws != nullptr, "lsm_loss_static: cudaGetSymbolAddress returned a null address");
    __atomic_store_n(&g_lsm_static_ptr[dev], ws, __ATOMIC_RELEASE);
  }
  cudaStream_t st = (cudaStream_t)stream;
  float* loss_rows = (float*)ws;
  int* correct_rows = (int*)(loss_rows + R);
  lsm_kernel<<<(unsigned)R, 256, 0, st>>>(logits, ldv, V, targets, ignore_id, smoothing, grad_scale, loss_rows, correct_rows,
                                          grad_f32, (bf16*)grad_bf16, (grad_f32 || grad_bf16) ? ldg : 0);
  B2_LAUNCH_CHECK();
  lsm_finish_kernel<<<1, 256, 0, st>>>(loss_rows, correct_rows, R, stats3);
  B2_LAUNCH_CHECK();
  return 0;
}
