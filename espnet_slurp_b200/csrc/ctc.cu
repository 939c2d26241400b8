// CTC loss, forward and gradient, in log space over the blank-extended label lattice.
//
// Replaces espnet2/asr/ctc.py:52-63,82-95 = log_softmax + torch.nn.CTCLoss(reduction="none",
// zero_infinity=True) (ATen LossCTC) on the logits of ctc_lo.  Lattice as spelled out in
// espnet/nets/pytorch_backend/gtn_ctc.py:26-36: S = 2L+1 states, even = blank, odd = label[(s-1)/2],
// arcs s->s, s-1->s and s-2->s iff s odd, s > 1 and label != previous label; start {0,1}, final
// {S-1,S-2}.
//
// Three kernels, logits read exactly twice and the gradient written once (12*T*B*V bytes fwd+bwd):
//  1. prep   (block per frame)      row logsumexp of the logits (fused log_softmax) and a compact
//                                   gather em[b,t,s] = logit[ext_s] - lse of the S lattice emissions;
//  2. lattice (2 warps / utterance) warp 0 runs the alpha recursion forward in time, warp 1 the beta
//                                   recursion backward, concurrently; each lane owns a contiguous chunk
//                                   of states in registers, neighbours come from warp shuffles, the
//                                   3-way logsumexp is shuffle-free inside a chunk; rows are renormalised
//                                   by their max every step (offsets accumulated in double), so stored
//                                   alpha~/beta~ stay O(1) and fp32 keeps its resolution at any T;
//  3. grad   (block per frame)      occupancy from alpha~ + beta~ - em, self-normalised per frame (no
//                                   dependence on the total likelihood), scattered into a shared-memory
//                                   row and fused with softmax: grad = scale * (softmax - occ).
#include <math.h>

#include "common.cuh"

namespace {

constexpr float NEG_INF = -INFINITY;

__device__ __forceinline__ float lse2(float a, float b) {
  const float m = fmaxf(a, b);
  if (m == NEG_INF) return NEG_INF;
  return m + logf(expf(a - m) + expf(b - m));
}
__device__ __forceinline__ float lse3(float a, float b, float c) {
  const float m = fmaxf(a, fmaxf(b, c));
  if (m == NEG_INF) return NEG_INF;
  return m + logf(expf(a - m) + expf(b - m) + expf(c - m));
}

__device__ __forceinline__ int ext_label(const int* __restrict__ lab, int s, int blank) { return (s & 1) ? lab[s >> 1] : blank; }

// ---- 1. prep: one block per (b, t) ------------------------------------------------------------
__global__ void __launch_bounds__(256) ctc_prep_kernel(const float* __restrict__ logits, long long ldv, int T, int V,
                                                       const int* __restrict__ hlens, const int* __restrict__ labels, int Lmax,
                                                       const int* __restrict__ llens, int blank, float* __restrict__ lse_out,
                                                       float* __restrict__ em, int Smax) {
  __shared__ float red[32];
  const int t = blockIdx.x, b = blockIdx.y;
  if (t >= hlens[b]) return;
  const float* row = logits + ((long long)b * T + t) * ldv;
  float mx = NEG_INF;
  for (int v = threadIdx.x; v < V; v += blockDim.x) mx = fmaxf(mx, row[v]);
  mx = warp_max(mx);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
  __syncthreads();
  mx = red[0];
  for (int w = 1; w < (blockDim.x >> 5); ++w) mx = fmaxf(mx, red[w]);
  float sum = 0.f;
  for (int v = threadIdx.x; v < V; v += blockDim.x) sum += expf(row[v] - mx);
  sum = block_sum(sum, red);
  const float lse = mx + logf(sum);
  if (threadIdx.x == 0) lse_out[(long long)b * T + t] = lse;
  const int S = 2 * llens[b] + 1;
  const int* lab = labels + (long long)b * Lmax;
  float* e = em + ((long long)b * T + t) * Smax;
  for (int s = threadIdx.x; s < S; s += blockDim.x) e[s] = row[ext_label(lab, s, blank)] - lse;
}

// ---- 2. lattice: block = 2 warps per utterance -------------------------------------------------
template <int C>  // states per lane; S <= 32*C
__global__ void __launch_bounds__(64) ctc_lattice_kernel(const float* __restrict__ em, int T, int Smax,
                                                         const int* __restrict__ hlens, const int* __restrict__ labels, int Lmax,
                                                         const int* __restrict__ llens, int blank, float* __restrict__ alpha,
                                                         float* __restrict__ beta, float* __restrict__ nll) {
  const int b = blockIdx.x, lane = threadIdx.x & 31, dir = threadIdx.x >> 5;  // dir 0 = alpha, 1 = beta
  const int Tb = hlens[b], L = llens[b], S = 2 * L + 1;
  const int* lab = labels + (long long)b * Lmax;
  const float* eb = em + (long long)b * T * Smax;
  float* out = (dir == 0 ? alpha : beta) + (long long)b * T * Smax;
  if (Tb <= 0 || Tb > T || L > Lmax || S > 32 * C) {
    if (dir == 0 && lane == 0) nll[b] = INFINITY;  // infeasible -> zeroed by the caller-visible flag path
    return;
  }
  const int s0 = lane * C;
  // skip-arc permission of the state a transition ENTERS: alpha uses skip[s] (s-2 -> s); beta uses skip[s+2]
  bool skip[C];
#pragma unroll
  for (int i = 0; i < C; ++i) {
    const int s = dir == 0 ? s0 + i : s0 + i + 2;
    skip[i] = (s < S) && (s & 1) && s > 1 && (lab[s >> 1] != lab[(s >> 1) - 1]);
  }
  float a[C], e_next[C];
  double offset = 0.0;
  const int t_first = dir == 0 ? 0 : Tb - 1;
  const int step = dir == 0 ? 1 : -1;
#pragma unroll
  for (int i = 0; i < C; ++i) {
    const int s = s0 + i;
    const bool init = dir == 0 ? (s < 2) : (s >= S - 2);
    a[i] = (s < S && init) ? eb[(long long)t_first * Smax + s] : NEG_INF;
  }
  for (int n = 0; n < Tb; ++n) {
    const int t = t_first + n * step;
    if (n > 0) {
      // neighbours across the chunk boundary
      float p1, p2;  // alpha: states s0-1, s0-2 ; beta: states s0+C, s0+C+1
      if (dir == 0) {
        p1 = __shfl_up_sync(0xffffffffu, a[C - 1], 1);
        p2 = C >= 2 ? __shfl_up_sync(0xffffffffu, a[C >= 2 ? C - 2 : 0], 1) : __shfl_up_sync(0xffffffffu, a[0], 2);
        if (lane == 0) p1 = p2 = NEG_INF;
        if (C < 2 && lane == 1) p2 = NEG_INF;
#pragma unroll
        for (int i = C - 1; i >= 0; --i) {
          const float m1 = i >= 1 ? a[i - 1] : p1;
          const float m2 = i >= 2 ? a[i - 2] : (i == 1 ? p1 : p2);
          const float m2b = (C >= 2 && i == 0) ? p2 : m2;
          const float v = lse3(a[i], m1, skip[i] ? m2b : NEG_INF);
          a[i] = (s0 + i < S) ? v + e_next[i] : NEG_INF;
        }
      } else {
        p1 = __shfl_down_sync(0xffffffffu, a[0], 1);
        p2 = C >= 2 ? __shfl_down_sync(0xffffffffu, a[C >= 2 ? 1 : 0], 1) : __shfl_down_sync(0xffffffffu, a[0], 2);
        if (lane == 31) p1 = p2 = NEG_INF;
        if (C < 2 && lane == 30) p2 = NEG_INF;
#pragma unroll
        for (int i = 0; i < C; ++i) {
          const float m1 = i + 1 < C ? a[i + 1 < C ? i + 1 : 0] : p1;
          const float m2 = i + 2 < C ? a[i + 2 < C ? i + 2 : 0] : (i + 2 == C ? p1 : p2);
          const float v = lse3(a[i], m1, skip[i] ? m2 : NEG_INF);
          a[i] = (s0 + i < S) ? v + e_next[i] : NEG_INF;
        }
      }
    }
    // prefetch the next frame's emissions while this frame is renormalised and stored
    const int tn = t + step;
    if (n + 1 < Tb) {
#pragma unroll
      for (int i = 0; i < C; ++i) e_next[i] = (s0 + i < S) ? eb[(long long)tn * Smax + s0 + i] : 0.f;
    }
    float m = NEG_INF;
#pragma unroll
    for (int i = 0; i < C; ++i) m = fmaxf(m, a[i]);
    m = warp_max(m);
    if (m != NEG_INF) {
      offset += (double)m;
#pragma unroll
      for (int i = 0; i < C; ++i) a[i] -= m;
    }
#pragma unroll
    for (int i = 0; i < C; ++i)
      if (s0 + i < S) out[(long long)t * Smax + s0 + i] = a[i];
  }
  if (dir == 0) {
    // log-likelihood = offsets + lse(alpha~(S-1), alpha~(S-2)) at the last frame
    float fin = NEG_INF;
#pragma unroll
    for (int i = 0; i < C; ++i)
      if (s0 + i == S - 1 || s0 + i == S - 2) fin = lse2(fin, a[i]);
    // combine the (at most two) lanes holding the final states
    float mx = warp_max(fin);
    float sum = (fin == NEG_INF || mx == NEG_INF) ? 0.f : expf(fin - mx);
    sum = warp_sum(sum);
    if (lane == 0) {
      const double ll = (mx == NEG_INF) ? -INFINITY : offset + (double)mx + log((double)sum);
      nll[b] = (float)(-ll);  // +inf when infeasible; zero_infinity is applied by ctc_finish_kernel
    }
  }
}

// nll_raw (+inf if infeasible) -> loss (0 if infeasible) + feasibility flag; loss_sum = sum_b loss_b (one block)
__global__ void ctc_finish_kernel(const float* __restrict__ nll_raw, int B, float* __restrict__ loss, int* __restrict__ feasible,
                                  float* __restrict__ loss_sum) {
  __shared__ float red[32];
  float s = 0.f;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    const float v = nll_raw[b];
    const bool ok = isfinite(v);
    feasible[b] = ok ? 1 : 0;
    loss[b] = ok ? v : 0.f;
    s += ok ? v : 0.f;
  }
  s = block_sum(s, red);
  if (threadIdx.x == 0 && loss_sum != nullptr) *loss_sum = s;
}

// ---- 3. grad: one block per (b, t) -------------------------------------------------------------
__global__ void __launch_bounds__(256) ctc_grad_kernel(const float* __restrict__ logits, long long ldv, int T, int V,
                                                       const int* __restrict__ hlens, const int* __restrict__ labels, int Lmax,
                                                       const int* __restrict__ llens, int blank, const float* __restrict__ lse_in,
                                                       const float* __restrict__ em, const float* __restrict__ alpha,
                                                       const float* __restrict__ beta, int Smax, const int* __restrict__ feasible,
                                                       float scale, float* __restrict__ g32, bf16* __restrict__ g16,
                                                       long long ldg) {
  extern __shared__ float occ[];  // [V]
  __shared__ float red[32];
  const int t = blockIdx.x, b = blockIdx.y;
  const long long r = (long long)b * T + t;
  const int Vp = (int)ldg;
  if (t >= hlens[b] || !feasible[b]) {
    for (int v = threadIdx.x; v < Vp; v += blockDim.x) {
      if (g32 != nullptr) g32[r * ldg + v] = 0.f;
      if (g16 != nullptr) g16[r * ldg + v] = __float2bfloat16(0.f);
    }
    return;
  }
  for (int v = threadIdx.x; v < V; v += blockDim.x) occ[v] = 0.f;
  const int S = 2 * llens[b] + 1;
  const int* lab = labels + (long long)b * Lmax;
  const float* ar = alpha + r * Smax;
  const float* br = beta + r * Smax;
  const float* er = em + r * Smax;
  // frame-local normaliser Z = lse_s(alpha~ + beta~ - em)
  float mx = NEG_INF;
  for (int s = threadIdx.x; s < S; s += blockDim.x) mx = fmaxf(mx, ar[s] + br[s] - er[s]);
  mx = warp_max(mx);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
  __syncthreads();
  mx = red[0];
  for (int w = 1; w < (blockDim.x >> 5); ++w) mx = fmaxf(mx, red[w]);
  float sum = 0.f;
  for (int s = threadIdx.x; s < S; s += blockDim.x) {
    const float z = ar[s] + br[s] - er[s];
    sum += (z == NEG_INF) ? 0.f : expf(z - mx);
  }
  sum = block_sum(sum, red);  // contains the barriers that also order the occ[] zeroing
  const float inv = sum > 0.f ? 1.f / sum : 0.f;
  for (int s = threadIdx.x; s < S; s += blockDim.x) {
    const float z = ar[s] + br[s] - er[s];
    if (z != NEG_INF) atomicAdd(&occ[ext_label(lab, s, blank)], expf(z - mx) * inv);
  }
  __syncthreads();
  const float lse = lse_in[r];
  const float* row = logits + r * ldv;
  for (int v = threadIdx.x; v < Vp; v += blockDim.x) {
    const float g = v < V ? scale * (expf(row[v] - lse) - occ[v]) : 0.f;
    if (g32 != nullptr) g32[r * ldg + v] = g;
    if (g16 != nullptr) g16[r * ldg + v] = __float2bfloat16(g);
  }
}

long long ws_floats(int B, int T, int Lmax) {
  const long long Smax = 2LL * Lmax + 1;
  return 3LL * B * T * Smax + (long long)B * T + B + 8;  // em, alpha, beta, lse, nll_raw (+ pad)
}

}  // namespace

extern "C" long long b200slu_ctc_workspace_size(int B, int T, int Lmax) { return ws_floats(B, T, Lmax) * 4 + (long long)B * 4; }

// Forward: per-utterance loss (0 where infeasible, zero_infinity) and their sum.  The workspace keeps
// lse / em / alpha~ / beta~ / feasibility for b200slu_ctc_bwd.
extern "C" int b200slu_ctc_fwd(const float* logits, long long ldv, int B, int T, int V, const int* hlens, const int* labels,
                               int Lmax, const int* label_lens, int blank, void* workspace, float* loss_per_utt, float* loss_sum,
                               void* stream) {
  B2_CHECK_ARG(B > 0 && T > 0 && V > 1 && Lmax >= 0 && ldv >= V && B <= 65535, "ctc_fwd: bad shape B=%d T=%d V=%d Lmax=%d", B, T, V,
               Lmax);
  B2_CHECK_ARG(blank >= 0 && blank < V, "ctc_fwd: blank out of range");
  const int Smax = 2 * Lmax + 1;
  B2_CHECK_ARG(Smax <= 32 * 20, "ctc_fwd: label length %d exceeds the supported 319", Lmax);
  cudaStream_t st = (cudaStream_t)stream;
  float* ws = (float*)workspace;
  const long long n = (long long)B * T * Smax;
  float *em = ws, *alpha = ws + n, *beta = ws + 2 * n, *lse = ws + 3 * n, *nll_raw = lse + (long long)B * T;
  int* feasible = (int*)(ws + ws_floats(B, T, Lmax));
  B2_SCOPE(FAM_CTC, st, 0, 4.0 * B * T * V + 20.0 * n);
  ctc_prep_kernel<<<dim3(T, B), 256, 0, st>>>(logits, ldv, T, V, hlens, labels, Lmax > 0 ? Lmax : 1, label_lens, blank, lse, em,
                                              Smax);
  B2_LAUNCH_CHECK();
#define LAT(C)                                                                                                              \
  ctc_lattice_kernel<C><<<B, 64, 0, st>>>(em, T, Smax, hlens, labels, Lmax > 0 ? Lmax : 1, label_lens, blank, alpha, beta, \
                                          nll_raw)
  if (Smax <= 64) LAT(2);
  else if (Smax <= 128) LAT(4);
  else if (Smax <= 256) LAT(8);
  else LAT(20);
#undef LAT
  B2_LAUNCH_CHECK();
  ctc_finish_kernel<<<1, 256, 0, st>>>(nll_raw, B, loss_per_utt, feasible, loss_sum);
  B2_LAUNCH_CHECK();
  return 0;
}

// Gradient of (scale * sum_b loss_b) w.r.t. the logits: scale * (softmax - occupancy) for t < hlens[b], 0 beyond and 0 for
// infeasible utterances.  Writes fp32 and/or bf16 rows of stride ldg (columns [V, ldg) zeroed).
extern "C" int b200slu_ctc_bwd(const float* logits, long long ldv, int B, int T, int V, const int* hlens, const int* labels,
                               int Lmax, const int* label_lens, int blank, const void* workspace, float scale, float* grad_f32,
                               void* grad_bf16, long long ldg, void* stream) {
  B2_CHECK_ARG(B > 0 && T > 0 && V > 1 && ldg >= V && B <= 65535, "ctc_bwd: bad shape");
  B2_CHECK_ARG(grad_f32 != nullptr || grad_bf16 != nullptr, "ctc_bwd: no output");
  B2_CHECK_ARG((size_t)V * 4 <= 200 * 1024, "ctc_bwd: vocabulary too large for the shared-memory occupancy row");
  const int Smax = 2 * Lmax + 1;
  cudaStream_t st = (cudaStream_t)stream;
  const float* ws = (const float*)workspace;
  const long long n = (long long)B * T * Smax;
  const float *em = ws, *alpha = ws + n, *beta = ws + 2 * n, *lse = ws + 3 * n;
  const int* feasible = (const int*)(ws + ws_floats(B, T, Lmax));
  static bool cfg = false;
  if (!cfg) {
    B2_CUDA(cudaFuncSetAttribute(ctc_grad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    cfg = true;
  }
  B2_SCOPE(FAM_CTC, st, 0, (double)B * T * V * (4.0 + (grad_f32 ? 4.0 : 0.0) + (grad_bf16 ? 2.0 : 0.0)) + 12.0 * n);
  ctc_grad_kernel<<<dim3(T, B), 256, (size_t)V * 4, st>>>(logits, ldv, T, V, hlens, labels, Lmax > 0 ? Lmax : 1, label_lens, blank,
                                                          lse, em, alpha, beta, Smax, feasible, scale, grad_f32, (bf16*)grad_bf16,
                                                          ldg);
  B2_LAUNCH_CHECK();
  return 0;
}
