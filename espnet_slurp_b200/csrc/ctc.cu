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
// The lattice runs in the log2 domain with ex2.approx / lg2.approx (one MUFU op each, rel. error 2^-22) and a FINITE
// stand-in for log(0): with -1e30 the 3-way logsumexp needs no special case (exp2 of the differences is 1 or 0, the
// result stays ~-1e30), and an all-impossible row shows up as a total log-likelihood below -1e29.
constexpr float NEG_BIG = -1.0e30f;
constexpr float LOG2E = 1.4426950408889634f, LN2 = 0.6931471805599453f;
__device__ __forceinline__ float ex2_(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float lg2_(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float lse2(float a, float b) {  // log2 domain
  const float m = fmaxf(a, b);
  return m + lg2_(ex2_(a - m) + ex2_(b - m));
}
__device__ __forceinline__ float lse3(float a, float b, float c) {  // log2 domain
  const float m = fmaxf(a, fmaxf(b, c));
  return m + lg2_(ex2_(a - m) + ex2_(b - m) + ex2_(c - m));
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
  // ONE pass over the row: per-thread running (max, sum of 2^((x - max) log2 e)), merged across the block afterwards
  float m = NEG_BIG, sm = 0.f;
  if (ldv % 4 == 0 && ((uintptr_t)logits % 16) == 0) {
    for (int v = threadIdx.x * 4; v < V; v += blockDim.x * 4) {
      float x[4];
      if (v + 4 <= V) {
        const float4 q = *reinterpret_cast<const float4*>(row + v);
        x[0] = q.x; x[1] = q.y; x[2] = q.z; x[3] = q.w;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) x[j] = v + j < V ? row[v + j] : NEG_BIG;
      }
      const float lm = fmaxf(fmaxf(x[0], x[1]), fmaxf(x[2], x[3]));
      if (lm > m) {
        sm *= ex2_((m - lm) * LOG2E);
        m = lm;
      }
      const float mb = m * LOG2E;
      sm += ex2_(fmaf(x[0], LOG2E, -mb)) + ex2_(fmaf(x[1], LOG2E, -mb)) + ex2_(fmaf(x[2], LOG2E, -mb)) + ex2_(fmaf(x[3], LOG2E, -mb));
    }
  } else {
    for (int v = threadIdx.x; v < V; v += blockDim.x) {
      const float x = row[v];
      if (x > m) {
        sm *= ex2_((m - x) * LOG2E);
        m = x;
      }
      sm += ex2_((x - m) * LOG2E);
    }
  }
  float mx = warp_max(m);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
  __syncthreads();
  mx = red[0];
  for (int w = 1; w < (blockDim.x >> 5); ++w) mx = fmaxf(mx, red[w]);
  float sum = sm * ex2_((m - mx) * LOG2E);  // threads that saw no element hold (NEG_BIG, 0) -> 0
  sum = block_sum(sum, red);
  const float lse = mx + logf(sum);
  if (threadIdx.x == 0) lse_out[(long long)b * T + t] = lse;
  const int S = 2 * llens[b] + 1;
  const int* lab = labels + (long long)b * Lmax;
  float* e = em + ((long long)b * T + t) * Smax;
  for (int s = threadIdx.x; s < S; s += blockDim.x) e[s] = (row[ext_label(lab, s, blank)] - lse) * LOG2E;  // log2 units
}

// ---- 2. lattice: block = 2 warps per utterance -------------------------------------------------
// max over the warp in one instruction: order-preserving float -> int map, redux.sync, map back
__device__ __forceinline__ float warp_max_fast(float x) {
  int i = __float_as_int(x);
  i ^= (i >> 31) & 0x7fffffff;
  i = __reduce_max_sync(0xffffffffu, i);
  i ^= (i >> 31) & 0x7fffffff;
  return __int_as_float(i);
}

template <int C, int D>  // C states per lane (S <= 32*C); D frames of emission prefetch
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
  float a[C];
  float e[D][C];  // emissions of the next D frames: the loads are D frames ahead of their use (L2 latency off the chain)
  double offset = 0.0;
  const int t_first = dir == 0 ? 0 : Tb - 1;
  const int step = dir == 0 ? 1 : -1;
#pragma unroll
  for (int i = 0; i < C; ++i) {
    const int s = s0 + i;
    const bool init = dir == 0 ? (s < 2) : (s >= S - 2);
    a[i] = (s < S && init) ? eb[(long long)t_first * Smax + s] : NEG_BIG;
  }
#pragma unroll
  for (int j = 0; j < D; ++j) {
    const int nn = 1 + j;
#pragma unroll
    for (int i = 0; i < C; ++i)
      e[j][i] = (nn < Tb && s0 + i < S) ? eb[(long long)(t_first + nn * step) * Smax + s0 + i] : 0.f;
  }
  // renormalise the row by its maximum (one redux instruction) and store it
  auto renorm_store = [&](int t) {
    float m = NEG_BIG;
#pragma unroll
    for (int i = 0; i < C; ++i) m = fmaxf(m, a[i]);
    m = warp_max_fast(m);
    offset += (double)m;  // an impossible row (m ~ -1e30) drives the total below -1e29: reported as infeasible
#pragma unroll
    for (int i = 0; i < C; ++i) a[i] = fmaxf(a[i] - m, NEG_BIG);
#pragma unroll
    for (int i = 0; i < C; ++i)
      if (s0 + i < S) out[(long long)t * Smax + s0 + i] = a[i];
  };
  renorm_store(t_first);
  for (int n0 = 1; n0 < Tb; n0 += D) {
#pragma unroll
    for (int j = 0; j < D; ++j) {
      const int n = n0 + j;
      if (n >= Tb) break;
      const int t = t_first + n * step;
      // neighbours across the chunk boundary
      float p1, p2;  // alpha: states s0-1, s0-2 ; beta: states s0+C, s0+C+1
      if (dir == 0) {
        p1 = __shfl_up_sync(0xffffffffu, a[C - 1], 1);
        p2 = C >= 2 ? __shfl_up_sync(0xffffffffu, a[C >= 2 ? C - 2 : 0], 1) : __shfl_up_sync(0xffffffffu, a[0], 2);
        if (lane == 0) p1 = p2 = NEG_BIG;
        if (C < 2 && lane == 1) p2 = NEG_BIG;
#pragma unroll
        for (int i = C - 1; i >= 0; --i) {
          const float m1 = i >= 1 ? a[i >= 1 ? i - 1 : 0] : p1;
          const float m2 = i >= 2 ? a[i >= 2 ? i - 2 : 0] : (i == 1 ? p1 : p2);
          const float v = lse3(a[i], m1, skip[i] ? m2 : NEG_BIG);
          a[i] = (s0 + i < S) ? v + e[j][i] : NEG_BIG;
        }
      } else {
        p1 = __shfl_down_sync(0xffffffffu, a[0], 1);
        p2 = C >= 2 ? __shfl_down_sync(0xffffffffu, a[C >= 2 ? 1 : 0], 1) : __shfl_down_sync(0xffffffffu, a[0], 2);
        if (lane == 31) p1 = p2 = NEG_BIG;
        if (C < 2 && lane == 30) p2 = NEG_BIG;
#pragma unroll
        for (int i = 0; i < C; ++i) {
          const float m1 = i + 1 < C ? a[i + 1 < C ? i + 1 : 0] : p1;
          const float m2 = i + 2 < C ? a[i + 2 < C ? i + 2 : 0] : (i + 2 == C ? p1 : p2);
          const float v = lse3(a[i], m1, skip[i] ? m2 : NEG_BIG);
          a[i] = (s0 + i < S) ? v + e[j][i] : NEG_BIG;
        }
      }
      // refill this slot with the emissions of frame n + D
      const int nf = n + D;
      if (nf < Tb) {
#pragma unroll
        for (int i = 0; i < C; ++i) e[j][i] = (s0 + i < S) ? eb[(long long)(t_first + nf * step) * Smax + s0 + i] : 0.f;
      }
      renorm_store(t);
    }
  }
  if (dir == 0) {
    // log-likelihood = offsets + lse(alpha~(S-1), alpha~(S-2)) at the last frame
    float fin = NEG_BIG;
#pragma unroll
    for (int i = 0; i < C; ++i)
      if (s0 + i == S - 1 || s0 + i == S - 2) fin = lse2(fin, a[i]);
    // combine the (at most two) lanes holding the final states
    float mx = warp_max(fin);
    float sum = ex2_(fin - mx);
    sum = warp_sum(sum);
    if (lane == 0) {
      const double ll2 = offset + (double)mx + log2((double)sum);  // log2 of the sequence likelihood
      nll[b] = (ll2 < -1.0e29) ? INFINITY : (float)(-ll2 * 0.6931471805599453);  // +inf when infeasible (zero_infinity later)
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
                                                       long long ldg, bool vec4) {
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
  for (int s = threadIdx.x; s < S; s += blockDim.x) sum += ex2_(ar[s] + br[s] - er[s] - mx);  // log2 domain; impossible states -> 0
  sum = block_sum(sum, red);  // contains the barriers that also order the occ[] zeroing
  const float inv = sum > 0.f ? 1.f / sum : 0.f;
  for (int s = threadIdx.x; s < S; s += blockDim.x) {
    const float o = ex2_(ar[s] + br[s] - er[s] - mx) * inv;
    if (o > 0.f) atomicAdd(&occ[ext_label(lab, s, blank)], o);
  }
  __syncthreads();
  const float lse2 = lse_in[r] * LOG2E;
  const float* row = logits + r * ldv;
  if (vec4) {  // rows 16-byte aligned, ldg % 4 == 0: 16-byte loads and stores
    for (int v = threadIdx.x * 4; v < Vp; v += blockDim.x * 4) {
      float g[4];
      if (v + 4 <= V) {
        const float4 x = *reinterpret_cast<const float4*>(row + v);
        g[0] = scale * (ex2_(fmaf(x.x, LOG2E, -lse2)) - occ[v]);
        g[1] = scale * (ex2_(fmaf(x.y, LOG2E, -lse2)) - occ[v + 1]);
        g[2] = scale * (ex2_(fmaf(x.z, LOG2E, -lse2)) - occ[v + 2]);
        g[3] = scale * (ex2_(fmaf(x.w, LOG2E, -lse2)) - occ[v + 3]);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) g[j] = v + j < V ? scale * (ex2_(fmaf(row[v + j], LOG2E, -lse2)) - occ[v + j]) : 0.f;
      }
      if (g32 != nullptr) *reinterpret_cast<float4*>(g32 + r * ldg + v) = make_float4(g[0], g[1], g[2], g[3]);
      if (g16 != nullptr)
        *reinterpret_cast<uint2*>(g16 + r * ldg + v) = make_uint2(pack_bf16x2(g[0], g[1]), pack_bf16x2(g[2], g[3]));
    }
  } else {
    for (int v = threadIdx.x; v < Vp; v += blockDim.x) {
      const float g = v < V ? scale * (ex2_(fmaf(row[v], LOG2E, -lse2)) - occ[v]) : 0.f;
      if (g32 != nullptr) g32[r * ldg + v] = g;
      if (g16 != nullptr) g16[r * ldg + v] = __float2bfloat16(g);
    }
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
  ctc_lattice_kernel<C, (C <= 8 ? 4 : 2)><<<B, 64, 0, st>>>(em, T, Smax, hlens, labels, Lmax > 0 ? Lmax : 1, label_lens, blank, alpha, beta, \
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
  const bool vec4 = ldv % 4 == 0 && ldg % 4 == 0 && ((uintptr_t)logits % 16) == 0 &&
                    (grad_f32 == nullptr || ((uintptr_t)grad_f32 % 16) == 0) && (grad_bf16 == nullptr || ((uintptr_t)grad_bf16 % 8) == 0);
  ctc_grad_kernel<<<dim3(T, B), 256, (size_t)V * 4, st>>>(logits, ldv, T, V, hlens, labels, Lmax > 0 ? Lmax : 1, label_lens, blank,
                                                          lse, em, alpha, beta, Smax, feasible, scale, grad_f32, (bf16*)grad_bf16,
                                                          ldg, vec4);
  B2_LAUNCH_CHECK();
  return 0;
}
