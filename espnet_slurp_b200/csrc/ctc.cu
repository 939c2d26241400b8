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
#include <stdlib.h>

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
// skip arc s-2 -> s allowed iff s is a label state past the first one whose label differs from the previous label
// (gtn_ctc.py:26-36); the only definition: both lattice kernels and the export below call it
__device__ __forceinline__ bool ctc_skip_into(const int* __restrict__ lab, int s, int S) {
  return (s < S) && (s & 1) && s > 1 && (lab[s >> 1] != lab[(s >> 1) - 1]);
}
// the lattice as the device code sees it, for the bit-exact test against gtn_ctc.py:26-36
__global__ void ctc_lattice_states_kernel(const int* __restrict__ labels, int Lmax, const int* __restrict__ llens, int blank,
                                          int Smax, int* __restrict__ ext_out, unsigned char* __restrict__ skip_out) {
  const int b = blockIdx.x;
  const int* lab = labels + (long long)b * Lmax;
  const int S = 2 * llens[b] + 1;
  for (int s = threadIdx.x; s < Smax; s += blockDim.x) {
    ext_out[(long long)b * Smax + s] = s < S ? ext_label(lab, s, blank) : -1;
    skip_out[(long long)b * Smax + s] = ctc_skip_into(lab, s, S) ? 1 : 0;
  }
}

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

// Small vocabularies (V <= 2048: the SLURP / LibriSpeech-BPE heads): ONE WARP per frame, 8 frames per block in flight, each
// warp walking frames grid-stride.  A 500-entry row is 2 KB -- half the threads of a 256-thread block had nothing to load
// and every frame paid three block barriers and a block launch; here the reductions are warp shuffles and a row's loads
// are all issued before the first one is used.  (V = 500, T = 1000, B = 64: prep + grad 265 -> see DESIGN.md section 5.)
constexpr int CTC_WPB = 8;      // warps (= frames in flight) per block
constexpr int CTC_WARP_V = 2048;       // prep: the warp form wins up to here (V = 2000: forward 417 -> 370 us at T = 1000)
constexpr int CTC_WARP_V_GRAD = 1024;  // grad: beyond this the block form is ahead (a warp then walks 16 chunks per lane in turn)
__global__ void __launch_bounds__(32 * CTC_WPB) ctc_prep_warp_kernel(const float* __restrict__ logits, long long ldv, int T, int V,
                                                                    const int* __restrict__ hlens, const int* __restrict__ labels,
                                                                    int Lmax, const int* __restrict__ llens, int blank,
                                                                    float* __restrict__ lse_out, float* __restrict__ em, int Smax,
                                                                    long long nframes) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (long long r = (long long)blockIdx.x * CTC_WPB + w; r < nframes; r += (long long)gridDim.x * CTC_WPB) {
    const int b = (int)(r / T), t = (int)(r - (long long)b * T);
    if (t >= hlens[b]) continue;
    const float* row = logits + r * ldv;
    float m = NEG_BIG, sm = 0.f;
    constexpr int U = 4;  // 16-byte loads in flight per lane and round (V <= 512: the whole row in one round)
    for (int v0 = lane * 4; v0 < V; v0 += 128 * U) {
      float x[U][4];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int v = v0 + u * 128;
        if (v + 4 <= V) {
          const float4 q = *reinterpret_cast<const float4*>(row + v);
          x[u][0] = q.x; x[u][1] = q.y; x[u][2] = q.z; x[u][3] = q.w;
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) x[u][j] = v + j < V ? row[v + j] : NEG_BIG;
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const float lm = fmaxf(fmaxf(x[u][0], x[u][1]), fmaxf(x[u][2], x[u][3]));
        if (lm > m) {
          sm *= ex2_((m - lm) * LOG2E);
          m = lm;
        }
        const float mb = m * LOG2E;
        sm += ex2_(fmaf(x[u][0], LOG2E, -mb)) + ex2_(fmaf(x[u][1], LOG2E, -mb)) + ex2_(fmaf(x[u][2], LOG2E, -mb)) +
              ex2_(fmaf(x[u][3], LOG2E, -mb));
      }
    }
    const float mx = warp_max(m);
    const float sum = warp_sum(sm * ex2_((m - mx) * LOG2E));  // lanes that saw no element hold (NEG_BIG, 0) -> 0
    const float lse = mx + logf(sum);
    if (lane == 0) lse_out[r] = lse;
    const int S = 2 * llens[b] + 1;
    const int* lab = labels + (long long)b * Lmax;
    float* e = em + r * Smax;
    for (int s = lane; s < S; s += 32) e[s] = (row[ext_label(lab, s, blank)] - lse) * LOG2E;  // log2 units
  }
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

// One direction of the recursion (DIR 0 = alpha, forward in time; 1 = beta, backward) for one utterance, run by one warp.
// The recursion is a single-warp latency chain, so what a frame costs is its instruction count: the direction is a
// template parameter (no per-frame branches), the main loop runs whole groups of D frames without bounds checks (the
// prefetch of frame n + D clamps its frame number instead of predicating the load), the < D tail frames use the
// emissions already held in registers, and dead states carry an emission of -1e30 so they need no select.
template <int C, int D, int DIR, bool PEEL>  // C states per lane (S <= 32*C); D frames of emission prefetch
__device__ __forceinline__ void lattice_run(const float* __restrict__ eb, int Smax, int Tb, int S, const int* __restrict__ lab,
                                            int lane, float* __restrict__ out, float* __restrict__ nll) {
  const int s0 = lane * C;
  // skip-arc permission of the state a transition ENTERS: alpha uses skip[s] (s-2 -> s); beta uses skip[s+2]
  bool skip[C], live[C];
#pragma unroll
  for (int i = 0; i < C; ++i) {
    const int s = DIR == 0 ? s0 + i : s0 + i + 2;
    skip[i] = ctc_skip_into(lab, s, S);
    live[i] = s0 + i < S;
  }
  const int t_first = DIR == 0 ? 0 : Tb - 1;
  const int step = DIR == 0 ? 1 : -1;
  const float* const e0 = eb + s0;  // + frame * Smax
  float* op = out + (long long)t_first * Smax + s0;
  const long long ostep = DIR == 0 ? Smax : -Smax;
  float a[C];
  float e[D][C];  // emissions of the next D frames: the loads are D frames ahead of their use (L2 latency off the chain)
  double offset = 0.0;
#pragma unroll
  for (int i = 0; i < C; ++i) {
    const int s = s0 + i;
    const bool init = DIR == 0 ? (s < 2) : (s >= S - 2);
    a[i] = (live[i] && init) ? e0[(long long)t_first * Smax + i] : NEG_BIG;
  }
  auto load_frame = [&](int n, float (&dst)[C]) {  // emissions of the n-th frame of the walk (clamped to the last one)
    const float* src = e0 + (t_first + min(n, Tb - 1) * step) * Smax;
#pragma unroll
    for (int i = 0; i < C; ++i) dst[i] = live[i] ? src[i] : NEG_BIG;
  };
#pragma unroll
  for (int j = 0; j < D; ++j) load_frame(1 + j, e[j]);
  // renormalise the row by its maximum (one redux instruction) and store it
  auto renorm_store = [&]() {
    float m = a[0];
#pragma unroll
    for (int i = 1; i < C; ++i) m = fmaxf(m, a[i]);
    m = warp_max_fast(m);
    offset += (double)m;  // an impossible row (m ~ -1e30) drives the total below -1e29: reported as infeasible
#pragma unroll
    for (int i = 0; i < C; ++i) a[i] = fmaxf(a[i] - m, NEG_BIG);
#pragma unroll
    for (int i = 0; i < C; ++i)
      if (live[i]) op[i] = a[i];
  };
  const bool edge = DIR == 0 ? lane == 0 : lane == 31;
  const bool edge2 = DIR == 0 ? lane == 1 : lane == 30;  // C == 1: the second neighbour comes from two lanes away
  auto frame = [&](const float (&em)[C]) {
    // neighbours across the chunk boundary -- alpha: states s0-1, s0-2 ; beta: states s0+C, s0+C+1
    float p1, p2;
    if (DIR == 0) {
      p1 = __shfl_up_sync(0xffffffffu, a[C - 1], 1);
      p2 = C >= 2 ? __shfl_up_sync(0xffffffffu, a[C >= 2 ? C - 2 : 0], 1) : __shfl_up_sync(0xffffffffu, a[0], 2);
    } else {
      p1 = __shfl_down_sync(0xffffffffu, a[0], 1);
      p2 = C >= 2 ? __shfl_down_sync(0xffffffffu, a[C >= 2 ? 1 : 0], 1) : __shfl_down_sync(0xffffffffu, a[0], 2);
    }
    p1 = edge ? NEG_BIG : p1;
    p2 = (edge || (C < 2 && edge2)) ? NEG_BIG : p2;
    if (DIR == 0) {
#pragma unroll
      for (int i = C - 1; i >= 0; --i) {
        const float m1 = i >= 1 ? a[i >= 1 ? i - 1 : 0] : p1;
        const float m2 = i >= 2 ? a[i >= 2 ? i - 2 : 0] : (i == 1 ? p1 : p2);
        a[i] = lse3(a[i], m1, skip[i] ? m2 : NEG_BIG) + em[i];
      }
    } else {
#pragma unroll
      for (int i = 0; i < C; ++i) {
        const float m1 = i + 1 < C ? a[i + 1 < C ? i + 1 : 0] : p1;
        const float m2 = i + 2 < C ? a[i + 2 < C ? i + 2 : 0] : (i + 2 == C ? p1 : p2);
        a[i] = lse3(a[i], m1, skip[i] ? m2 : NEG_BIG) + em[i];
      }
    }
  };
  renorm_store();
  int n = 1;
  if (PEEL) {
    for (; n + D <= Tb; n += D) {
#pragma unroll
      for (int j = 0; j < D; ++j) {
        frame(e[j]);
        load_frame(n + j + D, e[j]);  // refill this slot with the emissions of frame n + j + D
        op += ostep;
        renorm_store();
      }
    }
  }
  // PEEL: the last Tb - n < D frames.  Otherwise (many states per lane: the frame body is ~500 instructions and a second
  // copy of it does not pay) the whole walk, with the per-frame bound check.
  for (; n < Tb; n += D) {
#pragma unroll
    for (int j = 0; j < D; ++j) {
      if (n + j < Tb) {
        frame(e[j]);
        load_frame(n + j + D, e[j]);
        op += ostep;
        renorm_store();
      }
    }
  }
  if (DIR == 0) {
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
      *nll = (ll2 < -1.0e29) ? INFINITY : (float)(-ll2 * 0.6931471805599453);  // +inf when infeasible (zero_infinity later)
    }
  }
}

template <int C, int D>
__global__ void __launch_bounds__(64) ctc_lattice_kernel(const float* __restrict__ em, int T, int Smax,
                                                         const int* __restrict__ hlens, const int* __restrict__ labels, int Lmax,
                                                         const int* __restrict__ llens, int blank, float* __restrict__ alpha,
                                                         float* __restrict__ beta, float* __restrict__ nll) {
  const int b = blockIdx.x, lane = threadIdx.x & 31, dir = threadIdx.x >> 5;  // dir 0 = alpha, 1 = beta
  const int Tb = hlens[b], L = llens[b], S = 2 * L + 1;
  const int* lab = labels + (long long)b * Lmax;
  const float* eb = em + (long long)b * T * Smax;
  if (Tb <= 0 || Tb > T || L > Lmax || S > 32 * C) {
    if (dir == 0 && lane == 0) nll[b] = INFINITY;  // infeasible -> zeroed by the caller-visible flag path
    return;
  }
  if (dir == 0) lattice_run<C, D, 0, (C <= 8)>(eb, Smax, Tb, S, lab, lane, alpha + (long long)b * T * Smax, nll + b);
  else lattice_run<C, D, 1, (C <= 8)>(eb, Smax, Tb, S, lab, lane, beta + (long long)b * T * Smax, nll + b);
}

// Long label sequences (S > 128 states): W warps per direction share the lattice (C states per lane, 32 * W * C >= S), so
// the per-frame chain is C = 4..5 states long instead of 8..20.  A frame costs ONE named barrier: every warp publishes
// its un-normalised boundary states and its row maximum, and after the barrier each warp reads all W maxima (the row
// normaliser is global, so every warp subtracts the same value) and its neighbour's boundary states for the next frame
// (slots double-buffered by frame parity).
template <int C, int D, int DIR, int W>
__device__ __forceinline__ void lattice_run_mw(const float* __restrict__ eb, int Smax, int Tb, int S, const int* __restrict__ lab,
                                               int w, int lane, float* __restrict__ out, float* __restrict__ nll,
                                               float (*slot)[W][4] /* [2 parities][W][4] */, int bar_id) {
  static_assert(C >= 2, "multi-warp lattice: at least two states per lane");
  const int s0 = (w * 32 + lane) * C;
  bool skip[C], live[C];
#pragma unroll
  for (int i = 0; i < C; ++i) {
    const int s = DIR == 0 ? s0 + i : s0 + i + 2;
    skip[i] = ctc_skip_into(lab, s, S);
    live[i] = s0 + i < S;
  }
  const int t_first = DIR == 0 ? 0 : Tb - 1;
  const int step = DIR == 0 ? 1 : -1;
  const float* const e0 = eb + s0;
  float* op = out + (long long)t_first * Smax + s0;
  const long long ostep = DIR == 0 ? Smax : -Smax;
  float a[C], e[D][C];
  double offset = 0.0;
#pragma unroll
  for (int i = 0; i < C; ++i) {
    const int s = s0 + i;
    const bool init = DIR == 0 ? (s < 2) : (s >= S - 2);
    a[i] = (live[i] && init) ? e0[(long long)t_first * Smax + i] : NEG_BIG;
  }
  auto load_frame = [&](int n, float (&dst)[C]) {
    const float* src = e0 + (t_first + min(n, Tb - 1) * step) * Smax;
#pragma unroll
    for (int i = 0; i < C; ++i) dst[i] = live[i] ? src[i] : NEG_BIG;
  };
#pragma unroll
  for (int j = 0; j < D; ++j) load_frame(1 + j, e[j]);
  const int wn = DIR == 0 ? w - 1 : w + 1;                 // the warp whose boundary states this warp's edge lane needs
  const bool has_nb = DIR == 0 ? w > 0 : w < W - 1;
  const bool edge = DIR == 0 ? lane == 0 : lane == 31;
  const bool pub = DIR == 0 ? lane == 31 : lane == 0;      // the lane that owns the states the next warp needs
  float nb1 = NEG_BIG, nb2 = NEG_BIG;                       // neighbour warp's boundary states of the previous frame, normalised
  int par = 0;
  auto publish_renorm_store = [&]() {
    float m = a[0];
#pragma unroll
    for (int i = 1; i < C; ++i) m = fmaxf(m, a[i]);
    m = warp_max_fast(m);
    if (pub) {
      slot[par][w][0] = DIR == 0 ? a[C - 1] : a[0];
      slot[par][w][1] = DIR == 0 ? a[C - 2] : a[1];
      slot[par][w][2] = m;
    }
    asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "n"(W * 32) : "memory");
    float mg = slot[par][0][2];
#pragma unroll
    for (int v = 1; v < W; ++v) mg = fmaxf(mg, slot[par][v][2]);
    if (has_nb) {
      nb1 = fmaxf(slot[par][wn][0] - mg, NEG_BIG);
      nb2 = fmaxf(slot[par][wn][1] - mg, NEG_BIG);
    }
    offset += (double)mg;
#pragma unroll
    for (int i = 0; i < C; ++i) a[i] = fmaxf(a[i] - mg, NEG_BIG);
#pragma unroll
    for (int i = 0; i < C; ++i)
      if (live[i]) op[i] = a[i];
    par ^= 1;
  };
  auto frame = [&](const float (&em)[C]) {
    float p1, p2;
    if (DIR == 0) {
      p1 = __shfl_up_sync(0xffffffffu, a[C - 1], 1);
      p2 = __shfl_up_sync(0xffffffffu, a[C - 2], 1);
    } else {
      p1 = __shfl_down_sync(0xffffffffu, a[0], 1);
      p2 = __shfl_down_sync(0xffffffffu, a[1], 1);
    }
    p1 = edge ? nb1 : p1;
    p2 = edge ? nb2 : p2;
    if (DIR == 0) {
#pragma unroll
      for (int i = C - 1; i >= 0; --i) {
        const float m1 = i >= 1 ? a[i >= 1 ? i - 1 : 0] : p1;
        const float m2 = i >= 2 ? a[i >= 2 ? i - 2 : 0] : (i == 1 ? p1 : p2);
        a[i] = lse3(a[i], m1, skip[i] ? m2 : NEG_BIG) + em[i];
      }
    } else {
#pragma unroll
      for (int i = 0; i < C; ++i) {
        const float m1 = i + 1 < C ? a[i + 1 < C ? i + 1 : 0] : p1;
        const float m2 = i + 2 < C ? a[i + 2 < C ? i + 2 : 0] : (i + 2 == C ? p1 : p2);
        a[i] = lse3(a[i], m1, skip[i] ? m2 : NEG_BIG) + em[i];
      }
    }
  };
  publish_renorm_store();
  int n = 1;
  for (; n + D <= Tb; n += D) {
#pragma unroll
    for (int j = 0; j < D; ++j) {
      frame(e[j]);
      load_frame(n + j + D, e[j]);
      op += ostep;
      publish_renorm_store();
    }
  }
#pragma unroll
  for (int j = 0; j < D; ++j) {  // Tb is the same for every warp of the utterance: the barrier count stays matched
    if (n + j < Tb) {
      frame(e[j]);
      op += ostep;
      publish_renorm_store();
    }
  }
  if (DIR == 0) {
    float fin = NEG_BIG;
#pragma unroll
    for (int i = 0; i < C; ++i)
      if (s0 + i == S - 1 || s0 + i == S - 2) fin = lse2(fin, a[i]);
    const float mx = warp_max(fin);
    float sum = ex2_(fin - mx);
    sum = warp_sum(sum);
    if (lane == 0) {  // `par` is the buffer nobody reads any more
      slot[par][w][0] = mx;
      slot[par][w][1] = sum;
    }
    asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "n"(W * 32) : "memory");
    if (w == 0 && lane == 0) {
      float M = slot[par][0][0];
#pragma unroll
      for (int v = 1; v < W; ++v) M = fmaxf(M, slot[par][v][0]);
      float tot = 0.f;
#pragma unroll
      for (int v = 0; v < W; ++v) tot += slot[par][v][1] * ex2_(slot[par][v][0] - M);
      const double ll2 = offset + (double)M + log2((double)tot);
      *nll = (ll2 < -1.0e29) ? INFINITY : (float)(-ll2 * 0.6931471805599453);
    }
  }
}

template <int C, int D, int W>
__global__ void __launch_bounds__(64 * W) ctc_lattice_mw_kernel(const float* __restrict__ em, int T, int Smax,
                                                                const int* __restrict__ hlens, const int* __restrict__ labels,
                                                                int Lmax, const int* __restrict__ llens, int blank,
                                                                float* __restrict__ alpha, float* __restrict__ beta,
                                                                float* __restrict__ nll) {
  __shared__ float slot[2][2][W][4];  // [direction][frame parity][warp][boundary 1, boundary 2, row max, -]
  const int b = blockIdx.x, lane = threadIdx.x & 31, wi = threadIdx.x >> 5, dir = wi / W, w = wi % W;
  const int Tb = hlens[b], L = llens[b], S = 2 * L + 1;
  const int* lab = labels + (long long)b * Lmax;
  const float* eb = em + (long long)b * T * Smax;
  if (Tb <= 0 || Tb > T || L > Lmax || S > 32 * W * C) {
    if (threadIdx.x == 0) nll[b] = INFINITY;
    return;
  }
  if (dir == 0) lattice_run_mw<C, D, 0, W>(eb, Smax, Tb, S, lab, w, lane, alpha + (long long)b * T * Smax, nll + b, slot[0], 1);
  else lattice_run_mw<C, D, 1, W>(eb, Smax, Tb, S, lab, w, lane, beta + (long long)b * T * Smax, nll + b, slot[1], 2);
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
  // the blank's occupancy (every other state: up to 301 adds on one word) is a block reduction; the label states go
  // through shared-memory atomics (repeated labels collide there)
  float ob = 0.f;
  for (int s = threadIdx.x; s < S; s += blockDim.x) {
    const float o = ex2_(ar[s] + br[s] - er[s] - mx) * inv;
    if (s & 1) {
      if (o > 0.f) atomicAdd(&occ[lab[s >> 1]], o);
    } else {
      ob += o;
    }
  }
  ob = block_sum(ob, red);  // (its barriers also order the atomics above before the reads below)
  if (threadIdx.x == 0 && ob > 0.f) atomicAdd(&occ[blank], ob);
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

// The same for small vocabularies: one warp per frame, its occupancy row in its own slice of shared memory.  The blank's
// occupancy (every other lattice state) is a warp reduction; only the label states go through shared-memory atomics
// (repeated labels collide there -- with the blank included, a frame of a 300-label utterance queued 301 adds on one word).
__global__ void __launch_bounds__(32 * CTC_WPB) ctc_grad_warp_kernel(const float* __restrict__ logits, long long ldv, int T, int V,
                                                                    const int* __restrict__ hlens, const int* __restrict__ labels,
                                                                    int Lmax, const int* __restrict__ llens, int blank,
                                                                    const float* __restrict__ lse_in, const float* __restrict__ em,
                                                                    const float* __restrict__ alpha, const float* __restrict__ beta,
                                                                    int Smax, const int* __restrict__ feasible, float scale,
                                                                    float* __restrict__ g32, bf16* __restrict__ g16, long long ldg,
                                                                    long long nframes, int Vs) {
  extern __shared__ __align__(16) float occ_all[];  // [CTC_WPB][Vs]
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  float* occ = occ_all + (long long)w * Vs;
  const int Vp = (int)ldg;
  for (long long r = (long long)blockIdx.x * CTC_WPB + w; r < nframes; r += (long long)gridDim.x * CTC_WPB) {
    const int b = (int)(r / T), t = (int)(r - (long long)b * T);
    if (t >= hlens[b] || !feasible[b]) {
      for (int v = lane * 4; v < Vp; v += 128) {
        if (g32 != nullptr) *reinterpret_cast<float4*>(g32 + r * ldg + v) = make_float4(0.f, 0.f, 0.f, 0.f);
        if (g16 != nullptr) *reinterpret_cast<uint2*>(g16 + r * ldg + v) = make_uint2(0u, 0u);
      }
      continue;
    }
    for (int v = lane * 4; v < Vs; v += 128) *reinterpret_cast<float4*>(occ + v) = make_float4(0.f, 0.f, 0.f, 0.f);
    const int S = 2 * llens[b] + 1;
    const int* lab = labels + (long long)b * Lmax;
    const float* ar = alpha + r * Smax;
    const float* br = beta + r * Smax;
    const float* er = em + r * Smax;
    // the row of the logits is requested now; it is consumed after the lattice terms
    const float* row = logits + r * ldv;
    // frame-local normaliser Z = lse_s(alpha~ + beta~ - em)
    float mx = NEG_INF;
    for (int s = lane; s < S; s += 32) mx = fmaxf(mx, ar[s] + br[s] - er[s]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int s = lane; s < S; s += 32) sum += ex2_(ar[s] + br[s] - er[s] - mx);  // log2 domain; impossible states -> 0
    sum = warp_sum(sum);
    const float inv = sum > 0.f ? 1.f / sum : 0.f;
    __syncwarp();  // occ[] zeroed by every lane before anyone adds into it
    float ob = 0.f;
    for (int s = lane; s < S; s += 32) {
      const float o = ex2_(ar[s] + br[s] - er[s] - mx) * inv;
      if (s & 1) {
        if (o > 0.f) atomicAdd(&occ[lab[s >> 1]], o);
      } else {
        ob += o;
      }
    }
    ob = warp_sum(ob);
    if (lane == 0 && ob > 0.f) atomicAdd(&occ[blank], ob);
    __syncwarp();
    const float lse2 = lse_in[r] * LOG2E;
    for (int v = lane * 4; v < Vp; v += 128) {
      float g[4];
      if (v + 4 <= V) {
        const float4 x = *reinterpret_cast<const float4*>(row + v);
        const float4 oc = *reinterpret_cast<const float4*>(occ + v);
        g[0] = scale * (ex2_(fmaf(x.x, LOG2E, -lse2)) - oc.x);
        g[1] = scale * (ex2_(fmaf(x.y, LOG2E, -lse2)) - oc.y);
        g[2] = scale * (ex2_(fmaf(x.z, LOG2E, -lse2)) - oc.z);
        g[3] = scale * (ex2_(fmaf(x.w, LOG2E, -lse2)) - oc.w);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) g[j] = v + j < V ? scale * (ex2_(fmaf(row[v + j], LOG2E, -lse2)) - occ[v + j]) : 0.f;
      }
      if (g32 != nullptr) *reinterpret_cast<float4*>(g32 + r * ldg + v) = make_float4(g[0], g[1], g[2], g[3]);
      if (g16 != nullptr) *reinterpret_cast<uint2*>(g16 + r * ldg + v) = make_uint2(pack_bf16x2(g[0], g[1]), pack_bf16x2(g[2], g[3]));
    }
    __syncwarp();  // the row is read by every lane before the next frame clears it
  }
}

long long ws_floats(int B, int T, int Lmax) {
  const long long Smax = 2LL * Lmax + 1;
  return 3LL * B * T * Smax + (long long)B * T + B + 8;  // em, alpha, beta, lse, nll_raw (+ pad)
}

}  // namespace

extern "C" int b200slu_ctc_lattice_states(const int* labels, int Lmax, const int* label_lens, int B, int blank, int* ext_out,
                                          unsigned char* skip_out, void* stream) {
  B2_CHECK_ARG(B > 0 && Lmax >= 0 && ext_out != nullptr && skip_out != nullptr, "ctc_lattice_states: bad arguments");
  ctc_lattice_states_kernel<<<B, 128, 0, (cudaStream_t)stream>>>(labels, Lmax > 0 ? Lmax : 1, label_lens, blank, 2 * Lmax + 1,
                                                                 ext_out, skip_out);
  B2_LAUNCH_CHECK();
  return 0;
}

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
  static const int no_warp = getenv("B200SLU_CTC_NO_WARP") != nullptr;  // A/B: the block-per-frame kernels at any V
  if (!no_warp && V <= CTC_WARP_V && ldv % 4 == 0 && ((uintptr_t)logits % 16) == 0) {
    const long long nframes = (long long)B * T;
    const int grid = b2_grid_for(nframes * 32, 32 * CTC_WPB, 8);
    ctc_prep_warp_kernel<<<grid, 32 * CTC_WPB, 0, st>>>(logits, ldv, T, V, hlens, labels, Lmax > 0 ? Lmax : 1, label_lens, blank, lse,
                                                        em, Smax, nframes);
  } else {
    ctc_prep_kernel<<<dim3(T, B), 256, 0, st>>>(logits, ldv, T, V, hlens, labels, Lmax > 0 ? Lmax : 1, label_lens, blank, lse, em,
                                                Smax);
  }
  B2_LAUNCH_CHECK();
  // D = frames of emission prefetch: the recursion is a latency chain (~170 cycles per frame at C = 2), the emission
  // rows come from L2 / HBM (~1000-2500 cycles away once the logits have streamed through L2), so D * frame time has
  // to cover that latency.
  static const int force_d = getenv("B200SLU_CTC_D") ? atoi(getenv("B200SLU_CTC_D")) : 0;  // tuning experiments only
#define LAT2(C, D)                                                                                                          \
  ctc_lattice_kernel<C, D><<<B, 64, 0, st>>>(em, T, Smax, hlens, labels, Lmax > 0 ? Lmax : 1, label_lens, blank, alpha, beta, \
                                             nll_raw)
#define LAT(C, DDEF)                                   \
  do {                                                 \
    const int d_ = force_d > 0 ? force_d : DDEF;       \
    if (d_ >= 16 && C <= 4) LAT2(C, 16);               \
    else if (d_ >= 8 && C <= 8) LAT2(C, 8);            \
    else if (d_ >= 4) LAT2(C, 4);                      \
    else LAT2(C, 2);                                   \
  } while (0)
  static const int no_mw = getenv("B200SLU_CTC_NO_MW") != nullptr;  // tuning experiments only
#define LATMW(C, D, W)                                                                                                       \
  ctc_lattice_mw_kernel<C, D, W><<<B, 64 * W, 0, st>>>(em, T, Smax, hlens, labels, Lmax > 0 ? Lmax : 1, label_lens, blank, alpha, \
                                                       beta, nll_raw)
  if (Smax <= 64) LAT(2, 16);
  else if (Smax <= 128) LAT(4, 8);
  else if (no_mw && Smax <= 256) LAT(8, 4);
  else if (no_mw) LAT(20, 2);
  else if (Smax <= 256) LATMW(4, 8, 2);
  else LATMW(5, 8, 4);
#undef LATMW
#undef LAT
#undef LAT2
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
  static unsigned long long cfg = 0;
  if (!b2_device_done(cfg)) {
    B2_CUDA(cudaFuncSetAttribute(ctc_grad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    b2_device_mark(cfg);
  }
  B2_SCOPE(FAM_CTC, st, 0, (double)B * T * V * (4.0 + (grad_f32 ? 4.0 : 0.0) + (grad_bf16 ? 2.0 : 0.0)) + 12.0 * n);
  const bool vec4 = ldv % 4 == 0 && ldg % 4 == 0 && ((uintptr_t)logits % 16) == 0 &&
                    (grad_f32 == nullptr || ((uintptr_t)grad_f32 % 16) == 0) && (grad_bf16 == nullptr || ((uintptr_t)grad_bf16 % 8) == 0);
  static const int no_warp = getenv("B200SLU_CTC_NO_WARP") != nullptr;
  if (!no_warp && V <= CTC_WARP_V_GRAD && vec4) {
    static unsigned long long cfgw = 0;
    if (!b2_device_done(cfgw)) {
      B2_CUDA(cudaFuncSetAttribute(ctc_grad_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CTC_WPB * CTC_WARP_V_GRAD * 4));
      b2_device_mark(cfgw);
    }
    const long long nframes = (long long)B * T;
    const int Vs = (V + 3) / 4 * 4;
    const size_t smem = (size_t)CTC_WPB * Vs * 4;
    const int per_sm = smem <= 16 * 1024 ? 8 : 6;
    const int grid = b2_grid_for(nframes * 32, 32 * CTC_WPB, per_sm);
    ctc_grad_warp_kernel<<<grid, 32 * CTC_WPB, smem, st>>>(logits, ldv, T, V, hlens, labels, Lmax > 0 ? Lmax : 1, label_lens, blank, lse,
                                                           em, alpha, beta, Smax, feasible, scale, grad_f32, (bf16*)grad_bf16, ldg,
                                                           nframes, Vs);
  } else {
    ctc_grad_kernel<<<dim3(T, B), 256, (size_t)V * 4, st>>>(logits, ldv, T, V, hlens, labels, Lmax > 0 ? Lmax : 1, label_lens, blank,
                                                            lse, em, alpha, beta, Smax, feasible, scale, grad_f32, (bf16*)grad_bf16,
                                                            ldg, vec4);
  }
  B2_LAUNCH_CHECK();
  return 0;
}
