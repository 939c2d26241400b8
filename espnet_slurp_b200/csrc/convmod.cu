// Conformer ConvolutionModule middle: GLU -> depthwise Conv1d(K, same padding) -> BatchNorm1d (training:
// batch statistics over ALL B*T positions, padded frames included) -> Swish, forward and backward.
//
// Replaces espnet/nets/pytorch_backend/conformer/convolution.py:67-79 between the two pointwise convs
// (which run on the tcgen05 GEMM).  Channels-last [B, T, C] throughout, so the reference's two
// transposes (:67, :79) disappear.  Memory-bound; shared-memory staging of a (TT + K - 1) x 64 halo
// tile, 16-byte vector loads, per-channel statistics reduced in shared memory then one atomic per
// channel per block.  Algorithmic bytes per position (bf16): fwd pass A 2*(2C) + 2*C, pass B 2*C + 2*C.
#include <stdlib.h>

#include "common.cuh"

namespace {

constexpr int CC = 64;    // channels per block
constexpr int KMAX = 31;  // cnn_module_kernel upper bound (SLURP/LibriSpeech: 31; tiny fixtures: 7)
// Output frames per block are a template parameter TT (multiple of 32): the K - 1 halo frames are staged (and their GLU
// recomputed) once per TT outputs, so a longer tile cuts both the redundant loads and the per-block weight fetch.
// KT = compile-time kernel size (31 for every recipe; 0 = run-time K <= KMAX for the small test fixtures): with KT the
// tap loops carry no run-time predicates (they were 70 % of the instructions of the first version).

// Stage glu(pw1) -- and for the backward d_dw -- for frames [t0 - pad, t0 + TT + pad) x channels [c0, c0 + CC) into
// shared memory as fp32 (0 outside [0, T)).  All of a thread's 16-byte loads are issued before the first one is used:
// a tile costs one memory round trip, not one per item (the staging loop used to be a third of the backward's time).
template <int TT, bool BWD>
__device__ __forceinline__ void stage_tiles(const bf16* __restrict__ pw1, const bf16* __restrict__ d_dw, int b, int T, int C,
                                            int t0, int c0, int K, float (*sg)[CC], float (*sd)[CC]) {
  constexpr int UT = ((TT + KMAX - 1) * (CC / 8) + 255) / 256;  // items per thread of a 256-thread block
  constexpr int U = UT > 5 ? (UT + 1) / 2 : UT;                 // ... in two rounds when that is many registers in flight
  const int pad = (K - 1) / 2, nitems = (TT + K - 1) * (CC / 8);
#pragma unroll 1
  for (int base = 0; base < UT * 256; base += U * 256) {
  uint4 ra[U], rg[U], rd[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int it = base + threadIdx.x + u * 256;
    const int f = it / (CC / 8), cg = it - f * (CC / 8);
    const int t = t0 - pad + f, c = c0 + cg * 8;
    ra[u] = rg[u] = rd[u] = make_uint4(0u, 0u, 0u, 0u);
    if (it < nitems && t >= 0 && t < T && c < C) {
      const bf16* row = pw1 + ((long long)b * T + t) * (2 * C);
      ra[u] = *reinterpret_cast<const uint4*>(row + c);
      rg[u] = *reinterpret_cast<const uint4*>(row + C + c);
      if (BWD) rd[u] = *reinterpret_cast<const uint4*>(d_dw + ((long long)b * T + t) * C + c);
    }
  }
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int it = base + threadIdx.x + u * 256;
    if (it < nitems) {
      const int f = it / (CC / 8), cg = it - f * (CC / 8);
      float a[8], gt[8], g[8];
      unpack8_bf16(ra[u], a);
      unpack8_bf16(rg[u], gt);
#pragma unroll
      for (int q = 0; q < 8; ++q) g[q] = a[q] * sigmoidf_(gt[q]);  // 0 * sigmoid(0) = 0 outside the utterance
      // a quarter-warp stores 8 x 16 B at a 32-byte stride: lanes cg 4..7 write their upper half first so that each
      // store instruction touches 32 distinct banks (ncu showed a 2-way conflict on these stores)
      const int h = (cg >> 2) & 1;
      float4* dst = reinterpret_cast<float4*>(&sg[f][cg * 8]);
      const float4 lo = make_float4(g[0], g[1], g[2], g[3]), hi = make_float4(g[4], g[5], g[6], g[7]);
      dst[h] = h ? hi : lo;
      dst[h ^ 1] = h ? lo : hi;
      if (BWD) {
        unpack8_bf16(rd[u], g);
        float4* dd = reinterpret_cast<float4*>(&sd[f][cg * 8]);
        const float4 dlo = make_float4(g[0], g[1], g[2], g[3]), dhi = make_float4(g[4], g[5], g[6], g[7]);
        dd[h] = h ? dhi : dlo;
        dd[h ^ 1] = h ? dlo : dhi;
      }
    }
  }
  }
}

// Asynchronous form of the forward's staging: the raw bf16 halves (a | gate) of the NEXT tile are copied global -> shared
// with cp.async (no registers held, zero-fill outside [0, T)) while the current tile is being computed; a conversion pass
// then turns them into the fp32 GLU tile.  One HBM round trip per tile leaves the critical path.
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
template <int TT>
__device__ __forceinline__ void issue_raw_tile(const bf16* __restrict__ pw1, int b, int T, int C, int t0, int c0, int K,
                                               uint4* raw_a, uint4* raw_g) {
  constexpr int UT = ((TT + KMAX - 1) * (CC / 8) + 255) / 256;
  const int pad = (K - 1) / 2, nitems = (TT + K - 1) * (CC / 8);
#pragma unroll
  for (int u = 0; u < UT; ++u) {
    const int it = threadIdx.x + u * 256;
    if (it < nitems) {
      const int f = it / (CC / 8), cg = it - f * (CC / 8);
      const int t = t0 - pad + f, c = c0 + cg * 8;
      const bool ok = t >= 0 && t < T && c < C;
      const bf16* row = pw1 + ((long long)b * T + (ok ? t : 0)) * (2 * C) + (ok ? c : 0);
      cp_async16((uint32_t)__cvta_generic_to_shared(raw_a + it), row, ok ? 16 : 0);
      cp_async16((uint32_t)__cvta_generic_to_shared(raw_g + it), row + C, ok ? 16 : 0);
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int TT>
__device__ __forceinline__ void convert_raw_tile(const uint4* raw_a, const uint4* raw_g, int K, float (*sg)[CC]) {
  constexpr int UT = ((TT + KMAX - 1) * (CC / 8) + 255) / 256;
  const int nitems = (TT + K - 1) * (CC / 8);
#pragma unroll
  for (int u = 0; u < UT; ++u) {
    const int it = threadIdx.x + u * 256;
    if (it < nitems) {
      const int f = it / (CC / 8), cg = it - f * (CC / 8);
      float a[8], gt[8], g[8];
      unpack8_bf16(raw_a[it], a);
      unpack8_bf16(raw_g[it], gt);
#pragma unroll
      for (int q = 0; q < 8; ++q) g[q] = a[q] * sigmoidf_(gt[q]);  // 0 * sigmoid(0) = 0 outside the utterance
      const int h = (cg >> 2) & 1;  // see stage_tiles: every store instruction touches 32 distinct banks
      float4* dst = reinterpret_cast<float4*>(&sg[f][cg * 8]);
      const float4 lo = make_float4(g[0], g[1], g[2], g[3]), hi = make_float4(g[4], g[5], g[6], g[7]);
      dst[h] = h ? hi : lo;
      dst[h ^ 1] = h ? lo : hi;
    }
  }
}

// Each thread owns a PAIR of neighbouring channels and, per sub-tile, FR consecutive output frames: the K + FR - 1 staged
// inputs it needs are read from shared memory once (one 8-byte load per frame) into a sliding window, the taps of both
// channels stay in registers as pairs and every multiply-add is a packed fma.rn.f32x2 (FFMA2: the FMA pipe's rate of
// scalar code at half the issue slots -- tools/ffma2_rate.cu -- which is what these kernels ran out of: ncu had them at
// IPC 2.5 with the FMA pipe 39 % busy).  Block = 32 channel pairs x 8 frame lanes.
template <int TT>
struct DwGeom {
  static constexpr int FR = (TT % 64 == 0) ? 8 : 4;   // frames per thread and sub-tile
  static constexpr int SUB = 8 * FR;                  // frames per sub-tile (8 frame lanes)
  static_assert(TT % SUB == 0, "tile = whole sub-tiles");
};

template <int TT, int KT>
__global__ void __launch_bounds__(256, 2) glu_dw_fwd_kernel(const bf16* __restrict__ pw1, const float* __restrict__ w,
                                                            const float* __restrict__ bias, int B, int T, int C, int Krt,
                                                            bf16* __restrict__ out, float* __restrict__ sums) {
  constexpr int FR = DwGeom<TT>::FR, SUB = DwGeom<TT>::SUB;
  b2_pdl_entry();
  extern __shared__ __align__(16) float dyn_smem[];
  float (*sm)[CC] = reinterpret_cast<float (*)[CC]>(dyn_smem);                           // glu halo tile [TT+KMAX-1][CC]
  float (*red)[8][CC] = reinterpret_cast<float (*)[8][CC]>(dyn_smem + (TT + KMAX - 1) * CC);  // [2][8 frame lanes][CC]
  uint4* raw_a = reinterpret_cast<uint4*>(dyn_smem + (TT + KMAX - 1) * CC + 2 * 8 * CC);      // raw bf16 a half, next tile
  uint4* raw_g = raw_a + (TT + KMAX - 1) * (CC / 8);                                          // ... gate half
  const int K = KT > 0 ? KT : Krt;
  const int c0 = blockIdx.x * CC;
  const int cp = threadIdx.x & 31, fr = threadIdx.x >> 5;  // channel pair, frame lane
  const int c = c0 + 2 * cp;
  const bool cok = c < C;  // C is even: the pair is in or out together
  // The block's 64 x K taps are contiguous in w: fetched coalesced through shared memory (the halo tile's space).  A
  // per-thread gather w[c * K + k] touches 32 cache lines per load instruction -- 16 k line requests per block, which took
  // longer than the block's two tiles of arithmetic.
  {
    float* wt = dyn_smem;
    const int nw = min(CC, C - c0) * K;
    constexpr int WU = (CC * KMAX + 455) / 256;
    float tmp[WU];
#pragma unroll
    for (int u = 0; u < WU; ++u) {  // all loads in flight before the first store
      const int it = threadIdx.x + u * 256;
      tmp[u] = it < nw ? w[(long long)c0 * K + it] : 0.f;
    }
#pragma unroll
    for (int u = 0; u < WU; ++u) {
      const int it = threadIdx.x + u * 256;
      if (it < nw) wt[it] = tmp[u];
    }
    __syncthreads();
  }
  float2 wr[KMAX];
#pragma unroll
  for (int k = 0; k < KMAX; ++k)
    wr[k] = (cok && k < K) ? make_float2(dyn_smem[(2 * cp) * K + k], dyn_smem[(2 * cp + 1) * K + k]) : make_float2(0.f, 0.f);
  const float2 bs = cok ? make_float2(bias[c], bias[c + 1]) : make_float2(0.f, 0.f);
  const int ntt = (T + TT - 1) / TT;
  float2 s1 = make_float2(0.f, 0.f), s2 = make_float2(0.f, 0.f);
  // Each block walks its (utterance, time tile) units and keeps the BatchNorm partial sums in registers across them:
  // the atomics on `sums` are per block (grid.y per channel), not per tile -- same-address atomics retire serially.
  if ((int)blockIdx.y < B * ntt) issue_raw_tile<TT>(pw1, blockIdx.y / ntt, T, C, (blockIdx.y % ntt) * TT, c0, K, raw_a, raw_g);
  for (int unit = blockIdx.y; unit < B * ntt; unit += gridDim.y) {
    const int b = unit / ntt, t0 = (unit - b * ntt) * TT;
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();  // this tile's raw halves have landed (every thread's copies); the previous tile is fully consumed
    convert_raw_tile<TT>(raw_a, raw_g, K, sm);
    __syncthreads();
    {  // the raw buffers are free again: the next tile's copies fly under this tile's arithmetic
      const int nx = unit + gridDim.y;
      if (nx < B * ntt) issue_raw_tile<TT>(pw1, nx / ntt, T, C, (nx % ntt) * TT, c0, K, raw_a, raw_g);
    }
    if (cok) {
#pragma unroll 1
      for (int sub = 0; sub < TT / SUB; ++sub) {
        const int f0 = sub * SUB + fr * FR;
        if (t0 + f0 >= T) break;
        float2 acc[FR];
#pragma unroll
        for (int o = 0; o < FR; ++o) acc[o] = bs;
#pragma unroll
        for (int j = 0; j < FR + KMAX - 1; ++j) {
          if (KT > 0 ? (j < FR + KT - 1) : (j < FR + K - 1)) {
            const float2 v = *reinterpret_cast<const float2*>(&sm[f0 + j][2 * cp]);
#pragma unroll
            for (int o = 0; o < FR; ++o)
              if (j - o >= 0 && j - o < (KT > 0 ? KT : KMAX)) acc[o] = b2_fma2(wr[j - o], v, acc[o]);  // wr[k] = 0 for k >= K
          }
        }
#pragma unroll
        for (int o = 0; o < FR; ++o) {
          const int t = t0 + f0 + o;
          if (t < T) {
            *reinterpret_cast<uint32_t*>(out + ((long long)b * T + t) * C + c) = pack_bf16x2(acc[o].x, acc[o].y);
            s1.x += acc[o].x;
            s1.y += acc[o].y;
            s2 = b2_fma2(acc[o], acc[o], s2);
          }
        }
      }
    }
  }
  *reinterpret_cast<float2*>(&red[0][fr][2 * cp]) = s1;
  *reinterpret_cast<float2*>(&red[1][fr][2 * cp]) = s2;
  __syncthreads();
  if (threadIdx.x < CC && c0 + threadIdx.x < C && sums != nullptr) {
    float a = 0.f, q = 0.f;
#pragma unroll
    for (int l = 0; l < 8; ++l) {
      a += red[0][l][threadIdx.x];
      q += red[1][l][threadIdx.x];
    }
    atomicAdd(sums + c0 + threadIdx.x, a);
    atomicAdd(sums + C + c0 + threadIdx.x, q);
  }
}

// (One channel per thread, scalar FMAs: the channel-pair / FFMA2 form that helps the forward was measured SLOWER here --
// 87.7 against 81.5 us -- its two passes per tile re-walk the sub-tiles and the kernel is bound by the staging round trips
// and barriers, not by FMA issue.)
constexpr int FR1 = 8;
template <int TT>
struct DwBwdGeom {
  static constexpr int NP = TT <= 96 ? 4 : 1;  // private copies of the block's dW / bias partials (one per frame lane)
};
// Backward: grid (channel tiles, utterance chunks); each block walks all its (utterance, time tile) pairs and keeps the
// depthwise weight gradient in registers across them, so the global atomics are per block, not per tile.
template <int TT, int KT>
__global__ void __launch_bounds__(256) glu_dw_bwd_kernel(const bf16* __restrict__ d_dw, const bf16* __restrict__ pw1,
                                                         const float* __restrict__ w, int B, int T, int C, int Krt,
                                                         bf16* __restrict__ d_pw1, float* __restrict__ dw_accum,
                                                         float* __restrict__ dbias_accum, float* __restrict__ dpw1_colsum) {
  b2_pdl_entry();
  extern __shared__ __align__(16) float dyn_smem[];
  float (*sg)[CC] = reinterpret_cast<float (*)[CC]>(dyn_smem);                             // glu halo tile  [TT+KMAX-1][CC]
  float (*sd)[CC] = reinterpret_cast<float (*)[CC]>(dyn_smem + (TT + KMAX - 1) * CC);      // d_dw halo tile
  // dW partial (+ dbias); rows padded to an odd stride so that the 32 channels of a warp hit 32 different banks.  NP = 4: one
  // private copy per frame lane, so a tile's partials join the block's with plain read-modify-writes -- as shared-memory
  // atomics (a compare-and-swap loop per float, four lanes per address) that flush was 11 % of the kernel's instructions
  // and 28 % of its stall samples.  The 128-frame tile has no room for the copies (two blocks per SM) and keeps the atomics.
  constexpr int NP = DwBwdGeom<TT>::NP;
  float (*swg)[KMAX + 4] = reinterpret_cast<float (*)[KMAX + 4]>(dyn_smem + 2 * (TT + KMAX - 1) * CC);  // [NP * CC][KMAX + 4]
  // reversed taps [KMAX][CC]: swr[m][ch] = w[ch][K-1-m] (0 for m >= K); re-read per tile so that the staging phase has the
  // whole register file for its loads in flight
  float (*swr)[CC] = reinterpret_cast<float (*)[CC]>(dyn_smem + 2 * (TT + KMAX - 1) * CC + NP * CC * (KMAX + 4));
  const int K = KT > 0 ? KT : Krt;
  const int c0 = blockIdx.x * CC;
  const int pad = (K - 1) / 2;
  const int cl = threadIdx.x & (CC - 1), fr = threadIdx.x >> 6;
  const int c = c0 + cl;
  const int ntt = (T + TT - 1) / TT;
  for (int it = threadIdx.x; it < NP * CC * (KMAX + 4); it += blockDim.x) (&swg[0][0])[it] = 0.f;
  for (int it = threadIdx.x; it < KMAX * CC; it += blockDim.x) {
    const int m = it / CC, ch = it - m * CC;
    swr[m][ch] = (c0 + ch < C && m < K) ? w[(c0 + ch) * K + (K - 1 - m)] : 0.f;
  }
  for (int unit = blockIdx.y; unit < B * ntt; unit += gridDim.y) {
    const int b = unit / ntt, t0 = (unit - b * ntt) * TT;
    __syncthreads();  // previous tile fully consumed (and, the first time, swg / swr written)
    stage_tiles<TT, true>(pw1, d_dw, b, T, C, t0, c0, K, sg, sd);
    __syncthreads();
    if (c < C) {
     float wrev[KMAX], acc[KMAX];  // wrev[m] = w[K-1-m]: keeps every register-array index a compile-time constant
#pragma unroll
     for (int k = 0; k < KMAX; ++k) {
       wrev[k] = swr[k][cl];
       acc[k] = 0.f;
     }
     float accb = 0.f, acca = 0.f, accg = 0.f;  // column sums of d_dw (depthwise bias) and of the two halves of d_pw1 (pw1's bias)
#pragma unroll 1
     for (int sub = 0; sub < TT / (4 * FR1); ++sub) {
      const int f0 = sub * 4 * FR1 + fr * FR1;
      if (t0 + f0 >= T) break;
      float dg[FR1], dv[FR1];
      // the GLU inputs of the thread's own frames (for d_a, d_gate) are fetched now; their latency hides under the taps
      bf16 pa[FR1], pg[FR1];
#pragma unroll
      for (int o = 0; o < FR1; ++o) {
        const int t = min(t0 + f0 + o, T - 1);
        const bf16* row = pw1 + ((long long)b * T + t) * (2 * C);
        pa[o] = row[c];
        pg[o] = row[C + c];
      }
#pragma unroll
      for (int o = 0; o < FR1; ++o) {
        dg[o] = 0.f;
        dv[o] = sd[f0 + o + pad][cl];  // d_dw at the thread's own frames (0 beyond T by staging)
        accb += dv[o];
      }
#pragma unroll
      for (int j = 0; j < FR1 + KMAX - 1; ++j) {
        if (KT > 0 ? (j < FR1 + KT - 1) : (j < FR1 + K - 1)) {
          // d_glu[f0+o] = sum_k w[k] * d_dw[f0 + o - k + pad]  = sum over staged row j = o + (K-1) - k
          const float dj = sd[f0 + j][cl];
          const float gj = sg[f0 + j][cl];
#pragma unroll
          for (int o = 0; o < FR1; ++o) {
            // forward tap index of staged row j for output o: k = j - o (glu[f0+o+k] feeds out[f0+o])
            if (j - o >= 0 && j - o < (KT > 0 ? KT : KMAX)) acc[j - o] += dv[o] * gj;
          }
#pragma unroll
          for (int o = 0; o < FR1; ++o) {
            // backward: d_dw[f0+o - k + pad] sits in staged row j = o + (K-1) - k  ->  w[k] = wrev[j - o]
            if (j - o >= 0 && j - o < (KT > 0 ? KT : KMAX)) dg[o] += wrev[j - o] * dj;
          }
        }
      }
#pragma unroll
      for (int o = 0; o < FR1; ++o) {
        const int t = t0 + f0 + o;
        if (t < T) {
          const float a = __bfloat162float(pa[o]), gt = __bfloat162float(pg[o]);
          const float sgm = sigmoidf_(gt);
          bf16* drow = d_pw1 + ((long long)b * T + t) * (2 * C);
          const float da = dg[o] * sgm, dgate = dg[o] * a * sgm * (1.f - sgm);
          drow[c] = __float2bfloat16(da);
          drow[C + c] = __float2bfloat16(dgate);
          acca += da;
          accg += dgate;
        }
      }
     }
     // this tile's weight / bias gradient partials join the block's
     if (NP == 4) {
       float* mine = swg[fr * CC + cl];  // this thread's own row: no other thread touches it
#pragma unroll
       for (int k = 0; k < KMAX; ++k)
         if (k < K) mine[k] += acc[k];
       mine[KMAX] += accb;
       mine[KMAX + 1] += acca;
       mine[KMAX + 2] += accg;
     } else {  // 4 frame lanes per address
#pragma unroll
       for (int k = 0; k < KMAX; ++k)
         if (k < K) atomicAdd(&swg[cl][k], acc[k]);
       atomicAdd(&swg[cl][KMAX], accb);
       if (dpw1_colsum != nullptr) {
         atomicAdd(&swg[cl][KMAX + 1], acca);
         atomicAdd(&swg[cl][KMAX + 2], accg);
       }
     }
    }
  }
  __syncthreads();
  auto total = [&](int ch, int k) {
    float v = swg[ch][k];
#pragma unroll
    for (int q = 1; q < NP; ++q) v += swg[q * CC + ch][k];
    return v;
  };
  for (int it = threadIdx.x; it < CC * K; it += blockDim.x) {
    const int ch = it / K, k = it - ch * K;
    if (c0 + ch < C) atomicAdd(dw_accum + (long long)(c0 + ch) * K + k, total(ch, k));
  }
  if (threadIdx.x < CC && c0 + threadIdx.x < C) {
    atomicAdd(dbias_accum + c0 + threadIdx.x, total(threadIdx.x, KMAX));
    if (dpw1_colsum != nullptr) {  // bias gradient of pointwise_conv1 = column sums of d_pw1: no separate pass over it
      atomicAdd(dpw1_colsum + c0 + threadIdx.x, total(threadIdx.x, KMAX + 1));
      atomicAdd(dpw1_colsum + C + c0 + threadIdx.x, total(threadIdx.x, KMAX + 2));
    }
  }
}

// BatchNorm1d training statistics: mean / rstd from the fp32 sums, running-stat update (momentum,
// unbiased variance) exactly as torch.nn.BatchNorm1d does (convolution.py:45).
__global__ void bn_finalize_kernel(const float* __restrict__ sums, int C, float n, float eps, float momentum,
                                   float* __restrict__ mean_rstd, float* __restrict__ running_mean,
                                   float* __restrict__ running_var, long long* __restrict__ num_batches) {
  b2_pdl_entry();  // programmatic dependent launch: see common.cuh
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < C; c += gridDim.x * blockDim.x) {
    const float mean = sums[c] / n;
    const float var = fmaxf(sums[C + c] / n - mean * mean, 0.f);
    mean_rstd[c] = mean;
    mean_rstd[C + c] = rsqrtf(var + eps);
    if (running_mean != nullptr) {
      running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
      running_var[c] = (1.f - momentum) * running_var[c] + momentum * var * (n / fmaxf(n - 1.f, 1.f));
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && num_batches != nullptr) *num_batches += 1;
}

// A thread owns one group of 8 channels (its BatchNorm constants live in registers, folded to one FMA per element) and
// walks rows with BN_U independent 16-byte loads in flight; block = 64 channel groups x 4 row lanes, grid (C/512, rows).
constexpr int BN_U = 4;
__global__ void __launch_bounds__(256) bn_swish_fwd_kernel(const bf16* __restrict__ x, const float* __restrict__ mean_rstd,
                                                           const float* __restrict__ gamma, const float* __restrict__ beta,
                                                           long long M, int C, bf16* __restrict__ y) {
  b2_pdl_entry();  // programmatic dependent launch: see common.cuh
  const int cg = blockIdx.x * 64 + (threadIdx.x & 63), rl = threadIdx.x >> 6;
  if (cg * 8 >= C) return;
  const int c = cg * 8;
  float sc[8], sh[8];
  {
    float mu[8], rs[8], ga[8], be[8];
    load8_f32(mean_rstd + c, mu);
    load8_f32(mean_rstd + C + c, rs);
    load8_f32(gamma + c, ga);
    load8_f32(beta + c, be);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      sc[j] = rs[j] * ga[j];
      sh[j] = be[j] - mu[j] * sc[j];
    }
  }
  const long long step = (long long)gridDim.y * 4;
  for (long long r0 = (long long)blockIdx.y * 4 + rl; r0 < M; r0 += BN_U * step) {
    uint4 raw[BN_U];
#pragma unroll
    for (int u = 0; u < BN_U; ++u) {
      const long long r = r0 + u * step;
      raw[u] = r < M ? *reinterpret_cast<const uint4*>(x + r * C + c) : make_uint4(0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int u = 0; u < BN_U; ++u) {
      const long long r = r0 + u * step;
      if (r < M) {
        const uint32_t w4[4] = {raw[u].x, raw[u].y, raw[u].z, raw[u].w};
        float v[8];
#pragma unroll
        for (int q = 0; q < 4; ++q) {  // packed fp32 on channel pairs: half the issue slots of the scalar form
          const float2 z = b2_fma2(unpack_bf16x2(w4[q]), make_float2(sc[2 * q], sc[2 * q + 1]), make_float2(sh[2 * q], sh[2 * q + 1]));
          const float2 y2 = b2_mul2(z, b2_sigmoid2(z));
          v[2 * q] = y2.x;
          v[2 * q + 1] = y2.y;
        }
        store8_bf16(y + r * C + c, v);
      }
    }
  }
}

// sums2[c] += sum dz ; sums2[C + c] += sum dz * xhat   with dz = dy * swish'(gamma * xhat + beta)
__global__ void __launch_bounds__(512, 1) bn_swish_bwd_reduce_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x,
                                                                     const float* __restrict__ mean_rstd,
                                                                     const float* __restrict__ gamma,
                                                                     const float* __restrict__ beta, int M, int C,
                                                                     float* __restrict__ sums2) {
  b2_pdl_entry();  // programmatic dependent launch: see common.cuh
  // narrow column strips (common.cuh): few blocks per output column, one vector atomic per four columns
  __shared__ float red[2][CS_ROWS][CS_COLS + 1];
  const int tx = threadIdx.x & (CS_TX - 1), ty = threadIdx.x / CS_TX;
  const int g = blockIdx.x * CS_TX + tx;
  float a1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, a2[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (g * 8 < C) {
    float mu[8], rs[8], ga[8], be[8];
    load8_f32(mean_rstd + g * 8, mu);
    load8_f32(mean_rstd + C + g * 8, rs);
    load8_f32(gamma + g * 8, ga);
    load8_f32(beta + g * 8, be);
    constexpr int U = 4;
    const long long step = (long long)gridDim.y * CS_ROWS;
    for (long long r0 = (long long)blockIdx.y * CS_ROWS + ty; r0 < M; r0 += U * step) {
      uint4 rd[U], rx[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long r = r0 + u * step;
        rd[u] = r < M ? *reinterpret_cast<const uint4*>(dy + r * C + g * 8) : make_uint4(0u, 0u, 0u, 0u);  // dy = 0: no contribution
        rx[u] = r < M ? *reinterpret_cast<const uint4*>(x + r * C + g * 8) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint32_t wd[4] = {rd[u].x, rd[u].y, rd[u].z, rd[u].w}, wx[4] = {rx[u].x, rx[u].y, rx[u].z, rx[u].w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {  // packed fp32 on channel pairs (the kernel is issue-bound, not bandwidth-bound)
          const float2 d2 = unpack_bf16x2(wd[q]), x2 = unpack_bf16x2(wx[q]);
          const float2 xh = b2_mul2(b2_add2(x2, make_float2(-mu[2 * q], -mu[2 * q + 1])), make_float2(rs[2 * q], rs[2 * q + 1]));
          const float2 z = b2_fma2(make_float2(ga[2 * q], ga[2 * q + 1]), xh, make_float2(be[2 * q], be[2 * q + 1]));
          const float2 dz = b2_mul2(d2, b2_swish_grad2(z));
          const float2 s1 = b2_add2(make_float2(a1[2 * q], a1[2 * q + 1]), dz);
          const float2 s2 = b2_fma2(dz, xh, make_float2(a2[2 * q], a2[2 * q + 1]));
          a1[2 * q] = s1.x; a1[2 * q + 1] = s1.y;
          a2[2 * q] = s2.x; a2[2 * q + 1] = s2.y;
        }
      }
    }
  }
  colsum_block_reduce(a1, red[0], tx, ty, C, sums2);
  colsum_block_reduce(a2, red[1], tx, ty, C, sums2 + C);
}

__global__ void __launch_bounds__(256, 2) bn_swish_bwd_apply_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x,
                                                                 const float* __restrict__ mean_rstd,
                                                                 const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                 const float* __restrict__ sums2, long long M, int C,
                                                                 bf16* __restrict__ dx, float* __restrict__ dgamma,
                                                                 float* __restrict__ dbeta) {
  b2_pdl_entry();  // programmatic dependent launch: see common.cuh
  const int cg = blockIdx.x * 64 + (threadIdx.x & 63), rl = threadIdx.x >> 6;
  if (cg * 8 < C) {
    const int c = cg * 8;
    const float invn = 1.f / (float)M;
    float mu[8], rs[8], ga[8], be[8], k1[8], k2[8];
    load8_f32(mean_rstd + c, mu);
    load8_f32(mean_rstd + C + c, rs);
    load8_f32(gamma + c, ga);
    load8_f32(beta + c, be);
    load8_f32(sums2 + c, k1);
    load8_f32(sums2 + C + c, k2);
    float gr[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      k1[j] *= invn;
      k2[j] *= invn;
      gr[j] = ga[j] * rs[j];
    }
    const long long step = (long long)gridDim.y * 4;
    for (long long r0 = (long long)blockIdx.y * 4 + rl; r0 < M; r0 += BN_U * step) {
      uint4 rd[BN_U], rx[BN_U];
#pragma unroll
      for (int u = 0; u < BN_U; ++u) {
        const long long r = r0 + u * step;
        rd[u] = r < M ? *reinterpret_cast<const uint4*>(dy + r * C + c) : make_uint4(0u, 0u, 0u, 0u);
        rx[u] = r < M ? *reinterpret_cast<const uint4*>(x + r * C + c) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int u = 0; u < BN_U; ++u) {
        const long long r = r0 + u * step;
        if (r < M) {
          const uint32_t wd[4] = {rd[u].x, rd[u].y, rd[u].z, rd[u].w}, wx[4] = {rx[u].x, rx[u].y, rx[u].z, rx[u].w};
          float o[8];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float2 d2 = unpack_bf16x2(wd[q]), x2 = unpack_bf16x2(wx[q]);
            const float2 xh = b2_mul2(b2_add2(x2, make_float2(-mu[2 * q], -mu[2 * q + 1])), make_float2(rs[2 * q], rs[2 * q + 1]));
            const float2 z = b2_fma2(make_float2(ga[2 * q], ga[2 * q + 1]), xh, make_float2(be[2 * q], be[2 * q + 1]));
            const float2 dz = b2_mul2(d2, b2_swish_grad2(z));
            // ga * rs * (dz - k1 - xh * k2)
            const float2 t = b2_fma2(xh, make_float2(-k2[2 * q], -k2[2 * q + 1]), b2_add2(dz, make_float2(-k1[2 * q], -k1[2 * q + 1])));
            const float2 o2 = b2_mul2(make_float2(gr[2 * q], gr[2 * q + 1]), t);
            o[2 * q] = o2.x;
            o[2 * q + 1] = o2.y;
          }
          store8_bf16(dx + r * C + c, o);
        }
      }
    }
  }
  if (blockIdx.x == 0 && blockIdx.y == 0) {
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      dbeta[c] += sums2[c];
      dgamma[c] += sums2[C + c];
    }
  }
}

// Frames per tile for K = 31: every tile stages TT + 30 frames and pays two block barriers, so pick the tile length that
// stages the fewest frames for this T (T = 124: one 128-frame tile, 158 staged, instead of two 96-frame tiles, 252).
// blocks per SM the grids are sized for (two are resident: 128 registers, <= 97 KB of shared memory)
int dw_blocks_per_sm() {
  static const int v = getenv("B200SLU_DW_BPS") ? atoi(getenv("B200SLU_DW_BPS")) : 2;  // tuning only
  return v > 0 ? v : 2;
}

int dw_tile_frames(int T, bool allow192) {
  static const int forced = getenv("B200SLU_DW_TT") ? atoi(getenv("B200SLU_DW_TT")) : 0;  // tuning only
  if (forced == 32 || forced == 96 || forced == 128 || (forced == 192 && allow192)) return forced;
  if (T <= 64) return 32;
  const int c96 = ((T + 95) / 96) * (96 + 30), c128 = ((T + 127) / 128) * (128 + 30), c192 = ((T + 191) / 192) * (192 + 30);
  if (allow192 && c192 <= c128 && c192 <= c96) return 192;  // forward only: the backward's two tiles would not fit twice per SM
  return c128 <= c96 ? 128 : 96;
}

}  // namespace

extern "C" int b200slu_glu_dwconv_fwd(const void* pw1, const float* w, const float* bias, int B, int T, int C, int K, void* out,
                                      float* bn_sums, void* stream) {
  B2_CHECK_ARG(B > 0 && T > 0 && C % 8 == 0 && K % 2 == 1 && K <= KMAX, "glu_dwconv: C %% 8 == 0, odd K <= %d (C=%d K=%d)", KMAX,
               C, K);
  B2_CHECK_ARG(B <= 65535, "glu_dwconv: B too large");
  cudaStream_t st = (cudaStream_t)stream;
  B2_SCOPE(FAM_CONVMOD, st, 2.0 * B * T * C * K, 6.0 * B * T * C);
  // tile length: 96 frames (halo overhead 1.31x) unless the utterances are short
#define B2_FWD(TT_, KT_)                                                                                              \
  do {                                                                                                                \
    const int ctiles = (C + CC - 1) / CC, units = B * ((T + TT_ - 1) / TT_);                                          \
    int gy = (dw_blocks_per_sm() * b2_num_sms() + ctiles - 1) / ctiles;                                               \
    if (gy > units) gy = units;                                                                                       \
    const size_t smem = (size_t)(2 * (TT_ + KMAX - 1) * CC + 2 * 8 * CC) * sizeof(float);                             \
    auto kern = glu_dw_fwd_kernel<TT_, KT_>;                                                                          \
    static unsigned long long cfg = 0;                                                                                \
    if (!b2_device_done(cfg)) {                                                                                       \
      B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));                    \
      b2_device_mark(cfg);                                                                                            \
    }                                                                                                                 \
    B2_CUDA(b2_launch(kern, dim3(ctiles, gy), dim3(256), smem, st, (const bf16*)pw1, w, bias, B, T, C, K, (bf16*)out, bn_sums)); \
  } while (0)
  if (K == 31) {
    // the tiling with the least staged frames
    const int tt = dw_tile_frames(T, false);
    if (tt == 128) B2_FWD(128, 31);
    else if (tt == 96) B2_FWD(96, 31);
    else B2_FWD(32, 31);
  } else {
    B2_FWD(32, 0);
  }
#undef B2_FWD
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_glu_dwconv_bwd_b(const void* d_out, const void* pw1, const float* w, int B, int T, int C, int K, void* d_pw1,
                                        float* dw_accum, float* dbias_accum, float* dpw1_colsum_accum, void* stream);
extern "C" int b200slu_glu_dwconv_bwd(const void* d_out, const void* pw1, const float* w, int B, int T, int C, int K, void* d_pw1,
                                      float* dw_accum, float* dbias_accum, void* stream) {
  return b200slu_glu_dwconv_bwd_b(d_out, pw1, w, B, T, C, K, d_pw1, dw_accum, dbias_accum, nullptr, stream);
}

extern "C" int b200slu_glu_dwconv_bwd_b(const void* d_out, const void* pw1, const float* w, int B, int T, int C, int K, void* d_pw1,
                                        float* dw_accum, float* dbias_accum, float* dpw1_colsum_accum, void* stream) {
  B2_CHECK_ARG(B > 0 && T > 0 && C % 8 == 0 && K % 2 == 1 && K <= KMAX, "glu_dwconv_bwd: C %% 8 == 0, odd K <= %d", KMAX);
  B2_CHECK_ARG(B <= 65535, "glu_dwconv_bwd: B too large");
  cudaStream_t st = (cudaStream_t)stream;
  B2_SCOPE(FAM_CONVMOD, st, 4.0 * B * T * C * K, 10.0 * B * T * C);
#define B2_BWD(TT_, KT_)                                                                                                     \
  do {                                                                                                                       \
    const int ctiles = (C + CC - 1) / CC, units = B * ((T + TT_ - 1) / TT_);                                                 \
    int gy = (dw_blocks_per_sm() * b2_num_sms() + ctiles - 1) / ctiles;                                                      \
    if (gy > units) gy = units;                                                                                              \
    const size_t smem = (size_t)(2 * (TT_ + KMAX - 1) * CC + DwBwdGeom<TT_>::NP * CC * (KMAX + 4) + KMAX * CC) * sizeof(float);          \
    auto kern = glu_dw_bwd_kernel<TT_, KT_>;                                                                                 \
    static unsigned long long cfg = 0;                                                                                                 \
    if (!b2_device_done(cfg)) {                                                                                                              \
      B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));                           \
      b2_device_mark(cfg);                                                                                                            \
    }                                                                                                                        \
    B2_CUDA(b2_launch(kern, dim3(ctiles, gy), dim3(256), smem, st, (const bf16*)d_out, (const bf16*)pw1, w, B, T, C, K,       \
                      (bf16*)d_pw1, dw_accum, dbias_accum, dpw1_colsum_accum));                                              \
  } while (0)
  if (K == 31) {
    const int tt = dw_tile_frames(T, false);
    if (tt == 128) B2_BWD(128, 31);
    else if (tt == 96) B2_BWD(96, 31);
    else B2_BWD(32, 31);
  } else {
    B2_BWD(32, 0);
  }
#undef B2_BWD
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_bn_finalize(const float* sums, int C, long long n, float eps, float momentum, float* mean_rstd,
                                   float* running_mean, float* running_var, long long* num_batches_tracked, void* stream) {
  B2_CHECK_ARG(C > 0 && n > 0, "bn_finalize: empty");
  cudaStream_t st = (cudaStream_t)stream;
  B2_SCOPE(FAM_CONVMOD, st, 0, 24.0 * C);
  B2_CUDA(b2_launch(bn_finalize_kernel, dim3((C + 255) / 256), dim3(256), 0, st, sums, C, (float)n, eps, momentum, mean_rstd,
                    running_mean, running_var, num_batches_tracked));
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_bn_swish_fwd(const void* x, const float* mean_rstd, const float* gamma, const float* beta, long long M,
                                    int C, void* y, void* stream) {
  B2_CHECK_ARG(M > 0 && C % 8 == 0, "bn_swish_fwd: C %% 8 == 0");
  cudaStream_t st = (cudaStream_t)stream;
  B2_SCOPE(FAM_CONVMOD, st, 0, 4.0 * M * C);
  {
    dim3 grid((C + 511) / 512, 1);
    const long long gy = (M + 4 * BN_U - 1) / (4 * BN_U);
    const long long cap = (8LL * b2_num_sms() + grid.x - 1) / grid.x;
    grid.y = (unsigned)(gy < cap ? gy : cap);
    B2_CUDA(b2_launch(bn_swish_fwd_kernel, grid, dim3(256), 0, st, (const bf16*)x, mean_rstd, gamma, beta, M, C, (bf16*)y));
  }
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_bn_swish_bwd(const void* dy, const void* x, const float* mean_rstd, const float* gamma, const float* beta,
                                    long long M, int C, float* sums2_zeroed, void* dx, float* dgamma_accum, float* dbeta_accum,
                                    void* stream) {
  B2_CHECK_ARG(M > 0 && M < (1LL << 31) && C % 8 == 0, "bn_swish_bwd: C %% 8 == 0");
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid((C + CS_COLS - 1) / CS_COLS, 1);
  long long gy = (M + CS_ROWS * 4 - 1) / (CS_ROWS * 4);
  const int cap = (b2_num_sms() + grid.x - 1) / grid.x;  // one 512-thread block (128 registers) per SM
  grid.y = (unsigned)(gy < cap ? gy : cap);
  B2_SCOPE(FAM_CONVMOD, st, 0, 10.0 * M * C);
  B2_CUDA(b2_launch(bn_swish_bwd_reduce_kernel, grid, dim3(512), 0, st, (const bf16*)dy, (const bf16*)x, mean_rstd, gamma, beta, (int)M,
                    C, sums2_zeroed));
  B2_LAUNCH_CHECK();
  {
    dim3 ga((C + 511) / 512, 1);
    const long long gya = (M + 4 * BN_U - 1) / (4 * BN_U);
    const long long capa = (4LL * b2_num_sms() + ga.x - 1) / ga.x;
    ga.y = (unsigned)(gya < capa ? gya : capa);
    B2_CUDA(b2_launch(bn_swish_bwd_apply_kernel, ga, dim3(256), 0, st, (const bf16*)dy, (const bf16*)x, mean_rstd, gamma, beta,
                      sums2_zeroed, M, C, (bf16*)dx, dgamma_accum, dbeta_accum));
  }
  B2_LAUNCH_CHECK();
  return 0;
}
