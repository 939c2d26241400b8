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
  constexpr int U = ((TT + KMAX - 1) * (CC / 8) + 255) / 256;  // items per thread of a 256-thread block
  const int pad = (K - 1) / 2, nitems = (TT + K - 1) * (CC / 8);
  uint4 ra[U], rg[U], rd[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int it = threadIdx.x + u * 256;
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
    const int it = threadIdx.x + u * 256;
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

// Each thread owns one channel and, per 32-frame sub-tile, FR = 8 consecutive output frames: the K + 7 staged inputs
// it needs are read from shared memory once into a sliding register window (4.75 LDS per output instead of K),
// weights stay in registers for the whole tile.
constexpr int FR = 8;

template <int TT, int KT>
__global__ void __launch_bounds__(256) glu_dw_fwd_kernel(const bf16* __restrict__ pw1, const float* __restrict__ w,
                                                         const float* __restrict__ bias, int B, int T, int C, int Krt,
                                                         bf16* __restrict__ out, float* __restrict__ sums) {
  static_assert(TT % (4 * FR) == 0, "tile = whole sub-tiles of 4 frame lanes x FR frames");
  __shared__ __align__(16) float sm[TT + KMAX - 1][CC];
  __shared__ float red[2][4][CC];
  const int K = KT > 0 ? KT : Krt;
  const int c0 = blockIdx.x * CC;
  const int cl = threadIdx.x & (CC - 1), fr = threadIdx.x >> 6;  // 4 frame lanes x FR frames
  const int c = c0 + cl;
  float wr[KMAX];
#pragma unroll
  for (int k = 0; k < KMAX; ++k) wr[k] = (c < C && k < K) ? w[c * K + k] : 0.f;
  const float bs = c < C ? bias[c] : 0.f;
  const int ntt = (T + TT - 1) / TT;
  float s1 = 0.f, s2 = 0.f;
  // Each block walks its (utterance, time tile) units and keeps the BatchNorm partial sums in registers across them:
  // the atomics on `sums` are per block (grid.y per channel), not per tile -- same-address atomics retire serially.
  for (int unit = blockIdx.y; unit < B * ntt; unit += gridDim.y) {
  const int b = unit / ntt, t0 = (unit - b * ntt) * TT;
  __syncthreads();  // previous tile fully consumed
  stage_tiles<TT, false>(pw1, nullptr, b, T, C, t0, c0, K, sm, nullptr);
  __syncthreads();
  if (c < C) {
#pragma unroll 1
    for (int sub = 0; sub < TT / (4 * FR); ++sub) {
      const int f0 = sub * 4 * FR + fr * FR;
      if (t0 + f0 >= T) break;
      float acc[FR];
#pragma unroll
      for (int o = 0; o < FR; ++o) acc[o] = bs;
#pragma unroll
      for (int j = 0; j < FR + KMAX - 1; ++j) {
        if (KT > 0 ? (j < FR + KT - 1) : (j < FR + K - 1)) {
          const float v = sm[f0 + j][cl];
#pragma unroll
          for (int o = 0; o < FR; ++o)
            if (j - o >= 0 && j - o < (KT > 0 ? KT : KMAX)) acc[o] += wr[j - o] * v;  // wr[k] = 0 for k >= K
        }
      }
#pragma unroll
      for (int o = 0; o < FR; ++o) {
        const int t = t0 + f0 + o;
        if (t < T) {
          out[((long long)b * T + t) * C + c] = __float2bfloat16(acc[o]);
          s1 += acc[o];
          s2 += acc[o] * acc[o];
        }
      }
    }
  }
  }
  red[0][fr][cl] = s1;
  red[1][fr][cl] = s2;
  __syncthreads();
  if (fr == 0 && c < C && sums != nullptr) {
    atomicAdd(sums + c, red[0][0][cl] + red[0][1][cl] + red[0][2][cl] + red[0][3][cl]);
    atomicAdd(sums + C + c, red[1][0][cl] + red[1][1][cl] + red[1][2][cl] + red[1][3][cl]);
  }
}

// Backward: grid (channel tiles, utterance chunks); each block walks all its (utterance, time tile) pairs and keeps the
// depthwise weight gradient in registers across them, so the global atomics are per block, not per tile.
template <int TT, int KT>
__global__ void __launch_bounds__(256) glu_dw_bwd_kernel(const bf16* __restrict__ d_dw, const bf16* __restrict__ pw1,
                                                         const float* __restrict__ w, int B, int T, int C, int Krt,
                                                         bf16* __restrict__ d_pw1, float* __restrict__ dw_accum,
                                                         float* __restrict__ dbias_accum) {
  extern __shared__ __align__(16) float dyn_smem[];
  float (*sg)[CC] = reinterpret_cast<float (*)[CC]>(dyn_smem);                             // glu halo tile  [TT+KMAX-1][CC]
  float (*sd)[CC] = reinterpret_cast<float (*)[CC]>(dyn_smem + (TT + KMAX - 1) * CC);      // d_dw halo tile
  // dW partial (+ dbias); rows padded to an odd stride so that the 32 channels of a warp hit 32 different banks
  float (*swg)[KMAX + 2] = reinterpret_cast<float (*)[KMAX + 2]>(dyn_smem + 2 * (TT + KMAX - 1) * CC);
  // reversed taps [KMAX][CC]: swr[m][ch] = w[ch][K-1-m] (0 for m >= K); re-read per tile so that the staging phase has the
  // whole register file for its loads in flight
  float (*swr)[CC] = reinterpret_cast<float (*)[CC]>(dyn_smem + 2 * (TT + KMAX - 1) * CC + CC * (KMAX + 2));
  const int K = KT > 0 ? KT : Krt;
  const int c0 = blockIdx.x * CC;
  const int pad = (K - 1) / 2;
  const int cl = threadIdx.x & (CC - 1), fr = threadIdx.x >> 6;
  const int c = c0 + cl;
  const int ntt = (T + TT - 1) / TT;
  for (int it = threadIdx.x; it < CC * (KMAX + 2); it += blockDim.x) (&swg[0][0])[it] = 0.f;
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
     float accb = 0.f;
#pragma unroll 1
     for (int sub = 0; sub < TT / (4 * FR); ++sub) {
      const int f0 = sub * 4 * FR + fr * FR;
      if (t0 + f0 >= T) break;
      float dg[FR], dv[FR];
      // the GLU inputs of the thread's own frames (for d_a, d_gate) are fetched now; their latency hides under the taps
      bf16 pa[FR], pg[FR];
#pragma unroll
      for (int o = 0; o < FR; ++o) {
        const int t = min(t0 + f0 + o, T - 1);
        const bf16* row = pw1 + ((long long)b * T + t) * (2 * C);
        pa[o] = row[c];
        pg[o] = row[C + c];
      }
#pragma unroll
      for (int o = 0; o < FR; ++o) {
        dg[o] = 0.f;
        dv[o] = sd[f0 + o + pad][cl];  // d_dw at the thread's own frames (0 beyond T by staging)
        accb += dv[o];
      }
#pragma unroll
      for (int j = 0; j < FR + KMAX - 1; ++j) {
        if (KT > 0 ? (j < FR + KT - 1) : (j < FR + K - 1)) {
          // d_glu[f0+o] = sum_k w[k] * d_dw[f0 + o - k + pad]  = sum over staged row j = o + (K-1) - k
          const float dj = sd[f0 + j][cl];
          const float gj = sg[f0 + j][cl];
#pragma unroll
          for (int o = 0; o < FR; ++o) {
            // forward tap index of staged row j for output o: k = j - o (glu[f0+o+k] feeds out[f0+o])
            if (j - o >= 0 && j - o < (KT > 0 ? KT : KMAX)) acc[j - o] += dv[o] * gj;
          }
#pragma unroll
          for (int o = 0; o < FR; ++o) {
            // backward: d_dw[f0+o - k + pad] sits in staged row j = o + (K-1) - k  ->  w[k] = wrev[j - o]
            if (j - o >= 0 && j - o < (KT > 0 ? KT : KMAX)) dg[o] += wrev[j - o] * dj;
          }
        }
      }
#pragma unroll
      for (int o = 0; o < FR; ++o) {
        const int t = t0 + f0 + o;
        if (t < T) {
          const float a = __bfloat162float(pa[o]), gt = __bfloat162float(pg[o]);
          const float sgm = sigmoidf_(gt);
          bf16* drow = d_pw1 + ((long long)b * T + t) * (2 * C);
          drow[c] = __float2bfloat16(dg[o] * sgm);
          drow[C + c] = __float2bfloat16(dg[o] * a * sgm * (1.f - sgm));
        }
      }
     }
     // this tile's weight / bias gradient partials join the block's (4 frame lanes per address)
#pragma unroll
     for (int k = 0; k < KMAX; ++k)
       if (k < K) atomicAdd(&swg[cl][k], acc[k]);
     atomicAdd(&swg[cl][KMAX], accb);
    }
  }
  __syncthreads();
  for (int it = threadIdx.x; it < CC * K; it += blockDim.x) {
    const int ch = it / K, k = it - ch * K;
    if (c0 + ch < C) atomicAdd(dw_accum + (long long)(c0 + ch) * K + k, swg[ch][k]);
  }
  if (threadIdx.x < CC && c0 + threadIdx.x < C) atomicAdd(dbias_accum + c0 + threadIdx.x, swg[threadIdx.x][KMAX]);
}

// BatchNorm1d training statistics: mean / rstd from the fp32 sums, running-stat update (momentum,
// unbiased variance) exactly as torch.nn.BatchNorm1d does (convolution.py:45).
__global__ void bn_finalize_kernel(const float* __restrict__ sums, int C, float n, float eps, float momentum,
                                   float* __restrict__ mean_rstd, float* __restrict__ running_mean,
                                   float* __restrict__ running_var, long long* __restrict__ num_batches) {
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
        for (int q = 0; q < 4; ++q) {
          const float2 p2 = unpack_bf16x2(w4[q]);
          v[2 * q] = swishf_(fmaf(p2.x, sc[2 * q], sh[2 * q]));
          v[2 * q + 1] = swishf_(fmaf(p2.y, sc[2 * q + 1], sh[2 * q + 1]));
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
        for (int q = 0; q < 4; ++q) {
          const float2 d2 = unpack_bf16x2(wd[q]), x2 = unpack_bf16x2(wx[q]);
          const float dd[2] = {d2.x, d2.y}, xx[2] = {x2.x, x2.y};
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int j = 2 * q + h;
            const float xh = (xx[h] - mu[j]) * rs[j];
            const float dz = dd[h] * swish_gradf_(ga[j] * xh + be[j]);
            a1[j] += dz;
            a2[j] += dz * xh;
          }
        }
      }
    }
  }
  colsum_block_reduce(a1, red[0], tx, ty, C, sums2);
  colsum_block_reduce(a2, red[1], tx, ty, C, sums2 + C);
}

__global__ void __launch_bounds__(256) bn_swish_bwd_apply_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x,
                                                                 const float* __restrict__ mean_rstd,
                                                                 const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                 const float* __restrict__ sums2, long long M, int C,
                                                                 bf16* __restrict__ dx, float* __restrict__ dgamma,
                                                                 float* __restrict__ dbeta) {
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
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      k1[j] *= invn;
      k2[j] *= invn;
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
            const float dd[2] = {d2.x, d2.y}, xx[2] = {x2.x, x2.y};
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const int j = 2 * q + h;
              const float xh = (xx[h] - mu[j]) * rs[j];
              const float dz = dd[h] * swish_gradf_(fmaf(ga[j], xh, be[j]));
              o[j] = ga[j] * rs[j] * (dz - k1[j] - xh * k2[j]);
            }
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
int dw_tile_frames(int T) {
  static const int forced = getenv("B200SLU_DW_TT") ? atoi(getenv("B200SLU_DW_TT")) : 0;  // tuning only
  if (forced == 32 || forced == 96 || forced == 128) return forced;
  if (T <= 64) return 32;
  const int c96 = ((T + 95) / 96) * (96 + 30), c128 = ((T + 127) / 128) * (128 + 30);
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
    int gy = (4 * b2_num_sms() + ctiles - 1) / ctiles;                                                                \
    if (gy > units) gy = units;                                                                                       \
    glu_dw_fwd_kernel<TT_, KT_><<<dim3(ctiles, gy), 256, 0, st>>>((const bf16*)pw1, w, bias, B, T, C, K, (bf16*)out,  \
                                                                  bn_sums);                                           \
  } while (0)
  if (K == 31) {
    const int tt = dw_tile_frames(T);
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

extern "C" int b200slu_glu_dwconv_bwd(const void* d_out, const void* pw1, const float* w, int B, int T, int C, int K, void* d_pw1,
                                      float* dw_accum, float* dbias_accum, void* stream) {
  B2_CHECK_ARG(B > 0 && T > 0 && C % 8 == 0 && K % 2 == 1 && K <= KMAX, "glu_dwconv_bwd: C %% 8 == 0, odd K <= %d", KMAX);
  B2_CHECK_ARG(B <= 65535, "glu_dwconv_bwd: B too large");
  cudaStream_t st = (cudaStream_t)stream;
  B2_SCOPE(FAM_CONVMOD, st, 4.0 * B * T * C * K, 10.0 * B * T * C);
#define B2_BWD(TT_, KT_)                                                                                                     \
  do {                                                                                                                       \
    const int ctiles = (C + CC - 1) / CC, units = B * ((T + TT_ - 1) / TT_);                                                 \
    int gy = (2 * b2_num_sms() + ctiles - 1) / ctiles;                                                                       \
    if (gy > units) gy = units;                                                                                              \
    const size_t smem = (size_t)(2 * (TT_ + KMAX - 1) * CC + CC * (KMAX + 2) + KMAX * CC) * sizeof(float);                               \
    auto kern = glu_dw_bwd_kernel<TT_, KT_>;                                                                                 \
    static unsigned long long cfg = 0;                                                                                                 \
    if (!b2_device_done(cfg)) {                                                                                                              \
      B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));                           \
      b2_device_mark(cfg);                                                                                                            \
    }                                                                                                                        \
    kern<<<dim3(ctiles, gy), 256, smem, st>>>((const bf16*)d_out, (const bf16*)pw1, w, B, T, C, K, (bf16*)d_pw1, dw_accum,  \
                                              dbias_accum);                                                                  \
  } while (0)
  if (K == 31) {
    const int tt = dw_tile_frames(T);
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
  bn_finalize_kernel<<<(C + 255) / 256, 256, 0, st>>>(sums, C, (float)n, eps, momentum, mean_rstd, running_mean, running_var,
                                                      num_batches_tracked);
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
    bn_swish_fwd_kernel<<<grid, 256, 0, st>>>((const bf16*)x, mean_rstd, gamma, beta, M, C, (bf16*)y);
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
  bn_swish_bwd_reduce_kernel<<<grid, 512, 0, st>>>((const bf16*)dy, (const bf16*)x, mean_rstd, gamma, beta, (int)M, C,
                                                   sums2_zeroed);
  B2_LAUNCH_CHECK();
  {
    dim3 ga((C + 511) / 512, 1);
    const long long gya = (M + 4 * BN_U - 1) / (4 * BN_U);
    const long long capa = (4LL * b2_num_sms() + ga.x - 1) / ga.x;
    ga.y = (unsigned)(gya < capa ? gya : capa);
    bn_swish_bwd_apply_kernel<<<ga, 256, 0, st>>>((const bf16*)dy, (const bf16*)x, mean_rstd, gamma, beta, sums2_zeroed, M, C,
                                                  (bf16*)dx, dgamma_accum, dbeta_accum);
  }
  B2_LAUNCH_CHECK();
  return 0;
}
