// LayerNorm forward / backward over the last dimension, warp per row, fp32 statistics.
//
// Replaces espnet/nets/pytorch_backend/transformer/layer_norm.py:12-38 (eps = 1e-12 at :23) at the
// five call sites of conformer/encoder_layer.py:97-149, conformer_encoder.py:362-363 and the
// decoder's norm1/2/3 + after_norm (decoder_layer.py:63-134, transformer_decoder.py:139-141).
// The residual stream is fp32; the normalised output feeds a tensor-core GEMM, so the forward can
// emit bf16 and/or fp32, and the backward fuses (a) the residual-path gradient add and (b) the
// bf16 "branch gradient" alpha*dropout(dx) that the next dgrad GEMM consumes, together with (c) its column sums, which
// are the bias gradient of the branch's last Linear (saves the separate cast + column-sum pass over dx).
// Memory-bound: algorithmic bytes fwd = M*D*(4 + 2|4), bwd = M*D*(dy + 4 + 4 [+4] [+2]).
#include <stdlib.h>

#include "common.cuh"

namespace {

template <int NG>  // NG groups of 8 columns per lane: D <= NG*256
__global__ void __launch_bounds__(256) ln_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                     const float* __restrict__ beta, int M, int D, float eps,
                                                     bf16* __restrict__ y16, float* __restrict__ y32,
                                                     float* __restrict__ stats) {
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int ngroups = D >> 3;
  for (long long row = (long long)blockIdx.x * warps_per_block + (threadIdx.x >> 5); row < M;
       row += (long long)gridDim.x * warps_per_block) {
    const float* xr = x + row * D;
    float v[NG][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NG; ++i) {
      const int g = lane + 32 * i;
      if (g < ngroups) {
        load8_f32(xr + g * 8, v[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) s += v[i][j];
      }
    }
    const float mean = warp_sum(s) / (float)D;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < NG; ++i) {
      const int g = lane + 32 * i;
      if (g < ngroups) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float d = v[i][j] - mean;
          q += d * d;
        }
      }
    }
    const float rstd = rsqrtf(warp_sum(q) / (float)D + eps);
    if (lane == 0 && stats != nullptr) {
      stats[row * 2] = mean;
      stats[row * 2 + 1] = rstd;
    }
#pragma unroll
    for (int i = 0; i < NG; ++i) {
      const int g = lane + 32 * i;
      if (g < ngroups) {
        float ga[8], be[8], o[8];
        load8_f32(gamma + g * 8, ga);
        load8_f32(beta + g * 8, be);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * ga[j] + be[j];
        if (y16 != nullptr) store8_bf16(y16 + row * D + g * 8, o);
        if (y32 != nullptr) store8_f32(y32 + row * D + g * 8, o);
      }
    }
  }
}

template <int NG, bool DY_F32, bool GSUM>
__global__ void __launch_bounds__(NG <= 2 ? 384 : 256, 1) ln_bwd_kernel(const void* __restrict__ dy_, const float* __restrict__ x,
                                                     const float* __restrict__ gamma, const float* __restrict__ stats,
                                                     const float* __restrict__ dx_add, int M, int D, float* __restrict__ dx,
                                                     float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                     bf16* __restrict__ g16, float g_alpha, B2Dropout drop,
                                                     float* __restrict__ gsum) {
  extern __shared__ float sm[];  // [2][D] block partials of dgamma / dbeta (+ [D] column sums of g16 with GSUM)
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int ngroups = D >> 3;
  for (int i = threadIdx.x; i < (GSUM ? 3 : 2) * D; i += blockDim.x) sm[i] = 0.f;
  __syncthreads();
  float dg[NG][8], db[NG][8], dc[GSUM ? NG : 1][8];
#pragma unroll
  for (int i = 0; i < NG; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) dg[i][j] = db[i][j] = 0.f;
#pragma unroll
  for (int i = 0; i < (GSUM ? NG : 1); ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) dc[i][j] = 0.f;

  for (long long row = (long long)blockIdx.x * warps_per_block + (threadIdx.x >> 5); row < M;
       row += (long long)gridDim.x * warps_per_block) {
    const float mean = stats[row * 2], rstd = stats[row * 2 + 1];
    float xh[NG][8], dyg[NG][8];
    // the residual-path gradient is fetched with the row's other loads (one memory round trip per row, not two)
    float addv[NG <= 2 ? NG : 1][8];
    if (NG <= 2 && dx_add != nullptr) {
#pragma unroll
      for (int i = 0; i < (NG <= 2 ? NG : 1); ++i) {
        const int g = lane + 32 * i;
        if (g < ngroups) load8_f32(dx_add + row * D + g * 8, addv[i]);
      }
    }
    float c1 = 0.f, c2 = 0.f;
#pragma unroll
    for (int i = 0; i < NG; ++i) {
      const int g = lane + 32 * i;
      if (g < ngroups) {
        float xv[8], dyv[8], ga[8];
        load8_f32(x + row * D + g * 8, xv);
        if (DY_F32)
          load8_f32(reinterpret_cast<const float*>(dy_) + row * D + g * 8, dyv);
        else
          load8_bf16(reinterpret_cast<const bf16*>(dy_) + row * D + g * 8, dyv);
        load8_f32(gamma + g * 8, ga);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] = (xv[j] - mean) * rstd;
          dyg[i][j] = dyv[j] * ga[j];
          c1 += dyg[i][j];
          c2 += dyg[i][j] * xh[i][j];
          dg[i][j] += dyv[j] * xh[i][j];
          db[i][j] += dyv[j];
        }
      }
    }
    c1 = warp_sum(c1) / (float)D;
    c2 = warp_sum(c2) / (float)D;
#pragma unroll
    for (int i = 0; i < NG; ++i) {
      const int g = lane + 32 * i;
      if (g < ngroups) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * (dyg[i][j] - c1 - xh[i][j] * c2);
        if (dx_add != nullptr) {
          if (NG <= 2) {
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] += addv[NG <= 2 ? i : 0][j];
          } else {
            float a[8];
            load8_f32(dx_add + row * D + g * 8, a);
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] += a[j];
          }
        }
        store8_f32(dx + row * D + g * 8, o);
        if (g16 != nullptr) {
          dropout8(drop, (unsigned long long)row * D + g * 8, o);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] *= g_alpha;
          if (GSUM) {
#pragma unroll
            for (int j = 0; j < 8; ++j) dc[i][j] += o[j];
          }
          store8_bf16(g16 + row * D + g * 8, o);
        }
      }
    }
  }
  // block reduction of the parameter gradients, then one atomic per column per block
#pragma unroll
  for (int i = 0; i < NG; ++i) {
    const int g = lane + 32 * i;
    if (g < ngroups) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        atomicAdd(&sm[g * 8 + j], dg[i][j]);
        atomicAdd(&sm[D + g * 8 + j], db[i][j]);
        if (GSUM) atomicAdd(&sm[2 * D + g * 8 + j], dc[i][j]);
      }
    }
  }
  __syncthreads();
  // Same-address atomics from different SMs retire serially (tens of ns each): one 16-byte vector atomic per four
  // columns and block, and as few, fat blocks as keep the loads in flight (host side).  Summing the block partials
  // across a thread-block cluster through DSMEM first was measured slower (59 us against 45 us: the cluster's
  // co-scheduling and its two cluster barriers cost more than the atomics they save).
  const bool vec4 = (((uintptr_t)dgamma | (uintptr_t)dbeta | (GSUM ? (uintptr_t)gsum : 0)) & 15) == 0 && (D & 3) == 0;
  if (vec4) {
    for (int i = threadIdx.x; i < D / 4; i += blockDim.x) {
      atomicAdd(reinterpret_cast<float4*>(dgamma) + i, reinterpret_cast<const float4*>(sm)[i]);
      atomicAdd(reinterpret_cast<float4*>(dbeta) + i, reinterpret_cast<const float4*>(sm + D)[i]);
      if (GSUM) atomicAdd(reinterpret_cast<float4*>(gsum) + i, reinterpret_cast<const float4*>(sm + 2 * D)[i]);
    }
  } else {
    for (int i = threadIdx.x; i < D; i += blockDim.x) {
      atomicAdd(dgamma + i, sm[i]);
      atomicAdd(dbeta + i, sm[D + i]);
      if (GSUM) atomicAdd(gsum + i, sm[2 * D + i]);
    }
  }
}

}  // namespace

extern "C" int b200slu_layernorm_fwd(const float* x, const float* gamma, const float* beta, int M, int D, float eps,
                                     void* y_bf16, float* y_f32, float* stats, void* stream) {
  B2_CHECK_ARG(M > 0 && D > 0 && D % 8 == 0 && D <= 2048, "layernorm: need D %% 8 == 0 and D <= 2048 (M=%d D=%d)", M, D);
  B2_CHECK_ARG(y_bf16 != nullptr || y_f32 != nullptr, "layernorm: no output");
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = b2_grid_for((long long)M * 32, 256, 8);
  B2_SCOPE(FAM_LN, st, 0, (double)M * D * (4.0 + (y_bf16 ? 2.0 : 0.0) + (y_f32 ? 4.0 : 0.0)));
#define LN_F(NG) ln_fwd_kernel<NG><<<grid, 256, 0, st>>>(x, gamma, beta, M, D, eps, (bf16*)y_bf16, y_f32, stats)
  if (D <= 256) LN_F(1);
  else if (D <= 512) LN_F(2);
  else if (D <= 1024) LN_F(4);
  else LN_F(8);
#undef LN_F
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_layernorm_bwd(const void* dy, int dy_is_f32, const float* x, const float* gamma, const float* stats,
                                     const float* dx_add, int M, int D, float* dx, float* dgamma, float* dbeta,
                                     void* g_bf16, float g_alpha, float* g_colsum_accum, float drop_p, unsigned long long seed,
                                     unsigned int site, void* stream) {
  B2_CHECK_ARG(M > 0 && D > 0 && D % 8 == 0 && D <= 2048, "layernorm_bwd: need D %% 8 == 0 and D <= 2048 (M=%d D=%d)", M, D);
  B2_CHECK_ARG(g_colsum_accum == nullptr || g_bf16 != nullptr, "layernorm_bwd: g_colsum_accum needs g_bf16");
  cudaStream_t st = (cudaStream_t)stream;
  // one 512-thread block per SM (16 row-warps, <= 128 registers): the block count is the atomic count per column
  static const int bps = getenv("B200SLU_LN_BPS") ? atoi(getenv("B200SLU_LN_BPS")) : 1;  // tuning knobs
  static const int thr = getenv("B200SLU_LN_THREADS") ? atoi(getenv("B200SLU_LN_THREADS")) : 384;
  const int threads = D <= 512 ? thr : 256;  // wider rows hold more registers per lane
  const int grid = b2_grid_for((long long)M * 32, threads, (bps > 0 ? bps : 1) * (threads <= 256 ? 2 : 1));
  B2Dropout drop = b2_make_drop(drop_p, seed, site);
  const size_t smem = (g_colsum_accum ? 3 : 2) * (size_t)D * sizeof(float);
  B2_SCOPE(FAM_LN, st, 0,
           (double)M * D * ((dy_is_f32 ? 4.0 : 2.0) + 4.0 + 4.0 + (dx_add ? 4.0 : 0.0) + (g_bf16 ? 2.0 : 0.0)));
#define LN_B3(NG, F32, GS)                                                                                                \
  ln_bwd_kernel<NG, F32, GS><<<grid, threads, smem, st>>>(dy, x, gamma, stats, dx_add, M, D, dx, dgamma, dbeta, (bf16*)g_bf16, \
                                                          g_alpha, drop, g_colsum_accum)
#define LN_B(NG)                                                                                                          \
  do {                                                                                                                    \
    if (dy_is_f32) {                                                                                                      \
      if (g_colsum_accum) LN_B3(NG, true, true);                                                                          \
      else LN_B3(NG, true, false);                                                                                        \
    } else {                                                                                                              \
      if (g_colsum_accum) LN_B3(NG, false, true);                                                                         \
      else LN_B3(NG, false, false);                                                                                       \
    }                                                                                                                     \
  } while (0)
  if (D <= 256) LN_B(1);
  else if (D <= 512) LN_B(2);
  else if (D <= 1024) LN_B(4);
  else LN_B(8);
#undef LN_B
#undef LN_B3
  B2_LAUNCH_CHECK();
  return 0;
}
