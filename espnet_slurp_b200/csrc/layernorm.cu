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
#include "sm100_ptx.cuh"

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

// Forward with the rows streamed through a per-warp ring of shared-memory slots (1-D bulk copies on mbarriers, NBUF rows
// ahead), like ln_bwd_pipe_kernel below: 16 warps x 4 x 2 KB in flight per SM instead of one round trip per warp and row.
constexpr int LNF_WARPS = 16, LNF_NBUF = 4;
template <int NG>
__global__ void __launch_bounds__(LNF_WARPS * 32, 1) ln_fwd_pipe_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                                       const float* __restrict__ beta, int M, int D, float eps,
                                                                       bf16* __restrict__ y16, float* __restrict__ y32,
                                                                       float* __restrict__ stats) {
  extern __shared__ __align__(128) unsigned char smraw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int ngroups = D >> 3;
  const uint32_t slot = (uint32_t)D * 4u;
  const uint32_t sbase = smem_u32(smraw);
  const uint32_t bar0 = sbase + (uint32_t)(warp * LNF_NBUF) * 8u;
  const uint32_t ring = 1024u + (uint32_t)(warp * LNF_NBUF) * slot;  // barriers live in the first KB
  b2_pdl_launch_dependents();  // the next kernel may start launching; it waits for this one before touching memory
  if (lane == 0) {
    for (int b = 0; b < LNF_NBUF; ++b) mbar_init(bar0 + b * 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  b2_pdl_wait();  // barrier set-up overlapped the predecessor's tail; its results are needed from here on
  const long long W = (long long)gridDim.x * LNF_WARPS, gw = (long long)blockIdx.x * LNF_WARPS + warp;
  auto issue = [&](long long row, int b) {
    mbar_expect_tx(bar0 + b * 8, slot);
    bulk_g2s(sbase + ring + (uint32_t)b * slot, x + row * D, slot, bar0 + b * 8);
  };
  if (lane == 0) {
    for (int b = 0; b < LNF_NBUF; ++b)
      if (gw + b * W < M) issue(gw + b * W, b);
  }
  float ga[NG][8], be[NG][8];
#pragma unroll
  for (int i = 0; i < NG; ++i) {
    const int g = lane + 32 * i;
#pragma unroll
    for (int j = 0; j < 8; ++j) ga[i][j] = be[i][j] = 0.f;
    if (g < ngroups) {
      load8_f32(gamma + g * 8, ga[i]);
      load8_f32(beta + g * 8, be[i]);
    }
  }
  int b = 0;
  uint32_t ph = 0;
  for (long long row = gw; row < M; row += W) {
    mbar_wait(bar0 + b * 8, ph);
    const float* sl = reinterpret_cast<const float*>(smraw + ring + (uint32_t)b * slot);
    float v[NG][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NG; ++i) {
      const int g = lane + 32 * i;
#pragma unroll
      for (int j = 0; j < 8; ++j) v[i][j] = 0.f;
      if (g < ngroups) {
        load8_f32(sl + g * 8, v[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) s += v[i][j];
      }
    }
    __syncwarp();
    if (lane == 0 && row + (long long)LNF_NBUF * W < M) issue(row + (long long)LNF_NBUF * W, b);
    if (++b == LNF_NBUF) {
      b = 0;
      ph ^= 1u;
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
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * ga[i][j] + be[i][j];
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

// The same backward with the row loads decoupled from the arithmetic: every warp owns a ring of NBUF row slots in shared
// memory (x | dy | dx_add of one row each), filled by 1-D bulk copies that complete on the slot's mbarrier, issued NBUF rows
// ahead.  The register-only version above has each warp wait out a full HBM round trip per row with 12 warps per SM --
// about 30 KB in flight per SM against the ~45 KB that 6.5 TB/s x 1 us of latency asks for (measured 0.61 of the HBM
// peak); here 12 warps x NBUF x 5 KB are in flight whatever the arithmetic is doing.  D <= 512.
constexpr int LNP_WARPS = 12;
template <int NG, bool DY_F32, bool GSUM>
__global__ void __launch_bounds__(LNP_WARPS * 32, 1) ln_bwd_pipe_kernel(const void* __restrict__ dy_, const float* __restrict__ x,
                                                     const float* __restrict__ gamma, const float* __restrict__ stats,
                                                     const float* __restrict__ dx_add, int M, int D, float* __restrict__ dx,
                                                     float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                     bf16* __restrict__ g16, float g_alpha, B2Dropout drop,
                                                     float* __restrict__ gsum, int nbuf) {
  extern __shared__ __align__(128) unsigned char smraw[];
  float* sm = reinterpret_cast<float*>(smraw);  // [3][D] block partials, then the barriers, then the row slots
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int ngroups = D >> 3;
  const uint32_t bx = (uint32_t)D * 4u, bdy = (uint32_t)D * (DY_F32 ? 4u : 2u), badd = dx_add != nullptr ? (uint32_t)D * 4u : 0u;
  const uint32_t slot = bx + bdy + badd;
  const uint32_t bar0 = smem_u32(smraw) + 3u * D * 4u + (uint32_t)(warp * nbuf) * 8u;
  const uint32_t ring_off = (3u * D * 4u + (uint32_t)(LNP_WARPS * nbuf) * 8u + 127u) & ~127u;
  const uint32_t buf0 = ring_off + (uint32_t)(warp * nbuf) * slot;  // byte offset of this warp's first slot
  b2_pdl_launch_dependents();
  for (int i = threadIdx.x; i < 3 * D; i += blockDim.x) sm[i] = 0.f;
  if (lane == 0) {
    for (int b = 0; b < nbuf; ++b) mbar_init(bar0 + b * 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  b2_pdl_wait();
  const long long W = (long long)gridDim.x * LNP_WARPS, gw = (long long)blockIdx.x * LNP_WARPS + warp;
  const uint32_t sbase = smem_u32(smraw);
  auto issue = [&](long long row, int b) {  // one lane
    const uint32_t bar = bar0 + b * 8, dst = sbase + buf0 + (uint32_t)b * slot;
    mbar_expect_tx(bar, slot);
    bulk_g2s(dst, x + row * D, bx, bar);
    bulk_g2s(dst + bx, reinterpret_cast<const unsigned char*>(dy_) + row * (long long)bdy, bdy, bar);
    if (badd) bulk_g2s(dst + bx + bdy, dx_add + row * D, badd, bar);
  };
  if (lane == 0) {
    for (int b = 0; b < nbuf; ++b)
      if (gw + b * W < M) issue(gw + b * W, b);
  }
  float ga[NG][8];
#pragma unroll
  for (int i = 0; i < NG; ++i) {
    const int g = lane + 32 * i;
#pragma unroll
    for (int j = 0; j < 8; ++j) ga[i][j] = 0.f;
    if (g < ngroups) load8_f32(gamma + g * 8, ga[i]);
  }
  float dg[NG][8], db[NG][8], dc[GSUM ? NG : 1][8];
#pragma unroll
  for (int i = 0; i < NG; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) dg[i][j] = db[i][j] = 0.f;
#pragma unroll
  for (int i = 0; i < (GSUM ? NG : 1); ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) dc[i][j] = 0.f;

  float2 st_next = gw < M ? *reinterpret_cast<const float2*>(stats + gw * 2) : make_float2(0.f, 0.f);
  int b = 0;
  uint32_t ph = 0;
  for (long long row = gw; row < M; row += W) {
    const float mean = st_next.x, rstd = st_next.y;
    if (row + W < M) st_next = *reinterpret_cast<const float2*>(stats + (row + W) * 2);  // one row ahead of its use
    mbar_wait(bar0 + b * 8, ph);
    const unsigned char* sl = smraw + buf0 + (uint32_t)b * slot;
    float xh[NG][8], dyg[NG][8], addv[NG][8];
    float c1 = 0.f, c2 = 0.f;
#pragma unroll
    for (int i = 0; i < NG; ++i) {
      const int g = lane + 32 * i;
      if (g < ngroups) {
        float xv[8], dyv[8];
        load8_f32(reinterpret_cast<const float*>(sl) + g * 8, xv);
        if (DY_F32)
          load8_f32(reinterpret_cast<const float*>(sl + bx) + g * 8, dyv);
        else
          load8_bf16(reinterpret_cast<const bf16*>(sl + bx) + g * 8, dyv);
        if (badd) load8_f32(reinterpret_cast<const float*>(sl + bx + bdy) + g * 8, addv[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] = (xv[j] - mean) * rstd;
          dyg[i][j] = dyv[j] * ga[i][j];
          c1 += dyg[i][j];
          c2 += dyg[i][j] * xh[i][j];
          dg[i][j] += dyv[j] * xh[i][j];
          db[i][j] += dyv[j];
        }
      }
    }
    __syncwarp();  // every lane has read the slot: refill it with the row NBUF rows on
    if (lane == 0 && row + (long long)nbuf * W < M) issue(row + (long long)nbuf * W, b);
    if (++b == nbuf) {
      b = 0;
      ph ^= 1u;
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
        if (badd) {
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += addv[i][j];
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
  // Each warp's column partials go to its OWN (now idle) row slots -- every copy it issued has been consumed, and nbuf >= 2
  // slots of >= 6 D bytes hold the 3 x D floats -- with plain vector stores; the block then sums the 12 copies column-wise.
  // (As shared-memory atomics -- a compare-and-swap loop per float, twelve warps per address -- this was 14 % of the
  // kernel's stall samples.)
  {
    float* mine = reinterpret_cast<float*>(smraw + buf0);
#pragma unroll
    for (int i = 0; i < NG; ++i) {
      const int g = lane + 32 * i;
      if (g < ngroups) {
        store8_f32(mine + g * 8, dg[i]);
        store8_f32(mine + D + g * 8, db[i]);
        if (GSUM) store8_f32(mine + 2 * D + g * 8, dc[i]);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < (GSUM ? 3 : 2) * (D >> 2); i += blockDim.x) {
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int w = 0; w < LNP_WARPS; ++w) {
      const float4 v = reinterpret_cast<const float4*>(smraw + ring_off + (uint32_t)(w * nbuf) * slot)[i];
      t.x += v.x;
      t.y += v.y;
      t.z += v.z;
      t.w += v.w;
    }
    reinterpret_cast<float4*>(sm)[i] = t;
  }
  __syncthreads();
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

template <int NG, bool F32, bool GS>
int launch_ln_bwd_pipe(const void* dy, const float* x, const float* gamma, const float* stats, const float* dx_add, int M, int D,
                       float* dx, float* dgamma, float* dbeta, bf16* g16, float g_alpha, B2Dropout drop, float* gsum, int nbuf,
                       size_t smem, cudaStream_t st) {
  auto kern = ln_bwd_pipe_kernel<NG, F32, GS>;
  static unsigned long long configured = 0;  // per device
  if (!b2_device_done(configured)) {
    B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
    b2_device_mark(configured);
  }
  const long long warps = M;
  const int grid = (int)min((long long)b2_num_sms(), (warps + LNP_WARPS - 1) / LNP_WARPS);
  B2_CUDA(b2_launch(kern, dim3(grid), dim3(LNP_WARPS * 32), smem, st, dy, x, gamma, stats, dx_add, M, D, dx, dgamma, dbeta, g16, g_alpha,
                    drop, gsum, nbuf));
  return 0;
}

}  // namespace

extern "C" int b200slu_layernorm_fwd(const float* x, const float* gamma, const float* beta, int M, int D, float eps,
                                     void* y_bf16, float* y_f32, float* stats, void* stream) {
  B2_CHECK_ARG(M > 0 && D > 0 && D % 8 == 0 && D <= 2048, "layernorm: need D %% 8 == 0 and D <= 2048 (M=%d D=%d)", M, D);
  B2_CHECK_ARG(y_bf16 != nullptr || y_f32 != nullptr, "layernorm: no output");
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = b2_grid_for((long long)M * 32, 256, 8);
  B2_SCOPE(FAM_LN, st, 0, (double)M * D * (4.0 + (y_bf16 ? 2.0 : 0.0) + (y_f32 ? 4.0 : 0.0)));
  static const bool fpipe_off = getenv("B200SLU_LN_PIPE") && atoi(getenv("B200SLU_LN_PIPE")) == 0;  // tuning / A-B only
  if (!fpipe_off && D <= 512 && ((uintptr_t)x % 16) == 0 && M >= 4 * LNF_WARPS) {
    const size_t psmem = 1024 + (size_t)LNF_WARPS * LNF_NBUF * D * 4;
    const int pgrid = (int)min((long long)b2_num_sms(), ((long long)M + LNF_WARPS - 1) / LNF_WARPS);
    static unsigned long long configured = 0;  // per device
    if (!b2_device_done(configured)) {
      B2_CUDA(cudaFuncSetAttribute(ln_fwd_pipe_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
      B2_CUDA(cudaFuncSetAttribute(ln_fwd_pipe_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
      b2_device_mark(configured);
    }
    if (D <= 256)
      B2_CUDA(b2_launch(ln_fwd_pipe_kernel<1>, dim3(pgrid), dim3(LNF_WARPS * 32), psmem, st, x, gamma, beta, M, D, eps, (bf16*)y_bf16,
                        y_f32, stats));
    else
      B2_CUDA(b2_launch(ln_fwd_pipe_kernel<2>, dim3(pgrid), dim3(LNF_WARPS * 32), psmem, st, x, gamma, beta, M, D, eps, (bf16*)y_bf16,
                        y_f32, stats));
    B2_LAUNCH_CHECK();
    return 0;
  }
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
  // rows through a shared-memory ring (D <= 512, aligned rows): see ln_bwd_pipe_kernel
  static const bool pipe_off = getenv("B200SLU_LN_PIPE") && atoi(getenv("B200SLU_LN_PIPE")) == 0;  // tuning / A-B only
  if (!pipe_off && D <= 512 && ((uintptr_t)x % 16) == 0 && ((uintptr_t)dy % 16) == 0 && ((uintptr_t)dx_add % 16) == 0) {
    const size_t slot = (size_t)D * (4 + (dy_is_f32 ? 4 : 2) + (dx_add ? 4 : 0));
    const size_t fixed = 3 * (size_t)D * 4 + 1024;
    int nbuf = (int)((232448 - fixed) / (LNP_WARPS * slot));
    if (nbuf > 4) nbuf = 4;
    if (nbuf >= 2) {
      const size_t psmem = ((3 * (size_t)D * 4 + (size_t)LNP_WARPS * nbuf * 8 + 127) & ~(size_t)127) + (size_t)LNP_WARPS * nbuf * slot;
#define LN_P(NG, F32, GS)                                                                                                  \
  B2_TRY((launch_ln_bwd_pipe<NG, F32, GS>(dy, x, gamma, stats, dx_add, M, D, dx, dgamma, dbeta, (bf16*)g_bf16, g_alpha, drop, \
                                          g_colsum_accum, nbuf, psmem, st)))
#define LN_PP(NG)                                        \
  do {                                                   \
    if (dy_is_f32) {                                     \
      if (g_colsum_accum) LN_P(NG, true, true);          \
      else LN_P(NG, true, false);                        \
    } else {                                             \
      if (g_colsum_accum) LN_P(NG, false, true);         \
      else LN_P(NG, false, false);                       \
    }                                                    \
  } while (0)
      if (D <= 256) LN_PP(1);
      else LN_PP(2);
#undef LN_PP
#undef LN_P
      B2_LAUNCH_CHECK();
      return 0;
    }
  }
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
