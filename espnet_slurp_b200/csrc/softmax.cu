// Masked softmax over attention scores with rel_shift folded into the score indexing, fwd + bwd.
//
// Replaces forward_attention (espnet/nets/pytorch_backend/transformer/attention.py:64-96) and the
// rel_shift + add + scale of RelPositionMultiHeadedAttention.forward (:297-306, rel_shift :240-263)
// / LegacyRelPositionMultiHeadedAttention.forward (:196-207, rel_shift :145-165):
//   s[i,j] = (ac[i,j] + shift(bd)[i,j]) * scale ; masked keys (j >= len[b], or j > i when causal) get
//   probability exactly 0 (the reference fills finfo.min, softmaxes and re-zeroes) ; dropout.
//   latest: shift(bd)[i,j] = bd[i, T-1-i+j]                         (bd is T x (2T-1))
//   legacy: j<=i -> bd[i, T-1-(i-j)] ; j==i+1 -> 0 ; j>i+1 -> bd[i+1, j-i-2]   (bd is T x T)
// The (un-shifted) bd matrix is never re-laid-out: the shift is an index computation here, and in
// the backward the gradient is scattered straight back into the un-shifted layout.
// Score tensors are head-major [H, B, Tq, ld]; one warp owns one row; lanes own groups of 8 keys.
#include "common.cuh"

namespace {

struct SmArgs {
  int H, B, Tq, Tk;
  long long ld_ac, ld_bd, ld_p;
  int mode, causal;
  float scale;
  const int* lens;  // [B] valid keys per utterance
};

__device__ __forceinline__ float bd_term(const float* __restrict__ bd, long long hb, int i, int j, int T, long long ld_bd,
                                         int mode) {
  if (mode == B200SLU_RELPOS_NONE) return 0.f;
  const B2ShiftSrc src = b2_rel_shift_src(i, j, T, mode);
  return src.row < 0 ? 0.f : bd[(hb * T + src.row) * ld_bd + src.col];
}

template <int NG>
__global__ void __launch_bounds__(256) softmax_fwd_kernel(const float* __restrict__ ac, const float* __restrict__ bd, SmArgs a,
                                                          bf16* __restrict__ p, bf16* __restrict__ pd, B2Dropout drop) {
  const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  const long long rows = (long long)a.H * a.B * a.Tq;
  const int ngroups = (int)(a.ld_p >> 3);
  for (long long r = (long long)blockIdx.x * wpb + (threadIdx.x >> 5); r < rows; r += (long long)gridDim.x * wpb) {
    const long long hb = r / a.Tq;
    const int i = (int)(r - hb * a.Tq);
    const int b = (int)(hb % a.B);
    int nvalid = a.lens != nullptr ? a.lens[b] : a.Tk;
    nvalid = min(nvalid, a.Tk);
    if (a.causal) nvalid = min(nvalid, i + 1);
    float s[NG][8];
    float mx = -INFINITY;
#pragma unroll
    for (int k = 0; k < NG; ++k) {
      const int g = lane + 32 * k;
      if (g < ngroups) {
        float v[8];
        load8_f32(ac + r * a.ld_ac + g * 8, v);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const int j = g * 8 + e;
          if (j < nvalid) {
            s[k][e] = (v[e] + bd_term(bd, hb, i, j, a.Tq, a.ld_bd, a.mode)) * a.scale;
            mx = fmaxf(mx, s[k][e]);
          } else {
            s[k][e] = -INFINITY;
          }
        }
      }
    }
    mx = warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < NG; ++k) {
      const int g = lane + 32 * k;
      if (g < ngroups) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          s[k][e] = (s[k][e] == -INFINITY) ? 0.f : __expf(s[k][e] - mx);
          sum += s[k][e];
        }
      }
    }
    sum = warp_sum(sum);
    const float inv = sum > 0.f ? 1.f / sum : 0.f;
#pragma unroll
    for (int k = 0; k < NG; ++k) {
      const int g = lane + 32 * k;
      if (g < ngroups) {
#pragma unroll
        for (int e = 0; e < 8; ++e) s[k][e] *= inv;
        store8_bf16(p + r * a.ld_p + g * 8, s[k]);
        if (pd != p) {
          dropout8(drop, (unsigned long long)r * a.ld_p + g * 8, s[k]);
          store8_bf16(pd + r * a.ld_p + g * 8, s[k]);
        }
      }
    }
  }
}

template <int NG>
__global__ void __launch_bounds__(256) softmax_bwd_kernel(const float* __restrict__ dpd, const bf16* __restrict__ p, SmArgs a,
                                                          bf16* __restrict__ dac, bf16* __restrict__ dbd, B2Dropout drop) {
  extern __shared__ float sm[];  // [warps][ld_p] staged d(score) row for the un-shift scatter
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const long long rows = (long long)a.H * a.B * a.Tq;
  const int ngroups = (int)(a.ld_p >> 3);
  float* row_sm = sm + (size_t)w * a.ld_p;
  const int T = a.Tq;
  for (long long r = (long long)blockIdx.x * wpb + w; r < rows; r += (long long)gridDim.x * wpb) {
    const long long hb = r / a.Tq;
    const int i = (int)(r - hb * a.Tq);
    float pv[NG][8], dp[NG][8];
    float dot = 0.f;
#pragma unroll
    for (int k = 0; k < NG; ++k) {
      const int g = lane + 32 * k;
      if (g < ngroups) {
        load8_bf16(p + r * a.ld_p + g * 8, pv[k]);
        load8_f32(dpd + r * a.ld_p + g * 8, dp[k]);
        dropout8(drop, (unsigned long long)r * a.ld_p + g * 8, dp[k]);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          if (g * 8 + e >= a.Tk) dp[k][e] = 0.f;  // padding columns of dpd are unspecified
          dot += dp[k][e] * pv[k][e];
        }
      }
    }
    dot = warp_sum(dot);
#pragma unroll
    for (int k = 0; k < NG; ++k) {
      const int g = lane + 32 * k;
      if (g < ngroups) {
        float ds[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          ds[e] = pv[k][e] * (dp[k][e] - dot) * a.scale;
          if (dbd != nullptr) row_sm[g * 8 + e] = ds[e];
        }
        store8_bf16(dac + r * a.ld_p + g * 8, ds);
      }
    }
    if (dbd != nullptr) {
      __syncwarp();
      bf16* own = dbd + (hb * T + i) * a.ld_bd;
      if (a.mode == B200SLU_RELPOS_LATEST) {
        const int lo = b2_rel_shift_src(i, 0, T, B200SLU_RELPOS_LATEST).col;  // column of j = 0
        for (int n = lane; n < a.ld_bd; n += 32) {
          const int j = n - lo;
          own[n] = __float2bfloat16((j >= 0 && j < T) ? row_sm[j] : 0.f);
        }
      } else {  // legacy
        const int lo = b2_rel_shift_src(i, 0, T, B200SLU_RELPOS_LATEST).col;
        for (int n = lane; n < a.ld_bd; n += 32) {
          if (n >= lo && n < T) own[n] = __float2bfloat16(row_sm[n - lo]);           // j = n - lo <= i
          else if (n >= T || i == 0) own[n] = __float2bfloat16(0.f);               // padding / never-read part of row 0
        }
        if (i + 1 < T) {
          bf16* nxt = own + a.ld_bd;
          for (int j = i + 2 + lane; j < T; j += 32) nxt[j - i - 2] = __float2bfloat16(row_sm[j]);
        }
      }
      __syncwarp();
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Short rows (the SLURP utterances give T ~ 155 keys = 20 groups of 8): LPR lanes own a row and a warp works on
// 32 / LPR rows at once, so all 32 lanes are busy (20 of 32 with a warp per row), a warp has 4x the bytes in flight
// and the row reductions are log2(LPR) shuffle steps.  Modes NONE and LATEST (LEGACY keeps the warp-per-row kernels).
template <int LPR>
__device__ __forceinline__ float seg_max(float v) {
#pragma unroll
  for (int o = LPR / 2; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
template <int LPR>
__device__ __forceinline__ float seg_sum(float v) {
#pragma unroll
  for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <int LPR, int NGL>
__global__ void __launch_bounds__(256) softmax_fwd_rows_kernel(const float* __restrict__ ac, const float* __restrict__ bd,
                                                               SmArgs a, bf16* __restrict__ p, bf16* __restrict__ pd,
                                                               B2Dropout drop) {
  constexpr int RPW = 32 / LPR;
  const int lane = threadIdx.x & 31, sub = lane / LPR, l = lane % LPR, wpb = blockDim.x >> 5;
  const long long rows = (long long)a.H * a.B * a.Tq;
  const int ngroups = (int)(a.ld_p >> 3);
  B2DropKey dk = {0ull, 0u, 1.f};
  if (drop.p > 0.f) dk = dropout_key(drop);
  for (long long rw = (long long)blockIdx.x * wpb + (threadIdx.x >> 5); rw * RPW < rows; rw += (long long)gridDim.x * wpb) {
    const long long r = rw * RPW + sub;
    const bool live = r < rows;
    const long long hb = live ? r / a.Tq : 0;
    const int i = live ? (int)(r - hb * a.Tq) : 0;
    int nvalid = 0;
    if (live) {
      nvalid = a.lens != nullptr ? a.lens[(int)(hb % a.B)] : a.Tk;
      nvalid = min(nvalid, a.Tk);
      if (a.causal) nvalid = min(nvalid, i + 1);
    }
    // latest: the source of (i, j) is (i, src(i, 0).col + j): one contiguous run per row
    const float* bdr = (a.mode == B200SLU_RELPOS_LATEST)
                           ? bd + (hb * a.Tq + i) * a.ld_bd + b2_rel_shift_src(i, 0, a.Tq, B200SLU_RELPOS_LATEST).col
                           : nullptr;
    float s[NGL][8];
    float mx = -INFINITY;
#pragma unroll
    for (int k = 0; k < NGL; ++k) {
      const int g = l + LPR * k;
#pragma unroll
      for (int e = 0; e < 8; ++e) s[k][e] = -INFINITY;
      if (g < ngroups && g * 8 < nvalid) {
        float v[8], t[8];
        load8_f32(ac + r * a.ld_ac + g * 8, v);
#pragma unroll
        for (int e = 0; e < 8; ++e) t[e] = (bdr != nullptr && g * 8 + e < nvalid) ? bdr[g * 8 + e] : 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          if (g * 8 + e < nvalid) {
            s[k][e] = (v[e] + t[e]) * a.scale;
            mx = fmaxf(mx, s[k][e]);
          }
        }
      }
    }
    mx = seg_max<LPR>(mx);
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < NGL; ++k) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        s[k][e] = (s[k][e] == -INFINITY) ? 0.f : __expf(s[k][e] - mx);
        sum += s[k][e];
      }
    }
    sum = seg_sum<LPR>(sum);
    const float inv = sum > 0.f ? 1.f / sum : 0.f;
    if (live) {
#pragma unroll
      for (int k = 0; k < NGL; ++k) {
        const int g = l + LPR * k;
        if (g < ngroups) {
#pragma unroll
          for (int e = 0; e < 8; ++e) s[k][e] *= inv;
          store8_bf16(p + r * a.ld_p + g * 8, s[k]);
          if (pd != p) {
            if (drop.p > 0.f) {
              const uint32_t keep = dropout_keep8(dk, (unsigned long long)(r * a.ld_p + g * 8) >> 3);
#pragma unroll
              for (int e = 0; e < 8; ++e) s[k][e] = ((keep >> e) & 1u) ? s[k][e] * dk.inv : 0.f;
            }
            store8_bf16(pd + r * a.ld_p + g * 8, s[k]);
          }
        }
      }
    }
  }
}

template <int LPR, int NGL>
__global__ void __launch_bounds__(256) softmax_bwd_rows_kernel(const float* __restrict__ dpd, const bf16* __restrict__ p,
                                                               SmArgs a, bf16* __restrict__ dac, bf16* __restrict__ dbd,
                                                               B2Dropout drop) {
  extern __shared__ float sm[];  // [warps][32 / LPR][ld_p] staged d(score) rows for the un-shift scatter
  constexpr int RPW = 32 / LPR;
  const int lane = threadIdx.x & 31, sub = lane / LPR, l = lane % LPR, w = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const long long rows = (long long)a.H * a.B * a.Tq;
  const int ngroups = (int)(a.ld_p >> 3);
  const int T = a.Tq;
  float* row_sm = sm + ((size_t)w * RPW + sub) * a.ld_p;
  B2DropKey dk = {0ull, 0u, 1.f};
  if (drop.p > 0.f) dk = dropout_key(drop);
  for (long long rw = (long long)blockIdx.x * wpb + w; rw * RPW < rows; rw += (long long)gridDim.x * wpb) {
    const long long r = rw * RPW + sub;
    const bool live = r < rows;
    float pv[NGL][8], dp[NGL][8];
    float dot = 0.f;
#pragma unroll
    for (int k = 0; k < NGL; ++k) {
      const int g = l + LPR * k;
#pragma unroll
      for (int e = 0; e < 8; ++e) pv[k][e] = dp[k][e] = 0.f;
      if (live && g < ngroups) {
        load8_bf16(p + r * a.ld_p + g * 8, pv[k]);
        load8_f32(dpd + r * a.ld_p + g * 8, dp[k]);
      }
    }
#pragma unroll
    for (int k = 0; k < NGL; ++k) {
      const int g = l + LPR * k;
      if (live && g < ngroups) {
        if (drop.p > 0.f) {
          const uint32_t keep = dropout_keep8(dk, (unsigned long long)(r * a.ld_p + g * 8) >> 3);
#pragma unroll
          for (int e = 0; e < 8; ++e) dp[k][e] = ((keep >> e) & 1u) ? dp[k][e] * dk.inv : 0.f;
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          if (g * 8 + e >= a.Tk) dp[k][e] = 0.f;  // padding columns of dpd are unspecified
          dot += dp[k][e] * pv[k][e];
        }
      }
    }
    dot = seg_sum<LPR>(dot);
#pragma unroll
    for (int k = 0; k < NGL; ++k) {
      const int g = l + LPR * k;
      if (live && g < ngroups) {
        float ds[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) ds[e] = pv[k][e] * (dp[k][e] - dot) * a.scale;
        if (dbd != nullptr) {
          float4* dst = reinterpret_cast<float4*>(row_sm + g * 8);
          dst[0] = make_float4(ds[0], ds[1], ds[2], ds[3]);
          dst[1] = make_float4(ds[4], ds[5], ds[6], ds[7]);
        }
        store8_bf16(dac + r * a.ld_p + g * 8, ds);
      }
    }
    if (dbd != nullptr) {  // latest: dbd[i, n] = ds[i, n - (T-1-i)] inside the band, 0 outside; 8 outputs per lane and store
      __syncwarp();
      if (live) {
        const long long hb = r / a.Tq;
        const int i = (int)(r - hb * a.Tq);
        bf16* own = dbd + (hb * T + i) * a.ld_bd;
        const int lo = b2_rel_shift_src(i, 0, T, B200SLU_RELPOS_LATEST).col;
        for (int gq = l; gq < (int)(a.ld_bd >> 3); gq += LPR) {
          float o[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const int j = gq * 8 + e - lo;
            o[e] = (j >= 0 && j < T) ? row_sm[j] : 0.f;
          }
          store8_bf16(own + gq * 8, o);
        }
      }
      __syncwarp();
    }
  }
}

int check(const SmArgs& a) {
  B2_CHECK_ARG(a.H > 0 && a.B > 0 && a.Tq > 0 && a.Tk > 0, "softmax: empty problem");
  B2_CHECK_ARG(a.ld_p % 8 == 0 && a.ld_p >= a.Tk && a.ld_ac >= a.ld_p && a.ld_ac % 4 == 0,
               "softmax: ld_p %% 8 == 0, ld_p >= Tk, ld_ac >= ld_p (ld_p=%lld ld_ac=%lld Tk=%d)", a.ld_p, a.ld_ac, a.Tk);
  B2_CHECK_ARG(a.ld_p <= 2048, "softmax: at most 2048 keys per row (got ld %lld)", a.ld_p);
  if (a.mode == B200SLU_RELPOS_LATEST)
    B2_CHECK_ARG(a.Tq == a.Tk && a.ld_bd >= 2 * a.Tq - 1, "softmax: latest rel-pos needs Tq == Tk and ld_bd >= 2T-1");
  else if (a.mode == B200SLU_RELPOS_LEGACY)
    B2_CHECK_ARG(a.Tq == a.Tk && a.ld_bd >= a.Tq, "softmax: legacy rel-pos needs Tq == Tk and ld_bd >= T");
  else
    B2_CHECK_ARG(a.mode == B200SLU_RELPOS_NONE, "softmax: unknown rel-pos mode %d", a.mode);
  return 0;
}

// d(matrix_bd) (un-shifted layout) from d(scores) (shifted layout): the rel_shift scatter alone (latest mode), for
// callers that already hold d(scores) -- dbd[i, n] = ds[i, n - src(i, 0).col] inside the band, 0 outside.
__global__ void __launch_bounds__(256) rel_unshift_kernel(const bf16* __restrict__ ds, long long ld_p, long long rows, int T,
                                                          bf16* __restrict__ dbd, long long ld_bd) {
  const int gpr = (int)(ld_bd >> 3);
  const long long total = rows * gpr;
  for (long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x; w < total; w += (long long)gridDim.x * blockDim.x) {
    const long long r = w / gpr;
    const int g = (int)(w - r * gpr);
    const int i = (int)(r % T);
    const int lo = b2_rel_shift_src(i, 0, T, B200SLU_RELPOS_LATEST).col;
    const bf16* src = ds + r * ld_p;
    float o[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int j = g * 8 + e - lo;
      o[e] = (j >= 0 && j < T) ? __bfloat162float(src[j]) : 0.f;
    }
    store8_bf16(dbd + r * ld_bd + g * 8, o);
  }
}

}  // namespace

extern "C" int b200slu_softmax_fwd(const float* ac, long long ld_ac, const float* bd, long long ld_bd, const int* key_lens, int H,
                                   int B, int Tq, int Tk, int relpos_mode, int causal, float scale, void* p, void* p_drop,
                                   long long ld_p, float drop_p, unsigned long long seed, unsigned int site, void* stream) {
  SmArgs a{H, B, Tq, Tk, ld_ac, ld_bd, ld_p, relpos_mode, causal, scale, key_lens};
  B2_TRY(check(a));
  B2_CHECK_ARG(p != nullptr && p_drop != nullptr && (drop_p > 0.f) == (p != p_drop),
               "softmax_fwd: p_drop must alias p iff dropout is off");
  B2_CHECK_ARG(relpos_mode == B200SLU_RELPOS_NONE || bd != nullptr, "softmax_fwd: bd missing");
  cudaStream_t st = (cudaStream_t)stream;
  const long long rows = (long long)H * B * Tq;
  const int grid = b2_grid_for(rows * 32, 256, 8);
  B2_SCOPE(FAM_ATTN, st, 0, (double)rows * Tk * (4.0 + (relpos_mode ? 4.0 : 0.0) + 2.0 + (drop_p > 0.f ? 2.0 : 0.0)));
  B2Dropout drop = b2_make_drop(drop_p, seed, site);
#define SM_FR(LPR, NGL)                                                                                            \
  softmax_fwd_rows_kernel<LPR, NGL><<<b2_grid_for((rows + 32 / LPR - 1) / (32 / LPR) * 32, 256, 8), 256, 0, st>>>( \
      ac, bd, a, (bf16*)p, (bf16*)p_drop, drop)
  if (relpos_mode != B200SLU_RELPOS_LEGACY && ld_p <= 512) {
    if (ld_p <= 192) SM_FR(8, 3);
    else if (ld_p <= 256) SM_FR(8, 4);
    else SM_FR(16, 4);
    B2_LAUNCH_CHECK();
    return 0;
  }
#undef SM_FR
#define SM_F(NG) softmax_fwd_kernel<NG><<<grid, 256, 0, st>>>(ac, bd, a, (bf16*)p, (bf16*)p_drop, drop)
  if (ld_p <= 256) SM_F(1);
  else if (ld_p <= 512) SM_F(2);
  else if (ld_p <= 1024) SM_F(4);
  else SM_F(8);
#undef SM_F
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_softmax_bwd(const float* dp_drop, const void* p, long long ld_p, int H, int B, int Tq, int Tk,
                                   int relpos_mode, float scale, void* d_ac, void* d_bd, long long ld_bd, float drop_p,
                                   unsigned long long seed, unsigned int site, void* stream) {
  SmArgs a{H, B, Tq, Tk, ld_p, ld_bd, ld_p, relpos_mode, 0, scale, nullptr};
  B2_TRY(check(a));
  B2_CHECK_ARG(ld_bd % 8 == 0 || relpos_mode == B200SLU_RELPOS_NONE, "softmax_bwd: ld_bd %% 8 == 0");
  B2_CHECK_ARG((relpos_mode == B200SLU_RELPOS_NONE) == (d_bd == nullptr), "softmax_bwd: d_bd must be given iff rel-pos mode");
  cudaStream_t st = (cudaStream_t)stream;
  const long long rows = (long long)H * B * Tq;
  const int threads = 256;
  const size_t smem = d_bd != nullptr ? (size_t)(threads / 32) * ld_p * sizeof(float) : 0;
  const int grid = b2_grid_for(rows * 32, threads, 4);
  B2_SCOPE(FAM_ATTN, st, 0, (double)rows * Tk * (4.0 + 2.0 + 2.0) + (d_bd ? 2.0 * rows * ld_bd : 0.0));
  B2Dropout drop = b2_make_drop(drop_p, seed, site);
#define SM_BR(LPR, NGL)                                                                                                   \
  do {                                                                                                                    \
    const size_t smr = d_bd != nullptr ? (size_t)(threads / 32) * (32 / LPR) * ld_p * sizeof(float) : 0;                  \
    softmax_bwd_rows_kernel<LPR, NGL><<<b2_grid_for((rows + 32 / LPR - 1) / (32 / LPR) * 32, threads, 4), threads, smr, st>>>( \
        dp_drop, (const bf16*)p, a, (bf16*)d_ac, (bf16*)d_bd, drop);                                                      \
  } while (0)
  if (relpos_mode != B200SLU_RELPOS_LEGACY && ld_p <= 512) {
    if (ld_p <= 192) SM_BR(8, 3);
    else if (ld_p <= 256) SM_BR(8, 4);
    else SM_BR(16, 4);
    B2_LAUNCH_CHECK();
    return 0;
  }
#undef SM_BR
#define SM_B(NG)                                                                                                  \
  do {                                                                                                            \
    static unsigned long long cfg = 0;                                                                                      \
    if (!b2_device_done(cfg)) {                                                                                                   \
      B2_CUDA(cudaFuncSetAttribute(softmax_bwd_kernel<NG>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536)); \
      b2_device_mark(cfg);                                                                                                 \
    }                                                                                                             \
    softmax_bwd_kernel<NG><<<grid, threads, smem, st>>>(dp_drop, (const bf16*)p, a, (bf16*)d_ac, (bf16*)d_bd, drop); \
  } while (0)
  if (ld_p <= 256) SM_B(1);
  else if (ld_p <= 512) SM_B(2);
  else if (ld_p <= 1024) SM_B(4);
  else SM_B(8);
#undef SM_B
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_rel_unshift(const void* d_scores, long long ld_p, int H, int B, int T, int relpos_mode, void* d_bd,
                                   long long ld_bd, void* stream) {
  B2_CHECK_ARG(relpos_mode == B200SLU_RELPOS_LATEST, "rel_unshift: latest mode only");
  B2_CHECK_ARG(H > 0 && B > 0 && T > 0 && ld_p >= T && ld_bd >= 2 * T - 1 && ld_bd % 8 == 0, "rel_unshift: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  const long long rows = (long long)H * B * T;
  B2_SCOPE(FAM_ATTN, st, 0, (double)rows * (2.0 * T + 2.0 * ld_bd));
  rel_unshift_kernel<<<b2_grid_for(rows * (ld_bd >> 3), 256, 8), 256, 0, st>>>((const bf16*)d_scores, ld_p, rows, T, (bf16*)d_bd, ld_bd);
  B2_LAUNCH_CHECK();
  return 0;
}
