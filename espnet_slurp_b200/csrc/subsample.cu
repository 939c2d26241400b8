// Conv2dSubsampling front of the encoder (espnet/nets/pytorch_backend/transformer/subsampling.py:53-87):
//   conv.0 = Conv2d(1, C, 3, stride 2) + ReLU  -> direct memory-bound kernel, channels-last bf16 output
//   conv.2 = Conv2d(C, C, 3, stride 2) + ReLU  -> im2col here, contraction on the tcgen05 GEMM
//            (M = B*T*F2, K = 9*C with k = (kh, kw, c), N = C; ReLU in the GEMM epilogue)
//   out.0  = Linear(C*F2, d) on the GEMM directly over the [B*T, F2*C] view (weight K-dim permuted
//            once at cast time), so the reference's transpose(1,2).contiguous() copy (:84) never happens.
// Backward: col2im (+ ReLU mask of conv.0) gathers the <= 4 taps per input pixel; conv.0's weight
// gradient is a per-channel reduction over all output pixels.
#include "common.cuh"

namespace {

// x [B, Tin, F] f32 -> h1 [B, T1, F1, C] bf16 ;  grid (ceil(T1 / TR1), B).  Each thread keeps the 9 taps + bias of its
// 8 channels in registers and walks the output pixels of TR1 rows; the 2*TR1+1 input rows sit in shared memory
// (broadcast reads), every store is one 16-byte vector and a warp writes 512 contiguous bytes.
constexpr int TR1 = 8;
__global__ void __launch_bounds__(256) conv1_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                        const float* __restrict__ bias, int Tin, int F, int C, int T1, int F1,
                                                        bf16* __restrict__ h1) {
  extern __shared__ float sx[];  // [2*TR1+1][F]
  const int tb = blockIdx.x * TR1, b = blockIdx.y;
  const int nt = min(TR1, T1 - tb), nrows = 2 * nt + 1;
  for (int i = threadIdx.x; i < nrows * F; i += blockDim.x) sx[i] = x[((long long)b * Tin + 2 * tb) * F + i];
  __syncthreads();
  const int ncg = C >> 3;
  const int npl = blockDim.x / ncg;  // position lanes (4 for C = 512)
  for (int cg = threadIdx.x % ncg, pl = threadIdx.x / ncg; pl < npl && cg < ncg; cg += ncg) {
    float wr[9][8], bs[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      bs[j] = bias[cg * 8 + j];
#pragma unroll
      for (int k = 0; k < 9; ++k) wr[k][j] = w[(cg * 8 + j) * 9 + k];
    }
    for (int p = pl; p < nt * F1; p += npl) {
      const int tl = p / F1, f = p - tl * F1;
      float acc[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = bs[j];
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        const float xv = sx[(2 * tl + k / 3) * F + 2 * f + (k % 3)];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += xv * wr[k][j];
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = fmaxf(acc[j], 0.f);
      store8_bf16(h1 + (((long long)b * T1 + tb + tl) * F1 + f) * C + cg * 8, acc);
    }
  }
}

// dW[c][k] += sum_{b,t,f} dh1[b,t,f,c] * x[b,2t+kh,2f+kw] ; db[c] += sum dh1   (dh1 already ReLU-masked)
// grid (ceil(T1 / ROWS), B)
constexpr int ROWS = 16;
__global__ void __launch_bounds__(256) conv1_bwd_kernel(const bf16* __restrict__ dh1, const float* __restrict__ x, int Tin, int F,
                                                        int C, int T1, int F1, float* __restrict__ dw, float* __restrict__ db) {
  extern __shared__ float sm[];
  float* sacc = sm;             // [C][10]
  float* sx = sm + 10 * C;      // [3][F]
  const int b = blockIdx.y, tbeg = blockIdx.x * ROWS, tend = min(T1, tbeg + ROWS);
  const int ncg = C >> 3;
  const int cg = threadIdx.x % ncg, pl = threadIdx.x / ncg, npl = blockDim.x / ncg;
  for (int i = threadIdx.x; i < 10 * C; i += blockDim.x) sacc[i] = 0.f;
  float acc[10][8];
#pragma unroll
  for (int k = 0; k < 10; ++k)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[k][j] = 0.f;
  for (int t = tbeg; t < tend; ++t) {
    __syncthreads();
    for (int i = threadIdx.x; i < 3 * F; i += blockDim.x) {
      const int kh = i / F, f = i - kh * F;
      sx[i] = x[((long long)b * Tin + 2 * t + kh) * F + f];
    }
    __syncthreads();
    if (pl < npl) {
      constexpr int U = 5;  // independent 16-byte loads in flight per thread (the kernel is latency-bound otherwise)
      for (int f0 = pl; f0 < F1; f0 += npl * U) {
        uint4 raw[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int f = f0 + u * npl;
          raw[u] = f < F1 ? *reinterpret_cast<const uint4*>(dh1 + (((long long)b * T1 + t) * F1 + f) * C + cg * 8)
                          : make_uint4(0u, 0u, 0u, 0u);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int f = min(f0 + u * npl, F1 - 1);  // out-of-range slots carry zeros
          const float2 g0 = unpack_bf16x2(raw[u].x), g1 = unpack_bf16x2(raw[u].y), g2 = unpack_bf16x2(raw[u].z),
                       g3 = unpack_bf16x2(raw[u].w);
          const float g[8] = {g0.x, g0.y, g1.x, g1.y, g2.x, g2.y, g3.x, g3.y};
#pragma unroll
          for (int k = 0; k < 9; ++k) {
            const float xv = sx[(k / 3) * F + 2 * f + (k % 3)];
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[k][j] += g[j] * xv;
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[9][j] += g[j];
        }
      }
    }
  }
  if (pl < npl) {
#pragma unroll
    for (int k = 0; k < 10; ++k)
#pragma unroll
      for (int j = 0; j < 8; ++j) atomicAdd(&sacc[(cg * 8 + j) * 10 + k], acc[k][j]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 10 * C; i += blockDim.x) {
    const int c = i / 10, k = i - c * 10;
    if (k < 9) atomicAdd(dw + c * 9 + k, sacc[i]);
    else atomicAdd(db + c, sacc[i]);
  }
}

// col[(b,t,f), (kh,kw,c)] = h1[b, 2t+kh, 2f+kw, c]
__global__ void __launch_bounds__(256) im2col_kernel(const bf16* __restrict__ h1, int B, int T1, int F1, int C, int T2, int F2,
                                                     bf16* __restrict__ col) {
  const int ncg = C >> 3;
  const long long total = (long long)B * T2 * F2 * 9 * ncg;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % ncg);
    long long r = i / ncg;
    const int k = (int)(r % 9);
    r /= 9;  // output pixel (b, t, f)
    const int f = (int)(r % F2);
    const long long bt = r / F2;
    const int t = (int)(bt % T2);
    const long long b = bt / T2;
    const int kh = k / 3, kw = k - kh * 3;
    const uint4 v = *reinterpret_cast<const uint4*>(h1 + (((b * T1 + 2 * t + kh) * F1) + 2 * f + kw) * C + cg * 8);
    *reinterpret_cast<uint4*>(col + (r * 9 + k) * C + cg * 8) = v;
  }
}

// dh1[b,r,g,c] = (h1 > 0) * sum over taps (kh,kw) with 2t+kh = r, 2f+kw = g of dcol[(b,t,f),(kh,kw,c)]
__global__ void __launch_bounds__(256) col2im_relu_kernel(const bf16* __restrict__ dcol, const bf16* __restrict__ h1, int B, int T1,
                                                          int F1, int C, int T2, int F2, bf16* __restrict__ dh1) {
  const int ncg = C >> 3;
  const long long total = (long long)B * T1 * F1 * ncg;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % ncg);
    long long p = i / ncg;
    const int g = (int)(p % F1);
    const long long br = p / F1;
    const int r = (int)(br % T1);
    const long long b = br / T1;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
      const int tt = r - kh;
      if (tt < 0 || (tt & 1) || (tt >> 1) >= T2) continue;
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const int ff = g - kw;
        if (ff < 0 || (ff & 1) || (ff >> 1) >= F2) continue;
        float v[8];
        load8_bf16(dcol + ((((b * T2 + (tt >> 1)) * F2 + (ff >> 1)) * 9) + kh * 3 + kw) * C + cg * 8, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += v[j];
      }
    }
    float h[8];
    load8_bf16(h1 + p * C + cg * 8, h);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = h[j] > 0.f ? acc[j] : 0.f;
    store8_bf16(dh1 + p * C + cg * 8, acc);
  }
}

}  // namespace

extern "C" int b200slu_conv1_fwd(const float* x, const float* w, const float* bias, int B, int Tin, int F, int C, void* h1,
                                 void* stream) {
  B2_CHECK_ARG(B > 0 && Tin >= 3 && F >= 3 && C % 8 == 0 && B <= 65535, "conv1_fwd: bad shape B=%d Tin=%d F=%d C=%d", B, Tin, F, C);
  const int T1 = (Tin - 3) / 2 + 1, F1 = (F - 3) / 2 + 1;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem = (size_t)(2 * TR1 + 1) * F * sizeof(float);
  B2_CHECK_ARG(smem <= 48 * 1024 && C <= 2048, "conv1_fwd: F or C too large");
  B2_SCOPE(FAM_SUBSAMPLE, st, 18.0 * B * T1 * F1 * C, 4.0 * B * Tin * F + 2.0 * B * T1 * F1 * C);
  conv1_fwd_kernel<<<dim3((T1 + TR1 - 1) / TR1, B), 256, smem, st>>>(x, w, bias, Tin, F, C, T1, F1, (bf16*)h1);
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_conv1_bwd(const void* dh1, const float* x, int B, int Tin, int F, int C, float* dw_accum, float* db_accum,
                                 void* stream) {
  B2_CHECK_ARG(B > 0 && Tin >= 3 && F >= 3 && C % 8 == 0 && C <= 2048 && B <= 65535, "conv1_bwd: bad shape");
  const int T1 = (Tin - 3) / 2 + 1, F1 = (F - 3) / 2 + 1;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem = (size_t)(10 * C + 3 * F) * sizeof(float);
  B2_CHECK_ARG(smem <= 48 * 1024, "conv1_bwd: C too large");
  B2_SCOPE(FAM_SUBSAMPLE, st, 20.0 * B * T1 * F1 * C, 2.0 * B * T1 * F1 * C);
  conv1_bwd_kernel<<<dim3((T1 + ROWS - 1) / ROWS, B), 256, smem, st>>>((const bf16*)dh1, x, Tin, F, C, T1, F1, dw_accum,
                                                                       db_accum);
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_im2col3x3s2(const void* h1, int B, int T1, int F1, int C, void* col, void* stream) {
  B2_CHECK_ARG(B > 0 && T1 >= 3 && F1 >= 3 && C % 8 == 0, "im2col: bad shape");
  const int T2 = (T1 - 3) / 2 + 1, F2 = (F1 - 3) / 2 + 1;
  cudaStream_t st = (cudaStream_t)stream;
  const long long items = (long long)B * T2 * F2 * 9 * (C >> 3);
  B2_SCOPE(FAM_SUBSAMPLE, st, 0, 32.0 * items);
  im2col_kernel<<<b2_grid_for(items, 256, 16), 256, 0, st>>>((const bf16*)h1, B, T1, F1, C, T2, F2, (bf16*)col);
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_col2im3x3s2_relu(const void* dcol, const void* h1, int B, int T1, int F1, int C, void* dh1, void* stream) {
  B2_CHECK_ARG(B > 0 && T1 >= 3 && F1 >= 3 && C % 8 == 0, "col2im: bad shape");
  const int T2 = (T1 - 3) / 2 + 1, F2 = (F1 - 3) / 2 + 1;
  cudaStream_t st = (cudaStream_t)stream;
  const long long items = (long long)B * T1 * F1 * (C >> 3);
  B2_SCOPE(FAM_SUBSAMPLE, st, 0, 16.0 * B * T2 * F2 * 9 * (C >> 3) + 32.0 * items);
  col2im_relu_kernel<<<b2_grid_for(items, 256, 16), 256, 0, st>>>((const bf16*)dcol, (const bf16*)h1, B, T1, F1, C, T2, F2,
                                                                  (bf16*)dh1);
  B2_LAUNCH_CHECK();
  return 0;
}
