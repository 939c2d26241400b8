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
constexpr int TR1 = 16;
__global__ void __launch_bounds__(256) conv1_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                        const float* __restrict__ bias, int Tin, int F, int C, int T1, int F1,
                                                        bf16* __restrict__ h1, uint32_t* __restrict__ relu_bits) {
  extern __shared__ float sx[];  // [2*TR1+1][F]
  const int tb = blockIdx.x * TR1, b = blockIdx.y;
  const int nt = min(TR1, T1 - tb), nrows = 2 * nt + 1;
  for (int i = threadIdx.x; i < nrows * F; i += blockDim.x) sx[i] = x[((long long)b * Tin + 2 * tb) * F + i];
  __syncthreads();
  const int ncg = C >> 3;
  const int npl = blockDim.x / ncg;  // position lanes (4 for C = 512)
  for (int cg = threadIdx.x % ncg, pl = threadIdx.x / ncg; pl < npl && cg < ncg; cg += ncg) {
    // The kernel is issue-bound (72 FMAs per 16 bytes stored): the taps run as packed FFMA2 (fma.rn.f32x2, two
    // IEEE FMAs per instruction -- same rounding as the scalar form), channel pairs (2j, 2j+1) per register pair.
    float2 wr[9][4], bs[4];
    {
      // this thread's 8 x 9 taps are 72 consecutive floats of w: 18 vector loads instead of 72 strided scalar ones
      float wf[72];
      const float4* wsrc = reinterpret_cast<const float4*>(w + cg * 72);
#pragma unroll
      for (int q = 0; q < 18; ++q) {
        const float4 v = wsrc[q];
        wf[4 * q] = v.x; wf[4 * q + 1] = v.y; wf[4 * q + 2] = v.z; wf[4 * q + 3] = v.w;
      }
      const float4 b0 = *reinterpret_cast<const float4*>(bias + cg * 8), b1 = *reinterpret_cast<const float4*>(bias + cg * 8 + 4);
      bs[0] = make_float2(b0.x, b0.y); bs[1] = make_float2(b0.z, b0.w); bs[2] = make_float2(b1.x, b1.y); bs[3] = make_float2(b1.z, b1.w);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int k = 0; k < 9; ++k) wr[k][j] = make_float2(wf[(2 * j) * 9 + k], wf[(2 * j + 1) * 9 + k]);
      }
    }
    // pixel p = tl * F1 + f of the block's nt x F1 output rows; the walk steps (tl, f), the staged-input offset and the
    // output pointers instead of dividing per pixel (the division and the 64-bit index were 27 % of the instructions)
    int tl = pl / F1, f = pl - tl * F1;
    const float* sxp = sx + (2 * tl) * F + 2 * f;
    const long long pix0 = ((long long)b * T1 + tb) * F1;
    bf16* hp = h1 + (pix0 + pl) * C + cg * 8;
    unsigned char* bp = relu_bits != nullptr ? reinterpret_cast<unsigned char*>(relu_bits) + (pix0 + pl) * (C >> 3) + cg : nullptr;
    const long long hstep = (long long)npl * C;
    const int bstep = npl * (C >> 3), wrap = 2 * F - 2 * F1;
    for (int p = pl; p < nt * F1; p += npl) {
      float2 acc2[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) acc2[j] = bs[j];
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        const float xv = sxp[(k / 3) * F + (k % 3)];
        const float2 xx = make_float2(xv, xv);
#pragma unroll
        for (int j = 0; j < 4; ++j) acc2[j] = b2_fma2(xx, wr[k][j], acc2[j]);
      }
      uint32_t pk[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) pk[j] = pack_bf16x2(fmaxf(acc2[j].x, 0.f), fmaxf(acc2[j].y, 0.f));
      *reinterpret_cast<uint4*>(hp) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      hp += hstep;
      if (bp != nullptr) {
        // bit c of the pixel's C-bit row = (h1[c] > 0): what conv.2's dgrad needs of h1, at 1/16 of the bytes.  Taken
        // from the stored bf16 pairs (non-negative after the ReLU, so > 0 <=> != 0): adding 0x7fff to a half sets its top
        // bit iff it is nonzero and cannot carry into the neighbour.
        uint32_t x = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) x |= ((pk[j] + 0x7fff7fffu) & 0x80008000u) >> (15 - 2 * j);  // halves of pair j -> bits 2j, 2j+16
        // one byte per lane: a warp's 32 bytes are contiguous (little-endian words: byte cg % 4 of word cg / 4)
        *bp = (unsigned char)((x | (x >> 15)) & 0xffu);
        bp += bstep;
      }
      f += npl;
      sxp += 2 * npl;
      while (f >= F1) {  // next output row: two input rows down, back to the row's start
        f -= F1;
        sxp += wrap;
      }
    }
  }
}

// bits[p][c / 32] bit (c % 32) = (h1[p][c] > 0), for callers that hold only h1
__global__ void __launch_bounds__(256) relu_bits_kernel(const bf16* __restrict__ h1, long long words, uint32_t* __restrict__ bits) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (long long)gridDim.x * blockDim.x) {
    const uint4* src = reinterpret_cast<const uint4*>(h1 + i * 32);
    uint32_t m = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint4 v = src[q];
      const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f2 = unpack_bf16x2(w4[e]);
        m |= (f2.x > 0.f ? 1u : 0u) << (q * 8 + e * 2);
        m |= (f2.y > 0.f ? 1u : 0u) << (q * 8 + e * 2 + 1);
      }
    }
    bits[i] = m;
  }
}

// dW[c][k] += sum_{b,t,f} dh1[b,t,f,c] * x[b,2t+kh,2f+kw] ; db[c] += sum dh1   (dh1 already ReLU-masked)
// A reduction over B*T1*F1 pixels of a [pixels, C] matrix against 9 taps (+ a column of ones for the bias): a
// wgrad-shaped GEMM  [C, 16] = dh1^T [C, pixels] . patches [pixels, 16]  on the tensor cores (split-K over the
// pixels), bound by reading dh1 once from HBM.  (The CUDA-core version was bound by 80 FMAs per 16 bytes of dh1 and by
// 2560 blocks x 5120 same-address atomics: 1.6 ms against 0.25 ms of HBM time at 128 utterances.)
// patches[p, k] = bf16(x[b, 2t+kh, 2f+kw]) for k = kh*3+kw < 9, 1 for k = 9, 0 for k = 10..15.
__global__ void __launch_bounds__(256) conv1_patch_kernel(const float* __restrict__ x, int B, int Tin, int F, int T1, int F1,
                                                          bf16* __restrict__ patches) {
  const long long total = (long long)B * T1 * F1;
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (long long)gridDim.x * blockDim.x) {
    const int f = (int)(p % F1);
    const long long bt = p / F1;
    const int t = (int)(bt % T1);
    const long long b = bt / T1;
    const float* xp = x + (b * Tin + 2 * t) * F + 2 * f;
    float v[16];
#pragma unroll
    for (int k = 0; k < 9; ++k) v[k] = xp[(k / 3) * F + (k % 3)];
    v[9] = 1.f;
#pragma unroll
    for (int k = 10; k < 16; ++k) v[k] = 0.f;
    store8_bf16(patches + p * 16, *reinterpret_cast<float (*)[8]>(&v[0]));
    store8_bf16(patches + p * 16 + 8, *reinterpret_cast<float (*)[8]>(&v[8]));
  }
}

// dw[c*9 + k] += acc[c][k] (k < 9), db[c] += acc[c][9]
__global__ void __launch_bounds__(256) conv1_wgrad_scatter_kernel(const float* __restrict__ acc, int C, float* __restrict__ dw,
                                                                  float* __restrict__ db) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < C * 10) {
    const int c = i / 10, k = i - c * 10;
    if (k < 9) dw[c * 9 + k] += acc[c * 16 + k];
    else db[c] += acc[c * 16 + 9];
  }
}

// col[(b,t,f), (kh,kw,c)] = h1[b, 2t+kh, 2f+kw, c]
// grid (T2, B): a block copies one output row (b, t): F2 pixels x 9 taps x C channels.  A warp takes (pixel, tap)
// pairs -- each a contiguous C-channel run on both sides -- four at a time, so every lane has >= 4 independent 16-byte
// loads in flight and the index arithmetic is per warp, not per element.
__global__ void __launch_bounds__(256) im2col_kernel(const bf16* __restrict__ h1, int B, int T1, int F1, int C, int T2, int F2,
                                                     bf16* __restrict__ col) {
  const int t = blockIdx.x, b = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int ncg = C >> 3, npairs = F2 * 9;
  const bf16* src_row = h1 + ((long long)b * T1 + 2 * t) * F1 * C;
  bf16* dst_row = col + ((long long)b * T2 + t) * F2 * 9 * C;
  constexpr int U = 4;
  for (int p0 = warp; p0 < npairs; p0 += nwarps * U) {
    const bf16* src[U];
    bf16* dst[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int pk = min(p0 + u * nwarps, npairs - 1);
      const int f = pk / 9, k = pk - f * 9, kh = k / 3, kw = k - kh * 3;
      src[u] = src_row + ((long long)kh * F1 + 2 * f + kw) * C;
      dst[u] = dst_row + (long long)pk * C;
    }
    for (int cg = lane; cg < ncg; cg += 32) {
      uint4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) v[u] = *reinterpret_cast<const uint4*>(src[u] + cg * 8);
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (p0 + u * nwarps < npairs) *reinterpret_cast<uint4*>(dst[u] + cg * 8) = v[u];
    }
  }
}

// dh1[b,r,g,c] = (h1 > 0) * sum over taps (kh,kw) with 2t+kh = r, 2f+kw = g of dcol[(b,t,f),(kh,kw,c)]
// grid (T1, B): a block produces one row (b, r) of dh1; a warp takes pixels g.  Which taps contribute is uniform per
// block (kh) and per warp (kw); all (<= 4) tap loads and the ReLU-mask load of a lane are issued before any use.
__global__ void __launch_bounds__(256) col2im_relu_kernel(const bf16* __restrict__ dcol, const bf16* __restrict__ h1, int B, int T1,
                                                          int F1, int C, int T2, int F2, bf16* __restrict__ dh1) {
  const int r = blockIdx.x, b = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int ncg = C >> 3;
  // rows of dcol that feed this h1 row: r = 2t + kh
  // r even -> kh in {0, 2}, r odd -> kh = 1: two fixed slots with validity flags
  const int kh0 = (r & 1) ? 1 : 0;
  const int khs[2] = {kh0, 2}, ts[2] = {(r - kh0) >> 1, (r - 2) >> 1};
  const bool okt[2] = {ts[0] < T2, !(r & 1) && r >= 2 && ts[1] < T2};
  const long long row = ((long long)b * T1 + r) * F1;
  for (int g = warp; g < F1; g += nwarps) {
    // taps along f: g even -> kw in {0, 2}, g odd -> kw = 1; slot q = 2 * a + (0 | 1) with a the kh slot
    const int kw0 = (g & 1) ? 1 : 0, f0 = (g - kw0) >> 1, f1 = (g - 2) >> 1;
    const bool okf0 = f0 < F2, okf1 = !(g & 1) && g >= 2 && f1 < F2;
    const bf16* src[4];
    bool ok[4];
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      const long long base = ((long long)b * T2 + ts[a]) * F2;
      src[2 * a] = dcol + ((base + f0) * 9 + khs[a] * 3 + kw0) * C;
      src[2 * a + 1] = dcol + ((base + f1) * 9 + khs[a] * 3 + 2) * C;
      ok[2 * a] = okt[a] && okf0;
      ok[2 * a + 1] = okt[a] && okf1;
    }
    for (int cg = lane; cg < ncg; cg += 32) {
      uint4 v[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) v[q] = ok[q] ? *reinterpret_cast<const uint4*>(src[q] + cg * 8) : make_uint4(0u, 0u, 0u, 0u);
      const uint4 hv = *reinterpret_cast<const uint4*>(h1 + (row + g) * C + cg * 8);
      float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t w4[4] = {v[q].x, v[q].y, v[q].z, v[q].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 p2 = unpack_bf16x2(w4[e]);
          acc[2 * e] += p2.x;
          acc[2 * e + 1] += p2.y;
        }
      }
      const uint32_t h4[4] = {hv.x, hv.y, hv.z, hv.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 hh = unpack_bf16x2(h4[e]);
        if (!(hh.x > 0.f)) acc[2 * e] = 0.f;
        if (!(hh.y > 0.f)) acc[2 * e + 1] = 0.f;
      }
      store8_bf16(dh1 + (row + g) * C + cg * 8, acc);
    }
  }
}

}  // namespace

extern "C" int b200slu_conv1_fwd_bits(const float* x, const float* w, const float* bias, int B, int Tin, int F, int C, void* h1,
                                      unsigned int* relu_bits, void* stream);
extern "C" int b200slu_conv1_fwd(const float* x, const float* w, const float* bias, int B, int Tin, int F, int C, void* h1,
                                 void* stream) {
  return b200slu_conv1_fwd_bits(x, w, bias, B, Tin, F, C, h1, nullptr, stream);
}

extern "C" int b200slu_relu_bits(const void* h1, long long pixels, int C, unsigned int* relu_bits, void* stream) {
  B2_CHECK_ARG(pixels > 0 && C > 0 && C % 32 == 0 && ((uintptr_t)h1 % 16) == 0, "relu_bits: C %% 32 == 0, aligned input");
  const long long words = pixels * (C / 32);
  relu_bits_kernel<<<b2_grid_for(words, 256, 8), 256, 0, (cudaStream_t)stream>>>((const bf16*)h1, words, relu_bits);
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_conv1_fwd_bits(const float* x, const float* w, const float* bias, int B, int Tin, int F, int C, void* h1,
                                      unsigned int* relu_bits, void* stream) {
  B2_CHECK_ARG(B > 0 && Tin >= 3 && F >= 3 && C % 8 == 0 && B <= 65535, "conv1_fwd: bad shape B=%d Tin=%d F=%d C=%d", B, Tin, F, C);
  const int T1 = (Tin - 3) / 2 + 1, F1 = (F - 3) / 2 + 1;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem = (size_t)(2 * TR1 + 1) * F * sizeof(float);
  B2_CHECK_ARG(smem <= 48 * 1024 && C <= 2048, "conv1_fwd: F or C too large");
  B2_CHECK_ARG(((uintptr_t)w % 16) == 0 && ((uintptr_t)bias % 16) == 0 && ((uintptr_t)h1 % 16) == 0, "conv1_fwd: 16-byte aligned w, bias, h1");
  B2_SCOPE(FAM_SUBSAMPLE, st, 18.0 * B * T1 * F1 * C, 4.0 * B * Tin * F + 2.0 * B * T1 * F1 * C);
  B2_CHECK_ARG(relu_bits == nullptr || C % 32 == 0, "conv1_fwd: the ReLU bit rows need C %% 32 == 0");
  conv1_fwd_kernel<<<dim3((T1 + TR1 - 1) / TR1, B), 256, smem, st>>>(x, w, bias, Tin, F, C, T1, F1, (bf16*)h1, relu_bits);
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" long long b200slu_conv1_bwd_workspace_size(int B, int Tin, int F, int C) {
  if (B <= 0 || Tin < 3 || F < 3 || C <= 0) return 0;
  const long long T1 = (Tin - 3) / 2 + 1, F1 = (F - 3) / 2 + 1;
  return B * T1 * F1 * 16 * 2 + (long long)C * 16 * 4 + 256;
}

extern "C" int b200slu_conv1_bwd(const void* dh1, const float* x, int B, int Tin, int F, int C, float* dw_accum, float* db_accum,
                                 void* workspace, long long workspace_bytes, void* stream) {
  B2_CHECK_ARG(B > 0 && Tin >= 3 && F >= 3 && C % 8 == 0 && C <= 2048 && B <= 65535, "conv1_bwd: bad shape");
  const int T1 = (Tin - 3) / 2 + 1, F1 = (F - 3) / 2 + 1;
  const long long pixels = (long long)B * T1 * F1;
  B2_CHECK_ARG(workspace != nullptr && ((uintptr_t)workspace % 256) == 0 &&
                   workspace_bytes >= b200slu_conv1_bwd_workspace_size(B, Tin, F, C),
               "conv1_bwd: workspace of b200slu_conv1_bwd_workspace_size() bytes, 256-byte aligned, required");
  B2_CHECK_ARG(pixels < (1LL << 31), "conv1_bwd: too many pixels");
  cudaStream_t st = (cudaStream_t)stream;
  bf16* patches = (bf16*)workspace;
  float* acc = (float*)((char*)workspace + ((pixels * 32 + 255) / 256) * 256);
  {
    B2_SCOPE(FAM_SUBSAMPLE, st, 0, 4.0 * B * Tin * F + 32.0 * pixels);
    conv1_patch_kernel<<<b2_grid_for(pixels, 256, 8), 256, 0, st>>>(x, B, Tin, F, T1, F1, patches);
    B2_LAUNCH_CHECK();
    B2_CUDA(cudaMemsetAsync(acc, 0, (size_t)C * 16 * sizeof(float), st));
  }
  b200slu_gemm_t g = {};
  g.trans_a = 1;
  g.trans_b = 1;
  g.M = C;
  g.N = 16;
  g.K = (int)pixels;
  g.A = dh1;
  g.lda = C;
  g.B = patches;
  g.ldb = 16;
  g.C = acc;
  g.ldc = 16;
  g.batch_inner = g.batch_outer = 1;
  b200slu_epilogue_t e = {};
  e.alpha = 1.f;
  e.out_f32 = 1;
  e.accumulate = 1;
  B2_TRY(b2_gemm_ex(&g, &e, st));
  conv1_wgrad_scatter_kernel<<<(C * 10 + 255) / 256, 256, 0, st>>>(acc, C, dw_accum, db_accum);
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_im2col3x3s2(const void* h1, int B, int T1, int F1, int C, void* col, void* stream) {
  B2_CHECK_ARG(B > 0 && T1 >= 3 && F1 >= 3 && C % 8 == 0, "im2col: bad shape");
  const int T2 = (T1 - 3) / 2 + 1, F2 = (F1 - 3) / 2 + 1;
  cudaStream_t st = (cudaStream_t)stream;
  const long long items = (long long)B * T2 * F2 * 9 * (C >> 3);
  B2_SCOPE(FAM_SUBSAMPLE, st, 0, 32.0 * items);
  B2_CHECK_ARG(B <= 65535, "im2col: B too large");
  im2col_kernel<<<dim3(T2, B), 256, 0, st>>>((const bf16*)h1, B, T1, F1, C, T2, F2, (bf16*)col);
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_col2im3x3s2_relu(const void* dcol, const void* h1, int B, int T1, int F1, int C, void* dh1, void* stream) {
  B2_CHECK_ARG(B > 0 && T1 >= 3 && F1 >= 3 && C % 8 == 0, "col2im: bad shape");
  const int T2 = (T1 - 3) / 2 + 1, F2 = (F1 - 3) / 2 + 1;
  cudaStream_t st = (cudaStream_t)stream;
  const long long items = (long long)B * T1 * F1 * (C >> 3);
  B2_SCOPE(FAM_SUBSAMPLE, st, 0, 16.0 * B * T2 * F2 * 9 * (C >> 3) + 32.0 * items);
  B2_CHECK_ARG(B <= 65535, "col2im: B too large");
  col2im_relu_kernel<<<dim3(T1, B), 256, 0, st>>>((const bf16*)dcol, (const bf16*)h1, B, T1, F1, C, T2, F2, (bf16*)dh1);
  B2_LAUNCH_CHECK();
  return 0;
}
