// fp32 feature front-end: waveform -> log-mel (STFT fused), and SpecAug + utterance mean normalisation.
//
// logmel replaces DefaultFrontend.forward (espnet2/asr/frontend/default.py:82-117) = Stft
// (espnet2/layers/stft.py:83-170: torch.stft, centre reflect-pad n_fft/2, periodic Hann, onesided,
// frames past olens zeroed) -> power re^2+im^2 (default.py:110) -> LogMel (espnet2/layers/log_mel.py:
// 62-79: matmul with the mel matrix, clamp 1e-10, natural log, padded frames 0).  One block per frame:
// framing + window + radix-2 FFT in shared memory + power + mel + log, so the (B,257,T,2) spectrum
// and five elementwise passes never touch HBM (read 4*N*B, write 4*T_in*n_mels*B bytes).
//
// specaug_mvn replaces SpecAug.forward (espnet2/asr/specaug/specaug.py:89-96: TimeWarp
// espnet2/layers/time_warp.py:9-46,65-88 bicubic; MaskAlongAxis espnet2/layers/mask_along_axis.py:8-68
// freq then time, fill 0) followed by UtteranceMVN (espnet2/layers/utterance_mvn.py:45-88, means only:
// zero the padding, mean over valid frames, subtract from ALL frames).  The random draws stay on the
// host (same torch RNG calls, same order as the reference, Appendix A.15) and arrive as small arrays.
#include <math.h>
#include <stdlib.h>

#include "common.cuh"
#include "sm100_ptx.cuh"

namespace {

__global__ void __launch_bounds__(256) logmel_kernel(const float* __restrict__ wav, long long ldw, int N,
                                                     const int* __restrict__ lens, const float* __restrict__ melmat, int n_fft,
                                                     int log2n, int hop, int n_mels, int T, float* __restrict__ feats,
                                                     int* __restrict__ flens) {
  extern __shared__ float sm[];
  float* re = sm;               // [n_fft]
  float* im = sm + n_fft;       // [n_fft]
  float* twc = im + n_fft;      // [n_fft/2] cos
  float* tws = twc + n_fft / 2; // [n_fft/2] -sin
  const int t = blockIdx.x, b = blockIdx.y;
  const int pad = n_fft / 2;
  const int olen = (lens[b] + 2 * pad - n_fft) / hop + 1;
  if (t == 0 && threadIdx.x == 0 && flens != nullptr) flens[b] = olen;
  float* out = feats + ((long long)b * T + t) * n_mels;
  if (t >= olen) {
    for (int m = threadIdx.x; m < n_mels; m += blockDim.x) out[m] = 0.f;
    return;
  }
  const float* w = wav + (long long)b * ldw;
  for (int i = threadIdx.x; i < n_fft; i += blockDim.x) {
    int n = t * hop - pad + i;
    if (n < 0) n = -n;                    // reflect (no edge repeat), over the padded batch width N
    if (n >= N) n = 2 * (N - 1) - n;
    const float win = 0.5f - 0.5f * cospif(2.f * (float)i / (float)n_fft);
    const int rev = (int)(__brev((unsigned)i) >> (32 - log2n));
    re[rev] = w[n] * win;
    im[rev] = 0.f;
  }
  for (int k = threadIdx.x; k < n_fft / 2; k += blockDim.x) {
    float s, c;
    sincospif(-2.f * (float)k / (float)n_fft, &s, &c);
    twc[k] = c;
    tws[k] = s;
  }
  __syncthreads();
  for (int s = 1; s <= log2n; ++s) {
    const int half = 1 << (s - 1), stride = n_fft >> s;
    for (int j = threadIdx.x; j < n_fft / 2; j += blockDim.x) {
      const int grp = j >> (s - 1), pos = j & (half - 1);
      const int i0 = (grp << s) + pos, i1 = i0 + half;
      const float c = twc[pos * stride], sn = tws[pos * stride];
      const float xr = re[i1] * c - im[i1] * sn, xi = re[i1] * sn + im[i1] * c;
      const float ur = re[i0], ui = im[i0];
      re[i0] = ur + xr;
      im[i0] = ui + xi;
      re[i1] = ur - xr;
      im[i1] = ui - xi;
    }
    __syncthreads();
  }
  const int nbins = n_fft / 2 + 1;
  for (int k = threadIdx.x; k < nbins; k += blockDim.x) re[k] = re[k] * re[k] + im[k] * im[k];  // power, in place
  __syncthreads();
  for (int m = threadIdx.x; m < n_mels; m += blockDim.x) {
    float acc = 0.f;
    for (int k = 0; k < nbins; ++k) acc += re[k] * melmat[k * n_mels + m];
    out[m] = logf(fmaxf(acc, 1e-10f));
  }
}

// Warp-per-frame form of the same computation (n_fft <= 1024): the block-per-frame kernel above spends its time in
// 9 block-wide barriers per frame, 256 sincospi per frame and a dense 257 x 80 mel product run by 80 threads (0.83 ms
// for 80 k frames = 0.018 of the HBM roofline its bytes would allow).  Here a block of 8 warps computes the window, the
// twiddles and, per mel filter, the range of FFT bins with non-zero weight ONCE; then every warp walks its own frames:
// framing + window into its private shared-memory buffers (bit-reversed), radix-2 stages separated by __syncwarp only,
// power spectrum, and the mel product over the non-zero range of each filter (Slaney filters overlap two neighbours:
// ~6 bins per filter instead of 257; exact for any mel matrix, zero weights contribute exactly 0).
constexpr int LM_WARPS = 8;
// Shared memory is the bound of an in-place radix-2 FFT (10 shared accesses per butterfly), so the real input is
// transformed as ONE complex FFT of half the length (z[n] = x[2n] + i x[2n+1]; X[k] = E[k] + e^{-2 pi i k/N} O[k] with
// E, O from Z[k] and conj(Z[M-k])), and the twiddles of every stage sit contiguously (stage s reads entries
// half-1 .. 2 half-2: consecutive lanes, consecutive words, instead of a stride that folds 32 lanes onto 2 banks).
__global__ void __launch_bounds__(LM_WARPS * 32) logmel_warp_kernel(const float* __restrict__ wav, long long ldw, int N,
                                                                    const int* __restrict__ lens,
                                                                    const float* __restrict__ melmat, int n_fft, int log2n,
                                                                    int hop, int n_mels, int T, int B,
                                                                    float* __restrict__ feats, int* __restrict__ flens) {
  extern __shared__ float sm[];
  const int M = n_fft / 2, log2m = log2n - 1, nbins = M + 1, pad = n_fft / 2;
  float* win = sm;                   // [n_fft]
  float* twc = win + n_fft;          // [M] per-stage twiddles of the M-point FFT (M - 1 used): cos
  float* tws = twc + M;              // [M] -sin
  float* pc = tws + M;               // [M + 1] cos(2 pi k / n_fft)   (split of the real transform)
  float* ps = pc + (M + 1);          // [M + 1] -sin(2 pi k / n_fft)
  int* klo = reinterpret_cast<int*>(ps + (M + 1));     // [n_mels] first bin with a non-zero weight
  int* khi = klo + n_mels;                             // [n_mels] one past the last
  float* bufs = reinterpret_cast<float*>(khi + n_mels);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* re = bufs + (size_t)warp * (2 * M + nbins + 3);
  float* im = re + M;
  float* pw = im + M;                // [M + 1] power spectrum
  for (int i = threadIdx.x; i < n_fft; i += blockDim.x) win[i] = 0.5f - 0.5f * cospif(2.f * (float)i / (float)n_fft);
  for (int e = threadIdx.x; e < M - 1; e += blockDim.x) {
    // entry e belongs to the stage with half = 2^floor(log2(e + 1)), position pos = e - (half - 1): W = exp(-pi i pos / half)
    const int half = 1 << (31 - __clz(e + 1)), pos = e - (half - 1);
    float sn, c;
    sincospif(-(float)pos / (float)half, &sn, &c);
    twc[e] = c;
    tws[e] = sn;
  }
  for (int k = threadIdx.x; k <= M; k += blockDim.x) {
    float sn, c;
    sincospif(-2.f * (float)k / (float)n_fft, &sn, &c);
    pc[k] = c;
    ps[k] = sn;
  }
  for (int m = threadIdx.x; m < n_mels; m += blockDim.x) {
    int lo = nbins, hi = 0;
    for (int k = 0; k < nbins; ++k)
      if (melmat[k * n_mels + m] != 0.f) {
        lo = min(lo, k);
        hi = k + 1;
      }
    klo[m] = lo;
    khi[m] = hi;
  }
  if (blockIdx.x == 0 && flens != nullptr)
    for (int b = threadIdx.x; b < B; b += blockDim.x) flens[b] = (lens[b] + 2 * pad - n_fft) / hop + 1;
  __syncthreads();
  const long long frames = (long long)B * T, wstride = (long long)gridDim.x * LM_WARPS;
  for (long long f = (long long)blockIdx.x * LM_WARPS + warp; f < frames; f += wstride) {
    const int b = (int)(f / T), t = (int)(f - (long long)b * T);
    const int olen = (lens[b] + 2 * pad - n_fft) / hop + 1;
    float* out = feats + f * n_mels;
    if (t >= olen) {
      for (int m = lane; m < n_mels; m += 32) out[m] = 0.f;
      continue;
    }
    const float* w = wav + (long long)b * ldw;
    const int n0 = t * hop - pad;
    for (int i = lane; i < n_fft; i += 32) {
      int n = n0 + i;
      if (n < 0) n = -n;                    // reflect (no edge repeat), over the padded batch width N
      if (n >= N) n = 2 * (N - 1) - n;
      const int rev = (int)(__brev((unsigned)(i >> 1)) >> (32 - log2m));
      ((i & 1) ? im : re)[rev] = w[n] * win[i];
    }
    __syncwarp();
    for (int s = 1; s <= log2m; ++s) {
      const int half = 1 << (s - 1);
      const float *tc = twc + (half - 1), *ts = tws + (half - 1);
      for (int j = lane; j < M / 2; j += 32) {
        const int grp = j >> (s - 1), pos = j & (half - 1);
        const int i0 = (grp << s) + pos, i1 = i0 + half;
        const float c = tc[pos], sn = ts[pos];
        const float xr = re[i1] * c - im[i1] * sn, xi = re[i1] * sn + im[i1] * c;
        const float ur = re[i0], ui = im[i0];
        re[i0] = ur + xr;
        im[i0] = ui + xi;
        re[i1] = ur - xr;
        im[i1] = ui - xi;
      }
      __syncwarp();
    }
    for (int k = lane; k <= M; k += 32) {
      const int ka = k & (M - 1), kb = (M - k) & (M - 1);
      const float zr = re[ka], zi = im[ka], cr = re[kb], ci = -im[kb];   // Z[k], conj(Z[M - k])
      const float er = 0.5f * (zr + cr), ei = 0.5f * (zi + ci);          // E = (Z + Zc) / 2
      const float dr = 0.5f * (zr - cr), di = 0.5f * (zi - ci);          // (Z - Zc) / 2 ;  O = -i * that = (di, -dr)
      const float orr = di, oi = -dr;
      const float xr = er + pc[k] * orr - ps[k] * oi, xi = ei + pc[k] * oi + ps[k] * orr;
      pw[k] = xr * xr + xi * xi;
    }
    __syncwarp();
    for (int m = lane; m < n_mels; m += 32) {
      float acc = 0.f;
      for (int k = klo[m]; k < khi[m]; ++k) acc += pw[k] * melmat[k * n_mels + m];
      out[m] = logf(fmaxf(acc, 1e-10f));
    }
    __syncwarp();  // the buffers are rewritten by the next frame
  }
}

__device__ __forceinline__ void cubic_w(float t, float (&w)[4]) {
  const float A = -0.75f;
  const float x0 = t + 1.f, x1 = t, x2 = 1.f - t, x3 = 2.f - t;
  w[0] = ((A * x0 - 5.f * A) * x0 + 8.f * A) * x0 - 4.f * A;
  w[1] = ((A + 2.f) * x1 - (A + 3.f)) * x1 * x1 + 1.f;
  w[2] = ((A + 2.f) * x2 - (A + 3.f)) * x2 * x2 + 1.f;
  w[3] = ((A * x3 - 5.f * A) * x3 + 8.f * A) * x3 - 4.f * A;
}

// value of the time-warped utterance at (t, f): segment [seg0, seg0 + in_len) resized to out_len frames
__device__ __forceinline__ float warp_sample(const float* __restrict__ x, int F, int f, int seg0, int in_len, int out_len, int dst) {
  const float scale = (float)in_len / (float)out_len;
  const float src = ((float)dst + 0.5f) * scale - 0.5f;
  const float fl = floorf(src);
  float w[4];
  cubic_w(src - fl, w);
  const int i0 = (int)fl;
  float acc = 0.f;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    int idx = i0 - 1 + k;
    idx = max(0, min(in_len - 1, idx));
    acc += w[k] * x[(long long)(seg0 + idx) * F + f];
  }
  return acc;
}

constexpr int SA_CLUSTER = 8;
__global__ void __launch_bounds__(256) specaug_mvn_kernel(const float* __restrict__ x, int T, int F, const int* __restrict__ lens,
                                                          const int* __restrict__ center, const int* __restrict__ warped,
                                                          const int* __restrict__ fmask, int nf, const int* __restrict__ tmask,
                                                          int nt, int do_mvn, float* __restrict__ y) {
  // One utterance = one thread-block CLUSTER of SA_CLUSTER CTAs (one CTA per utterance left 128 blocks walking 626 frames
  // each: 209 us for 20 MB): the CTAs split the frames, each reduces its share of the per-bin sums in its own shared
  // memory, and after a cluster barrier every CTA adds up all shares through distributed shared memory.
  extern __shared__ float sm[];  // [nfl][F] partial sums, then [F] this CTA's share, then [F] means
  const int b = blockIdx.x / SA_CLUSTER, crank = (int)cluster_ctarank();
  const int len = min(lens[b], T);
  const float* xb = x + (long long)b * T * F;
  float* yb = y + (long long)b * T * F;
  const int nfl = blockDim.x / F;
  const int fl = threadIdx.x / F, f = threadIdx.x - fl * F;
  const bool active = fl < nfl;
  const int c = center != nullptr ? center[b] : -1, wp = center != nullptr ? warped[b] : 0;
  const bool keep_pad = center == nullptr && !do_mvn;
  bool fdrop = false;
  if (active)
    for (int k = 0; k < nf; ++k) {
      const int pos = fmask[(b * nf + k) * 2], ml = fmask[(b * nf + k) * 2 + 1];
      fdrop = fdrop || (f >= pos && f < pos + ml);
    }
  float sum = 0.f;
  if (active) {
    for (int t = crank * nfl + fl; t < T; t += nfl * SA_CLUSTER) {
      float v = 0.f;
      // without a warp stage (center == nullptr) and without the MVN the reference never touches the padding except through
      // the masks, which run over the padded axes; TimeWarp re-pads with 0 and UtteranceMVN zeroes the padding itself
      if (t < len || keep_pad) {
        if (c >= 0) v = t < wp ? warp_sample(xb, F, f, 0, c, wp, t) : warp_sample(xb, F, f, c, len - c, len - wp, t - wp);
        else v = xb[(long long)t * F + f];
        bool drop = fdrop;
        for (int k = 0; k < nt; ++k) {
          const int pos = tmask[(b * nt + k) * 2], ml = tmask[(b * nt + k) * 2 + 1];
          drop = drop || (t >= pos && t < pos + ml);
        }
        if (drop) v = 0.f;
      }
      yb[(long long)t * F + f] = v;
      sum += v;
    }
    sm[fl * F + f] = sum;
  }
  __syncthreads();
  if (threadIdx.x < F) {
    float s = 0.f;
    for (int k = 0; k < nfl; ++k) s += sm[k * F + threadIdx.x];
    sm[nfl * F + threadIdx.x] = s;
  }
  cluster_sync_all();  // every CTA of the utterance has published its share (also a block barrier)
  if (do_mvn && threadIdx.x < F) {
    float s = 0.f;
    const uint32_t mine = smem_u32(sm + nfl * F + threadIdx.x);
    for (int rk = 0; rk < SA_CLUSTER; ++rk) {
      float v;
      asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(mapa_rank(mine, (uint32_t)rk)) : "memory");
      s += v;
    }
    sm[(nfl + 1) * F + threadIdx.x] = s / (float)len;
  }
  cluster_sync_all();  // nobody leaves (or reuses its shares) while a peer may still be reading them
  if (!do_mvn) return;
  if (active) {
    const float mean = sm[(nfl + 1) * F + f];
    for (int t = crank * nfl + fl; t < T; t += nfl * SA_CLUSTER) yb[(long long)t * F + f] -= mean;
  }
}

// GlobalMVN (espnet2/layers/global_mvn.py:66-94): y = (x - mean[f]) on the valid frames, 0 on the padding, then / std[f].
// One thread per 4 contiguous bins (F % 4 == 0: 16-byte accesses), grid-stride over (utterance, frame) rows.
__global__ void __launch_bounds__(256) global_mvn_kernel(const float* __restrict__ x, int B, int T, int F,
                                                         const int* __restrict__ lens, const float* __restrict__ mean,
                                                         const float* __restrict__ inv_std, float* __restrict__ y) {
  const int f4 = F >> 2;
  const long long n = (long long)B * T * f4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int fq = (int)(i % f4);
    const long long row = i / f4;
    const int t = (int)(row % T), b = (int)(row / T);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (t < lens[b]) {
      v = *reinterpret_cast<const float4*>(x + row * F + fq * 4);
      if (mean != nullptr) {
        const float4 m = *reinterpret_cast<const float4*>(mean + fq * 4);
        v.x -= m.x; v.y -= m.y; v.z -= m.z; v.w -= m.w;
      }
      if (inv_std != nullptr) {
        const float4 s = *reinterpret_cast<const float4*>(inv_std + fq * 4);
        v.x *= s.x; v.y *= s.y; v.z *= s.z; v.w *= s.w;
      }
    }
    *reinterpret_cast<float4*>(y + row * F + fq * 4) = v;
  }
}

}  // namespace

extern "C" int b200slu_global_mvn_fwd(const float* x, int B, int T, int F, const int* lens, const float* mean, const float* inv_std,
                                      float* y, void* stream) {
  B2_CHECK_ARG(B > 0 && T > 0 && F > 0 && F % 4 == 0, "global_mvn: need F %% 4 == 0 (F=%d)", F);
  B2_CHECK_ARG(((uintptr_t)x % 16) == 0 && ((uintptr_t)y % 16) == 0 && (mean == nullptr || ((uintptr_t)mean % 16) == 0) &&
                   (inv_std == nullptr || ((uintptr_t)inv_std % 16) == 0),
               "global_mvn: operands must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  B2_SCOPE(FAM_FRONTEND, st, 0, 8.0 * B * T * F);
  global_mvn_kernel<<<b2_grid_for((long long)B * T * (F / 4), 256, 8), 256, 0, st>>>(x, B, T, F, lens, mean, inv_std, y);
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_logmel_fwd(const float* wav, long long ldw, int B, int N, const int* lens, const float* melmat, int n_fft,
                                  int hop, int n_mels, float* feats, int* feat_lens, void* stream) {
  int log2n = 0;
  while ((1 << log2n) < n_fft) ++log2n;
  B2_CHECK_ARG(B > 0 && B <= 65535 && N > n_fft / 2 && (1 << log2n) == n_fft && n_fft >= 4 && n_fft <= 4096 && hop > 0 &&
                   n_mels > 0,
               "logmel: need power-of-two 4 <= n_fft <= 4096 and N > n_fft/2 (N=%d n_fft=%d)", N, n_fft);
  const int T = N / hop + 1;
  cudaStream_t st = (cudaStream_t)stream;
  B2_SCOPE(FAM_FRONTEND, st, (double)B * T * (5.0 * n_fft * log2n + 2.0 * (n_fft / 2 + 1) * n_mels),
           4.0 * B * N + 4.0 * B * T * n_mels);
  static const bool old_kernel = getenv("B200SLU_LOGMEL_BLOCK") != nullptr;  // tuning: the block-per-frame kernel
  if (n_fft <= 1024 && !old_kernel) {
    const size_t smw = (size_t)(3 * n_fft + 2 + 2 * n_mels + LM_WARPS * (n_fft + n_fft / 2 + 4)) * sizeof(float);
    const long long frames = (long long)B * T;
    long long blocks = (frames + LM_WARPS - 1) / LM_WARPS;
    const long long cap = (long long)b2_num_sms() * (smw <= 48 * 1024 ? 4 : 2);
    if (blocks > cap) blocks = cap;
    static unsigned long long cfg = 0;
    if (smw > 48 * 1024 && !b2_device_done(cfg)) {
      B2_CUDA(cudaFuncSetAttribute(logmel_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
      b2_device_mark(cfg);
    }
    logmel_warp_kernel<<<(unsigned)blocks, LM_WARPS * 32, smw, st>>>(wav, ldw, N, lens, melmat, n_fft, log2n, hop, n_mels, T, B, feats,
                                                                     feat_lens);
    B2_LAUNCH_CHECK();
    return 0;
  }
  const size_t smem = (size_t)3 * n_fft * sizeof(float);
  logmel_kernel<<<dim3(T, B), 256, smem, st>>>(wav, ldw, N, lens, melmat, n_fft, log2n, hop, n_mels, T, feats, feat_lens);
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_specaug_mvn_fwd(const float* x, int B, int T, int F, const int* lens, const int* warp_center,
                                       const int* warp_warped, const int* freq_masks, int n_freq_masks, const int* time_masks,
                                       int n_time_masks, int do_mvn, float* y, void* stream) {
  B2_CHECK_ARG(B > 0 && T > 0 && F > 0 && F <= 256, "specaug_mvn: need F <= 256 (F=%d)", F);
  B2_CHECK_ARG((warp_center == nullptr) == (warp_warped == nullptr), "specaug_mvn: warp draws must come in pairs");
  cudaStream_t st = (cudaStream_t)stream;
  const int nfl = 256 / F;
  const size_t smem = (size_t)(nfl + 2) * F * sizeof(float);
  B2_SCOPE(FAM_FRONTEND, st, 0, 8.0 * B * T * F);
  {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)B * SA_CLUSTER);
    cfg.blockDim = dim3(256);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = SA_CLUSTER;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    B2_CUDA(cudaLaunchKernelEx(&cfg, specaug_mvn_kernel, x, T, F, lens, warp_center, warp_warped, freq_masks, n_freq_masks,
                               time_masks, n_time_masks, do_mvn, y));
  }
  return 0;
}
