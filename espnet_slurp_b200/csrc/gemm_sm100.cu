// bf16 GEMM on 5th-gen tensor cores: TMA (SWIZZLE_128B) -> smem ring -> tcgen05.mma
// (accumulators in TMEM) -> tcgen05.ld epilogue with fused bias / activation /
// dropout / residual / activation-gradient / split-K atomics.
//
// Replaces every cuBLAS/cuDNN call the reference's nn.Linear / k=1 nn.Conv1d /
// Conv2d(im2col) sites make: attention.py:33-36,232;
// positionwise_feed_forward.py:25-26; convolution.py:28-53; subsampling.py:56-65;
// espnet2/asr/ctc.py:33; espnet2/asr/decoder/transformer_decoder.py:85.
//
// Operand majorness is chosen per operand (4 instantiations of one kernel):
//   trans_a=0: A is [M,K] row-major (K contiguous)      trans_a=1: A is [K,M] row-major (M contiguous)
//   trans_b=0: B is [N,K] row-major (K contiguous)      trans_b=1: B is [K,N] row-major (N contiguous)
//   (0,0) fwd   Y = X W^T        (0,1) dgrad  dX = dY W      (1,1) wgrad  dW = dY^T X
// transposed operands use MN-major UMMA descriptors so nothing is ever transposed in HBM.
// Two batch levels (inner, outer) with independent strides per operand (stride 0 = broadcast)
// cover the per-(utterance, head) attention contractions: the TMA maps are 4-D.
//
// CTA = one 128 x BN output tile.  Warp roles: 0 = TMA producer, 1 = TMEM owner + MMA
// issuer (one elected lane), 2..5 = epilogue (TMEM lane group = warp_id % 4).
#include <stdlib.h>

#include "common.cuh"
#include "sm100_ptx.cuh"

namespace {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 B = one swizzle span

// Implicit-GEMM view of Conv2dSubsampling's second convolution (subsampling.py:56-65: 3x3, stride 2, C -> C channels):
// one operand's k-blocks (CONV = 1: A, rows = pixels) or column boxes (CONV = 2: B, wgrad) come from a TMA im2col map over
// the channels-last activation instead of a materialised [pixels, 9 C] matrix.
struct ConvArgs {
  int pw, ph;          // pixel grid (per image, w fastest) the GEMM's pixel dimension enumerates
  int mult;            // tensor-coordinate step between neighbouring pixels (the map's traversal stride)
  int base_w, base_h;  // tensor coordinates of pixel (0, 0) (the map's lower corner)
  int cb, C;           // channel blocks of 64 per tap; channels (= columns per tap of the [C, 9 C] weight matrix)
  unsigned char off_w[9], off_h[9], wtap[9];  // per tap of THIS GEMM: im2col offsets and the weight matrix's tap index
  // scatter epilogue (dgrad parity classes): pixel (n, hq, wq) of the GEMM is pixel (n, hq*o_mult + o_h0, wq*o_mult + o_w0)
  // of the [*, OH, OW, C] output; the same position of `aux` gates it (ReLU mask)
  int o_mult, o_w0, o_h0, OW, OH;
};

struct EpiArgs {
  const float* bias;
  const float* res;
  const bf16* aux;
  bf16* pre_out;
  void* out;
  long long ldc, ldres, ldaux;
  long long c_bs_inner, c_bs_outer;  // output batch strides (elements)
  int a_bc_inner, a_bc_outer, b_bc_inner, b_bc_outer;  // 1 = operand indexed by that batch level, 0 = broadcast
  float alpha;
  int act, aux_mode, out_f32, accumulate;
  uint32_t flags;  // EF_* summary of the fields above (what the generic kernel branches on)
  int debug;  // tuning experiments only (B200SLU_GEMM_DEBUG): 1 = no stores, 2 = no epilogue, 4 = no TMA loads,
              // 16 = CTA 0 records phase cycle counts into dbg_out (printed by the host after a synchronize)
  long long* dbg_out;
  B2Dropout drop;
  ConvArgs cv;
  float* colsum;  // EF_COLSUM: fp32 [N] += column sums of the bf16 output
};

// Epilogue feature bits.  A kernel instantiated with EF != EF_GENERIC has its epilogue resolved at compile time: the
// K = 512 GEMMs of the encoder leave ~250 issue cycles per 16-column slice per warp, which a chain of
// load-constant / test / branch decisions per feature does not fit into (measured: 840 cycles per slice).
enum : uint32_t {
  EF_BIAS = 1u, EF_SWISH = 2u, EF_RELU = 4u, EF_DACT = 8u, EF_DROP = 16u, EF_ALPHA = 32u, EF_AUX_MUL = 64u,
  EF_AUX_SWISH = 128u, EF_AUX_RELU = 256u, EF_RES = 512u, EF_F32 = 1024u, EF_ACC = 2048u, EF_RES_TMA = 4096u, EF_SCATTER = 8192u, EF_COLSUM = 16384u, EF_AUX_TMA = 32768u, EF_GENERIC = 1u << 31
};
// Epilogue warps: 2 per TMEM lane quadrant for the register-heavy epilogues (168 registers each).
// 16 warps (4 per quadrant, 64 columns each) halve the epilogue time of the light epilogues (bias-only: 6500 -> 3300
// cycles per tile) but their 64 KB of staging costs a stage of the operand ring, and the ring depth is what bounds the
// main loop (TMA round trip): measured 26.2 us against 23.9 us on the 9920 x 2048 x 512 forward.  Kept switchable.
// Also measured on the CTA-pair kernels (32 KB stages leave room for both): conv2 dgrad 1651 us with 16 warps against
// 1530 us with 8.  Kept switchable.
constexpr uint32_t EF_LIGHT = 0;  // EF_BIAS | EF_RELU | EF_F32 | EF_ACC | EF_AUX_MUL | EF_AUX_RELU | EF_ALPHA;
__host__ __device__ constexpr int epi_warps(uint32_t ef, int bn, int cg) {
  return (EF_LIGHT != 0 && cg == 2 && bn == 256 && (ef & ~EF_LIGHT) == 0) ? 16 : 8;
}
__host__ __device__ constexpr int num_threads(uint32_t ef, int bn, int cg) { return 64 + 32 * epi_warps(ef, bn, cg); }

// Staging boxes per epilogue warp for the TMA stores.  Two when the kernel writes two outputs per box (saved
// derivative), when it flushes a box every 32 columns (fp32 output, not the one-epilogue-per-long-K wgrad) or in the
// generic kernel; otherwise one, and the shared memory goes to a deeper operand ring: the main loop is bound by the
// TMA round trip (free stage -> L2 -> full stage ~ 2000+ cycles against 512 cycles of MMA per stage), so bytes in
// flight are what buys tensor-pipe utilisation.
__host__ __device__ constexpr int stg_bufs(uint32_t ef, int bn, int cg) {
  return epi_warps(ef, bn, cg) == 16 ? 1 : ((ef & (EF_DACT | EF_GENERIC | EF_AUX_TMA)) || ((ef & EF_F32) && !(ef & EF_ACC))) ? 2 : 1;
}
template <int BN, int CG>
__host__ __device__ constexpr int num_stages(uint32_t ef) {
  // 227 KB minus staging, barriers and the bias slice, in stages of (128 + BN / CG) x 64 bf16
  return (232448 - epi_warps(ef, BN, CG) * stg_bufs(ef, BN, CG) * 4096 - 2 * BN * 4 - 256) / ((BM + BN / CG) * BK * 2);
}

// CG = CTAs per MMA (tcgen05 cta_group): with CG = 2 a CTA pair owns a 256 x BN tile, each CTA holding its own 128
// rows of A and HALF of B's rows -- 32 KB instead of 48 KB of L2 traffic per CTA and k-block, which is what bounds
// the K = 512 .. 2048 GEMMs of the step (L2 -> SM tops out near 6300 B/clk chip-wide).
template <int BN, int STAGES, int STGB, int CG, int EW = 8>
struct SmemLayout {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = (BN / CG) * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int NUM_BARS = 2 * STAGES + 4 + 2 * EW;  // ring full / empty, accumulator full / empty, residual boxes per epilogue warp
  // output staging for the TMA stores: per epilogue warp STGB boxes of 32 rows x 128 B (SWIZZLE_128B), 1024-B aligned
  static constexpr int STG_OFF = STAGES * STAGE_BYTES;
  static constexpr int STG_BYTES = EW * STGB * 4096;
  static constexpr int BAR_OFF = STG_OFF + STG_BYTES;
  static constexpr int BIAS_OFF = BAR_OFF + NUM_BARS * 8 + 16;       // per-tile bias slice, double-buffered by tile parity
  static constexpr int BIAS_BYTES = 2 * BN * 4;
  static constexpr int TOTAL = BIAS_OFF + BIAS_BYTES;                // the dynamic segment is declared 1024-B aligned
  static_assert(TOTAL <= 232448, "shared memory budget");
};

// ----------------------------------------------------------------------------- fused epilogue
// Per-thread, per-tile constants of the epilogue (hoisted out of the column loop).
struct EpiRow {
  const float* res;   // residual row (or nullptr)
  const bf16* aux;    // activation-derivative row (or nullptr)
  unsigned long long e0;  // flat element index of (row, 0) for the dropout counter
};

// The arithmetic of 8 consecutive columns.  b = bias, x = aux (bf16 widened), r = residual; d receives the saved
// derivative.  F is a compile-time constant in the specialised kernels and a register in the generic one.
template <uint32_t EF>
__device__ __forceinline__ void epi_math8(uint32_t rt_flags, const EpiArgs& ep, const B2DropKey& dk, float alpha,
                                          unsigned long long e0, int col, float (&v)[8], const float (&b)[8],
                                          const float (&x)[8], const float (&r)[8], float (&d)[8]) {
  // Packed fp32 throughout (FADD2 / FMUL2 / FFMA2 on neighbouring columns): the FMA pipe's rate of scalar code at half the
  // issue slots, and issue slots are what the K = 512 epilogues run out of (two epilogue warps per scheduler).
  const uint32_t F = (EF & EF_GENERIC) ? rt_flags : EF;
  const float2 one = make_float2(1.f, 1.f), mone = make_float2(-1.f, -1.f);
  // the residual add folds into the keep multiply (one FFMA2) when nothing sits between them
  const bool res_in_drop = (F & EF_DROP) && !(F & EF_DACT) && (F & EF_RES) && !(F & (EF_AUX_MUL | EF_AUX_SWISH | EF_AUX_RELU));
  float2 V[4], D[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) V[q] = make_float2(v[2 * q], v[2 * q + 1]);
  if (F & EF_BIAS) {
#pragma unroll
    for (int q = 0; q < 4; ++q) V[q] = b2_add2(V[q], make_float2(b[2 * q], b[2 * q + 1]));
  }
  if (F & EF_SWISH) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 sg = b2_sigmoid2(V[q]);
      const float2 t = b2_mul2(V[q], sg);                 // swish
      D[q] = b2_fma2(t, b2_fma2(sg, mone, one), sg);      // swish' = sg + swish * (1 - sg)
      V[q] = t;
    }
  } else if (F & EF_RELU) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      D[q] = make_float2(V[q].x > 0.f ? 1.f : 0.f, V[q].y > 0.f ? 1.f : 0.f);
      V[q] = make_float2(fmaxf(V[q].x, 0.f), fmaxf(V[q].y, 0.f));
    }
  } else {
#pragma unroll
    for (int q = 0; q < 4; ++q) D[q] = one;
  }
  if (F & EF_DROP) {
    // alpha rides on the keep multiplier (the host clears EF_ALPHA when EF_DROP is set)
    float mul[8];
    if (F & EF_DACT) {
      dropout_mul8(dk, (e0 + (unsigned long long)col) >> 3, dk.inv, mul);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 m = make_float2(mul[2 * q], mul[2 * q + 1]);
        D[q] = b2_mul2(D[q], m);
        V[q] = b2_mul2(V[q], m);
      }
      if (alpha != 1.f) {
        const float2 al = make_float2(alpha, alpha);
#pragma unroll
        for (int q = 0; q < 4; ++q) V[q] = b2_mul2(V[q], al);
      }
    } else {
      dropout_mul8(dk, (e0 + (unsigned long long)col) >> 3, dk.inv * alpha, mul);
      if (res_in_drop) {
#pragma unroll
        for (int q = 0; q < 4; ++q)
          V[q] = b2_fma2(V[q], make_float2(mul[2 * q], mul[2 * q + 1]), make_float2(r[2 * q], r[2 * q + 1]));
      } else {
#pragma unroll
        for (int q = 0; q < 4; ++q) V[q] = b2_mul2(V[q], make_float2(mul[2 * q], mul[2 * q + 1]));
      }
    }
  } else if (F & EF_ALPHA) {
    const float2 al = make_float2(alpha, alpha);
#pragma unroll
    for (int q = 0; q < 4; ++q) V[q] = b2_mul2(V[q], al);
  }
  if (F & EF_AUX_MUL) {
#pragma unroll
    for (int q = 0; q < 4; ++q) V[q] = b2_mul2(V[q], make_float2(x[2 * q], x[2 * q + 1]));
  } else if (F & EF_AUX_SWISH) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 xx = make_float2(x[2 * q], x[2 * q + 1]);
      const float2 sg = b2_sigmoid2(xx);
      // swish'(x) = sg * (1 + x * (1 - sg))   (swish_gradf_)
      V[q] = b2_mul2(V[q], b2_mul2(sg, b2_fma2(xx, b2_fma2(sg, mone, one), one)));
    }
  } else if (F & EF_AUX_RELU) {
#pragma unroll
    for (int q = 0; q < 4; ++q) V[q] = make_float2(x[2 * q] > 0.f ? V[q].x : 0.f, x[2 * q + 1] > 0.f ? V[q].y : 0.f);
  }
  if ((F & EF_RES) && !res_in_drop) {
#pragma unroll
    for (int q = 0; q < 4; ++q) V[q] = b2_add2(V[q], make_float2(r[2 * q], r[2 * q + 1]));
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    v[2 * q] = V[q].x;
    v[2 * q + 1] = V[q].y;
    d[2 * q] = D[q].x;
    d[2 * q + 1] = D[q].y;
  }
}

// Residual / saved-derivative operands of one 16-column slice, fetched one slice ahead of their use.
struct EpiPre {
  float r[2][8];
  uint4 x[2];
};
template <uint32_t EF>
__device__ __forceinline__ void epi_prefetch(uint32_t rt_flags, const EpiRow& er, int col0, int N, EpiPre& p) {
  const uint32_t F = (EF & EF_GENERIC) ? rt_flags : EF;
  if (col0 + 16 <= N) {
#pragma unroll
    for (int g = 0; g < 2; ++g) {
      if ((F & (EF_AUX_MUL | EF_AUX_SWISH | EF_AUX_RELU)) && !(F & EF_AUX_TMA)) p.x[g] = *reinterpret_cast<const uint4*>(er.aux + col0 + g * 8);
      if ((F & EF_RES) && !(F & EF_RES_TMA)) load8_f32(er.res + col0 + g * 8, p.r[g]);
    }
  }
}

// One 16-column slice (index sl within the warp's columns) of one accumulator row: fused arithmetic, then the results
// go to this lane's row of the staging boxes in shared memory (SWIZZLE_128B: 16-byte chunk c of row r sits at
// r*128 + ((c ^ (r & 7)) << 4), which is also bank-conflict-free for the 32 lanes writing 32 rows).  The boxes leave
// through TMA bulk stores, which clip rows >= M and columns >= N; only the operand LOADS of a ragged last slice
// (N % 16 != 0) need guarding.  bias_s: this slice's 16 bias values in shared memory (zero-filled past N).
// rt_flags already has the split-K gating applied (bias / residual on the first split only).
template <uint32_t EF>
__device__ __forceinline__ void epilogue_slice(uint32_t rt_flags, const EpiArgs& ep, const B2DropKey& dk, const EpiRow& er,
                                               float alpha, const float* bias_s, const EpiPre& pre, const uint32_t (&acc)[16],
                                               int col0, int N, int sl, int lane, uint32_t stg_out, uint32_t stg_d,
                                               uint32_t keep16 = 0xffffu) {
  const uint32_t F = (EF & EF_GENERIC) ? rt_flags : EF;
  const uint32_t rowoff = (uint32_t)lane * 128u, sw = (uint32_t)(lane & 7);
  const bool whole = col0 + 16 <= N;
#pragma unroll
  for (int g = 0; g < 2; ++g) {
    const int c = col0 + g * 8;
    float w[8], d[8], b[8] = {}, x[8] = {}, r[8] = {};
    if (F & EF_BIAS) load8_f32(bias_s + g * 8, b);
    if (F & EF_RES_TMA) {
      // the residual box was TMA-loaded into the very staging box the results go to: same lane, same 16-byte chunks
      const uint32_t ch = (uint32_t)((sl & 1) * 4 + g * 2);
      lds128(stg_out + rowoff + ((ch ^ sw) << 4), r[0], r[1], r[2], r[3]);
      lds128(stg_out + rowoff + (((ch + 1) ^ sw) << 4), r[4], r[5], r[6], r[7]);
    }
    if (F & EF_AUX_TMA) {
      // the saved-derivative box was TMA-loaded into the very staging box the results go to: same lane, same 16-byte chunk
      const uint32_t ch = (uint32_t)((sl & 3) * 2 + g);
      uint32_t a0, a1, a2, a3;
      lds128u(stg_out + rowoff + ((ch ^ sw) << 4), a0, a1, a2, a3);
      const float2 x0 = unpack_bf16x2(a0), x1 = unpack_bf16x2(a1), x2 = unpack_bf16x2(a2), x3 = unpack_bf16x2(a3);
      x[0] = x0.x; x[1] = x0.y; x[2] = x1.x; x[3] = x1.y; x[4] = x2.x; x[5] = x2.y; x[6] = x3.x; x[7] = x3.y;
    } else if (whole) {
      if (F & (EF_AUX_MUL | EF_AUX_SWISH | EF_AUX_RELU)) {
        const float2 x0 = unpack_bf16x2(pre.x[g].x), x1 = unpack_bf16x2(pre.x[g].y), x2 = unpack_bf16x2(pre.x[g].z),
                     x3 = unpack_bf16x2(pre.x[g].w);
        x[0] = x0.x; x[1] = x0.y; x[2] = x1.x; x[3] = x1.y; x[4] = x2.x; x[5] = x2.y; x[6] = x3.x; x[7] = x3.y;
      }
      if ((F & EF_RES) && !(F & EF_RES_TMA)) {
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = pre.r[g][j];
      }
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const bool ok = c + j < N;
        if (F & (EF_AUX_MUL | EF_AUX_SWISH | EF_AUX_RELU)) x[j] = ok ? __bfloat162float(er.aux[c + j]) : 0.f;
        if ((F & EF_RES) && !(F & EF_RES_TMA)) r[j] = ok ? er.res[c + j] : 0.f;
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) w[j] = __uint_as_float(acc[g * 8 + j]);
    epi_math8<EF>(rt_flags, ep, dk, alpha, er.e0, c, w, b, x, r, d);
    if (F & EF_SCATTER) {  // ReLU gate of the layer below, one bit per column
#pragma unroll
      for (int j = 0; j < 8; ++j) w[j] = ((keep16 >> (g * 8 + j)) & 1u) ? w[j] : 0.f;
    }
    if (F & EF_DACT) {
      const uint32_t ch = (uint32_t)((sl & 3) * 2 + g);
      sts128(stg_d + rowoff + ((ch ^ sw) << 4), pack_bf16x2(d[0], d[1]), pack_bf16x2(d[2], d[3]), pack_bf16x2(d[4], d[5]),
             pack_bf16x2(d[6], d[7]));
    }
    if (F & EF_F32) {
      const uint32_t ch = (uint32_t)((sl & 1) * 4 + g * 2);
      sts128(stg_out + rowoff + ((ch ^ sw) << 4), __float_as_uint(w[0]), __float_as_uint(w[1]), __float_as_uint(w[2]),
             __float_as_uint(w[3]));
      sts128(stg_out + rowoff + (((ch + 1) ^ sw) << 4), __float_as_uint(w[4]), __float_as_uint(w[5]), __float_as_uint(w[6]),
             __float_as_uint(w[7]));
    } else {
      const uint32_t ch = (uint32_t)((sl & 3) * 2 + g);
      sts128(stg_out + rowoff + ((ch ^ sw) << 4), pack_bf16x2(w[0], w[1]), pack_bf16x2(w[2], w[3]), pack_bf16x2(w[4], w[5]),
             pack_bf16x2(w[6], w[7]));
    }
  }
}

// Persistent: grid = min(#work units, #SMs); unit u = (batch, tile, k-split), tiles n-fastest.
template <int BN, int STAGES, bool TA, bool TB, uint32_t EF, int CG, int CONV = 0>
__global__ void __launch_bounds__(num_threads(EF, BN, CG), 1)
gemm_bf16_tcgen05(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                  const __grid_constant__ CUtensorMap map_c, const __grid_constant__ CUtensorMap map_d, int M, int N, int K, int tiles_n, int tiles_per_batch, int batch_inner, int splits, int kb_per_split,
                  int num_units, EpiArgs ep) {
  constexpr int STGB = stg_bufs(EF, BN, CG);
  using L = SmemLayout<BN, STAGES, STGB, CG, epi_warps(EF, BN, CG)>;
  constexpr int TM = BM * CG;  // rows of one work unit (this CTA owns rows [cta_rank * BM, +BM) of it)
  const uint32_t cta_rank = CG == 2 ? cluster_ctarank() : 0u;
  const int first_unit = blockIdx.x / CG, unit_step = gridDim.x / CG;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  const uint32_t sbase = smem_u32(smem);
  if (sbase & 1023u) __trap();  // SWIZZLE_128B operand tiles and staging boxes need 1024-B alignment
  const uint32_t bar_full = sbase + L::BAR_OFF;
  const uint32_t bar_empty = bar_full + STAGES * 8;
  const uint32_t bar_acc_full = bar_empty + STAGES * 8;   // [2]
  const uint32_t bar_acc_empty = bar_acc_full + 16;       // [2]
  const uint32_t bar_res = bar_acc_empty + 16;            // [epilogue warps][2]: residual box landed in staging buffer b
  uint32_t* tmem_slot = (uint32_t*)(smem + L::BAR_OFF + L::NUM_BARS * 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_kb_total = (K + BK - 1) / BK;
  const int units_per_batch = tiles_per_batch * splits;

  b2_pdl_launch_dependents();  // the next kernel of the stream may start launching; it waits for us before reading
  if (threadIdx.x == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_c) : "memory");
    if (((EF & EF_GENERIC) ? ep.flags : EF) & (EF_DACT | EF_RES_TMA)) asm volatile("prefetch.tensormap [%0];" ::"l"(&map_d) : "memory");
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(bar_full + s * 8, 1);
      mbar_init(bar_empty + s * 8, 1);
    }
    for (int b = 0; b < 2 * epi_warps(EF, BN, CG); ++b) mbar_init(bar_res + b * 8, 1);
    for (int b = 0; b < 2; ++b) {
      mbar_init(bar_acc_full + b * 8, 1);
      mbar_init(bar_acc_empty + b * 8, epi_warps(EF, BN, CG) * CG);  // CG = 2: both CTAs' epilogue warps release the leader
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // TMEM: two accumulator buffers of BN fp32 columns x 128 lanes (in each CTA of a pair)
    if (CG == 1) {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(2 * BN)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(2 * BN)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
  }
  tcgen05_fence_before();
  if (CG == 1) __syncthreads();
  else cluster_sync_all();  // the peer's barriers are initialised before anything signals them
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  b2_pdl_wait();  // barrier init / TMEM allocation / descriptor prefetch overlapped the predecessor; its data is needed from here on

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int i = 0;  // running k-block counter (ring position)
      for (int u = first_unit; u < num_units; u += unit_step) {
        const int batch = u / units_per_batch, r = u - batch * units_per_batch;
        const int bo = batch / batch_inner, bi = batch - bo * batch_inner;
        // CONV = 2 orders its units k-chunk-major: the tiles running at any moment then read the same pixels (through
        // different taps) and L2 serves all but the first
        const int tile = CONV == 2 ? r % tiles_per_batch : r / splits, sp = CONV == 2 ? r / tiles_per_batch : r - tile * splits;
        const int tile_m = tile / tiles_n, tile_n = tile - tile_m * tiles_n;
        const int kb0 = sp * kb_per_split;
        const int nkb = min(kb_per_split, num_kb_total - kb0);
        const int abi = bi * ep.a_bc_inner, abo = bo * ep.a_bc_outer, bbi = bi * ep.b_bc_inner, bbo = bo * ep.b_bc_outer;
        // CONV: everything that needs an integer division is worked out once per unit; the k-block loop only steps
        // counters.  (One lane issues all loads of the CTA: four divisions per k-block made it the slowest stage of the
        // pipeline -- 960 cycles per k-block against 512 of MMA.)
        int pn0 = 0, pwq = 0, phq = 0;  // pixel (image, row, column in the pixel grid): CONV = 1 this CTA's first row, CONV = 2 the k-block's first pixel
        int tj = 0, cbi = 0;            // CONV = 1: tap and channel block of the k-block
        int btap[BN / CG / 64], bch[BN / CG / 64];  // CONV = 2: tap and first channel of this CTA's 64-column boxes of B
        if (CONV != 0) {
          const int p0 = CONV == 1 ? tile_m * TM + (int)cta_rank * BM : kb0 * BK, per = ep.cv.pw * ep.cv.ph;
          pn0 = p0 / per;
          const int r2 = p0 - pn0 * per;
          phq = r2 / ep.cv.pw;
          pwq = r2 - phq * ep.cv.pw;
          if (CONV == 1) {
            tj = kb0 / ep.cv.cb;
            cbi = kb0 - tj * ep.cv.cb;
          } else {
#pragma unroll
            for (int b = 0; b < BN / CG / 64; ++b) {
              const int ncol = tile_n * BN + (int)cta_rank * (BN / CG) + b * 64;
              btap[b] = ncol / ep.cv.C;
              bch[b] = ncol - btap[b] * ep.cv.C;
            }
          }
        }
        for (int kb = 0; kb < nkb; ++kb, ++i) {
          const int s = i % STAGES, it = i / STAGES;
          const bool rec = (ep.debug & 16) && blockIdx.x == 0;
          const long long c0 = rec ? clock64() : 0;
          mbar_wait(bar_empty + s * 8, (it & 1) ^ 1);
          if (rec) ep.dbg_out[7] += clock64() - c0;  // producer: waiting for a free stage
          const uint32_t sa = sbase + s * L::STAGE_BYTES, sb = sa + L::A_BYTES;
          // the full barrier of a pair lives in the leader CTA and counts the bytes of both CTAs' loads
          const uint32_t fb = CG == 2 ? mapa_rank(bar_full + s * 8, 0) : bar_full + s * 8;
          if ((EF & EF_GENERIC) && (ep.debug & 4)) {
            if (cta_rank == 0) mbar_arrive(bar_full + s * 8);
            continue;
          }
          if (cta_rank == 0) mbar_expect_tx(bar_full + s * 8, CG * L::STAGE_BYTES);
          const int k = (kb0 + kb) * BK;
          const int arow = tile_m * TM + (int)cta_rank * BM;               // this CTA's rows of A
          constexpr int BNL = BN / CG;                                     // ... and its share of B's rows
          const int brow = tile_n * BN + (int)cta_rank * BNL;
          auto load = [&](uint32_t dst, const CUtensorMap* m, int c0, int c1, int c2, int c3) {
            if (CG == 2) tma_load_4d_pair(dst, m, fb, c0, c1, c2, c3);
            else tma_load_4d(dst, m, fb, c0, c1, c2, c3);
          };
          auto load_px = [&](uint32_t dst, const CUtensorMap* m, int c, int w, int h, int n, int tap) {
            if (CG == 2) tma_load_im2col_4d_pair(dst, m, fb, c, w, h, n, ep.cv.off_w[tap], ep.cv.off_h[tap]);
            else tma_load_im2col_4d(dst, m, fb, c, w, h, n, ep.cv.off_w[tap], ep.cv.off_h[tap]);
          };
          if (CONV == 1) {
            // k-block = (tap, 64 channels): 128 pixels x 64 channels of the activation, shifted by the tap; the weight
            // operand's coordinates follow the tap (forward: k = (tap, cin) as stored; dgrad: column block (tap, cin))
            static_assert(CONV != 1 || !TA, "im2col A operand is K-major");
            const int c0 = cbi * 64;
            load_px(sa, &map_a, c0, ep.cv.base_w + pwq * ep.cv.mult, ep.cv.base_h + phq * ep.cv.mult, pn0, tj);
            const int wt = (int)ep.cv.wtap[tj] * ep.cv.C;
            if (!TB) {
              load(sb, &map_b, wt + c0, brow, 0, 0);
            } else {
#pragma unroll
              for (int b = 0; b < BNL / 64; ++b) load(sb + b * (64 * BK * 2), &map_b, brow + b * 64 + wt, c0, 0, 0);
            }
            if (++cbi == ep.cv.cb) {
              cbi = 0;
              ++tj;
            }
            continue;
          }
          if (CONV == 2) {
            // wgrad: k = pixels.  A = dy^T as usual; B's 64-column boxes are (tap, 64 channels) of 64 consecutive pixels
            static_assert(CONV != 2 || (TA && TB), "im2col B operand is MN-major");
#pragma unroll
            for (int a = 0; a < BM / 64; ++a) load(sa + a * (64 * BK * 2), &map_a, arow + a * 64, k, 0, 0);
            const int ph = ep.cv.base_h + phq * ep.cv.mult, pw = ep.cv.base_w + pwq * ep.cv.mult;
#pragma unroll
            for (int b = 0; b < BNL / 64; ++b) load_px(sb + b * (64 * BK * 2), &map_b, bch[b], pw, ph, pn0, btap[b]);
            pwq += BK;  // the next k-block starts 64 pixels on
            while (pwq >= ep.cv.pw) {
              pwq -= ep.cv.pw;
              if (++phq == ep.cv.ph) {
                phq = 0;
                ++pn0;
              }
            }
            continue;
          }
          if (!TA) {
            load(sa, &map_a, k, arow, abi, abo);   // box {64 k, 128 rows}
          } else {
            // box {64 mn, 64 k-rows}; one box per 64-wide MN atom
#pragma unroll
            for (int a = 0; a < BM / 64; ++a) load(sa + a * (64 * BK * 2), &map_a, arow + a * 64, k, abi, abo);
          }
          if (!TB) {
            load(sb, &map_b, k, brow, bbi, bbo);   // box {64 k, BNL rows}
          } else {
#pragma unroll
            for (int b = 0; b < BNL / 64; ++b) load(sb + b * (64 * BK * 2), &map_b, brow + b * 64, k, bbi, bbo);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // One lane runs the whole loop: a warp-wide poll + __syncwarp per k-block costs 14 % of the MMA rate
    // (tools/mma_rate2.cu: 146 vs 128 cycles per 128x256x16 MMA).
    if (lane == 0 && cta_rank == 0) {
      constexpr uint32_t idesc = make_idesc(BN, TA, TB, TM);
      int i = 0, t = 0;
      const bool rec = (ep.debug & 16) && blockIdx.x == 0;
      long long w_full = 0, w_acc = 0, t_begin = rec ? clock64() : 0;
      for (int u = first_unit; u < num_units; u += unit_step, ++t) {
        const int r = u % units_per_batch;
        const int sp = CONV == 2 ? r / tiles_per_batch : r % splits;
        const int kb0 = sp * kb_per_split;
        const int nkb = min(kb_per_split, num_kb_total - kb0);
        const int buf = t & 1;
        long long c0 = rec ? clock64() : 0;
        mbar_wait(bar_acc_empty + buf * 8, ((t >> 1) & 1) ^ 1);  // epilogue drained this buffer
        if (rec) w_acc += clock64() - c0;
        tcgen05_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BN;
        for (int kb = 0; kb < nkb; ++kb, ++i) {
          const int s = i % STAGES, it = i / STAGES;
          c0 = rec ? clock64() : 0;
          mbar_wait(bar_full + s * 8, it & 1);
          if (rec) w_full += clock64() - c0;
          tcgen05_fence_after();
          const uint32_t sa = sbase + s * L::STAGE_BYTES, sb = sa + L::A_BYTES;
#pragma unroll
          for (int kk = 0; kk < BK / 16; ++kk) {
            // K-major: 8-row groups 1024 B apart; advance 32 B per K=16 inside the swizzle span.
            // MN-major: 64-element MN atoms (64*BK*2 B apart = LBO), 8-row K groups 1024 B apart (SBO).
            const uint64_t da = TA ? make_smem_desc(sa + kk * 2048, 64 * BK * 2, 1024) : make_smem_desc(sa + kk * 32, 16, 1024);
            const uint64_t db = TB ? make_smem_desc(sb + kk * 2048, 64 * BK * 2, 1024) : make_smem_desc(sb + kk * 32, 16, 1024);
            if (CG == 2) umma_bf16_pair(d_tmem, da, db, idesc, (kb | kk) != 0);
            else umma_bf16(d_tmem, da, db, idesc, (kb | kk) != 0);
          }
          if (CG == 2) {
            umma_commit_pair(bar_empty + s * 8);                          // both producers get their stage back
            if (kb == nkb - 1) umma_commit_pair(bar_acc_full + buf * 8);  // both CTAs' epilogues
          } else {
            umma_commit(bar_empty + s * 8);                          // frees the smem stage once these MMAs retire
            if (kb == nkb - 1) umma_commit(bar_acc_full + buf * 8);  // accumulator complete
          }
        }
      }
      if (rec) {
        ep.dbg_out[0] = clock64() - t_begin;  // MMA issue loop, total
        ep.dbg_out[1] = w_full;               //   of which waiting for operands
        ep.dbg_out[2] = w_acc;                //   of which waiting for a drained accumulator
        ep.dbg_out[3] = i;                    // k-blocks issued
        ep.dbg_out[4] = t;                    // tiles
      }
    }
    __syncwarp();
  } else {
    // ===================== epilogue (8 warps: 4 lane quadrants x 2 column halves) =====================
    constexpr uint32_t EFC = EF;
    constexpr int CGROUPS = epi_warps(EF, BN, CG) / 4;
    constexpr int WCOLS = BN / CGROUPS;  // columns per warp
    constexpr int NPAIR = WCOLS / 32;    // pairs of 16-column slices per warp
    const int ew = warp - 2;
    const int lane_grp = warp & 3;       // TMEM lanes [32*lane_grp, +32) -- hardware rule: warp_id % 4
    const int cgrp = ew >> 2;            // columns [cgrp*WCOLS, +WCOLS) of the tile
    float* const bias_base = reinterpret_cast<float*>(smem + L::BIAS_OFF) + cgrp * WCOLS;
    const uint32_t stg0 = sbase + L::STG_OFF + ew * (STGB * 4096), stg1 = stg0 + (STGB - 1) * 4096;
    B2DropKey dk = {0ull, 0u, 1.f};
    if (((EFC & EF_GENERIC) ? ep.flags : EFC) & EF_DROP) dk = dropout_key(ep.drop);
    const float alpha = ep.alpha;
    const float* __restrict__ bias = ep.bias;
    const int debug = (EFC & EF_GENERIC) ? ep.debug : 0;
    int t = 0;
    uint32_t nflush = 0;  // boxes of the main output flushed so far by this warp (selects the staging buffer)
    for (int u = first_unit; u < num_units; u += unit_step, ++t) {
      const int batch = u / units_per_batch, r = u - batch * units_per_batch;
      const int bo = batch / batch_inner, bi = batch - bo * batch_inner;
      const int tile = CONV == 2 ? r % tiles_per_batch : r / splits, sp = CONV == 2 ? r / tiles_per_batch : r - tile * splits;
      const int tile_m = tile / tiles_n, tile_n = tile - tile_m * tiles_n;
      const int buf = t & 1;
      const int row0 = tile_m * TM + (int)cta_rank * BM + lane_grp * 32, row = row0 + lane;
      const bool row_ok = row < M;
      const int colw = tile_n * BN + cgrp * WCOLS;
      // split-K: bias / residual belong to the first split only
      const uint32_t rtf = (sp == 0) ? ep.flags : (ep.flags & ~(EF_BIAS | EF_RES));
      const uint32_t F = (EFC & EF_GENERIC) ? rtf : EFC;
      // with a saved-derivative output the two boxes are (out, dact); otherwise they double-buffer the output
      const bool two_out = (F & EF_DACT) != 0;
      EpiRow er;
      er.res = ep.res + (long long)row * ep.ldres;
      er.aux = ep.aux + (long long)row * ep.ldaux;
      er.e0 = (unsigned long long)row * (unsigned long long)N;
      int spix = -1;  // EF_SCATTER: output pixel of this lane's row
      if (!(EFC & EF_GENERIC) && (EFC & EF_SCATTER) && row_ok) {
        const int per = ep.cv.pw * ep.cv.ph, pn = row / per, r2 = row - pn * per, hq = r2 / ep.cv.pw, wq = r2 - hq * ep.cv.pw;
        spix = (pn * ep.cv.OH + hq * ep.cv.o_mult + ep.cv.o_h0) * ep.cv.OW + wq * ep.cv.o_mult + ep.cv.o_w0;
      }
      // this warp's bias slice -> shared memory and the first slice's operands -> registers, both before the wait
      // for the accumulator so their latency hides behind the MMAs of the tile
      // The four warps of a column half share one bias slice (tile parity selects the buffer; the named barrier
      // below keeps the four within one tile of each other, so buffer t&1 is no longer read when tile t+2 fills it).
      const float* bias_s = bias_base + (t & 1) * BN;
      const bool has_bias = (ep.flags & EF_BIAS) != 0;  // uniform over the CTA's tiles (F drops it on later k-splits)
      if (((EFC & EF_GENERIC) || (EFC & EF_BIAS)) && has_bias) {
        static_assert(WCOLS == 128 || WCOLS == 64, "bias staging assumes 4 warps x 32 lanes cover the half tile");
        const int c = lane_grp * 32 + lane;
        if (c < WCOLS) bias_base[(t & 1) * BN + c] = (colw + c < N) ? bias[colw + c] : 0.f;
      }
      // Residual through TMA (EF_RES_TMA): a row-per-lane 16-byte global load touches 32 cache lines per instruction and
      // the L1 tag stage made the residual read of a 128 x 256 fp32 tile cost more than the tile's MMAs.  The residual box
      // of a 32-column pair is bulk-loaded into the staging box its results will leave from (pair p of this warp: buffer
      // p & 1, barrier phase (p >> 1) & 1), one pair ahead of its use; the first pair of a tile before the accumulator wait.
      constexpr bool RT = !(EFC & EF_GENERIC) && (EFC & EF_RES_TMA) != 0;
      static_assert(!RT || (STGB == 2 && (EFC & EF_F32) && !(EFC & (EF_DACT | EF_ACC))), "TMA residual: fp32 output, two staging boxes");
      const uint32_t rbar = bar_res + ew * 16;
      // The saved-derivative operand (EF_AUX_TMA) takes the same road, in 64-column bf16 boxes (one per two slice pairs).
      constexpr bool AT = !(EFC & EF_GENERIC) && (EFC & EF_AUX_TMA) != 0;
      static_assert(!AT || (STGB == 2 && !(EFC & (EF_F32 | EF_DACT | EF_ACC | EF_RES))), "TMA saved derivative: bf16 output, two staging boxes");
      if ((RT || AT) && lane == 0) {
        bulk_wait_read<1>();  // the store that last read this buffer (two boxes ago) is done with it
        const uint32_t bsel = nflush & 1u;
        mbar_expect_tx(rbar + bsel * 8, 4096);
        tma_load_4d(bsel ? stg1 : stg0, &map_d, rbar + bsel * 8, colw, row0, 0, 0);
      }
      // EF_SCATTER: this lane's ReLU bits for the warp's 128 columns (one 16-byte load per tile), requested before the
      // accumulator wait.  (Reading the mask as bf16 rows of h1 -- 64 KB per tile in flight per SM -- stalled the epilogue
      // warps at load issue for as long as the tile's arithmetic took.)
      constexpr bool SC = !(EFC & EF_GENERIC) && (EFC & EF_SCATTER) != 0;
      static_assert(!SC || WCOLS == 128, "scatter epilogue: 128 columns = 4 mask words per warp");
      uint4 keep = make_uint4(0u, 0u, 0u, 0u);
      if (SC && spix >= 0 && colw < N)
        keep = ld_nc_u4(reinterpret_cast<const uint32_t*>(ep.aux) + (long long)spix * (ep.ldaux >> 5) + (colw >> 5));
      EpiPre pa, pb;
      if (!(EFC & EF_GENERIC) && row_ok) epi_prefetch<EFC>(rtf, er, colw, N, pa);
      if (((EFC & EF_GENERIC) || (EFC & EF_BIAS)) && has_bias) asm volatile("bar.sync %0, 128;" ::"r"(1 + cgrp) : "memory");
      const bool rec = (ep.debug & 16) && blockIdx.x == 0 && ew == 0 && lane == 0;
      long long c0 = rec ? clock64() : 0;
      mbar_wait(bar_acc_full + buf * 8, (t >> 1) & 1);
      if (rec) { long long c1 = clock64(); ep.dbg_out[5] += c1 - c0; c0 = c1; }
      tcgen05_fence_after();
      // TMEM -> registers is double-buffered in 16-column slices: the load of slice i+1 is in flight while slice i
      // runs the epilogue math.  The loop walks slice PAIRS and is not unrolled, so the (long) body exists twice only.
      const uint32_t tbase = tmem_base + ((uint32_t)(lane_grp * 32) << 16) + buf * BN + cgrp * WCOLS;
      uint32_t ra[16], rb[16];
      tmem_ld16(tbase, ra);
#pragma unroll 1
      for (int cp = 0; cp < NPAIR; ++cp) {
        const int c0c = colw + cp * 32, c1c = c0c + 16, c2c = c0c + 32;
        const int sl0 = cp * 2, sl1 = sl0 + 1;
        // A box of the main output spans 4 slices (bf16, 64 columns) or 2 slices (fp32, 32 columns).
        const bool box_start = (F & EF_F32) ? true : ((cp & 1) == 0);
        const bool box_end = (F & EF_F32) ? true : ((cp & 1) == 1);
        const uint32_t so = (two_out || STGB == 1) ? stg0 : ((nflush & 1) ? stg1 : stg0);
        long long q0 = rec ? clock64() : 0;
        if (RT) {
          if (lane == 0 && cp + 1 < NPAIR) {  // next pair's residual -> the other buffer, once its last store has read it
            bulk_wait_read<0>();
            const uint32_t bn = (nflush + 1u) & 1u;
            mbar_expect_tx(rbar + bn * 8, 4096);
            tma_load_4d(bn ? stg1 : stg0, &map_d, rbar + bn * 8, c2c, row0, 0, 0);
          }
          mbar_wait(rbar + (nflush & 1u) * 8, (nflush >> 1) & 1u);
        } else if (AT) {
          if (box_start) {
            if (lane == 0 && cp + 2 < NPAIR) {  // next box's saved derivative -> the other buffer, once its last store has read it
              bulk_wait_read<0>();
              const uint32_t bn = (nflush + 1u) & 1u;
              mbar_expect_tx(rbar + bn * 8, 4096);
              tma_load_4d(bn ? stg1 : stg0, &map_d, rbar + bn * 8, c0c + 64, row0, 0, 0);
            }
            mbar_wait(rbar + (nflush & 1u) * 8, (nflush >> 1) & 1u);
          }
        } else if (box_start && !(!(EFC & EF_GENERIC) && (EFC & EF_SCATTER) != 0)) {
          // the bulk store that last read this staging box must have finished reading it
          if (lane == 0) {
            if (two_out || STGB == 1) bulk_wait_read<0>();
            else bulk_wait_read<1>();
          }
          __syncwarp();
        }
        if (rec) { const long long q1 = clock64(); ep.dbg_out[8] += q1 - q0; q0 = q1; }
        tmem_ld_wait();
        if (rec) { const long long q1 = clock64(); ep.dbg_out[9] += q1 - q0; q0 = q1; }
        tmem_ld16(tbase + cp * 32 + 16, rb);
        if (row_ok && c0c < N && !(debug & 2)) {
          if (EFC & EF_GENERIC) epi_prefetch<EFC>(rtf, er, c0c, N, pa);
          else if (c1c < N) epi_prefetch<EFC>(rtf, er, c1c, N, pb);
          const uint32_t kw = SC ? (cp == 0 ? keep.x : cp == 1 ? keep.y : cp == 2 ? keep.z : keep.w) : 0u;
          epilogue_slice<EFC>(rtf, ep, dk, er, alpha, bias_s + cp * 32, pa, ra, c0c, N, sl0, lane, so, stg1, kw & 0xffffu);
        }
        tmem_ld_wait();
        if (cp + 1 < NPAIR) tmem_ld16(tbase + cp * 32 + 32, ra);
        if (row_ok && c1c < N && !(debug & 2)) {
          if (EFC & EF_GENERIC) epi_prefetch<EFC>(rtf, er, c1c, N, pb);
          else if (cp + 1 < NPAIR && c2c < N) epi_prefetch<EFC>(rtf, er, c2c, N, pa);
          const uint32_t kw = SC ? (cp == 0 ? keep.x : cp == 1 ? keep.y : cp == 2 ? keep.z : keep.w) : 0u;
          epilogue_slice<EFC>(rtf, ep, dk, er, alpha, bias_s + cp * 32 + 16, pb, rb, c1c, N, sl1, lane, so, stg1, kw >> 16);
        }
        if (rec) { const long long q1 = clock64(); ep.dbg_out[10] += q1 - q0; q0 = q1; }
        if (SC) {
          if (box_end) {
            // the 32 x 64 box leaves row by row: 8 lanes move one row's 128 bytes (4 rows per pass), each row to the pixel its
            // lane computed above
            __syncwarp();
            const int bcol = c0c - 32 + (lane & 7) * 8;
            bf16* __restrict__ outp = reinterpret_cast<bf16*>(ep.out);
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              const int rr = q * 4 + (lane >> 3);
              const int px = __shfl_sync(0xffffffffu, spix, rr);
              uint32_t v[4];
              lds128u(so + rr * 128 + (((lane & 7) ^ (rr & 7)) << 4), v[0], v[1], v[2], v[3]);
              if (px >= 0 && bcol < N) *reinterpret_cast<uint4*>(outp + (long long)px * ep.ldc + bcol) = make_uint4(v[0], v[1], v[2], v[3]);
            }
            __syncwarp();  // the box is rewritten by the next slices
          }
        } else if (box_end && !(debug & 1)) {
          fence_async_smem();  // generic-proxy writes of every lane -> visible to the async proxy
          __syncwarp();
          const int bcol = (F & EF_F32) ? c0c : c0c - 32;  // first column of the box
          if (F & EF_COLSUM) {
            // column sums of the staged 32 x 64 bf16 box: lane l owns columns 2l, 2l+1 (16-byte chunk l / 4 of every row,
            // XOR-swizzled with the row: the 32 lanes of a load hit 32 different banks); rows past M hold stale data
            const int nrows = min(32, M - row0);
            const uint32_t cbase = so + (uint32_t)(lane & 3) * 4u;
            float s0 = 0.f, s1 = 0.f;
#pragma unroll
            for (int rr = 0; rr < 32; ++rr) {
              if (rr < nrows) {
                uint32_t w;
                asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w) : "r"(cbase + (uint32_t)rr * 128u + ((((uint32_t)lane >> 2) ^ (uint32_t)(rr & 7)) << 4)));
                const float2 f2 = unpack_bf16x2(w);
                s0 += f2.x;
                s1 += f2.y;
              }
            }
            const int col = bcol + 2 * lane;
            if (col < N && nrows > 0) {
              atomicAdd(ep.colsum + col, s0);
              atomicAdd(ep.colsum + col + 1, s1);
            }
          }
          if (lane == 0) {
            if (bcol < N && row0 < M) {
              if (F & EF_ACC) tma_reduce_add_4d(&map_c, so, bcol, row0, bi, bo);
              else tma_store_4d(&map_c, so, bcol, row0, bi, bo);
              if (two_out) tma_store_4d(&map_d, stg1, bcol, row0, bi, bo);
            }
            bulk_commit();  // also when nothing was stored: keeps the group count in step with the staging buffers
          }
          if (!two_out) ++nflush;
        }
        if (rec) { const long long q1 = clock64(); ep.dbg_out[11] += q1 - q0; q0 = q1; }
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CG == 2) mbar_arrive_cluster(mapa_rank(bar_acc_empty + buf * 8, 0));
        else mbar_arrive(bar_acc_empty + buf * 8);
      }
      if (rec) ep.dbg_out[6] += clock64() - c0;  // epilogue warp 0: busy cycles (dbg_out[5] = waiting for accumulators)
    }
    if (lane == 0) bulk_wait_read<0>();  // the staging boxes are no longer being read when the CTA exits
  }
  tcgen05_fence_before();
  if (CG == 1) __syncthreads();
  else cluster_sync_all();  // neither CTA's shared memory / TMEM goes away while the other may still touch it
  if (warp == 1) {
    tcgen05_fence_after();
    if (CG == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(2 * BN) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(2 * BN) : "memory");
  }
}

// ----------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// 4-D tensor [n_outer, n_inner, rows, cols] of bf16 (esize 2) or fp32 (esize 4): row stride ld, batch strides
// s_inner / s_outer (elements; 0 = broadcast -> that dimension gets extent 1).  box = {box_cols, box_rows, 1, 1}.
int make_map(CUtensorMap* m, const void* base, long long rows, long long cols, long long ld, long long n_inner,
             long long s_inner, long long n_outer, long long s_outer, int box_cols, int box_rows, int esize = 2) {
  EncodeTiledFn enc = get_encode();
  if (enc == nullptr) {
    b200slu_set_error("cuTensorMapEncodeTiled driver entry point unavailable");
    return B200SLU_ECUDA;
  }
  if (s_inner == 0) n_inner = 1;
  if (s_outer == 0) n_outer = 1;
  if (n_inner == 1) s_inner = ld;  // any legal stride; the coordinate is always 0
  if (n_outer == 1) s_outer = ld;
  cuuint64_t dims[4] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)n_inner, (cuuint64_t)n_outer};
  cuuint64_t strides[3] = {(cuuint64_t)ld * esize, (cuuint64_t)s_inner * esize, (cuuint64_t)s_outer * esize};
  cuuint32_t box[4] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(m, esize == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4,
                   const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    b200slu_set_error("cuTensorMapEncodeTiled failed (%d): rows=%lld cols=%lld ld=%lld esize=%d batch=(%lld x %lld, strides %lld %lld) base=%p",
                      (int)r, rows, cols, ld, esize, n_inner, n_outer, s_inner, s_outer, base);
    return B200SLU_ECUDA;
  }
  return 0;
}

template <int BN, int STAGES, bool TA, bool TB, uint32_t EF, int CG, int CONV = 0>
int launch(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc, const CUtensorMap& md, int M, int N, int K, int nbatch, int batch_inner, int splits,
           int kb_per, const EpiArgs& ep, cudaStream_t st) {
  using L = SmemLayout<BN, STAGES, stg_bufs(EF, BN, CG), CG, epi_warps(EF, BN, CG)>;
  auto kern = gemm_bf16_tcgen05<BN, STAGES, TA, TB, EF, CG, CONV>;
  static unsigned long long configured = 0;  // per device
  if (!b2_device_done(configured)) {
    B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
    b2_device_mark(configured);
  }
  const int tiles_n = (N + BN - 1) / BN, tiles_m = (M + BM * CG - 1) / (BM * CG);
  const long long units = (long long)tiles_n * tiles_m * splits * nbatch;
  B2_CHECK_ARG(units < (1LL << 30), "gemm: too many work units (%lld)", units);
  const int slots = b2_num_sms() / CG;  // CTAs (CG = 1) or CTA pairs (CG = 2: one pair per TPC)
  const int grid = (units < slots ? (int)units : slots) * CG;
  EpiArgs epl = ep;
  long long* dbg_buf = nullptr;
  if (ep.debug & 16) {
    dbg_buf = b2_debug_scratch();
    B2_CHECK_ARG(dbg_buf != nullptr, "gemm: debug scratch unavailable");
    B2_CUDA(cudaMemsetAsync(dbg_buf, 0, 12 * sizeof(long long), st));
    epl.dbg_out = dbg_buf;
  }
  if (CG == 1) {
    B2_CUDA(b2_launch(kern, dim3(grid), dim3(num_threads(EF, BN, CG)), (size_t)L::TOTAL, st, ma, mb, mc, md, M, N, K, tiles_n,
                      tiles_n * tiles_m, batch_inner, splits, kb_per, (int)units, epl));
  } else {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(num_threads(EF, BN, CG));
    cfg.dynamicSmemBytes = (size_t)L::TOTAL;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = b2_pdl_enabled() ? 2 : 1;
    B2_CUDA(cudaLaunchKernelEx(&cfg, kern, ma, mb, mc, md, M, N, K, tiles_n, tiles_n * tiles_m, batch_inner, splits, kb_per,
                               (int)units, epl));
  }
  if (ep.debug & 16) {
    long long h[12];
    B2_CUDA(cudaStreamSynchronize(st));
    B2_CUDA(cudaMemcpy(h, dbg_buf, sizeof(h), cudaMemcpyDeviceToHost));
    fprintf(stderr,
            "[gemm dbg] M=%d N=%d K=%d BN=%d CG=%d TA=%d TB=%d splits=%d grid=%d conv=%d | CTA0: tiles=%lld kblocks=%lld mma_loop=%lld cyc "
            "(wait operands %lld, wait accumulator %lld) | epi warp0: wait acc %lld, busy %lld [staging free %lld, tmem wait %lld, "
            "slices %lld, flush %lld] | producer wait stage %lld | EF=%u CONV=%d\n",
            M, N, K, BN, CG, (int)TA, (int)TB, splits, grid, CONV, h[4], h[3], h[0], h[1], h[2], h[5], h[6], h[8], h[9], h[10], h[11], h[7],
            (unsigned)EF, CONV);
  }
  return 0;
}

}  // namespace

static int gemm_debug_env() {
  static const int dbg = getenv("B200SLU_GEMM_DEBUG") ? atoi(getenv("B200SLU_GEMM_DEBUG")) : 0;
  return dbg;
}

// Internal C++ entry (used by the C ABI below).
int b2_gemm_ex(const b200slu_gemm_t* g, const b200slu_epilogue_t* e, cudaStream_t st) {
  B2_CHECK_ARG(g != nullptr, "gemm: null descriptor");
  const int M = g->M, N = g->N, K = g->K;
  const int nbi = g->batch_inner > 0 ? g->batch_inner : 1, nbo = g->batch_outer > 0 ? g->batch_outer : 1;
  const int nbatch = nbi * nbo;
  B2_CHECK_ARG(M > 0 && N > 0 && K > 0, "gemm: empty problem M=%d N=%d K=%d", M, N, K);
  B2_CHECK_ARG(((uintptr_t)g->A % 16) == 0 && ((uintptr_t)g->B % 16) == 0 && g->lda % 8 == 0 && g->ldb % 8 == 0,
               "gemm: operands must be 16-byte aligned with row strides multiple of 8 elements (lda=%lld ldb=%lld)", g->lda,
               g->ldb);
  B2_CHECK_ARG(g->a_stride_inner % 8 == 0 && g->a_stride_outer % 8 == 0 && g->b_stride_inner % 8 == 0 &&
                   g->b_stride_outer % 8 == 0,
               "gemm: batch strides must be multiples of 8 elements");
  EpiArgs ep{};
  ep.out = g->C;
  ep.ldc = g->ldc;
  ep.c_bs_inner = g->c_stride_inner;
  ep.c_bs_outer = g->c_stride_outer;
  ep.a_bc_inner = g->a_stride_inner != 0;
  ep.a_bc_outer = g->a_stride_outer != 0;
  ep.b_bc_inner = g->b_stride_inner != 0;
  ep.b_bc_outer = g->b_stride_outer != 0;
  ep.alpha = 1.f;
  if (e != nullptr) {
    ep.bias = e->bias;
    ep.res = e->residual;
    ep.ldres = e->ld_residual;
    ep.aux = (const bf16*)e->act_grad_input;
    ep.ldaux = e->ld_act_grad_input;
    ep.aux_mode = e->act_grad;
    ep.pre_out = (bf16*)e->act_grad_out;
    ep.alpha = e->alpha;
    ep.act = e->act;
    ep.out_f32 = e->out_f32;
    ep.accumulate = e->accumulate;
    ep.drop = b2_make_drop(e->dropout_p, e->dropout_seed, e->dropout_site);
    ep.colsum = e->out_colsum_accum;
    ep.debug = gemm_debug_env();
  }
  {
    uint32_t f = 0;
    if (ep.bias != nullptr) f |= EF_BIAS;
    if (ep.act == B200SLU_ACT_SWISH) f |= EF_SWISH;
    if (ep.act == B200SLU_ACT_RELU) f |= EF_RELU;
    if (ep.pre_out != nullptr) f |= EF_DACT;
    if (ep.drop.p > 0.f) f |= EF_DROP;
    if (ep.alpha != 1.f && !(f & EF_DROP)) f |= EF_ALPHA;  // with dropout, alpha rides on the keep multiplier
    if (ep.aux != nullptr)
      f |= ep.aux_mode == B200SLU_ACT_MUL ? EF_AUX_MUL : ep.aux_mode == B200SLU_ACT_SWISH ? EF_AUX_SWISH : EF_AUX_RELU;
    if (ep.res != nullptr) f |= EF_RES;
    if (ep.out_f32) f |= EF_F32;
    if (ep.accumulate) f |= EF_ACC;
    if (ep.colsum != nullptr) f |= EF_COLSUM;
    // residual boxes through TMA (specialised fp32-output kernels only): whole 32-column boxes, 16-byte aligned rows
    static const bool res_tma_off = getenv("B200SLU_GEMM_RES_TMA") && atoi(getenv("B200SLU_GEMM_RES_TMA")) == 0;  // tuning only
    if ((f & EF_RES) && (f & EF_F32) && !(f & (EF_ACC | EF_DACT)) && nbatch == 1 && N % 32 == 0 && ep.ldres % 4 == 0 &&
        ((uintptr_t)ep.res % 16) == 0 && !res_tma_off)
      f |= EF_RES_TMA;
    // saved-derivative boxes through TMA (specialised bf16-output dgrad kernels only): whole 64-column boxes
    static const bool aux_tma_off = getenv("B200SLU_GEMM_AUX_TMA") && atoi(getenv("B200SLU_GEMM_AUX_TMA")) == 0;  // tuning only
    if ((f & (EF_AUX_MUL | EF_AUX_RELU)) && !(f & (EF_F32 | EF_ACC | EF_DACT | EF_RES)) && nbatch == 1 && N % 64 == 0 && ep.ldaux % 8 == 0 &&
        ((uintptr_t)ep.aux % 16) == 0 && !aux_tma_off)
      f |= EF_AUX_TMA;
    ep.flags = f;
  }
  B2_CHECK_ARG(e == nullptr || e->act == B200SLU_ACT_NONE || e->act == B200SLU_ACT_SWISH || e->act == B200SLU_ACT_RELU,
               "gemm: act must be NONE / SWISH / RELU");
  B2_CHECK_ARG(nbatch == 1 || (ep.res == nullptr && ep.aux == nullptr && ep.pre_out == nullptr && !(ep.drop.p > 0.f)),
               "gemm: batched mode supports only bias / activation / alpha epilogues");
  B2_CHECK_ARG(!(ep.drop.p > 0.f) || N % 8 == 0, "gemm: dropout epilogue needs N %% 8 == 0");
  B2_CHECK_ARG(!ep.accumulate || ep.out_f32, "gemm: accumulate needs fp32 output");
  B2_CHECK_ARG(ep.colsum == nullptr || (!ep.out_f32 && N % 64 == 0 && nbatch == 1 && ep.pre_out == nullptr),
               "gemm: out_colsum_accum needs an unbatched bf16 output with N %% 64 == 0 and no act_grad_out");
  B2_CHECK_ARG(ep.pre_out == nullptr || (!ep.out_f32 && g->ldc % 8 == 0 && ((uintptr_t)ep.pre_out % 16) == 0),
               "gemm: act_grad_out needs a bf16 output and shares its row stride");
  const int vec = ep.out_f32 ? 4 : 8;
  B2_CHECK_ARG(g->ldc % vec == 0 && ((uintptr_t)g->C % 16) == 0 && g->c_stride_inner % vec == 0 && g->c_stride_outer % vec == 0,
               "gemm: output must be 16-byte aligned, ldc / batch strides %% %d == 0", vec);
  B2_CHECK_ARG(ep.res == nullptr || (ep.ldres % 4 == 0 && ((uintptr_t)ep.res % 16) == 0), "gemm: residual alignment");
  B2_CHECK_ARG(ep.aux == nullptr || (ep.ldaux % 8 == 0 && ((uintptr_t)ep.aux % 16) == 0), "gemm: act_grad_input alignment");
  B2_CHECK_ARG(ep.bias == nullptr || ((uintptr_t)ep.bias % 16) == 0, "gemm: bias alignment");

  const int num_kb = (K + BK - 1) / BK;
  // Tile width and split-K factor from a small cost model (units of one k-block of a 128-wide tile):
  //  * a tcgen05 MMA with N = 128 retires no faster than one with N = 256 (tools/mma_rate2.cu: 114-127 vs 128 cycles),
  //    so the wide tile does twice the work per tensor-pipe cycle; it pays 1.15x per k-block for the larger L2 -> SM load;
  //  * every tile adds its epilogue / pipeline turn-around;
  //  * the persistent grid runs ceil(units / SMs) waves, each as long as its longest unit.
  //  Split-K only for accumulating outputs (wgrad: K = B*T is long, the output is small).
  //  * a CTA pair (cta_group::2, 256 x 256 tile) loads 32 KB instead of 48 KB per CTA and k-block and its ring is 5-6
  //    stages deep: measured 8 % faster per k-block once K >= 1024 (long main loops), slower below (the pair's
  //    cross-CTA barrier latencies are paid per tile); unbatched GEMMs only.
  int bn = 128, splits = 1, cg = 1;
  {
    static const int force_bn = getenv("B200SLU_GEMM_BN") ? atoi(getenv("B200SLU_GEMM_BN")) : 0;  // tuning experiments only
    static const int force_cg = getenv("B200SLU_GEMM_CG") ? atoi(getenv("B200SLU_GEMM_CG")) : 0;
    static const int wide_min = getenv("B200SLU_GEMM_WIDE_MIN") ? atoi(getenv("B200SLU_GEMM_WIDE_MIN")) : 129;  // N in (128, 256): one wide tile beats 128 + a sliver
    const long long sms = b2_num_sms();
    double best = 1e300;
    for (int cand = 0; cand < 3; ++cand) {
      const int cbn = cand == 0 ? 128 : 256, ccg = cand == 2 ? 2 : 1;
      if (cbn == 256 && N < wide_min) continue;
      // pairs pay off on long main loops, or when a pair runs so many tiles that its start-up / drain is amortised
      // (conv2 dgrad, K = 512, 358 tiles per pair: 1530 us against 1697 us)
      const long long pair_tiles = ((M + 2 * BM - 1) / (2 * BM)) * ((N + 255) / 256);
      if (ccg == 2 && (nbatch != 1 || M <= BM ||
                       (force_cg != 2 && num_kb / (ep.accumulate ? 8 : 1) < 16 && pair_tiles < 32 * (sms / 2))))
        continue;
      if ((force_bn == 128 || (force_bn == 256 && N >= 256)) && cbn != force_bn) continue;
      if (force_cg != 0 && cbn == 256 && ccg != force_cg && !(force_cg == 2 && (nbatch != 1 || M <= BM))) continue;
      const long long tm = (M + BM * ccg - 1) / (BM * ccg);
      const long long tiles = tm * ((N + cbn - 1) / cbn) * nbatch;
      const long long slots = sms / ccg;
      const double w = ccg == 2 ? 1.05 : cbn == 256 ? 1.15 : 1.0, c = cbn == 256 ? 4.0 : 2.0;
      const int max_s = ep.accumulate ? max(1, num_kb / 4) : 1;
      for (int sp = 1; sp <= max_s && sp <= 64; ++sp) {
        const int kbs = (num_kb + sp - 1) / sp;
        const long long waves = (tiles * sp + slots - 1) / slots;
        const double cost = (double)waves * (kbs * w + c);
        if (cost < best * 0.999) {
          best = cost;
          bn = cbn;
          cg = ccg;
          splits = sp;
        }
      }
    }
  }
  const int kb_per = (num_kb + splits - 1) / splits;
  splits = (num_kb + kb_per - 1) / kb_per;
  if (ep.flags & EF_AUX_TMA) {  // implemented by the two specialised dgrad kernels below only
    const uint32_t base = ep.flags & ~EF_AUX_TMA;
    const bool ok = bn == 256 && !g->trans_a && g->trans_b && (ep.debug & ~16) == 0 &&
                    (base == EF_AUX_MUL || base == (EF_AUX_MUL | EF_COLSUM) || base == EF_AUX_RELU);
    if (!ok) ep.flags = base;
  }
  if (ep.flags & EF_RES_TMA) {  // only the specialised kernels below implement it; anything else keeps the register path
    const uint32_t base = ep.flags & ~EF_RES_TMA;
    const bool ok = bn == 256 && !g->trans_a && (ep.debug & ~16) == 0 &&
                    ((!g->trans_b && base == (EF_BIAS | EF_DROP | EF_RES | EF_F32)) || (g->trans_b && base == (EF_RES | EF_F32)));
    if (!ok) ep.flags = base;
  }

  CUtensorMap ma, mb;
  const double flops = 2.0 * M * N * (double)K * nbatch;
  B2_SCOPE(FAM_GEMM, st, flops, nbatch * (2.0 * ((double)M * K + (double)N * K) + (ep.out_f32 ? 4.0 : 2.0) * M * N));
  if (!g->trans_a)
    B2_TRY(make_map(&ma, g->A, M, K, g->lda, nbi, g->a_stride_inner, nbo, g->a_stride_outer, BK, BM));
  else
    B2_TRY(make_map(&ma, g->A, K, M, g->lda, nbi, g->a_stride_inner, nbo, g->a_stride_outer, 64, BK));
  if (!g->trans_b)
    B2_TRY(make_map(&mb, g->B, N, K, g->ldb, nbi, g->b_stride_inner, nbo, g->b_stride_outer, BK, bn / cg));
  else
    B2_TRY(make_map(&mb, g->B, K, N, g->ldb, nbi, g->b_stride_inner, nbo, g->b_stride_outer, 64, BK));
  // output boxes: 32 rows x 128 bytes, one per epilogue warp and flush
  CUtensorMap mc, md;
  B2_TRY(make_map(&mc, g->C, M, N, g->ldc, nbi, g->c_stride_inner, nbo, g->c_stride_outer, ep.out_f32 ? 32 : 64, 32,
                  ep.out_f32 ? 4 : 2));
  if (ep.pre_out != nullptr)
    B2_TRY(make_map(&md, ep.pre_out, M, N, g->ldc, 1, 0, 1, 0, 64, 32, 2));
  else if (ep.flags & EF_RES_TMA)
    B2_TRY(make_map(&md, ep.res, M, N, ep.ldres, 1, 0, 1, 0, 32, 32, 4));  // residual boxes share the second-output slot
  else if (ep.flags & EF_AUX_TMA)
    B2_TRY(make_map(&md, ep.aux, M, N, ep.ldaux, 1, 0, 1, 0, 64, 32, 2));  // ... and so do the saved-derivative boxes
  else
    md = mc;
  // Specialised epilogues for the hot GEMMs of the train step; anything else runs the generic kernel.
#define B2_ARGS ma, mb, mc, md, M, N, K, nbatch, nbi, splits, kb_per, ep, st
#define B2_L(BN_, ST_, TA_, TB_, EF_)                                                          \
  do {                                                                                         \
    if (BN_ == 256 && cg == 2) return launch<256, num_stages<256, 2>(EF_), TA_, TB_, EF_, 2>(B2_ARGS); \
    return launch<BN_, num_stages<BN_, 1>(EF_), TA_, TB_, EF_, 1>(B2_ARGS);                     \
  } while (0)
#define B2_CASE(BN_, ST_, TA_, TB_, EF_) \
  if (ep.flags == (EF_)) B2_L(BN_, ST_, TA_, TB_, (EF_))
  const bool dbg_generic = (ep.debug & ~16) != 0;
  if (bn == 256) {
    if (!g->trans_a && !g->trans_b) {  // forward:  y = x W^T
      if (!dbg_generic) {
        B2_CASE(256, 3, false, false, EF_BIAS);
        B2_CASE(256, 3, false, false, EF_BIAS | EF_SWISH | EF_DACT | EF_DROP);
        B2_CASE(256, 3, false, false, EF_BIAS | EF_RELU | EF_DACT | EF_DROP);
        B2_CASE(256, 3, false, false, EF_BIAS | EF_DROP | EF_RES | EF_F32);
        B2_CASE(256, 3, false, false, EF_BIAS | EF_DROP | EF_RES | EF_F32 | EF_RES_TMA);
        B2_CASE(256, 3, false, false, EF_BIAS | EF_RELU);
        B2_CASE(256, 3, false, false, EF_BIAS | EF_F32);
        B2_CASE(256, 3, false, false, EF_BIAS | EF_DROP | EF_F32);
        B2_CASE(256, 3, false, false, EF_F32);
        B2_CASE(256, 3, false, false, 0u);
      }
      B2_L(256, 3, false, false, EF_GENERIC);
    }
    if (!g->trans_a && g->trans_b) {  // dgrad:  dx = dy W
      if (!dbg_generic) {
        B2_CASE(256, 3, false, true, 0u);
        B2_CASE(256, 3, false, true, EF_AUX_MUL);
        B2_CASE(256, 3, false, true, EF_AUX_MUL | EF_COLSUM);
        B2_CASE(256, 3, false, true, EF_AUX_MUL | EF_AUX_TMA);
        B2_CASE(256, 3, false, true, EF_AUX_RELU | EF_AUX_TMA);
        B2_CASE(256, 3, false, true, EF_AUX_MUL | EF_COLSUM | EF_AUX_TMA);
        B2_CASE(256, 3, false, true, EF_AUX_RELU);
        B2_CASE(256, 3, false, true, EF_RES | EF_F32);
        B2_CASE(256, 3, false, true, EF_RES | EF_F32 | EF_RES_TMA);
        B2_CASE(256, 3, false, true, EF_F32);
      }
      B2_L(256, 3, false, true, EF_GENERIC);
    }
    if (g->trans_a && g->trans_b) {  // wgrad:  dW += dy^T x
      if (!dbg_generic) B2_CASE(256, 3, true, true, EF_F32 | EF_ACC);
      B2_L(256, 3, true, true, EF_GENERIC);
    }
    B2_L(256, 3, true, false, EF_GENERIC);
  }
  // narrow tiles: the batched attention contractions (N = T or d_k) and small-N heads
  if (!g->trans_a && !g->trans_b) {
    if (!dbg_generic) {
      B2_CASE(128, 4, false, false, EF_F32);
      B2_CASE(128, 4, false, false, 0u);
      B2_CASE(128, 4, false, false, EF_BIAS);
    }
    B2_L(128, 4, false, false, EF_GENERIC);
  }
  if (!g->trans_a && g->trans_b) {
    if (!dbg_generic) B2_CASE(128, 4, false, true, 0u);
    B2_L(128, 4, false, true, EF_GENERIC);
  }
  if (g->trans_a && g->trans_b) {
    if (!dbg_generic) {
      B2_CASE(128, 4, true, true, EF_F32 | EF_ACC);
      B2_CASE(128, 4, true, true, 0u);
    }
    B2_L(128, 4, true, true, EF_GENERIC);
  }
  B2_L(128, 4, true, false, EF_GENERIC);
#undef B2_CASE
#undef B2_L
#undef B2_ARGS
}

// ----------------------------------------------------------------------------- conv.2 as implicit GEMM
namespace {

typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// im2col map over a channels-last bf16 tensor [Nimg, H, W, C]: base pixels run over [lower, dim - 1 + upper] per axis in
// steps of `stride`; a load moves `pixels` consecutive base pixels x 64 channels.
int make_im2col_map(CUtensorMap* m, const void* base, int Nimg, int H, int W, int C, int lower_w, int lower_h, int upper_w,
                    int upper_h, int stride, int pixels) {
  static EncodeIm2colFn enc = nullptr;
  if (enc == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      enc = (EncodeIm2colFn)p;
  }
  if (enc == nullptr) {
    b200slu_set_error("cuTensorMapEncodeIm2col driver entry point unavailable");
    return B200SLU_ECUDA;
  }
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)Nimg};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  int lower[2] = {lower_w, lower_h}, upper[2] = {upper_w, upper_h};
  cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, lower, upper, 64,
                   (cuuint32_t)pixels, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    b200slu_set_error("cuTensorMapEncodeIm2col failed (%d): N=%d H=%d W=%d C=%d corners (%d,%d)/(%d,%d) stride %d", (int)r, Nimg, H,
                      W, C, lower_w, lower_h, upper_w, upper_h, stride);
    return B200SLU_ECUDA;
  }
  // Drivers up to CUDA 13.1 encode im2col maps of tensors smaller than 128 KB with a flag the hardware path then trips
  // over (same correction as CUTLASS' make_im2col_tma_copy_desc applies).
  int drv = 0;
  if (cudaDriverGetVersion(&drv) == cudaSuccess && drv <= 13010 && (long long)Nimg * H * W * C * 2 < 131072)
    reinterpret_cast<uint64_t*>(m)[1] &= ~(1ull << 21);
  return 0;
}

struct Conv2Geom {
  int B, T1, F1, C, T2, F2;
  long long M;  // output pixels
};

int conv2_geom(int B, int T1, int F1, int C, Conv2Geom* g) {
  B2_CHECK_ARG(B > 0 && T1 >= 3 && F1 >= 3 && C > 0 && C % 256 == 0 && C <= 4096, "conv2: B > 0, T1, F1 >= 3, C a multiple of 256 (got B=%d T1=%d F1=%d C=%d)",
               B, T1, F1, C);
  g->B = B; g->T1 = T1; g->F1 = F1; g->C = C;
  g->T2 = (T1 - 3) / 2 + 1;
  g->F2 = (F1 - 3) / 2 + 1;
  g->M = (long long)B * g->T2 * g->F2;
  B2_CHECK_ARG((long long)B * T1 * F1 < (1LL << 30) && T1 < 32768 && F1 < 32768, "conv2: tensor too large for 32-bit pixel indices");
  return 0;
}

}  // namespace

extern "C" int b200slu_conv2_implicit_supported(int B, int T1, int F1, int C) {
  return B > 0 && T1 >= 3 && F1 >= 3 && C > 0 && C % 256 == 0 && C <= 4096 && (long long)B * T1 * F1 < (1LL << 30) && T1 < 32768 &&
         F1 < 32768;
}

// h2[(b,t,f), co] = relu(bias[co] + sum_{kt,kf,ci} h1[b, 2t+kt, 2f+kf, ci] * w[co, (kt*3+kf)*C + ci])
extern "C" int b200slu_conv2_fwd(const void* h1, int B, int T1, int F1, int C, const void* w, const float* bias, void* h2,
                                 void* stream) {
  Conv2Geom g;
  B2_TRY(conv2_geom(B, T1, F1, C, &g));
  cudaStream_t st = (cudaStream_t)stream;
  B2_CHECK_ARG(((uintptr_t)h1 % 16) == 0 && ((uintptr_t)w % 16) == 0 && ((uintptr_t)h2 % 16) == 0 && ((uintptr_t)bias % 16) == 0,
               "conv2_fwd: 16-byte aligned operands");
  const int M = (int)g.M, N = C, K = 9 * C;
  CUtensorMap ma, mb, mc;
  B2_TRY(make_im2col_map(&ma, h1, B, T1, F1, C, 0, 0, -2, -2, 2, BM));
  B2_TRY(make_map(&mb, w, N, K, K, 1, 0, 1, 0, BK, 128));
  B2_TRY(make_map(&mc, h2, M, N, N, 1, 0, 1, 0, 64, 32, 2));
  EpiArgs ep{};
  ep.out = h2; ep.ldc = N; ep.alpha = 1.f; ep.bias = bias; ep.act = B200SLU_ACT_RELU;
  ep.flags = EF_BIAS | EF_RELU;
  ep.debug = gemm_debug_env() & 16;
  ep.cv.pw = g.F2; ep.cv.ph = g.T2; ep.cv.mult = 2; ep.cv.base_w = 0; ep.cv.base_h = 0; ep.cv.cb = C / 64; ep.cv.C = C;
  for (int t = 0; t < 9; ++t) {
    ep.cv.off_h[t] = (unsigned char)(t / 3);
    ep.cv.off_w[t] = (unsigned char)(t % 3);
    ep.cv.wtap[t] = (unsigned char)t;
  }
  B2_SCOPE(FAM_GEMM, st, 2.0 * M * N * (double)K, 2.0 * ((double)B * T1 * F1 * C + (double)N * K + (double)M * N));
  constexpr uint32_t EF = EF_BIAS | EF_RELU;
  return launch<256, num_stages<256, 2>(EF), false, false, EF, 2, 1>(ma, mb, mc, mc, M, N, K, 1, 1, 1, K / BK, ep, st);
}

// dw[co, (kt*3+kf)*C + ci] += sum_{b,t,f} dh2[(b,t,f), co] * h1[b, 2t+kt, 2f+kf, ci]      (fp32, accumulated)
extern "C" int b200slu_conv2_wgrad(const void* dh2, const void* h1, int B, int T1, int F1, int C, float* dw_accum, void* stream) {
  Conv2Geom g;
  B2_TRY(conv2_geom(B, T1, F1, C, &g));
  cudaStream_t st = (cudaStream_t)stream;
  B2_CHECK_ARG(((uintptr_t)h1 % 16) == 0 && ((uintptr_t)dh2 % 16) == 0 && ((uintptr_t)dw_accum % 16) == 0,
               "conv2_wgrad: 16-byte aligned operands");
  const int M = C, N = 9 * C, K = (int)g.M;
  CUtensorMap ma, mb, mc;
  B2_TRY(make_map(&ma, dh2, K, M, C, 1, 0, 1, 0, 64, BK));
  B2_TRY(make_im2col_map(&mb, h1, B, T1, F1, C, 0, 0, -2, -2, 2, BK));
  B2_TRY(make_map(&mc, dw_accum, M, N, N, 1, 0, 1, 0, 32, 32, 4));
  EpiArgs ep{};
  ep.out = dw_accum; ep.ldc = N; ep.alpha = 1.f; ep.out_f32 = 1; ep.accumulate = 1;
  ep.flags = EF_F32 | EF_ACC;
  ep.debug = gemm_debug_env() & 16;
  ep.cv.pw = g.F2; ep.cv.ph = g.T2; ep.cv.mult = 2; ep.cv.base_w = 0; ep.cv.base_h = 0; ep.cv.cb = C / 64; ep.cv.C = C;
  for (int t = 0; t < 9; ++t) {
    ep.cv.off_h[t] = (unsigned char)(t / 3);
    ep.cv.off_w[t] = (unsigned char)(t % 3);
    ep.cv.wtap[t] = (unsigned char)t;
  }
  // split K (= pixels) so that the (C/256) x (9C/256) pair tiles fill the machine in one wave; the units run split-major,
  // so the tiles in flight read the same pixels through different taps.  (Short k-chunks, tried for L2 locality, only add
  // reduce-add epilogues: 1551 us with 48-block chunks against 1451 us.)
  const int num_kb = (K + BK - 1) / BK;
  const long long tiles = (long long)((M + 255) / 256) * (N / 256), slots = b2_num_sms() / 2;
  int splits = 1;
  double best = 1e300;
  for (int sp = 1; sp <= 64 && sp <= max(1, num_kb / 4); ++sp) {
    const int kbs = (num_kb + sp - 1) / sp;
    const double cost = (double)((tiles * sp + slots - 1) / slots) * (kbs * 1.05 + 4.0);
    if (cost < best * 0.999) {
      best = cost;
      splits = sp;
    }
  }
  const int kb_per = (num_kb + splits - 1) / splits;
  splits = (num_kb + kb_per - 1) / kb_per;
  B2_SCOPE(FAM_GEMM, st, 2.0 * M * N * (double)K, 2.0 * ((double)B * T1 * F1 * C + (double)M * K) + 4.0 * M * N);
  constexpr uint32_t EF = EF_F32 | EF_ACC;
  return launch<256, num_stages<256, 2>(EF), true, true, EF, 2, 2>(ma, mb, mc, mc, M, N, K, 1, 1, splits, kb_per, ep, st);
}

// dh1[b, r, g, ci] = (h1[b, r, g, ci] > 0) * sum over taps with 2t+kt = r, 2f+kf = g of dh2[(b,t,f), co] * w[co, (kt*3+kf)*C + ci]
// Gather form, one GEMM per parity class (r & 1, g & 1): 4 / 2 / 2 / 1 taps, no scatter-add and no [pixels, 9 C] matrix.
extern "C" int b200slu_conv2_dgrad_relu(const void* dh2, const void* w, const unsigned int* relu_bits, int B, int T1, int F1, int C,
                                        void* dh1, void* stream) {
  Conv2Geom g;
  B2_TRY(conv2_geom(B, T1, F1, C, &g));
  cudaStream_t st = (cudaStream_t)stream;
  B2_CHECK_ARG(((uintptr_t)relu_bits % 16) == 0 && ((uintptr_t)dh2 % 16) == 0 && ((uintptr_t)w % 16) == 0 && ((uintptr_t)dh1 % 16) == 0,
               "conv2_dgrad: 16-byte aligned operands");
  CUtensorMap mb, mc;
  B2_TRY(make_map(&mb, w, C, 9 * C, 9 * C, 1, 0, 1, 0, 64, BK));  // [K = co, N = (tap, ci)], N contiguous
  B2_TRY(make_map(&mc, dh1, (long long)B * T1 * F1, C, C, 1, 0, 1, 0, 64, 32, 2));  // not stored through (scatter epilogue)
  constexpr uint32_t EF = EF_SCATTER;
  for (int cls = 0; cls < 4; ++cls) {
    const int pt = cls >> 1, pf = cls & 1;
    const int Hc = (T1 - pt + 1) / 2, Wc = (F1 - pf + 1) / 2;  // pixels r = 2a + pt < T1, g = 2b + pf < F1
    if (Hc <= 0 || Wc <= 0) continue;
    const int lh = pt == 0 ? -1 : 0, lw = pf == 0 ? -1 : 0;
    CUtensorMap ma;
    B2_TRY(make_im2col_map(&ma, dh2, B, g.T2, g.F2, C, lw, lh, Wc - g.F2 + lw, Hc - g.T2 + lh, 1, BM));
    EpiArgs ep{};
    ep.out = dh1; ep.ldc = C; ep.alpha = 1.f; ep.aux = (const bf16*)relu_bits; ep.ldaux = C;  // bit rows, C bits each
    ep.flags = EF;
    ep.debug = gemm_debug_env() & 16;
    ep.cv.pw = Wc; ep.cv.ph = Hc; ep.cv.mult = 1; ep.cv.base_w = lw; ep.cv.base_h = lh; ep.cv.cb = C / 64; ep.cv.C = C;
    ep.cv.o_mult = 2; ep.cv.o_w0 = pf; ep.cv.o_h0 = pt; ep.cv.OW = F1; ep.cv.OH = T1;
    // taps of the class: kt in {0, 2} (pt = 0: source row a - kt/2 = base row + (1 - kt/2)) or {1} (source row a)
    int nt = 0;
    for (int it = 0; it < (pt == 0 ? 2 : 1); ++it)
      for (int jf = 0; jf < (pf == 0 ? 2 : 1); ++jf) {
        const int kt = pt == 0 ? 2 * it : 1, kf = pf == 0 ? 2 * jf : 1;
        ep.cv.off_h[nt] = (unsigned char)(pt == 0 ? 1 - it : 0);
        ep.cv.off_w[nt] = (unsigned char)(pf == 0 ? 1 - jf : 0);
        ep.cv.wtap[nt] = (unsigned char)(kt * 3 + kf);
        ++nt;
      }
    const long long Mc = (long long)B * Hc * Wc;
    const int K = nt * C;
    B2_SCOPE(FAM_GEMM, st, 2.0 * Mc * C * (double)K, 2.0 * ((double)B * g.T2 * g.F2 * C + (double)C * K + 2.0 * Mc * C));
    B2_TRY((launch<256, num_stages<256, 2>(EF), false, true, EF, 2, 1>(ma, mb, mc, mc, (int)Mc, C, K, 1, 1, 1, K / BK, ep, st)));
  }
  return 0;
}

extern "C" int b200slu_gemm_bf16_ex(const b200slu_gemm_t* g, const b200slu_epilogue_t* epilogue, void* stream) {
  return b2_gemm_ex(g, epilogue, (cudaStream_t)stream);
}

extern "C" int b200slu_gemm_bf16(int trans, int M, int N, int K, const void* A, long long lda, const void* B, long long ldb,
                                 void* C, long long ldc, const b200slu_epilogue_t* epilogue, void* stream) {
  b200slu_gemm_t g{};
  g.trans_a = g.trans_b = trans;
  g.M = M; g.N = N; g.K = K;
  g.A = A; g.lda = lda;
  g.B = B; g.ldb = ldb;
  g.C = C; g.ldc = ldc;
  g.batch_inner = g.batch_outer = 1;
  return b2_gemm_ex(&g, epilogue, (cudaStream_t)stream);
}
