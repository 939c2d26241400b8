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

namespace {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 B = one swizzle span
constexpr int NUM_EPI_WARPS = 8;
constexpr int NUM_THREADS = 64 + 32 * NUM_EPI_WARPS;

// ----------------------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug traps (kernel error) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) __trap();
  }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accum)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout), SWIZZLE_128B, version 1.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);            // [0,14)  start address
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;   // [16,30) leading byte offset
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;   // [32,46) stride byte offset
  d |= (uint64_t)1 << 46;                              // [46,48) descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                              // [61,64) SWIZZLE_128B
  return d;
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): bf16 x bf16 -> f32, M=128, N=n.
__host__ __device__ constexpr uint32_t make_idesc(int n, int a_mn_major, int b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(n >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
}

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
  B2Dropout drop;
};

template <int BN, int STAGES>
struct SmemLayout {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_OFF = STAGES * STAGE_BYTES;
  static constexpr int NUM_BARS = 2 * STAGES + 4;
  static constexpr int TOTAL = BAR_OFF + NUM_BARS * 8 + 16 + 1024;  // + alignment slack
};

// One 32-column chunk of one accumulator row: fused epilogue + store.
__device__ __forceinline__ void epilogue_chunk(const EpiArgs& ep, float (&v)[32], int row, int col0, int N, bool first_split,
                                               long long out_off) {
  const bool full = (col0 + 32 <= N);
  if (ep.bias != nullptr && first_split) {
    if (full) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        float4 b = *reinterpret_cast<const float4*>(ep.bias + col0 + j);
        v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < N) v[j] += ep.bias[col0 + j];
    }
  }
  // activation (+ its derivative when the backward wants it saved) and dropout share one pass; the saved derivative
  // already carries the dropout multiplier, so the dgrad epilogue is a single multiply (no MUFU, no hash)
  float dv[32];
  const bool want_d = ep.pre_out != nullptr;
  if (ep.act == B200SLU_ACT_SWISH) {
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const float sg = sigmoidf_(v[j]);
      if (want_d) dv[j] = sg * (1.f + v[j] * (1.f - sg));
      v[j] *= sg;
    }
  } else if (ep.act == B200SLU_ACT_RELU) {
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      if (want_d) dv[j] = v[j] > 0.f ? 1.f : 0.f;
      v[j] = fmaxf(v[j], 0.f);
    }
  } else if (want_d) {
#pragma unroll
    for (int j = 0; j < 32; ++j) dv[j] = 1.f;
  }
  if (ep.drop.p > 0.f) {
    const float inv = 1.f / (1.f - ep.drop.p);
    const unsigned long long e0 = (unsigned long long)row * (unsigned long long)N + col0;  // N % 8 == 0 (host check)
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      uint32_t keep = dropout_keep8(ep.drop, (e0 >> 3) + g);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float m = ((keep >> j) & 1u) ? inv : 0.f;
        v[g * 8 + j] *= m;
        if (want_d) dv[g * 8 + j] *= m;
      }
    }
  }
  if (want_d) {
    bf16* p = ep.pre_out + (long long)row * ep.ldc + col0;
    if (full) {
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        uint4 u = make_uint4(pack_bf16x2(dv[j], dv[j + 1]), pack_bf16x2(dv[j + 2], dv[j + 3]),
                             pack_bf16x2(dv[j + 4], dv[j + 5]), pack_bf16x2(dv[j + 6], dv[j + 7]));
        *reinterpret_cast<uint4*>(p + j) = u;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < N) p[j] = __float2bfloat16(dv[j]);
    }
  }
  if (ep.alpha != 1.f) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] *= ep.alpha;
  }
  if (ep.aux != nullptr) {
    const bf16* a = ep.aux + (long long)row * ep.ldaux + col0;
    if (full) {
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        uint4 u = *reinterpret_cast<const uint4*>(a + j);
        uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float2 x = unpack_bf16x2(w[q]);
          if (ep.aux_mode == B200SLU_ACT_MUL) {
            v[j + 2 * q] *= x.x;
            v[j + 2 * q + 1] *= x.y;
          } else if (ep.aux_mode == B200SLU_ACT_SWISH) {
            v[j + 2 * q] *= swish_gradf_(x.x);
            v[j + 2 * q + 1] *= swish_gradf_(x.y);
          } else {
            v[j + 2 * q] = x.x > 0.f ? v[j + 2 * q] : 0.f;
            v[j + 2 * q + 1] = x.y > 0.f ? v[j + 2 * q + 1] : 0.f;
          }
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        if (col0 + j < N) {
          float x = __bfloat162float(a[j]);
          v[j] *= (ep.aux_mode == B200SLU_ACT_MUL) ? x : (ep.aux_mode == B200SLU_ACT_SWISH) ? swish_gradf_(x) : (x > 0.f ? 1.f : 0.f);
        }
      }
    }
  }
  if (ep.res != nullptr) {
    const float* rp = ep.res + (long long)row * ep.ldres + col0;
    if (full) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        float4 b = *reinterpret_cast<const float4*>(rp + j);
        v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < N) v[j] += rp[j];
    }
  }
  if (ep.out_f32) {
    float* o = reinterpret_cast<float*>(ep.out) + out_off + (long long)row * ep.ldc + col0;
    if (ep.accumulate) {
      if (full) {  // split-K partial sums: 16-byte vector reductions (4x fewer L2 atomic requests than scalar RED)
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o + j), "f"(v[j]), "f"(v[j + 1]), "f"(v[j + 2]),
                       "f"(v[j + 3])
                       : "memory");
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (col0 + j < N) atomicAdd(o + j, v[j]);
      }
    } else if (full) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < N) o[j] = v[j];
    }
  } else {
    bf16* o = reinterpret_cast<bf16*>(ep.out) + out_off + (long long)row * ep.ldc + col0;
    if (full) {
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        uint4 u = make_uint4(pack_bf16x2(v[j], v[j + 1]), pack_bf16x2(v[j + 2], v[j + 3]),
                             pack_bf16x2(v[j + 4], v[j + 5]), pack_bf16x2(v[j + 6], v[j + 7]));
        *reinterpret_cast<uint4*>(o + j) = u;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < N) o[j] = __float2bfloat16(v[j]);
    }
  }
}

// Persistent: grid = min(#work units, #SMs); unit u = (batch, tile, k-split), tiles n-fastest.
template <int BN, int STAGES, bool TA, bool TB>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_bf16_tcgen05(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                  int M, int N, int K, int tiles_n, int tiles_per_batch, int batch_inner, int splits, int kb_per_split,
                  int num_units, EpiArgs ep) {
  using L = SmemLayout<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sbase = smem_u32(smem);
  const uint32_t bar_full = sbase + L::BAR_OFF;
  const uint32_t bar_empty = bar_full + STAGES * 8;
  const uint32_t bar_acc_full = bar_empty + STAGES * 8;   // [2]
  const uint32_t bar_acc_empty = bar_acc_full + 16;       // [2]
  uint32_t* tmem_slot = (uint32_t*)(smem + L::BAR_OFF + L::NUM_BARS * 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_kb_total = (K + BK - 1) / BK;
  const int units_per_batch = tiles_per_batch * splits;

  if (threadIdx.x == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(bar_full + s * 8, 1);
      mbar_init(bar_empty + s * 8, 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(bar_acc_full + b * 8, 1);
      mbar_init(bar_acc_empty + b * 8, NUM_EPI_WARPS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // TMEM: two accumulator buffers of BN fp32 columns x 128 lanes
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(2 * BN)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int i = 0;  // running k-block counter (ring position)
      for (int u = blockIdx.x; u < num_units; u += gridDim.x) {
        const int batch = u / units_per_batch, r = u - batch * units_per_batch;
        const int bo = batch / batch_inner, bi = batch - bo * batch_inner;
        const int tile = r / splits, sp = r - tile * splits;
        const int tile_m = tile / tiles_n, tile_n = tile - tile_m * tiles_n;
        const int kb0 = sp * kb_per_split;
        const int nkb = min(kb_per_split, num_kb_total - kb0);
        const int abi = bi * ep.a_bc_inner, abo = bo * ep.a_bc_outer, bbi = bi * ep.b_bc_inner, bbo = bo * ep.b_bc_outer;
        for (int kb = 0; kb < nkb; ++kb, ++i) {
          const int s = i % STAGES, it = i / STAGES;
          mbar_wait(bar_empty + s * 8, (it & 1) ^ 1);
          const uint32_t sa = sbase + s * L::STAGE_BYTES, sb = sa + L::A_BYTES;
          const uint32_t fb = bar_full + s * 8;
          mbar_expect_tx(fb, L::STAGE_BYTES);
          const int k = (kb0 + kb) * BK;
          if (!TA) {
            tma_load_4d(sa, &map_a, fb, k, tile_m * BM, abi, abo);   // box {64 k, 128 rows}
          } else {
            // box {64 mn, 64 k-rows}; one box per 64-wide MN atom
#pragma unroll
            for (int a = 0; a < BM / 64; ++a) tma_load_4d(sa + a * (64 * BK * 2), &map_a, fb, tile_m * BM + a * 64, k, abi, abo);
          }
          if (!TB) {
            tma_load_4d(sb, &map_b, fb, k, tile_n * BN, bbi, bbo);   // box {64 k, BN rows}
          } else {
#pragma unroll
            for (int b = 0; b < BN / 64; ++b) tma_load_4d(sb + b * (64 * BK * 2), &map_b, fb, tile_n * BN + b * 64, k, bbi, bbo);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = make_idesc(BN, TA, TB);
    int i = 0, t = 0;
    for (int u = blockIdx.x; u < num_units; u += gridDim.x, ++t) {
      const int r = u % units_per_batch;
      const int sp = r % splits;
      const int kb0 = sp * kb_per_split;
      const int nkb = min(kb_per_split, num_kb_total - kb0);
      const int buf = t & 1;
      mbar_wait(bar_acc_empty + buf * 8, ((t >> 1) & 1) ^ 1);  // epilogue drained this buffer
      tcgen05_fence_after();
      const uint32_t d_tmem = tmem_base + buf * BN;
      for (int kb = 0; kb < nkb; ++kb, ++i) {
        const int s = i % STAGES, it = i / STAGES;
        mbar_wait(bar_full + s * 8, it & 1);
        tcgen05_fence_after();
        if (lane == 0) {
          const uint32_t sa = sbase + s * L::STAGE_BYTES, sb = sa + L::A_BYTES;
#pragma unroll
          for (int kk = 0; kk < BK / 16; ++kk) {
            // K-major: 8-row groups 1024 B apart; advance 32 B per K=16 inside the swizzle span.
            // MN-major: 64-element MN atoms (64*BK*2 B apart = LBO), 8-row K groups 1024 B apart (SBO).
            const uint64_t da = TA ? make_smem_desc(sa + kk * 2048, 64 * BK * 2, 1024) : make_smem_desc(sa + kk * 32, 16, 1024);
            const uint64_t db = TB ? make_smem_desc(sb + kk * 2048, 64 * BK * 2, 1024) : make_smem_desc(sb + kk * 32, 16, 1024);
            umma_bf16(d_tmem, da, db, idesc, (kb | kk) != 0);
          }
          umma_commit(bar_empty + s * 8);                       // frees the smem stage once these MMAs retire
          if (kb == nkb - 1) umma_commit(bar_acc_full + buf * 8);  // accumulator complete
        }
        __syncwarp();
      }
    }
  } else {
    // ===================== epilogue (8 warps: 4 lane groups x 2 column halves) =====================
    const int ew = warp - 2;
    const int lane_grp = warp & 3;      // TMEM lanes [32*lane_grp, +32) -- hardware rule: warp_id % 4
    const int half = ew >> 2;           // columns [half*BN/2, +BN/2)
    int t = 0;
    for (int u = blockIdx.x; u < num_units; u += gridDim.x, ++t) {
      const int batch = u / units_per_batch, r = u - batch * units_per_batch;
      const int bo = batch / batch_inner, bi = batch - bo * batch_inner;
      const int tile = r / splits, sp = r - tile * splits;
      const int tile_m = tile / tiles_n, tile_n = tile - tile_m * tiles_n;
      const long long out_off = (long long)bo * ep.c_bs_outer + (long long)bi * ep.c_bs_inner;
      const int buf = t & 1;
      mbar_wait(bar_acc_full + buf * 8, (t >> 1) & 1);
      tcgen05_fence_after();
      const int row = tile_m * BM + lane_grp * 32 + lane;
      const bool row_ok = row < M;
      // TMEM -> registers is double-buffered: the load of chunk i+1 is in flight while chunk i runs the epilogue math
      constexpr int NCHUNK = BN / 64;  // 32-column chunks per column half
      const uint32_t tbase = tmem_base + ((uint32_t)(lane_grp * 32) << 16) + buf * BN + half * (BN / 2);
      uint32_t r32[2][32];
      tmem_ld32(tbase, r32[0]);
#pragma unroll
      for (int ci = 0; ci < NCHUNK; ++ci) {
        tmem_ld_wait();
        if (ci + 1 < NCHUNK) tmem_ld32(tbase + (ci + 1) * 32, r32[(ci + 1) & 1]);
        const int col0 = tile_n * BN + half * (BN / 2) + ci * 32;
        if (row_ok && col0 < N) {
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r32[ci & 1][j]);
          epilogue_chunk(ep, v, row, col0, N, sp == 0, out_off);
        }
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_acc_empty + buf * 8);
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(2 * BN) : "memory");
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

// 4-D bf16 tensor [n_outer, n_inner, rows, cols]: row stride ld, batch strides s_inner / s_outer (elements;
// 0 = broadcast -> that dimension gets extent 1).  box = {box_cols, box_rows, 1, 1}.
int make_map(CUtensorMap* m, const void* base, long long rows, long long cols, long long ld, long long n_inner,
             long long s_inner, long long n_outer, long long s_outer, int box_cols, int box_rows) {
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
  cuuint64_t strides[3] = {(cuuint64_t)ld * 2, (cuuint64_t)s_inner * 2, (cuuint64_t)s_outer * 2};
  cuuint32_t box[4] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    b200slu_set_error("cuTensorMapEncodeTiled failed (%d): rows=%lld cols=%lld ld=%lld batch=(%lld x %lld, strides %lld %lld) base=%p",
                      (int)r, rows, cols, ld, n_inner, n_outer, s_inner, s_outer, base);
    return B200SLU_ECUDA;
  }
  return 0;
}

template <int BN, int STAGES, bool TA, bool TB>
int launch(const CUtensorMap& ma, const CUtensorMap& mb, int M, int N, int K, int nbatch, int batch_inner, int splits,
           int kb_per, const EpiArgs& ep, cudaStream_t st) {
  using L = SmemLayout<BN, STAGES>;
  auto kern = gemm_bf16_tcgen05<BN, STAGES, TA, TB>;
  static bool configured = false;
  if (!configured) {
    B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
    configured = true;
  }
  const int tiles_n = (N + BN - 1) / BN, tiles_m = (M + BM - 1) / BM;
  const long long units = (long long)tiles_n * tiles_m * splits * nbatch;
  B2_CHECK_ARG(units < (1LL << 30), "gemm: too many work units (%lld)", units);
  const int grid = units < b2_num_sms() ? (int)units : b2_num_sms();
  kern<<<grid, NUM_THREADS, L::TOTAL, st>>>(ma, mb, M, N, K, tiles_n, tiles_n * tiles_m, batch_inner, splits, kb_per,
                                            (int)units, ep);
  B2_LAUNCH_CHECK();
  return 0;
}

}  // namespace

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
  }
  B2_CHECK_ARG(nbatch == 1 || (ep.res == nullptr && ep.aux == nullptr && ep.pre_out == nullptr && !(ep.drop.p > 0.f)),
               "gemm: batched mode supports only bias / activation / alpha epilogues");
  B2_CHECK_ARG(!(ep.drop.p > 0.f) || N % 8 == 0, "gemm: dropout epilogue needs N %% 8 == 0");
  B2_CHECK_ARG(!ep.accumulate || ep.out_f32, "gemm: accumulate needs fp32 output");
  const int vec = ep.out_f32 ? 4 : 8;
  B2_CHECK_ARG(g->ldc % vec == 0 && ((uintptr_t)g->C % 16) == 0 && g->c_stride_inner % vec == 0 && g->c_stride_outer % vec == 0,
               "gemm: output must be 16-byte aligned, ldc / batch strides %% %d == 0", vec);
  B2_CHECK_ARG(ep.res == nullptr || (ep.ldres % 4 == 0 && ((uintptr_t)ep.res % 16) == 0), "gemm: residual alignment");
  B2_CHECK_ARG(ep.aux == nullptr || (ep.ldaux % 8 == 0 && ((uintptr_t)ep.aux % 16) == 0), "gemm: act_grad_input alignment");
  B2_CHECK_ARG(ep.bias == nullptr || ((uintptr_t)ep.bias % 16) == 0, "gemm: bias alignment");

  const int num_kb = (K + BK - 1) / BK;
  int splits = 1;
  // Tile width: a 128x128 single-CTA tile reads 32 KB of shared memory per 256 MMA cycles (= the 128 B/clk port limit);
  // 128x256 reads 48 KB per 512 cycles and lets the tensor pipe run ahead.  Take the wide tile when it does not cost
  // more in wave quantisation than it gains (relative per-tile costs measured: profiles/r1_gemm_shapes_*.txt).
  int bn = 128;
  {
    const long long sms = b2_num_sms();
    const long long tm = (M + BM - 1) / BM;
    const long long t128 = tm * ((N + 127) / 128) * nbatch, t256 = tm * ((N + 255) / 256) * nbatch;
    const double c128 = (double)((t128 + sms - 1) / sms) * 1.0, c256 = (double)((t256 + sms - 1) / sms) * 1.45;
    if (N >= 256 && !ep.accumulate && c256 < c128) bn = 256;
    static const int force_bn = getenv("B200SLU_GEMM_BN") ? atoi(getenv("B200SLU_GEMM_BN")) : 0;  // tuning experiments only
    if (force_bn == 128 || (force_bn == 256 && N >= 256 && !ep.accumulate)) bn = force_bn;
  }
  const long long tiles128 = (long long)((M + BM - 1) / BM) * ((N + 127) / 128) * nbatch;
  if (ep.accumulate) {
    // split-K (wgrad: K = B*T is long, the output is small): target >= 2 CTAs per SM
    const int want = (int)((2LL * b2_num_sms() + tiles128 - 1) / tiles128);
    splits = max(1, min(want, num_kb / 4 > 0 ? num_kb / 4 : 1));
  }
  const int kb_per = (num_kb + splits - 1) / splits;
  splits = (num_kb + kb_per - 1) / kb_per;

  CUtensorMap ma, mb;
  const double flops = 2.0 * M * N * (double)K * nbatch;
  B2_SCOPE(FAM_GEMM, st, flops, nbatch * (2.0 * ((double)M * K + (double)N * K) + (ep.out_f32 ? 4.0 : 2.0) * M * N));
  if (!g->trans_a)
    B2_TRY(make_map(&ma, g->A, M, K, g->lda, nbi, g->a_stride_inner, nbo, g->a_stride_outer, BK, BM));
  else
    B2_TRY(make_map(&ma, g->A, K, M, g->lda, nbi, g->a_stride_inner, nbo, g->a_stride_outer, 64, BK));
  if (!g->trans_b)
    B2_TRY(make_map(&mb, g->B, N, K, g->ldb, nbi, g->b_stride_inner, nbo, g->b_stride_outer, BK, bn));
  else
    B2_TRY(make_map(&mb, g->B, K, N, g->ldb, nbi, g->b_stride_inner, nbo, g->b_stride_outer, 64, BK));
#define B2_GEMM_DISPATCH(BN_, ST_)                                                                                    \
  do {                                                                                                                \
    if (!g->trans_a && !g->trans_b) return launch<BN_, ST_, false, false>(ma, mb, M, N, K, nbatch, nbi, splits, kb_per, ep, st); \
    if (!g->trans_a && g->trans_b) return launch<BN_, ST_, false, true>(ma, mb, M, N, K, nbatch, nbi, splits, kb_per, ep, st);   \
    if (g->trans_a && !g->trans_b) return launch<BN_, ST_, true, false>(ma, mb, M, N, K, nbatch, nbi, splits, kb_per, ep, st);   \
    return launch<BN_, ST_, true, true>(ma, mb, M, N, K, nbatch, nbi, splits, kb_per, ep, st);                       \
  } while (0)
  if (bn == 256) B2_GEMM_DISPATCH(256, 4);
  B2_GEMM_DISPATCH(128, 6);
#undef B2_GEMM_DISPATCH
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
