// Fused relative-position multi-head attention for short sequences (keys <= 192, d_k = 64), forward.
//
// Replaces, in one kernel per layer, the chain the reference runs as separate torch ops
// (espnet/nets/pytorch_backend/transformer/attention.py): matrix_ac = (q+u) k^T and matrix_bd = (q+v) p^T (:297-302),
// rel_shift (:240-263), (ac + bd) / sqrt(d_k) (:304-306), masked softmax + dropout (forward_attention :78-91) and
// attn . v (:93-95).  The same kernel with relpos_mode = NONE is MultiHeadedAttention.forward (:98-114) of the decoder
// (causal self-attention, cross-attention over the encoder memory).
//
// One CTA = one (utterance, head, 128-query tile).  All operands of the tile arrive by TMA (SWIZZLE_128B); scores live
// in tensor memory only:
//   TMEM columns [0, TKP)          ac  = QU[128 x 64] . K[TKP x 64]^T                       (tcgen05.mma, M = 128)
//   TMEM columns [TKP, TKP + WIN)  bd  = QV[128 x 64] . Pwin[WIN x 64]^T, WIN = TKP + 128: the window of relative
//                                  positions rows i0 .. i0+127 can reach; P row c0 + n with c0 = T-1-i0-127
// rel_shift is index arithmetic on that window: out[i0+r, j] = bd[r, 127 - r + j].  A TMEM lane (= query row) is read
// by the thread of the same number, tcgen05.ld gives every lane of a warp the same columns, so the per-row offset is
// applied by a round trip through a private shared-memory row (16-byte stores, scalar loads at the shifted offset:
// bank-conflict-free with a 52-float row stride).  8 warps: lane quadrant = warp % 4, column half = warp / 4; three
// passes over the warp's columns (max, exp + sum, normalise) keep the scores in TMEM between passes (tcgen05.st), so a
// thread never holds more than one 16-column chunk.  The normalised, dropped-out probabilities are written (a) to HBM
// (bf16, head-major [H,B,Tq,ld_p], saved for the backward) and (b) into shared memory as the K-major A operand of the
// second contraction O = P . V (V is the MN-major B operand, exactly as TMA delivered it).
// Shared memory is recycled between phases (Q/K/P-window are dead once the score MMAs retire), which keeps a T = 155
// tile at 110 KB: two CTAs per SM, so one CTA's TMA loads and TMEM allocation wait overlap the other's compute.
#include <stdlib.h>

#include <cuda_fp16.h>

#include "common.cuh"
#include "sm100_ptx.cuh"

namespace {

constexpr int DK = 64;        // head dimension (one 128-byte swizzle span of bf16)
constexpr int QT = 128;       // query rows per CTA = TMEM lanes
constexpr int CH = 16;        // score columns per softmax chunk
constexpr int RING_LD = 84;   // floats per private shift-ring row: 64 window columns, 16 mirrored ones (reads never wrap), 4 pad
constexpr int NCH = 6;        // softmax chunks per warp (half a row of at most 192 keys)
constexpr int MAX_TK = 192;   // ac (TKP) + bd (TKP + 128) columns must fit the 512 TMEM columns
constexpr int NTHREADS = 256;

// one full 32-byte sector per thread (sm_100 256-bit store)
__device__ __forceinline__ void stg256(void* p, const uint32_t (&v)[8]) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]),
               "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
               : "memory");
}

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct AttnArgs {
  int B, H, Tq, Tk, TKP, WIN, rel, causal;
  float scale_log2e;
  const int* lens;
  bf16 *p, *pd, *ctx;
  long long ld_p, ldctx;
  // key chunk of a longer key axis (b200slu_relattn_fwd_chunk): this launch sees keys [key_off, key_off + Tk) of every
  // utterance.  key_off shifts the rel-pos window, the valid length and the dropout counter; p already points at column
  // key_off (p_cols = columns left in the row); lse2 receives log2 of each row's partition sum over the chunk.
  int key_off, p_cols;
  float* lse2;
  uint32_t off_k, off_p, off_v, off_bar, tmem_cols;
  B2Dropout drop;
  long long* dbg;  // tuning only (B200SLU_ATTN_DEBUG=1): per-CTA phase time stamps
};

// group barrier: the 256 threads of one tile (barrier 0 = the whole CTA in the one-tile kernel)
__device__ __forceinline__ void group_sync(int bar_id) { asm volatile("bar.sync %0, 256;" ::"r"(bar_id) : "memory"); }

// Softmax of one 128-query tile whose scores sit in tensor memory at t_s (ac at column 0, the bd window at column TKP):
// leaves P in HBM and the (dropped-out) K-major A operand of P.V at the start of the tile's shared-memory region.
__device__ __forceinline__ void softmax_tile(const AttnArgs& a, uint8_t* smem, uint32_t sbase, int tid, int bar_id, uint32_t t_s,
                                             float* rowstat, int i0, int h, int b, int len_b, const B2DropKey& dk,
                                             uint32_t s_release) {
  // ---------------------------------------------------------------- softmax: ONE pass over tensor memory
  // TMEM -> register reads run at 64 B/clk/SM (16 B/clk per scheduler), so every score column is read exactly once
  // (+ 31 window columns per warp for the shift).  A thread keeps its half row as e = exp2(s - m_c) per 16-column
  // chunk (m_c = the chunk's own maximum), packed fp16x2 (e <= 1; the final bf16 rounding of p is 4x coarser); the
  // chunk references are reconciled when the row maximum and sum are known.
  const int warp = tid >> 5, lane = tid & 31;  // warp within the group of 8
  const int TKP = a.TKP;
  const bool dropping = a.drop.p > 0.f;
  const int q = warp & 3, hs = warp >> 2;
  const int r = q * 32 + lane, i = i0 + r;
  const bool row_ok = i < a.Tq;
  const uint32_t tlane = t_s + ((uint32_t)(q * 32) << 16);
  const int nch = TKP / CH, n0 = (nch + 1) / 2;
  const int cbeg = hs ? n0 : 0, nmine = hs ? nch - n0 : n0;  // this warp's chunks: [cbeg, cbeg + nmine), nmine <= NCH
  const int jbeg = cbeg * CH;
  const int nvalid = a.causal ? min(len_b, i + 1 - a.key_off) : len_b;
  // shift ring: 64 window columns per thread; positions 64..79 mirror 0..15 so that the 16 lane-shifted scalar loads of a
  // chunk are base + immediate (no wrap); the 84-float row pitch keeps 16-byte stores and shifted loads conflict-free
  const uint32_t ring = sbase + (uint32_t)(warp * 32 + lane) * (RING_LD * 4);
  const float* ring_g = reinterpret_cast<const float*>(smem) + (warp * 32 + lane) * RING_LD;
  const int nb = 96 - 32 * q + jbeg;  // window column of ring position 0: row r, key jbeg + x reads position 31 - lane + x
  const float kap = a.scale_log2e;

  uint32_t E[NCH][CH / 2];
  float mc[NCH], sc[NCH];
  uint32_t ta[16], tb[16];
  if (nmine > 0) {
    if (a.rel) {
      tmem_ld16(tlane + TKP + nb, ta);
      tmem_ld16(tlane + TKP + nb + 16, tb);
      tmem_ld_wait();
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        sts128(ring + t * 16, ta[4 * t], ta[4 * t + 1], ta[4 * t + 2], ta[4 * t + 3]);
        sts128(ring + 64 + t * 16, tb[4 * t], tb[4 * t + 1], tb[4 * t + 2], tb[4 * t + 3]);
      }
      tmem_ld16(tlane + TKP + nb + 32, tb);
    }
    tmem_ld16(tlane + jbeg, ta);
  }
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    mc[c] = -INFINITY;
    sc[c] = 0.f;
#pragma unroll
    for (int e = 0; e < CH / 2; ++e) E[c][e] = 0u;
    if (c < nmine) {
      const int j0 = jbeg + c * CH;
      float sv[CH];
      tmem_ld_wait();
#pragma unroll
      for (int jj = 0; jj < CH; ++jj) sv[jj] = __uint_as_float(ta[jj]);
      if (a.rel) {
        const uint32_t wr = ring + (uint32_t)(((32 + 16 * c) & 63) * 4);
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          sts128(wr + t * 16, tb[4 * t], tb[4 * t + 1], tb[4 * t + 2], tb[4 * t + 3]);
          if ((c & 3) == 2) sts128(wr + 256 + t * 16, tb[4 * t], tb[4 * t + 1], tb[4 * t + 2], tb[4 * t + 3]);
        }
      }
      if (c + 1 < nmine) {  // next chunk's TMEM reads fly while this one is processed
        if (a.rel) tmem_ld16(tlane + TKP + nb + 48 + 16 * c, tb);
        tmem_ld16(tlane + j0 + CH, ta);
      }
      if (a.rel) {
        const float* rd = ring_g + ((31 - lane + 16 * c) & 63);
#pragma unroll
        for (int jj = 0; jj < CH; ++jj) sv[jj] += rd[jj];
      }
      // scores stay raw (unscaled) until the exponent: e = 2^(kap * s - kap * m) is one FFMA + one MUFU per element
      if (j0 + CH > nvalid) {  // only chunks that straddle the valid length pay for masking
#pragma unroll
        for (int jj = 0; jj < CH; ++jj) sv[jj] = (j0 + jj < nvalid) ? sv[jj] : -INFINITY;
      }
      float m = -INFINITY;
#pragma unroll
      for (int jj = 0; jj < CH; ++jj) m = fmaxf(m, sv[jj]);
      m *= kap;  // kap > 0
      mc[c] = m;
      const float mneg = (m == -INFINITY) ? 0.f : -m;
      float sum = 0.f;
#pragma unroll
      for (int jj = 0; jj < CH; jj += 2) {
        const float e0 = ex2_approx(fmaf(sv[jj], kap, mneg)), e1 = ex2_approx(fmaf(sv[jj + 1], kap, mneg));
        sum += e0 + e1;
        const __half2 hv = __floats2half2_rn(e0, e1);
        E[c][jj >> 1] = *reinterpret_cast<const uint32_t*>(&hv);
      }
      sc[c] = sum;
    }
  }
  float mw = -INFINITY;
#pragma unroll
  for (int c = 0; c < NCH; ++c) mw = fmaxf(mw, mc[c]);
  float sw = 0.f;
#pragma unroll
  for (int c = 0; c < NCH; ++c) sw += (mc[c] == -INFINITY) ? 0.f : sc[c] * exp2f(mc[c] - mw);
  rowstat[hs * QT + r] = mw;
  rowstat[2 * QT + hs * QT + r] = sw;
  tcgen05_fence_before();
  group_sync(bar_id);  // also: every warp is done with its shift ring, which the P operand below overwrites
  if (s_release != 0u && tid == 0) mbar_arrive(s_release);  // the score columns may be overwritten (two-group kernel)
  const float mo = rowstat[(hs ^ 1) * QT + r], so = rowstat[2 * QT + (hs ^ 1) * QT + r];
  const float mrow = fmaxf(mw, mo);
  float tot = 0.f;
  if (mw != -INFINITY) tot += sw * exp2f(mw - mrow);
  if (mo != -INFINITY) tot += so * exp2f(mo - mrow);
  const float inv = tot > 0.f ? 1.f / tot : 0.f;
  // chunked keys: log2 of the row's partition sum over THIS chunk (-inf when no key of the chunk is valid), from which the
  // combine pass rescales p and the context rows to the softmax over all chunks
  if (a.lse2 != nullptr && hs == 0 && row_ok)
    a.lse2[((long long)h * a.B + b) * a.Tq + i] = tot > 0.f ? mrow + log2f(tot) : -INFINITY;

  // p = e * exp2(m_c - m) / sum -> HBM (bf16, one 32-byte sector per thread and chunk); dropout(p) -> K-major A operand
  const long long rg = ((long long)h * a.B + b) * a.Tq + i;  // row of the head-major [H,B,Tq,ld_p] tensors
  bf16* prow = a.p + rg * a.ld_p;
  bf16* pdrow = a.pd != nullptr ? a.pd + rg * a.ld_p : nullptr;
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    if (c < nmine) {
      const int j0 = jbeg + c * CH;
      const float f = (mc[c] == -INFINITY) ? 0.f : exp2f(mc[c] - mrow) * inv;
      float v[CH];
#pragma unroll
      for (int e = 0; e < CH / 2; ++e) {
        const float2 ev = __half22float2(*reinterpret_cast<const __half2*>(&E[c][e]));
        v[2 * e] = ev.x * f;
        v[2 * e + 1] = ev.y * f;
      }
      uint32_t pk[CH / 2];
#pragma unroll
      for (int e = 0; e < CH / 2; ++e) pk[e] = pack_bf16x2(v[2 * e], v[2 * e + 1]);
      const bool full = row_ok && j0 + CH <= a.p_cols, wide = (a.ld_p & 15) == 0;  // row pitch a multiple of 32 bytes
      auto store16 = [&](bf16* dst, const uint32_t (&w)[CH / 2]) {
        if (full && wide) {
          stg256(dst, w);
        } else if (row_ok && j0 < a.p_cols) {
          *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
          if (full) *reinterpret_cast<uint4*>(dst + 8) = make_uint4(w[4], w[5], w[6], w[7]);
        }
      };
      store16(prow + j0, pk);
      if (dropping) {
        // the A operand is p with the dropped lanes zeroed; the 1 / (1 - rate) factor is applied to the context rows
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          uint32_t w[4];
          dropout_lanes8(dk, (unsigned long long)(rg * a.ld_p + a.key_off + j0 + g * 8) >> 3, w);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const uint32_t mlo = (w[e] & 0xffffu) >= dk.thr ? 0x0000ffffu : 0u, mhi = (w[e] >> 16) >= dk.thr ? 0xffff0000u : 0u;
            pk[g * 4 + e] &= (mlo | mhi);
          }
        }
        if (pdrow != nullptr) {  // only for the unfused backward: dropout(p) with its scale, as b200slu_softmax_fwd writes it
          uint32_t pq[CH / 2];
#pragma unroll
          for (int e = 0; e < CH / 2; ++e) {
            const float2 x = unpack_bf16x2(pk[e]);
            pq[e] = pack_bf16x2(x.x * dk.inv, x.y * dk.inv);
          }
          store16(pdrow + j0, pq);
        }
      }
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        const int jc = j0 + g * 8;
        const uint32_t chunk = (uint32_t)((jc & 63) >> 3);
        sts128(sbase + (uint32_t)(jc >> 6) * (QT * 128) + (uint32_t)r * 128u + ((chunk ^ (uint32_t)(r & 7)) << 4), pk[g * 4],
               pk[g * 4 + 1], pk[g * 4 + 2], pk[g * 4 + 3]);
      }
    }
  }
  fence_async_smem();
  tcgen05_fence_before();
  group_sync(bar_id);

}

__global__ void __launch_bounds__(NTHREADS, 2)
relattn_fwd_kernel(const __grid_constant__ CUtensorMap map_qu, const __grid_constant__ CUtensorMap map_qv,
                   const __grid_constant__ CUtensorMap map_k, const __grid_constant__ CUtensorMap map_v,
                   const __grid_constant__ CUtensorMap map_pos, const AttnArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = smem_u32(smem);
  if (sbase & 1023u) __trap();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nqt = (a.Tq + QT - 1) / QT;
  const int qt = blockIdx.x % nqt, hb = blockIdx.x / nqt, h = hb % a.H, b = hb / a.H;
  const int i0 = qt * QT;
  const int TKP = a.TKP, WIN = a.WIN;
  const uint32_t bar_qk = sbase + a.off_bar, bar_v = bar_qk + 8, bar_s = bar_qk + 16, bar_o = bar_qk + 24;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + a.off_bar + 32);
  float* rowstat = reinterpret_cast<float*>(smem + a.off_bar + 64);  // [2 halves][128 rows] max, then the same for sums
  const uint32_t s_qu = sbase, s_qv = sbase + QT * 128, s_k = sbase + a.off_k, s_p = sbase + a.off_p, s_v = sbase + a.off_v;

  long long ts[8];
  ts[0] = clock64();
  b2_pdl_launch_dependents();
  if (threadIdx.x == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_qu) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_k) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_v) : "memory");
    if (a.rel) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_qv) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_pos) : "memory");
    }
    mbar_init(bar_qk, 1);
    mbar_init(bar_v, 1);
    mbar_init(bar_s, 1);
    mbar_init(bar_o, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    // every operand of the tile is requested before this CTA queues for tensor memory
    b2_pdl_wait();
    mbar_expect_tx(bar_qk, (uint32_t)(QT * 128 * (a.rel ? 2 : 1) + TKP * 128 + (a.rel ? WIN * 128 : 0)));
    tma_load_3d(s_qu, &map_qu, bar_qk, h * DK, i0, b);
    tma_load_3d(s_k, &map_k, bar_qk, h * DK, 0, b);
    if (a.rel) {
      tma_load_3d(s_qv, &map_qv, bar_qk, h * DK, i0, b);
      // first position row the tile can reach = source column of (last query row of the tile, key 0); may be negative:
      // out-of-range rows are zero-filled
      const int c0 = b2_rel_shift_src(i0 + QT - 1, 0, a.Tq, B200SLU_RELPOS_LATEST).col + a.key_off;
      tma_load_3d(s_p, &map_pos, bar_qk, h * DK, c0, 0);
      tma_load_3d(s_p + (WIN / 2) * 128, &map_pos, bar_qk, h * DK, c0 + WIN / 2, 0);
    }
    mbar_expect_tx(bar_v, (uint32_t)(TKP * 128));
    tma_load_3d(s_v, &map_v, bar_v, h * DK, 0, b);
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(a.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // global scalars of the softmax (valid length, dropout key): fetched here so their latency hides behind the TMA loads
  b2_pdl_wait();
  const int len_b = a.lens != nullptr ? max(0, min(a.lens[b] - a.key_off, a.Tk)) : a.Tk;
  B2DropKey dk = {0ull, 0u, 1.f};
  const bool dropping = a.drop.p > 0.f;
  if (dropping) dk = dropout_key(a.drop);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  ts[1] = clock64();

  // ---------------------------------------------------------------- scores: ac (and bd) into tensor memory
  if (threadIdx.x == 0) {
    mbar_wait(bar_qk, 0);
    tcgen05_fence_after();
    const uint32_t id_ac = make_idesc(TKP, 0, 0, QT);
#pragma unroll
    for (int kk = 0; kk < DK / 16; ++kk)
      umma_bf16(tmem_base, make_smem_desc(s_qu + kk * 32, 16, 1024), make_smem_desc(s_k + kk * 32, 16, 1024), id_ac, kk != 0);
    if (a.rel) {
      const uint32_t id_bd = make_idesc(WIN / 2, 0, 0, QT);
      for (int half = 0; half < 2; ++half) {
#pragma unroll
        for (int kk = 0; kk < DK / 16; ++kk)
          umma_bf16(tmem_base + TKP + half * (WIN / 2), make_smem_desc(s_qv + kk * 32, 16, 1024),
                    make_smem_desc(s_p + half * (WIN / 2) * 128 + kk * 32, 16, 1024), id_bd, kk != 0);
      }
    }
    umma_commit(bar_s);
  }
  mbar_wait(bar_s, 0);
  tcgen05_fence_after();
  ts[2] = clock64();

  softmax_tile(a, smem, sbase, threadIdx.x, 0, tmem_base, rowstat, i0, h, b, len_b, dk, 0u);
  const int q = warp & 3, hs = warp >> 2;
  const int r = q * 32 + lane, i = i0 + r;
  const bool row_ok = i < a.Tq;
  const uint32_t tlane = tmem_base + ((uint32_t)(q * 32) << 16);
  ts[3] = ts[4] = clock64();
  ts[5] = clock64();
  // ---------------------------------------------------------------- context: O = P . V
  const uint32_t t_o = tmem_base + TKP;  // the bd window is dead
  if (threadIdx.x == 0) {
    mbar_wait(bar_v, 0);
    tcgen05_fence_after();
    const uint32_t id_o = make_idesc(DK, 0, 1, QT);
    for (int ks = 0; ks < TKP / 16; ++ks)
      umma_bf16(t_o, make_smem_desc(sbase + (uint32_t)(ks >> 2) * (QT * 128) + (ks & 3) * 32, 16, 1024),
                make_smem_desc(s_v + ks * 2048, 64 * DK * 2, 1024), id_o, ks != 0);
    umma_commit(bar_o);
  }
  mbar_wait(bar_o, 0);
  tcgen05_fence_after();
  ts[6] = clock64();
  {
    uint32_t o[32];
    tmem_ld32(tlane + TKP + hs * 32, o);
    tmem_ld_wait();
    if (row_ok) {
      bf16* dst = a.ctx + ((long long)b * a.Tq + i) * a.ldctx + h * DK + hs * 32;
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(o[g * 8 + e]) * dk.inv;
        store8_bf16(dst + g * 8, v);
      }
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (a.dbg != nullptr && threadIdx.x == 0 && blockIdx.x < 512) {
    ts[7] = clock64();
    for (int t = 0; t < 8; ++t) a.dbg[blockIdx.x * 8 + t] = ts[t];
  }
  if (warp == 1) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(a.tmem_cols) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Two tiles per SM in flight.  The one-tile kernel above lets two CTAs share an SM, but its 448 score columns make tensor
// memory exclusive, so the CTAs take turns and every phase runs as a latency chain of 8 warps.  Here ONE persistent CTA
// of 16 warps owns all 512 columns and runs two groups of 8 warps, each with its own shared-memory region, on alternate
// tiles.  Only two things are shared and handed back and forth through mbarriers: the score columns [0, 448) (from the
// score MMAs until the softmax pass has read them) and the context columns [448, 512) (from the P.V MMAs until the
// epilogue has read them).  A group's finalise / P.V / epilogue therefore overlaps the other group's loads / score MMAs /
// softmax pass.
constexpr int PIPE_O_COL = 448;
constexpr int PIPE_CTL_BYTES = 128 + 2 * 4 * QT * 4;

__global__ void __launch_bounds__(2 * NTHREADS, 1)
relattn_fwd_pipe_kernel(const __grid_constant__ CUtensorMap map_qu, const __grid_constant__ CUtensorMap map_qv,
                        const __grid_constant__ CUtensorMap map_k, const __grid_constant__ CUtensorMap map_v,
                        const __grid_constant__ CUtensorMap map_pos, const AttnArgs a, int num_tiles) {
  extern __shared__ __align__(1024) uint8_t smem_all[];
  const uint32_t sall = smem_u32(smem_all);
  if (sall & 1023u) __trap();
  const int g = threadIdx.x >> 8, tid = threadIdx.x & 255, warp = tid >> 5, lane = tid & 31;
  const uint32_t region = a.off_bar;  // one tile's operands (the one-tile kernel's layout), a multiple of 1024 bytes
  uint8_t* smem = smem_all + (size_t)g * region;
  const uint32_t sbase = sall + g * region, ctl = sall + 2 * region;
  const uint32_t bar_qk = ctl + g * 32, bar_v = bar_qk + 8, bar_s = bar_qk + 16, bar_o = bar_qk + 24;
  const uint32_t s_free = ctl + 64, o_free = ctl + 80;  // [2] each: "group g may use the score / context columns"
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem_all + 2 * (size_t)region + 96);
  float* rowstat = reinterpret_cast<float*>(smem_all + 2 * (size_t)region + 128) + g * 4 * QT;
  const int TKP = a.TKP, WIN = a.WIN;
  const uint32_t s_qu = sbase, s_qv = sbase + QT * 128, s_k = sbase + a.off_k, s_p = sbase + a.off_p, s_v = sbase + a.off_v;
  const int nqt = (a.Tq + QT - 1) / QT;

  b2_pdl_launch_dependents();
  if (threadIdx.x == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_qu) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_k) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_v) : "memory");
    if (a.rel) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_qv) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_pos) : "memory");
    }
    for (int i = 0; i < 12; ++i) mbar_init(ctl + i * 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    mbar_arrive(s_free);  // group 0 goes first
    mbar_arrive(o_free);
  }
  if ((threadIdx.x >> 5) == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  b2_pdl_wait();
  B2DropKey dk = {0ull, 0u, 1.f};
  if (a.drop.p > 0.f) dk = dropout_key(a.drop);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t t_s = tmem_base, t_o = tmem_base + PIPE_O_COL;

  const int n_c = (int)blockIdx.x < num_tiles ? (num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;  // this CTA's tiles
  const int n_iter = (n_c + 1) / 2;
  const int q = warp & 3, hs = warp >> 2;
  int kdone = 0;  // tiles this group has run: parity of its own barriers
  for (int it = 0; it < n_iter; ++it) {
    const int j = 2 * it + g;
    const uint32_t tokpar = (uint32_t)(it & 1);
    if (j >= n_c) {  // odd tile count: keep the hand-over going
      if (tid == 0) {
        mbar_wait(s_free + g * 8, tokpar);
        mbar_arrive(s_free + (g ^ 1) * 8);
        mbar_wait(o_free + g * 8, tokpar);
        mbar_arrive(o_free + (g ^ 1) * 8);
      }
      continue;
    }
    const int tile = (int)blockIdx.x + j * (int)gridDim.x;
    const int qt = tile % nqt, hb = tile / nqt, h = hb % a.H, b = hb / a.H;
    const int i0 = qt * QT;
    const uint32_t par = (uint32_t)(kdone & 1);
    if (tid == 0) {
      mbar_expect_tx(bar_qk, (uint32_t)(QT * 128 * (a.rel ? 2 : 1) + TKP * 128 + (a.rel ? WIN * 128 : 0)));
      tma_load_3d(s_qu, &map_qu, bar_qk, h * DK, i0, b);
      tma_load_3d(s_k, &map_k, bar_qk, h * DK, 0, b);
      if (a.rel) {
        tma_load_3d(s_qv, &map_qv, bar_qk, h * DK, i0, b);
        const int c0 = b2_rel_shift_src(i0 + QT - 1, 0, a.Tq, B200SLU_RELPOS_LATEST).col + a.key_off;
        tma_load_3d(s_p, &map_pos, bar_qk, h * DK, c0, 0);
        tma_load_3d(s_p + (WIN / 2) * 128, &map_pos, bar_qk, h * DK, c0 + WIN / 2, 0);
      }
      mbar_expect_tx(bar_v, (uint32_t)(TKP * 128));
      tma_load_3d(s_v, &map_v, bar_v, h * DK, 0, b);
    }
    const int len_b = a.lens != nullptr ? max(0, min(a.lens[b] - a.key_off, a.Tk)) : a.Tk;
    if (tid == 0) {
      mbar_wait(bar_qk, par);
      mbar_wait(s_free + g * 8, tokpar);  // the other group's softmax pass has read its scores
      tcgen05_fence_after();
      const uint32_t id_ac = make_idesc(TKP, 0, 0, QT);
#pragma unroll
      for (int kk = 0; kk < DK / 16; ++kk)
        umma_bf16(t_s, make_smem_desc(s_qu + kk * 32, 16, 1024), make_smem_desc(s_k + kk * 32, 16, 1024), id_ac, kk != 0);
      if (a.rel) {
        const uint32_t id_bd = make_idesc(WIN / 2, 0, 0, QT);
        for (int half = 0; half < 2; ++half) {
#pragma unroll
          for (int kk = 0; kk < DK / 16; ++kk)
            umma_bf16(t_s + TKP + half * (WIN / 2), make_smem_desc(s_qv + kk * 32, 16, 1024),
                      make_smem_desc(s_p + half * (WIN / 2) * 128 + kk * 32, 16, 1024), id_bd, kk != 0);
        }
      }
      umma_commit(bar_s);
    }
    mbar_wait(bar_s, par);
    tcgen05_fence_after();
    softmax_tile(a, smem, sbase, tid, 1 + g, t_s, rowstat, i0, h, b, len_b, dk, s_free + (g ^ 1) * 8);
    if (tid == 0) {
      mbar_wait(bar_v, par);
      mbar_wait(o_free + g * 8, tokpar);  // the other group's epilogue has read its context rows
      tcgen05_fence_after();
      const uint32_t id_o = make_idesc(DK, 0, 1, QT);
      for (int ks = 0; ks < TKP / 16; ++ks)
        umma_bf16(t_o, make_smem_desc(sbase + (uint32_t)(ks >> 2) * (QT * 128) + (ks & 3) * 32, 16, 1024),
                  make_smem_desc(s_v + ks * 2048, 64 * DK * 2, 1024), id_o, ks != 0);
      umma_commit(bar_o);
    }
    mbar_wait(bar_o, par);
    tcgen05_fence_after();
    uint32_t o[32];
    tmem_ld32(t_o + ((uint32_t)(q * 32) << 16) + hs * 32, o);
    tmem_ld_wait();
    tcgen05_fence_before();
    fence_async_smem();  // this tile's ring / P operand writes are ordered before the next tile's TMA loads into the same bytes
    group_sync(1 + g);
    if (tid == 0) mbar_arrive(o_free + (g ^ 1) * 8);
    const int i = i0 + q * 32 + lane;
    if (i < a.Tq) {
      bf16* dst = a.ctx + ((long long)b * a.Tq + i) * a.ldctx + h * DK + hs * 32;
#pragma unroll
      for (int gq = 0; gq < 4; ++gq) {
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(o[gq * 8 + e]) * dk.inv;
        store8_bf16(dst + gq * 8, v);
      }
    }
    ++kdone;
  }
  tcgen05_fence_before();
  __syncthreads();
  if ((threadIdx.x >> 5) == 1) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Backward, keys <= 192: ONE kernel per layer for everything between the gradient of the context and the gradients
// of q_u / k / v (reference: autograd through attention.py:78-95 and the two matmuls of :297-302 / :111-113):
//
//   dPd = dO . V^T                    tcgen05, 128 queries x TKP keys into tensor memory           (never leaves the SM)
//   dS  = p (m dPd - delta) scale     delta_i = <dO_i, O_i> (= sum_j dP_ij p_ij, the flash-attention identity), m = the
//                                     element's dropout multiplier regenerated from (seed, site, index)
//   dQ_u = dS . K     dK += dS^T . Q_u     dV += (m p)^T . dO          tcgen05 out of shared memory
//
// p (bf16, what the forward saved) arrives by TMA straight into the shared-memory tile that becomes the A operand of
// the dV contraction: the threads turn it into dropout(p) IN PLACE and write dS into a second tile; both tiles are
// SWIZZLE_128B [128 query rows x 64 keys] blocks, read K-major (A = dS for dQ_u) and MN-major (A = dS^T / pd^T for
// dK / dV) through the UMMA descriptors -- no transposed copy exists anywhere.  dP, dS, dropout(p) and (rel-pos modes)
// d(matrix_bd) never touch HBM.  One CTA owns one (utterance, head) at a time and walks its query tiles, so dK / dV
// accumulate in tensor memory across the tiles; thread = TMEM lane = query row (or key row in the dK / dV epilogue).
// Rel-pos (latest): d(matrix_bd) is dS un-shifted, d(bd)[r, 127 - r + j] = dS[r, j] in the coordinates of the position
// window the query tile can reach (the forward's window).  Once dQ_u / dK / dV are issued the threads rebuild dS as that
// SKEWED tile [128 x WIN] over the dead operand tiles: a row moves right by s = 127 - r elements = (s / 8) 16-byte
// units (placement) + (s % 8) elements (a funnel shift over two neighbouring units in registers: two select levels for
// the 4-byte part, one PRMT for the odd element).  The same tile is then the K-major A operand of
// dQ_v = d(bd) . P_window and the MN-major A operand of dP_window = d(bd)^T . Q_v; dP rows go to the fp32 gradient of
// the projected positions with vector reductions (red.global.add.v4.f32).  Nothing of shape T x T is written to HBM.
constexpr int BWD_DK_COL = 192;   // dK: two key tiles x 64 columns
constexpr int BWD_DV_COL = 320;   // dV: two key tiles x 64 columns
constexpr int BWD_DQV_COL = 448;  // dQ_v (rel-pos); d(pos) tiles reuse columns [0, 192) once dQ_u has been read out
constexpr int BLK = QT * 128;     // bytes of one [128 rows x 64 bf16] swizzled block

struct AttnBwdArgs {
  int B, H, Tq, Tk, TKP;
  float scale;
  const bf16* ctx;  // forward output O, merged heads [B*Tq, ldctx]
  long long ldctx;
  bf16 *dq, *dk, *dv, *dac;
  long long lddq, lddkv, ld_p;
  bf16* dqv;        // rel-pos: gradient of q_v, rows like dq
  float* dpos;      // rel-pos: fp32 gradient of the projected positions [2T-1, lddpos], accumulated
  long long lddpos;
  int WIN;
  uint32_t off_pd, off_ds, off_k, off_v, off_do, off_qu, off_pos, off_stg, off_bar;
  B2Dropout drop;
  long long* dbg;  // tuning only (B200SLU_ATTN_BWD_DEBUG): CTA 0's cycles per phase, summed over its query tiles
  int dbgflags;    // tuning only (B200SLU_ATTN_BWD_FLAGS): 1 = no d(pos) reduce, 2 = no load prefetch
  // Key chunks of a longer key axis (b200slu_relattn_bwd_chunks): work unit = (chunk, utterance, head); chunk c covers keys
  // [c * chunk_keys, +min(chunk_keys, Tk - c * chunk_keys)) of the kv_rows keys per utterance (Tk = kv_rows then) and
  // writes its dq_u / dq_v contribution to slab c (dq_slab elements apart).  The key offset shifts the K / V / p tile
  // coordinates, the rel-pos window and the dropout counter.  One chunk: chunk_keys = Tk, the plain kernel.
  int nchunk, chunk_keys, kv_rows, key_begin, key_end;  // chunk c covers keys key_begin + c * chunk_keys ... (< key_end)
  float *dk_colsum, *dv_colsum;  // optional fp32 [H*64] accumulators: column sums of dk / dv = the bias gradients of the K / V
                                 // projections (saves the column-sum pass over dqkv)
  long long dq_slab;
};

// Column sums over a warp's 32 rows of a [32 rows (lanes)] x [32 columns (registers)] tile: a butterfly reduce-scatter,
// 31 shuffles; lane l returns the sum of column l.
__device__ __forceinline__ float warp_colsum32(float (&v)[32], int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < off; ++i) {
      const float send = up ? v[i] : v[i + off], keep = up ? v[i + off] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0];
}

template <bool REL>
__global__ void __launch_bounds__(NTHREADS, 1)
relattn_bwd_kernel(const __grid_constant__ CUtensorMap map_do, const __grid_constant__ CUtensorMap map_qu,
                   const __grid_constant__ CUtensorMap map_k, const __grid_constant__ CUtensorMap map_v,
                   const __grid_constant__ CUtensorMap map_p, const __grid_constant__ CUtensorMap map_qv,
                   const __grid_constant__ CUtensorMap map_pos, const AttnBwdArgs a, int num_units) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = smem_u32(smem);
  if (sbase & 1023u) __trap();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t s_pd = sbase + a.off_pd, s_ds = sbase + a.off_ds, s_k = sbase + a.off_k, s_v = sbase + a.off_v,
                 s_do = sbase + a.off_do, s_qu = sbase + a.off_qu;
  const uint32_t s_pos = sbase + a.off_pos, s_qv = s_k, s_dbd = sbase;  // Q_v lands where K was; the skewed tile over pd / dO / Q_u / V
  const uint32_t bar_kv = sbase + a.off_bar, bar_q = bar_kv + 8, bar_s = bar_kv + 16, bar_c = bar_kv + 24;
  const uint32_t bar_qv = bar_kv + 32, bar_c2 = bar_kv + 40;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + a.off_bar + 48);
  float* rowstat = reinterpret_cast<float*>(smem + a.off_bar + 64);  // [2 halves][128 rows] partial delta
  const int TKP = a.TKP, nblk = (TKP + 63) / 64, nkt = (TKP + QT - 1) / QT;
  const int nqt = (a.Tq + QT - 1) / QT;

  b2_pdl_launch_dependents();
  if (tid == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_do) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_qu) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_k) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_v) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_p) : "memory");
    if (REL) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_qv) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_pos) : "memory");
    }
    for (int i = 0; i < 6; ++i) mbar_init(bar_kv + i * 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  b2_pdl_wait();
  B2DropKey dk = {0ull, 0u, 1.f};
  const bool dropping = a.drop.p > 0.f;
  if (dropping) dk = dropout_key(a.drop);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int q = warp & 3, hs = warp >> 2;
  const int r = q * 32 + lane;
  const uint32_t tlane = tmem_base + ((uint32_t)(q * 32) << 16);
  const int nch = TKP / CH, n0 = (nch + 1) / 2;
  const int cbeg = hs ? n0 : 0, nmine = hs ? nch - n0 : n0;
  uint32_t n_q = 0;  // completed query tiles: the parity of every barrier (each completes once per tile)
  long long acc_t[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const bool rec = a.dbg != nullptr && blockIdx.x == 0 && tid == 32;
#define BWD_STAMP(k_)                     \
  if (rec) {                              \
    const long long now_ = clock64();     \
    acc_t[k_] += now_ - t_prev;           \
    t_prev = now_;                        \
  }
  long long t_prev = rec ? clock64() : 0;

  // All operand loads of one query tile (thread 0).  K and V come per tile too: their regions are recycled (Q_v lands on K,
  // the skewed tile on V).  Issued one tile AHEAD, as soon as the last MMA that reads those regions has completed, so the
  // load latency hides behind the previous tile's read-out.
  auto issue_loads = [&](int unit_, int qt_) {
    const int bh_ = unit_ % (a.B * a.H), koff_ = a.key_begin + (unit_ / (a.B * a.H)) * a.chunk_keys;
    const int h_ = bh_ % a.H, b_ = bh_ / a.H, i0_ = qt_ * QT;
    mbar_expect_tx(bar_kv, (uint32_t)(2 * TKP * 128));
    tma_load_3d(s_k, &map_k, bar_kv, h_ * DK, koff_, b_);
    tma_load_3d(s_v, &map_v, bar_kv, h_ * DK, koff_, b_);
    mbar_expect_tx(bar_q, (uint32_t)(2 * BLK + nblk * BLK + (REL ? a.WIN * 128 : 0)));
    tma_load_3d(s_do, &map_do, bar_q, h_ * DK, i0_, b_);
    tma_load_3d(s_qu, &map_qu, bar_q, h_ * DK, i0_, b_);
    for (int kb = 0; kb < nblk; ++kb) tma_load_3d(s_pd + kb * BLK, &map_p, bar_q, koff_ + kb * 64, i0_, h_ * a.B + b_);
    if (REL) {
      const int c0_ = b2_rel_shift_src(i0_ + QT - 1, 0, a.Tq, B200SLU_RELPOS_LATEST).col + koff_;
      tma_load_3d(s_pos, &map_pos, bar_q, h_ * DK, c0_, 0);
      tma_load_3d(s_pos + (a.WIN / 2) * 128, &map_pos, bar_q, h_ * DK, c0_ + a.WIN / 2, 0);
    }
  };
  auto issue_next = [&](int unit_, int qt_) {
    if (a.dbgflags & 2) return;
    if (qt_ + 1 < nqt) issue_loads(unit_, qt_ + 1);
    else if (unit_ + (int)gridDim.x < num_units) issue_loads(unit_ + (int)gridDim.x, 0);
  };
  if (tid == 0 && (int)blockIdx.x < num_units) issue_loads(blockIdx.x, 0);
  const uint32_t s_stg = sbase + a.off_stg;

  for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x) {
    const int bh = unit % (a.B * a.H), chunk = unit / (a.B * a.H);
    const int h = bh % a.H, b = bh / a.H;
    const int key_off = a.key_begin + chunk * a.chunk_keys, tk_c = min(a.chunk_keys, a.key_end - key_off);  // this unit's keys
    for (int qt = 0; qt < nqt; ++qt) {
      const int i0 = qt * QT, i = i0 + r;
      const bool row_ok = i < a.Tq;
      const bool warp_live = i0 + q * 32 < a.Tq;  // warp-uniform: any valid query row in this warp's 32 lanes
      const uint32_t par = n_q & 1u;
      if ((a.dbgflags & 2) && tid == 0 && !(unit == (int)blockIdx.x && qt == 0)) issue_loads(unit, qt);
      // first position row of the window this tile reaches (the forward's window); may be negative: zero-filled
      const int c0 = REL ? b2_rel_shift_src(i0 + QT - 1, 0, a.Tq, B200SLU_RELPOS_LATEST).col + key_off : 0;
      // this thread's half (32 columns) of its row of the forward output, for delta
      uint4 cx[4];
      if (row_ok) {
        const uint4* src = reinterpret_cast<const uint4*>(a.ctx + ((long long)b * a.Tq + i) * a.ldctx + h * DK + hs * 32);
#pragma unroll
        for (int t = 0; t < 4; ++t) cx[t] = __ldg(src + t);
      } else {
#pragma unroll
        for (int t = 0; t < 4; ++t) cx[t] = make_uint4(0u, 0u, 0u, 0u);
      }
      if (tid == 0) {
        mbar_wait(bar_kv, par);
        mbar_wait(bar_q, par);
        tcgen05_fence_after();
        const uint32_t id_s = make_idesc(TKP, 0, 0, QT);
#pragma unroll
        for (int kk = 0; kk < DK / 16; ++kk)
          umma_bf16(tmem_base, make_smem_desc(s_do + kk * 32, 16, 1024), make_smem_desc(s_v + kk * 32, 16, 1024), id_s, kk != 0);
        umma_commit(bar_s);
      }
      mbar_wait(bar_q, par);
      BWD_STAMP(0)
      float dpart = 0.f;
      {
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const uint32_t chunk = (uint32_t)(hs * 4 + t);
          float d0, d1, d2, d3;
          lds128(s_do + (uint32_t)r * 128u + ((chunk ^ (uint32_t)(r & 7)) << 4), d0, d1, d2, d3);
          const uint32_t dw[4] = {__float_as_uint(d0), __float_as_uint(d1), __float_as_uint(d2), __float_as_uint(d3)};
          const uint32_t cw[4] = {cx[t].x, cx[t].y, cx[t].z, cx[t].w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 x = unpack_bf16x2(dw[e]), y = unpack_bf16x2(cw[e]);
            dpart = fmaf(x.x, y.x, dpart);
            dpart = fmaf(x.y, y.y, dpart);
          }
        }
      }
      rowstat[hs * QT + r] = dpart;
      group_sync(0);
      const float delta = rowstat[r] + rowstat[QT + r];
      mbar_wait(bar_s, par);
      tcgen05_fence_after();
      BWD_STAMP(1)

      // ------------------------------------------------------------ dS and dropout(p), 16 keys at a time
      const long long rg = ((long long)h * a.B + b) * a.Tq + i;
      bf16* dacrow = a.dac != nullptr ? a.dac + rg * a.ld_p : nullptr;
      const float sc = a.scale;
      uint32_t ta[16], tb[16];
      if (!warp_live) {
        // no valid query row in this warp (tail tile): p arrived as zeros (TMA fill), so dropout(p) is in place already;
        // dS = 0
        for (int cc = 0; cc < nmine; ++cc) {
          const int j0 = (cbeg + cc) * CH;
          const uint32_t blk = (uint32_t)(j0 >> 6) * BLK + (uint32_t)r * 128u, ch0 = (uint32_t)((j0 & 63) >> 3);
          sts128(s_ds + blk + (((ch0) ^ (uint32_t)(r & 7)) << 4), 0u, 0u, 0u, 0u);
          sts128(s_ds + blk + (((ch0 + 1u) ^ (uint32_t)(r & 7)) << 4), 0u, 0u, 0u, 0u);
        }
      }
      if (warp_live && nmine > 0) tmem_ld16(tlane + cbeg * CH, ta);
#pragma unroll 1
      for (int c = 0; warp_live && c < nmine; c += 2) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int cc = c + half;
          if (cc < nmine) {
            const int j0 = (cbeg + cc) * CH;
            tmem_ld_wait();
            if (cc + 1 < nmine) tmem_ld16(tlane + j0 + CH, half ? ta : tb);
            const uint32_t(&dp)[16] = half ? tb : ta;
            const uint32_t blk = (uint32_t)(j0 >> 6) * BLK + (uint32_t)r * 128u;
            const uint32_t ch0 = (uint32_t)((j0 & 63) >> 3);
            const uint32_t ad0 = blk + (((ch0) ^ (uint32_t)(r & 7)) << 4), ad1 = blk + (((ch0 + 1u) ^ (uint32_t)(r & 7)) << 4);
            float p0, p1, p2, p3, p4, p5, p6, p7;
            lds128(s_pd + ad0, p0, p1, p2, p3);
            lds128(s_pd + ad1, p4, p5, p6, p7);
            const uint32_t pw[8] = {__float_as_uint(p0), __float_as_uint(p1), __float_as_uint(p2), __float_as_uint(p3),
                                    __float_as_uint(p4), __float_as_uint(p5), __float_as_uint(p6), __float_as_uint(p7)};
            float mul[16];
            if (dropping) {
#pragma unroll
              for (int g = 0; g < 2; ++g) {
                float m8[8];
                dropout_mul8(dk, (unsigned long long)(rg * a.ld_p + key_off + j0 + g * 8) >> 3, dk.inv, m8);
#pragma unroll
                for (int e = 0; e < 8; ++e) mul[g * 8 + e] = m8[e];
              }
            } else {
#pragma unroll
              for (int e = 0; e < 16; ++e) mul[e] = 1.f;
            }
            uint32_t pdw[8], dsw[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const float2 pv = unpack_bf16x2(pw[e]);
              const float g0 = __uint_as_float(dp[2 * e]), g1 = __uint_as_float(dp[2 * e + 1]);
              const float m0 = mul[2 * e], m1 = mul[2 * e + 1];
              pdw[e] = pack_bf16x2(pv.x * m0, pv.y * m1);
              // p == 0 (masked key, padded row) gives exactly 0 whatever the accumulator holds there
              const float s0 = pv.x != 0.f ? pv.x * (fmaf(g0, m0, -delta)) * sc : 0.f;
              const float s1 = pv.y != 0.f ? pv.y * (fmaf(g1, m1, -delta)) * sc : 0.f;
              dsw[e] = pack_bf16x2(s0, s1);
            }
            sts128(s_pd + ad0, pdw[0], pdw[1], pdw[2], pdw[3]);
            sts128(s_pd + ad1, pdw[4], pdw[5], pdw[6], pdw[7]);
            sts128(s_ds + ad0, dsw[0], dsw[1], dsw[2], dsw[3]);
            sts128(s_ds + ad1, dsw[4], dsw[5], dsw[6], dsw[7]);
            if (dacrow != nullptr && row_ok && j0 < a.ld_p) {
              if (j0 + CH <= a.ld_p && (a.ld_p & 15) == 0) {
                stg256(dacrow + j0, dsw);
              } else {
                *reinterpret_cast<uint4*>(dacrow + j0) = make_uint4(dsw[0], dsw[1], dsw[2], dsw[3]);
                if (j0 + CH <= a.ld_p) *reinterpret_cast<uint4*>(dacrow + j0 + 8) = make_uint4(dsw[4], dsw[5], dsw[6], dsw[7]);
              }
            }
          }
        }
      }
      fence_async_smem();
      tcgen05_fence_before();
      group_sync(0);
      BWD_STAMP(2)

      // ------------------------------------------------------------ dQ_u, dK, dV
      if (tid == 0) {
        tcgen05_fence_after();
        const uint32_t id_q = make_idesc(DK, 0, 1, QT);  // A = dS (K-major), B = K rows as MN-major
        for (int ks = 0; ks < TKP / 16; ++ks)
          umma_bf16(tmem_base, make_smem_desc(s_ds + (uint32_t)(ks >> 2) * BLK + (ks & 3) * 32, 16, 1024),
                    make_smem_desc(s_k + ks * 2048, 64 * DK * 2, 1024), id_q, ks != 0);
        const uint32_t id_t = make_idesc(DK, 1, 1, QT);  // A = dS^T / pd^T (MN-major), B = Q_u / dO rows (MN-major)
        // the contraction runs over the tile's query rows: a last tile with few valid rows (27 of 128 at T = 155) issues
        // only the k-steps that hold any (each tcgen05.mma costs ~115 cycles of issue whatever its shape)
        const int kkn = (min(QT, a.Tq - i0) + 15) >> 4;
        for (int m = 0; m < nkt; ++m) {
          for (int kk = 0; kk < kkn; ++kk) {
            const uint32_t acc = (qt != 0 || kk != 0) ? 1u : 0u;
            umma_bf16(tmem_base + BWD_DK_COL + m * DK, make_smem_desc(s_ds + (uint32_t)m * 2 * BLK + kk * 2048, BLK, 1024),
                      make_smem_desc(s_qu + kk * 2048, 64 * DK * 2, 1024), id_t, acc);
            umma_bf16(tmem_base + BWD_DV_COL + m * DK, make_smem_desc(s_pd + (uint32_t)m * 2 * BLK + kk * 2048, BLK, 1024),
                      make_smem_desc(s_do + kk * 2048, 64 * DK * 2, 1024), id_t, acc);
          }
        }
        umma_commit(bar_c);
      }
      mbar_wait(bar_c, par);
      tcgen05_fence_after();
      auto drain_dkv = [&]() {
        for (int m = 0; m < nkt; ++m) {
          const int j = m * QT + r;
          uint32_t ok_[32], ov_[32];
          tmem_ld32(tlane + BWD_DK_COL + m * DK + hs * 32, ok_);
          tmem_ld32(tlane + BWD_DV_COL + m * DK + hs * 32, ov_);
          tmem_ld_wait();
          if (j < tk_c) {
            const long long row = ((long long)b * a.kv_rows + key_off + j) * a.lddkv + h * DK + hs * 32;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              float v[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(ok_[g * 8 + e]);
              store8_bf16(a.dk + row + g * 8, v);
#pragma unroll
              for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(ov_[g * 8 + e]);
              store8_bf16(a.dv + row + g * 8, v);
            }
          }
          if (a.dk_colsum != nullptr) {  // warp-uniform
            float ck[32], cv[32];
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              ck[e] = j < tk_c ? __uint_as_float(ok_[e]) : 0.f;
              cv[e] = j < tk_c ? __uint_as_float(ov_[e]) : 0.f;
            }
            const float sk = warp_colsum32(ck, lane), sv = warp_colsum32(cv, lane);
            atomicAdd(a.dk_colsum + h * DK + hs * 32 + lane, sk);
            atomicAdd(a.dv_colsum + h * DK + hs * 32 + lane, sv);
          }
        }
      };
      BWD_STAMP(3)
      if (REL && tid == 0) {  // K is dead: Q_v (B operand of the d(pos) contraction) takes its place while the tile is skewed
        mbar_expect_tx(bar_qv, (uint32_t)BLK);
        tma_load_3d(s_qv, &map_qv, bar_qv, h * DK, i0, b);
      }
      if (!REL && tid == 0) issue_next(unit, qt);  // every operand region is dead: the next tile's loads fly during the read-out
      {
        uint32_t o[32];
        tmem_ld32(tlane + hs * 32, o);
        tmem_ld_wait();
        if (row_ok) {
          bf16* dst = a.dq + chunk * a.dq_slab + ((long long)b * a.Tq + i) * a.lddq + h * DK + hs * 32;
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(o[g * 8 + e]);
            store8_bf16(dst + g * 8, v);
          }
        }
      }
      if (REL) {
        // ---------------------------------------------------------- d(bd): dS skewed into window coordinates
        const int WU = a.WIN >> 3, NU = TKP >> 3;   // 16-byte units per window row / per dS row
        const int sft = QT - 1 - r, ua = sft >> 3, bw = (sft >> 1) & 3;
        const uint32_t psel = (sft & 1) ? 0x5432u : 0x7654u;
        const bool s1 = bw & 1, s2 = bw & 2;
        const int ubeg = hs ? WU / 2 : 0, uend = hs ? WU : WU / 2;
        const uint32_t rsw = (uint32_t)(r & 7), rbase = (uint32_t)r * 128u;
        auto load_unit = [&](int k, uint32_t (&w)[4]) {
          if (k >= 0 && k < NU) {
            float f0, f1, f2, f3;
            lds128(s_ds + (uint32_t)(k >> 3) * BLK + rbase + ((((uint32_t)k & 7u) ^ rsw) << 4), f0, f1, f2, f3);
            w[0] = __float_as_uint(f0); w[1] = __float_as_uint(f1); w[2] = __float_as_uint(f2); w[3] = __float_as_uint(f3);
          } else {
            w[0] = w[1] = w[2] = w[3] = 0u;
          }
        };
        if (!warp_live) {
          for (int u = ubeg; u < uend; ++u)
            sts128(s_dbd + (uint32_t)(u >> 3) * BLK + rbase + ((((uint32_t)u & 7u) ^ rsw) << 4), 0u, 0u, 0u, 0u);
        }
        uint32_t y[8];
        if (warp_live) {
          uint32_t t4[4];
          load_unit(ubeg - ua - 1, t4);
#pragma unroll
          for (int e = 0; e < 4; ++e) y[4 + e] = t4[e];
        }
#pragma unroll 2
        for (int u = ubeg; warp_live && u < uend; ++u) {
          uint32_t t4[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) y[e] = y[4 + e];
          load_unit(u - ua, t4);
#pragma unroll
          for (int e = 0; e < 4; ++e) y[4 + e] = t4[e];
          uint32_t w1[7], z[5];
#pragma unroll
          for (int e = 0; e < 7; ++e) w1[e] = s1 ? y[e] : y[e + 1];
#pragma unroll
          for (int e = 0; e < 5; ++e) z[e] = s2 ? w1[e] : w1[e + 2];
          sts128(s_dbd + (uint32_t)(u >> 3) * BLK + rbase + ((((uint32_t)u & 7u) ^ rsw) << 4), __byte_perm(z[0], z[1], psel),
                 __byte_perm(z[1], z[2], psel), __byte_perm(z[2], z[3], psel), __byte_perm(z[3], z[4], psel));
        }
        fence_async_smem();
        tcgen05_fence_before();
        group_sync(0);
        BWD_STAMP(4)
        const int nmt = (a.WIN + QT - 1) / QT;
        if (tid == 0) {
          mbar_wait(bar_qv, par);
          tcgen05_fence_after();
          const uint32_t id_q = make_idesc(DK, 0, 1, QT);  // dQ_v = d(bd) . P_window
          for (int ks = 0; ks < a.WIN / 16; ++ks)
            umma_bf16(tmem_base + BWD_DQV_COL, make_smem_desc(s_dbd + (uint32_t)(ks >> 2) * BLK + (ks & 3) * 32, 16, 1024),
                      make_smem_desc(s_pos + ks * 2048, 64 * DK * 2, 1024), id_q, ks != 0);
          const uint32_t id_t = make_idesc(DK, 1, 1, QT);  // dP_window = d(bd)^T . Q_v
          const int kkn = (min(QT, a.Tq - i0) + 15) >> 4;
          for (int m = 0; m < nmt; ++m) {
            for (int kk = 0; kk < kkn; ++kk)
              umma_bf16(tmem_base + m * DK, make_smem_desc(s_dbd + (uint32_t)m * 2 * BLK + kk * 2048, BLK, 1024),
                        make_smem_desc(s_qv + kk * 2048, 64 * DK * 2, 1024), id_t, kk != 0);
          }
          umma_commit(bar_c2);
        }
        if (qt == nqt - 1) drain_dkv();  // final since the first MMA batch: read out under the second one
        mbar_wait(bar_c2, par);
        tcgen05_fence_after();
        if (tid == 0) issue_next(unit, qt);  // every operand region is dead: the next tile's loads fly during the read-out
        BWD_STAMP(5)
        {
          uint32_t o[32];
          tmem_ld32(tlane + BWD_DQV_COL + hs * 32, o);
          tmem_ld_wait();
          if (row_ok) {
            bf16* dst = a.dqv + chunk * a.dq_slab + ((long long)b * a.Tq + i) * a.lddq + h * DK + hs * 32;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              float v[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(o[g * 8 + e]);
              store8_bf16(dst + g * 8, v);
            }
          }
        }
        // d(pos): a warp's [32 window rows x 32 floats] slice goes through a private 4 KB staging patch (128-byte rows,
        // 16-byte chunks XORed with the row like every tile here), then leaves with 8 lanes per row: one reduction
        // instruction covers four whole 128-byte lines instead of 32 scattered 16-byte pieces (row-per-lane
        // red.global.add.v4 measured 6200 cycles per tile).  Only a __syncwarp between the two halves.
        const int rows_here = min(QT, a.Tq - i0);
        const int wlo = QT - rows_here, whi = QT - 1 + tk_c;  // window rows [wlo, whi) can be non-zero
        const uint32_t patch = s_stg + (uint32_t)warp * 4096u;
        for (int m = 0; m < nmt; ++m) {
          const int w0 = m * QT + q * 32;  // first window row of this warp's slice
          if (w0 >= whi || w0 + 32 <= wlo) continue;  // warp-uniform: nothing of the band in this slice
          uint32_t o[32];
          tmem_ld32(tlane + m * DK + hs * 32, o);
          tmem_ld_wait();
          __syncwarp();  // the previous slice has been read out of the patch
#pragma unroll
          for (int g = 0; g < 8; ++g)
            sts128(patch + (uint32_t)lane * 128u + ((((uint32_t)g) ^ (uint32_t)(lane & 7)) << 4), o[g * 4], o[g * 4 + 1], o[g * 4 + 2],
                   o[g * 4 + 3]);
          __syncwarp();
          const int ch = lane & 7;
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int rr = it * 4 + (lane >> 3), w = w0 + rr, n = c0 + w;
            float f0, f1, f2, f3;
            lds128(patch + (uint32_t)rr * 128u + ((((uint32_t)ch) ^ (uint32_t)(rr & 7)) << 4), f0, f1, f2, f3);
            if (w >= wlo && w < whi && n >= 0 && n < 2 * a.Tq - 1)
              atomicAdd(reinterpret_cast<float4*>(a.dpos + (long long)n * a.lddpos + h * DK + hs * 32 + ch * 4),
                        make_float4(f0, f1, f2, f3));
          }
        }
      }
      if (!REL && qt == nqt - 1) drain_dkv();
      ++n_q;
      tcgen05_fence_before();
      fence_async_smem();  // this tile's generic writes to the p / dS tiles are ordered before the next TMA loads into them
      group_sync(0);
      BWD_STAMP(6)
    }
  }
  if (rec) {
    for (int t = 0; t < 7; ++t) a.dbg[t] = acc_t[t];
    a.dbg[7] = (long long)n_q;
  }
#undef BWD_STAMP
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

// ----------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
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

// bf16 [n2, n1, cols] (row stride ld, slab stride s2, elements); box {64, box_rows, 1}; SWIZZLE_128B; zero fill out of range
int make_map3(CUtensorMap* m, const void* base, long long cols, long long n1, long long ld, long long n2, long long s2,
              int box_rows) {
  EncodeTiledFn enc = get_encode();
  if (enc == nullptr) {
    b200slu_set_error("cuTensorMapEncodeTiled driver entry point unavailable");
    return B200SLU_ECUDA;
  }
  if (n2 <= 1) {
    n2 = 1;
    s2 = ld * n1;
  }
  cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)n1, (cuuint64_t)n2};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)s2 * 2};
  cuuint32_t box[3] = {64, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    b200slu_set_error("relattn: cuTensorMapEncodeTiled failed (%d): cols=%lld n1=%lld ld=%lld n2=%lld s2=%lld box_rows=%d base=%p",
                      (int)r, cols, n1, ld, n2, s2, box_rows, base);
    return B200SLU_ECUDA;
  }
  return 0;
}

}  // namespace

extern "C" int b200slu_relattn_supported(int Tq, int Tk, int dk, int relpos_mode) {
  if (dk != DK || Tq < 1 || Tk < 1 || Tk > MAX_TK) return 0;
  if (relpos_mode == B200SLU_RELPOS_LATEST) return Tq == Tk;
  return relpos_mode == B200SLU_RELPOS_NONE;
}

extern "C" int b200slu_relattn_chunk_supported(int Tq, int Tk_chunk, int dk, int relpos_mode) {
  return dk == DK && Tq >= 1 && Tk_chunk >= 1 && Tk_chunk <= MAX_TK &&
         (relpos_mode == B200SLU_RELPOS_LATEST || relpos_mode == B200SLU_RELPOS_NONE);
}

// chunk form: keys [key_off, key_off + Tk) of kv_rows keys per utterance; k, v, p, p_drop point at key / column key_off
static int relattn_fwd_impl(const void* qu, const void* qv, long long ldq, const void* k, const void* v, long long ldkv,
                            const void* pos, long long ldpos, const int* key_lens, int B, int H, int Tq, int Tk,
                            int relpos_mode, int causal, float scale, void* p, void* p_drop, long long ld_p, void* ctx,
                            long long ldctx, float drop_p, unsigned long long seed, unsigned int site, int key_off, int kv_rows,
                            float* lse2, void* stream) {
  B2_CHECK_ARG(B > 0 && H > 0, "relattn_fwd: empty problem");
  B2_CHECK_ARG(b200slu_relattn_chunk_supported(Tq, Tk, DK, relpos_mode) && key_off >= 0 && key_off % 16 == 0 &&
                   kv_rows >= key_off + Tk && ld_p >= key_off + Tk,
               "relattn_fwd: supports d_k = 64, <= %d keys per launch, rel-pos mode latest or none, key_off %% 16 == 0 (got Tq=%d "
               "Tk=%d mode=%d key_off=%d kv_rows=%d)", MAX_TK, Tq, Tk, relpos_mode, key_off, kv_rows);
  const bool rel = relpos_mode == B200SLU_RELPOS_LATEST;
  B2_CHECK_ARG(qu != nullptr && k != nullptr && v != nullptr && p != nullptr && ctx != nullptr, "relattn_fwd: null operand");
  B2_CHECK_ARG(!rel || (qv != nullptr && pos != nullptr), "relattn_fwd: rel-pos mode needs q_v and the projected positions");
  B2_CHECK_ARG(drop_p > 0.f || p_drop == p || p_drop == nullptr, "relattn_fwd: without dropout p_drop must alias p (or be NULL)");
  B2_CHECK_ARG(ldq % 8 == 0 && ldkv % 8 == 0 && ldctx % 8 == 0 && ld_p % 8 == 0 && ld_p >= Tk && (!rel || ldpos % 8 == 0),
               "relattn_fwd: row strides must be multiples of 8 elements, ld_p >= Tk");
  B2_CHECK_ARG(relpos_mode != B200SLU_RELPOS_LATEST || kv_rows == Tq, "relattn_fwd: rel-pos mode attends over the queries' own axis");
  B2_CHECK_ARG(((uintptr_t)qu % 16) == 0 && ((uintptr_t)k % 16) == 0 && ((uintptr_t)v % 16) == 0 && ((uintptr_t)ctx % 16) == 0 &&
                   ((uintptr_t)p % 32) == 0 && ((uintptr_t)p_drop % 32) == 0 && (!rel || (((uintptr_t)qv % 16) == 0 && ((uintptr_t)pos % 16) == 0)),
               "relattn_fwd: operands must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  AttnArgs a{};
  a.B = B; a.H = H; a.Tq = Tq; a.Tk = Tk;
  a.TKP = (Tk + 15) / 16 * 16;
  a.WIN = a.TKP + QT;
  a.rel = rel; a.causal = causal;
  a.scale_log2e = scale * 1.4426950408889634f;
  a.lens = key_lens;
  a.p = (bf16*)p; a.pd = (p_drop != p) ? (bf16*)p_drop : nullptr; a.ctx = (bf16*)ctx;
  a.ld_p = ld_p; a.ldctx = ldctx;
  a.key_off = key_off; a.p_cols = (int)(ld_p - key_off); a.lse2 = lse2;
  a.off_k = 2 * QT * 128;
  a.off_p = a.off_k + a.TKP * 128;
  a.off_v = a.off_p + (rel ? a.WIN * 128 : 0);
  const uint32_t recycled = rel ? (uint32_t)(NTHREADS * RING_LD * 4) : 0u;            // scratch rows of pass 1
  const uint32_t pm = (uint32_t)((a.TKP + 63) / 64) * (QT * 128);                      // A operand of P.V
  const uint32_t need = (recycled > pm ? recycled : pm);
  if (a.off_v < need) a.off_v = (need + 1023u) / 1024u * 1024u;
  a.off_bar = a.off_v + a.TKP * 128;
  const int cols = rel ? a.TKP + a.WIN : a.TKP + DK;
  a.tmem_cols = cols <= 32 ? 32 : cols <= 64 ? 64 : cols <= 128 ? 128 : cols <= 256 ? 256 : 512;
  a.drop = b2_make_drop(drop_p, seed, site);
  const size_t smem = (size_t)a.off_bar + 64 + 4 * QT * sizeof(float);

  CUtensorMap mqu, mqv, mk, mv, mp;
  const int d = H * DK;
  B2_TRY(make_map3(&mqu, qu, d, Tq, ldq, B, (long long)Tq * ldq, QT));
  B2_TRY(make_map3(&mk, k, d, Tk, ldkv, B, (long long)kv_rows * ldkv, a.TKP));
  B2_TRY(make_map3(&mv, v, d, Tk, ldkv, B, (long long)kv_rows * ldkv, a.TKP));
  if (rel) {
    B2_TRY(make_map3(&mqv, qv, d, Tq, ldq, B, (long long)Tq * ldq, QT));
    // projected positions [2T-1, d], shared by all utterances: rank 2 (encoded as rank 3 with one slab)
    B2_TRY(make_map3(&mp, pos, d, 2 * Tq - 1, ldpos, 1, 0, a.WIN / 2));
  } else {
    mqv = mqu;
    mp = mk;
  }
  static unsigned long long configured = 0;  // per device
  if (!b2_device_done(configured)) {
    B2_CUDA(cudaFuncSetAttribute(relattn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    B2_CUDA(cudaFuncSetAttribute(relattn_fwd_pipe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    b2_device_mark(configured);
  }
  const int nqt = (Tq + QT - 1) / QT;
  const double tiles = (double)B * H;
  B2_SCOPE(FAM_ATTN, st, tiles * (4.0 * Tq * Tk * DK + (rel ? 2.0 * Tq * (2.0 * Tq - 1) * DK : 0.0)),
           tiles * (2.0 * DK * (2.0 * Tq + 2.0 * Tk) + 2.0 * Tq * Tk * (drop_p > 0.f ? 2.0 : 1.0)));
  static const bool dbg = getenv("B200SLU_ATTN_DEBUG") != nullptr;
  long long* dbg_buf = nullptr;
  if (dbg) {
    dbg_buf = b2_debug_scratch();
    B2_CHECK_ARG(dbg_buf != nullptr, "relattn: debug scratch unavailable");
    B2_CUDA(cudaMemsetAsync(dbg_buf, 0, 512 * 8 * sizeof(long long), st));
    a.dbg = dbg_buf;
  }
  static const int force_pipe = getenv("B200SLU_ATTN_PIPE") ? atoi(getenv("B200SLU_ATTN_PIPE")) : -1;  // tuning: 0 / 1
  const size_t smem_pipe = 2 * (size_t)a.off_bar + PIPE_CTL_BYTES;
  const int num_tiles = B * H * nqt;
  // plain mode needs <= 256 columns: two one-tile CTAs already run side by side on an SM (measured 74 vs 78 us)
  if (!dbg && force_pipe != 0 && (rel || force_pipe == 1) && smem_pipe <= 232448 && (!rel || a.TKP + a.WIN <= PIPE_O_COL)) {
    const int sms = b2_num_sms();
    const int grid = (num_tiles + 1) / 2 < sms ? (num_tiles + 1) / 2 : sms;
    B2_CUDA(b2_launch(relattn_fwd_pipe_kernel, dim3((unsigned)grid), dim3(2 * NTHREADS), smem_pipe, st, mqu, mqv, mk, mv, mp, a,
                      num_tiles));
    return 0;
  }
  B2_CUDA(b2_launch(relattn_fwd_kernel, dim3((unsigned)(B * H * nqt)), dim3(NTHREADS), smem, st, mqu, mqv, mk, mv, mp, a));
  if (dbg) {
    static long long h[512 * 8];
    B2_CUDA(cudaStreamSynchronize(st));
    B2_CUDA(cudaMemcpy(h, dbg_buf, sizeof(h), cudaMemcpyDeviceToHost));
    double acc[7] = {0};
    int n = 0;
    for (int c = 0; c < 512 && c < B * H * nqt; ++c, ++n)
      for (int t = 0; t < 7; ++t) acc[t] += (double)(h[c * 8 + t + 1] - h[c * 8 + t]);
    fprintf(stderr, "[relattn dbg] Tq=%d Tk=%d rel=%d smem=%zu tmem=%u | mean cycles over %d CTAs: alloc+sync %.0f, loads+score MMA %.0f, pass1 %.0f, "
            "pass2 %.0f, pass3 %.0f, PV MMA %.0f, epilogue %.0f\n", Tq, Tk, (int)rel, smem, a.tmem_cols, n, acc[0] / n, acc[1] / n, acc[2] / n,
            acc[3] / n, acc[4] / n, acc[5] / n, acc[6] / n);
  }
  return 0;
}

extern "C" int b200slu_relattn_fwd(const void* qu, const void* qv, long long ldq, const void* k, const void* v, long long ldkv,
                                   const void* pos, long long ldpos, const int* key_lens, int B, int H, int Tq, int Tk,
                                   int relpos_mode, int causal, float scale, void* p, void* p_drop, long long ld_p, void* ctx,
                                   long long ldctx, float drop_p, unsigned long long seed, unsigned int site, void* stream) {
  B2_CHECK_ARG(b200slu_relattn_supported(Tq, Tk, DK, relpos_mode),
               "relattn_fwd: supports d_k = 64, keys <= %d, rel-pos mode latest (Tq == Tk) or none (got Tq=%d Tk=%d mode=%d)",
               MAX_TK, Tq, Tk, relpos_mode);
  return relattn_fwd_impl(qu, qv, ldq, k, v, ldkv, pos, ldpos, key_lens, B, H, Tq, Tk, relpos_mode, causal, scale, p, p_drop, ld_p,
                          ctx, ldctx, drop_p, seed, site, 0, Tk, nullptr, stream);
}

extern "C" int b200slu_relattn_fwd_chunk(const void* qu, const void* qv, long long ldq, const void* k, const void* v,
                                         long long ldkv, const void* pos, long long ldpos, const int* key_lens, int B, int H, int Tq,
                                         int Tk_chunk, int relpos_mode, int causal, float scale, void* p, long long ld_p,
                                         void* ctx_chunk, long long ldctx, float drop_p, unsigned long long seed, unsigned int site,
                                         int key_off, int kv_rows, float* lse2, void* stream) {
  B2_CHECK_ARG(lse2 != nullptr, "relattn_fwd_chunk: lse2 output required");
  return relattn_fwd_impl(qu, qv, ldq, k, v, ldkv, pos, ldpos, key_lens, B, H, Tq, Tk_chunk, relpos_mode, causal, scale, p, p, ld_p,
                          ctx_chunk, ldctx, drop_p, seed, site, key_off, kv_rows, lse2, stream);
}

extern "C" int b200slu_relattn_bwd_supported(int Tq, int Tk, int dk, int relpos_mode) {
  if (dk != DK || Tq < 1 || Tk < 1 || Tk > MAX_TK) return 0;
  if (relpos_mode == B200SLU_RELPOS_LATEST) return Tq == Tk;
  return relpos_mode == B200SLU_RELPOS_NONE;
}

static int relattn_bwd_impl(const void* d_ctx, long long lddo, const void* q_u, long long ldq, const void* k, const void* v,
                            long long ldkv, const void* p, long long ld_p, const void* ctx, long long ldctx, int B, int H,
                            int Tq, int Tk, int relpos_mode, float scale, void* dq_u, long long lddq, void* dk_out,
                            void* dv_out, long long lddkv, void* d_ac, float drop_p, unsigned long long seed,
                            unsigned int site, const void* q_v, const void* pos, long long ldpos, void* dq_v,
                            float* dpos_accum, long long lddpos, int chunk_keys, long long dq_slab, int key_begin,
                            int key_end, float* dk_colsum, float* dv_colsum, void* stream) {
  // Tk = all keys of an utterance; this launch covers keys [key_begin, key_end) in chunks of chunk_keys (one work unit
  // each; [0, Tk) in one chunk for the plain form)
  B2_CHECK_ARG(B > 0 && H > 0, "relattn_bwd: empty problem");
  const int nchunk = chunk_keys > 0 && key_begin >= 0 && key_begin < key_end && key_end <= Tk
                         ? (key_end - key_begin + chunk_keys - 1) / chunk_keys : 0;
  B2_CHECK_ARG(chunk_keys >= 1 && nchunk >= 1 && b200slu_relattn_chunk_supported(Tq, chunk_keys, DK, relpos_mode) &&
                   ((nchunk == 1 && key_begin == 0 && key_end == Tk) || (chunk_keys % 64 == 0 && key_begin % 64 == 0 && d_ac == nullptr && dq_slab % 8 == 0)) &&
                   ld_p >= Tk,
               "relattn_bwd: supports d_k = 64, <= %d keys per chunk (whole 64-key blocks when there are several), rel-pos mode latest "
               "or none (got Tq=%d Tk=%d mode=%d chunk_keys=%d)", MAX_TK, Tq, Tk, relpos_mode, chunk_keys);
  B2_CHECK_ARG(relpos_mode != B200SLU_RELPOS_LATEST || Tk == Tq, "relattn_bwd: rel-pos mode attends over the queries' own axis");
  const bool rel = relpos_mode == B200SLU_RELPOS_LATEST;
  B2_CHECK_ARG(d_ctx && q_u && k && v && p && ctx && dq_u && dk_out && dv_out, "relattn_bwd: null operand");
  B2_CHECK_ARG(!rel || (q_v && pos && dq_v && dpos_accum), "relattn_bwd: rel-pos mode needs q_v, pos, dq_v and dpos_accum");
  B2_CHECK_ARG(lddo % 8 == 0 && ldq % 8 == 0 && ldkv % 8 == 0 && ldctx % 8 == 0 && lddq % 8 == 0 && lddkv % 8 == 0 && ld_p % 8 == 0 &&
                   ld_p >= Tk && (!rel || (ldpos % 8 == 0 && lddpos % 4 == 0)),
               "relattn_bwd: row strides must be multiples of 8 elements, ld_p >= Tk");
  B2_CHECK_ARG(((uintptr_t)d_ctx % 16) == 0 && ((uintptr_t)q_u % 16) == 0 && ((uintptr_t)k % 16) == 0 && ((uintptr_t)v % 16) == 0 &&
                   ((uintptr_t)p % 16) == 0 && ((uintptr_t)ctx % 16) == 0 && ((uintptr_t)dq_u % 16) == 0 &&
                   ((uintptr_t)dk_out % 16) == 0 && ((uintptr_t)dv_out % 16) == 0 && ((uintptr_t)d_ac % 32) == 0 &&
                   (!rel || (((uintptr_t)q_v % 16) == 0 && ((uintptr_t)pos % 16) == 0 && ((uintptr_t)dq_v % 16) == 0 &&
                             ((uintptr_t)dpos_accum % 16) == 0)),
               "relattn_bwd: operands must be 16-byte aligned (d_ac: 32)");
  cudaStream_t st = (cudaStream_t)stream;
  AttnBwdArgs a{};
  a.B = B; a.H = H; a.Tq = Tq; a.Tk = Tk;
  a.TKP = (chunk_keys + 15) / 16 * 16;
  a.WIN = a.TKP + QT;
  a.scale = scale;
  a.ctx = (const bf16*)ctx; a.ldctx = ldctx;
  a.dq = (bf16*)dq_u; a.dk = (bf16*)dk_out; a.dv = (bf16*)dv_out; a.dac = (bf16*)d_ac;
  a.lddq = lddq; a.lddkv = lddkv; a.ld_p = ld_p;
  a.dqv = (bf16*)dq_v; a.dpos = dpos_accum; a.lddpos = lddpos;
  a.nchunk = nchunk; a.chunk_keys = chunk_keys; a.kv_rows = Tk; a.dq_slab = dq_slab; a.key_begin = key_begin; a.key_end = key_end;
  B2_CHECK_ARG((dk_colsum == nullptr) == (dv_colsum == nullptr), "relattn_bwd: dk_colsum and dv_colsum come together");
  a.dk_colsum = dk_colsum; a.dv_colsum = dv_colsum;
  // [pd tile | dO | Q_u | V] [dS tile] [K, later Q_v] [position window].  The first group is dead once dQ_u / dK / dV are
  // issued: the skewed d(bd) tile is built over it (dS, its source, stays).  An M = 128 MN-major read over the last key /
  // window tile may run past the blocks that exist into the next region: never faulting (bounded below), never stored.
  const uint32_t nblk = (uint32_t)(a.TKP + 63) / 64, vpad = (uint32_t)(a.TKP * 128 + 1023) / 1024 * 1024;
  a.off_pd = 0;
  a.off_do = nblk * BLK;
  a.off_qu = a.off_do + BLK;
  a.off_v = a.off_qu + BLK;
  a.off_ds = a.off_v + vpad;
  a.off_k = a.off_ds + nblk * BLK;
  a.off_pos = a.off_k + (rel ? (vpad > (uint32_t)BLK ? vpad : (uint32_t)BLK) : vpad);
  a.off_stg = a.off_pos + (rel ? (uint32_t)(a.WIN * 128 + 1023) / 1024 * 1024 : 0u);
  // d(pos) staging (8 warps x 4 KB): the dS tile is dead by then and, from three key blocks on, large enough to host it
  a.off_bar = a.off_stg;
  if (rel) {
    if (nblk >= 3) a.off_stg = a.off_ds;
    else a.off_bar = a.off_stg + 2 * BLK;
  }
  a.drop = b2_make_drop(drop_p, seed, site);
  size_t smem = (size_t)a.off_bar + 64 + 2 * QT * sizeof(float);
  const size_t nkt = (size_t)(a.TKP + QT - 1) / QT, nmt = (size_t)(a.WIN + QT - 1) / QT;
  size_t reach = (size_t)a.off_ds + nkt * 2 * BLK;  // what the widest MN-major reads touch
  if (rel && nmt * 2 * BLK > reach) reach = nmt * 2 * BLK;
  if (smem < reach) smem = reach;
  B2_CHECK_ARG(smem <= 232448, "relattn_bwd: shared memory %zu", smem);
  B2_CHECK_ARG(!rel || (size_t)((a.WIN + 63) / 64) * BLK <= a.off_ds, "relattn_bwd: skewed tile overlaps its source");

  CUtensorMap mdo, mqu, mk, mv, mp, mqv, mpos;
  const int d = H * DK;
  B2_TRY(make_map3(&mdo, d_ctx, d, Tq, lddo, B, (long long)Tq * lddo, QT));
  B2_TRY(make_map3(&mqu, q_u, d, Tq, ldq, B, (long long)Tq * ldq, QT));
  B2_TRY(make_map3(&mk, k, d, Tk, ldkv, B, (long long)Tk * ldkv, a.TKP));
  B2_TRY(make_map3(&mv, v, d, Tk, ldkv, B, (long long)Tk * ldkv, a.TKP));
  B2_TRY(make_map3(&mp, p, ld_p, Tq, ld_p, (long long)H * B, (long long)Tq * ld_p, QT));
  if (rel) {
    B2_TRY(make_map3(&mqv, q_v, d, Tq, ldq, B, (long long)Tq * ldq, QT));
    B2_TRY(make_map3(&mpos, pos, d, 2 * Tq - 1, ldpos, 1, 0, a.WIN / 2));
  } else {
    mqv = mqu;
    mpos = mk;
  }
  static unsigned long long configured = 0;  // per device
  if (!b2_device_done(configured)) {
    B2_CUDA(cudaFuncSetAttribute(relattn_bwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    B2_CUDA(cudaFuncSetAttribute(relattn_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    b2_device_mark(configured);
  }
  const int units = B * H * nchunk, sms = b2_num_sms();
  B2_SCOPE(FAM_ATTN, st, (double)B * H * (8.0 * Tq * Tk * DK + (rel ? 4.0 * Tq * (2.0 * Tq - 1) * DK : 0.0)),
           (double)B * H * (nchunk * 2.0 * DK * 3.0 * Tq + 2.0 * DK * 2.0 * Tk + 2.0 * Tq * Tk + 2.0 * DK * (nchunk * Tq + 2.0 * Tk) +
                            (rel ? 2.0 * DK * 2.0 * Tq * nchunk : 0.0) + (d_ac ? 2.0 * Tq * Tk : 0.0)));
  const dim3 grid((unsigned)(units < sms ? units : sms));
  static const bool dbg = getenv("B200SLU_ATTN_BWD_DEBUG") != nullptr && atoi(getenv("B200SLU_ATTN_BWD_DEBUG")) != 0;
  if (dbg) {
    a.dbg = b2_debug_scratch();
    B2_CHECK_ARG(a.dbg != nullptr, "relattn_bwd: debug scratch unavailable");
  }
  static const int dbgflags = getenv("B200SLU_ATTN_BWD_FLAGS") ? atoi(getenv("B200SLU_ATTN_BWD_FLAGS")) : 0;
  a.dbgflags = dbgflags;
  if (rel) B2_CUDA(b2_launch(relattn_bwd_kernel<true>, grid, dim3(NTHREADS), smem, st, mdo, mqu, mk, mv, mp, mqv, mpos, a, units));
  else B2_CUDA(b2_launch(relattn_bwd_kernel<false>, grid, dim3(NTHREADS), smem, st, mdo, mqu, mk, mv, mp, mqv, mpos, a, units));
  if (dbg) {
    long long hbuf[8];
    B2_CUDA(cudaStreamSynchronize(st));
    B2_CUDA(cudaMemcpy(hbuf, a.dbg, sizeof(hbuf), cudaMemcpyDeviceToHost));
    const double n = hbuf[7] > 0 ? (double)hbuf[7] : 1.0;
    fprintf(stderr, "[relattn_bwd dbg] Tq=%d Tk=%d rel=%d smem=%zu | CTA 0, %lld query tiles, mean cycles per tile: loads %.0f, dP MMA %.0f, "
            "dS pass %.0f, dQu/dK/dV MMA %.0f, dQu drain + skew %.0f, dQv/dPos MMA %.0f, drains %.0f\n", Tq, Tk, (int)rel, smem, hbuf[7],
            hbuf[0] / n, hbuf[1] / n, hbuf[2] / n, hbuf[3] / n, hbuf[4] / n, hbuf[5] / n, hbuf[6] / n);
  }
  return 0;
}

extern "C" int b200slu_relattn_bwd(const void* d_ctx, long long lddo, const void* q_u, long long ldq, const void* k, const void* v,
                                   long long ldkv, const void* p, long long ld_p, const void* ctx, long long ldctx, int B, int H,
                                   int Tq, int Tk, int relpos_mode, float scale, void* dq_u, long long lddq, void* dk_out,
                                   void* dv_out, long long lddkv, void* d_ac, float drop_p, unsigned long long seed,
                                   unsigned int site, const void* q_v, const void* pos, long long ldpos, void* dq_v,
                                   float* dpos_accum, long long lddpos, void* stream) {
  B2_CHECK_ARG(b200slu_relattn_bwd_supported(Tq, Tk, DK, relpos_mode),
               "relattn_bwd: supports d_k = 64, keys <= %d, rel-pos mode latest (Tq == Tk) or none (got Tq=%d Tk=%d mode=%d)", MAX_TK,
               Tq, Tk, relpos_mode);
  return relattn_bwd_impl(d_ctx, lddo, q_u, ldq, k, v, ldkv, p, ld_p, ctx, ldctx, B, H, Tq, Tk, relpos_mode, scale, dq_u, lddq, dk_out,
                          dv_out, lddkv, d_ac, drop_p, seed, site, q_v, pos, ldpos, dq_v, dpos_accum, lddpos, Tk, 0, 0, Tk, nullptr, nullptr, stream);
}

extern "C" int b200slu_relattn_bwd_chunks(const void* d_ctx, long long lddo, const void* q_u, long long ldq, const void* k,
                                          const void* v, long long ldkv, const void* p, long long ld_p, const void* ctx,
                                          long long ldctx, int B, int H, int Tq, int Tk, int chunk_keys, int relpos_mode,
                                          float scale, void* dq_u_slabs, long long lddq, long long dq_slab, void* dk_out,
                                          void* dv_out, long long lddkv, float drop_p, unsigned long long seed, unsigned int site,
                                          const void* q_v, const void* pos, long long ldpos, void* dq_v_slabs, float* dpos_accum,
                                          long long lddpos, int key_begin, int key_end, float* dk_colsum_accum,
                                          float* dv_colsum_accum, void* stream) {
  return relattn_bwd_impl(d_ctx, lddo, q_u, ldq, k, v, ldkv, p, ld_p, ctx, ldctx, B, H, Tq, Tk, relpos_mode, scale, dq_u_slabs, lddq,
                          dk_out, dv_out, lddkv, nullptr, drop_p, seed, site, q_v, pos, ldpos, dq_v_slabs, dpos_accum, lddpos,
                          chunk_keys, dq_slab, key_begin, key_end, dk_colsum_accum, dv_colsum_accum, stream);
}

// ----------------------------------------------------------------------------- key chunks -> one softmax
namespace {

constexpr int MAX_CHUNKS = 16;

// One warp per (head, utterance, query) row.  alpha_c = 2^(lse2_c - LSE2) with LSE2 = log2 sum_c 2^lse2_c rescales what each
// chunk launch normalised over its own keys: p[row, chunk c's columns] *= alpha_c in place (bf16), ctx = sum_c alpha_c ctx_c.
__global__ void __launch_bounds__(256) relattn_combine_kernel(bf16* __restrict__ p, long long ld_p, int Tk, int chunk_keys, int nchunk,
                                                              const float* __restrict__ lse2, long long rows,
                                                              const bf16* __restrict__ ctx_chunks, long long chunk_stride,
                                                              long long ldctx, bf16* __restrict__ ctx, long long ldout, int B, int H,
                                                              int Tq) {
  const int lane = threadIdx.x & 31;
  for (long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows;
       row += (long long)gridDim.x * (blockDim.x >> 5)) {
    float al[MAX_CHUNKS];
    float mx = -INFINITY;
#pragma unroll
    for (int c = 0; c < MAX_CHUNKS; ++c) {
      al[c] = c < nchunk ? lse2[(long long)c * rows + row] : -INFINITY;
      mx = fmaxf(mx, al[c]);
    }
    float tot = 0.f;
#pragma unroll
    for (int c = 0; c < MAX_CHUNKS; ++c) {
      al[c] = (al[c] == -INFINITY) ? 0.f : exp2f(al[c] - mx);
      tot += al[c];
    }
    const float inv = tot > 0.f ? 1.f / tot : 0.f;
#pragma unroll
    for (int c = 0; c < MAX_CHUNKS; ++c) al[c] *= inv;
    // p row: groups of 8 columns; a group never straddles a chunk (chunk_keys % 8 == 0)
    bf16* pr = p + row * ld_p;
    for (int g = lane; g * 8 < Tk; g += 32) {
      const int c = (g * 8) / chunk_keys;
      float v[8], a = 0.f;
#pragma unroll
      for (int cc = 0; cc < MAX_CHUNKS; ++cc) a = cc == c ? al[cc] : a;
      load8_bf16(pr + g * 8, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] *= a;
      store8_bf16(pr + g * 8, v);
    }
    // context: row = (h, b, i) -> [b * Tq + i, h * 64 .. +64), 8 lanes x 8 channels
    const long long hb = row / Tq;
    const int i = (int)(row - hb * Tq), h = (int)(hb / B), b = (int)(hb - (long long)h * B);
    if (lane < 8) {
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      const long long off = ((long long)b * Tq + i) * ldctx + h * 64 + lane * 8;
#pragma unroll
      for (int c = 0; c < MAX_CHUNKS; ++c) {
        if (c < nchunk && al[c] != 0.f) {
          float v[8];
          load8_bf16(ctx_chunks + (long long)c * chunk_stride + off, v);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] += al[c] * v[j];
        }
      }
      store8_bf16(ctx + ((long long)b * Tq + i) * ldout + h * 64 + lane * 8, acc);
    }
  }
}

// dst[r, :] = sum_c src[c][r, :] (fp32 sum, bf16 in and out), rows of `cols` columns (cols % 8 == 0)
__global__ void __launch_bounds__(256) sum_chunks_kernel(const bf16* __restrict__ src, long long chunk_stride, long long ld_src,
                                                         int nchunk, long long rows, int cols, bf16* __restrict__ dst,
                                                         long long ld_dst) {
  const int gpr = cols >> 3;
  const long long total = rows * gpr;
  for (long long it = (long long)blockIdx.x * blockDim.x + threadIdx.x; it < total; it += (long long)gridDim.x * blockDim.x) {
    const long long r = it / gpr;
    const int g = (int)(it - r * gpr);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int c = 0; c < nchunk; ++c) {
      float v[8];
      load8_bf16(src + (long long)c * chunk_stride + r * ld_src + g * 8, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += v[j];
    }
    store8_bf16(dst + r * ld_dst + g * 8, acc);
  }
}

}  // namespace

extern "C" int b200slu_relattn_combine(void* p, long long ld_p, int Tk, int chunk_keys, int nchunk, const float* lse2,
                                       const void* ctx_chunks, long long chunk_stride, long long ldctx, void* ctx, long long ldout,
                                       int B, int H, int Tq, void* stream) {
  B2_CHECK_ARG(B > 0 && H > 0 && Tq > 0 && Tk > 0 && nchunk >= 1 && nchunk <= MAX_CHUNKS && chunk_keys % 8 == 0 &&
                   (long long)nchunk * chunk_keys >= Tk && ld_p % 8 == 0 && ldctx % 8 == 0 && ldout % 8 == 0 && chunk_stride % 8 == 0,
               "relattn_combine: 1..%d chunks of a multiple of 8 keys covering Tk, strides multiples of 8 elements", MAX_CHUNKS);
  B2_CHECK_ARG(((uintptr_t)p % 16) == 0 && ((uintptr_t)ctx_chunks % 16) == 0 && ((uintptr_t)ctx % 16) == 0, "relattn_combine: alignment");
  cudaStream_t st = (cudaStream_t)stream;
  const long long rows = (long long)H * B * Tq;
  B2_SCOPE(FAM_ATTN, st, 0, (double)rows * (4.0 * Tk + 4.0 * nchunk + 2.0 * 64 * (nchunk + 1)));
  relattn_combine_kernel<<<b2_grid_for(rows * 32, 256, 8), 256, 0, st>>>((bf16*)p, ld_p, Tk, chunk_keys, nchunk, lse2, rows,
                                                                         (const bf16*)ctx_chunks, chunk_stride, ldctx, (bf16*)ctx,
                                                                         ldout, B, H, Tq);
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_sum_chunks_bf16(const void* src, long long chunk_stride, long long ld_src, int nchunk, long long rows,
                                       int cols, void* dst, long long ld_dst, void* stream) {
  B2_CHECK_ARG(nchunk >= 1 && rows > 0 && cols > 0 && cols % 8 == 0 && ld_src % 8 == 0 && ld_dst % 8 == 0 && chunk_stride % 8 == 0 &&
                   ((uintptr_t)src % 16) == 0 && ((uintptr_t)dst % 16) == 0,
               "sum_chunks: cols / strides multiples of 8 elements, 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  B2_SCOPE(FAM_ELEMWISE, st, 0, 2.0 * rows * cols * (nchunk + 1));
  sum_chunks_kernel<<<b2_grid_for(rows * (cols >> 3), 256, 8), 256, 0, st>>>((const bf16*)src, chunk_stride, ld_src, nchunk, rows, cols,
                                                                             (bf16*)dst, ld_dst);
  B2_LAUNCH_CHECK();
  return 0;
}
