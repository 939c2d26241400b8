// Small memory-bound helpers around the GEMMs: fp32->bf16 cast with scale/dropout, bias-gradient
// column sums, the rel-pos bias add (attention.py:289-291) and the decoder token embedding
// (transformer_decoder.py:72-77, embedding.py:82-94).  All are single-pass, 16-byte vectorised.
#include <stdlib.h>

#include "common.cuh"

namespace {

__global__ void __launch_bounds__(256) cast_kernel(const float* __restrict__ x, bf16* __restrict__ y, long long n, float alpha,
                                                   B2Dropout drop) {
  b2_pdl_entry();  // programmatic dependent launch: see common.cuh
  const long long ng = n >> 3;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < ng; g += (long long)gridDim.x * blockDim.x) {
    float v[8];
    load8_f32(x + g * 8, v);
    dropout8(drop, (unsigned long long)g * 8, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] *= alpha;
    store8_bf16(y + g * 8, v);
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 7)) {
    const long long i = (ng << 3) + threadIdx.x;
    y[i] = __float2bfloat16(x[i] * alpha);
  }
}

// y = bf16(alpha * dropout(x)) and colsum[n] += sum_m alpha * dropout(x)[m, n] in one pass over x (the backward of
// "x + dropout(branch)" feeds both the dgrad/wgrad GEMMs (y) and the bias gradient of the branch's last Linear).
__global__ void __launch_bounds__(512, 2) cast_colsum_kernel(const float* __restrict__ x, bf16* __restrict__ y, int M,
                                                                      int N, float alpha, B2Dropout drop,
                                                                      float* __restrict__ out) {
  b2_pdl_entry();  // programmatic dependent launch: see common.cuh
  __shared__ float red[CS_ROWS][CS_COLS + 1];
  const int tx = threadIdx.x & (CS_TX - 1), ty = threadIdx.x / CS_TX;
  const int g = blockIdx.x * CS_TX + tx;  // column group of 8 (N % 8 == 0)
  const bool ok = g * 8 < N;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (ok) {
    B2DropKey dk = {0ull, 0u, 1.f};
    if (drop.p > 0.f) dk = dropout_key(drop);
    constexpr int U = 3;  // fp32 rows are two 16-byte loads per thread (4 rows would spill at 64 registers)
    const long long step = (long long)gridDim.y * CS_ROWS;
    for (long long r0 = (long long)blockIdx.y * CS_ROWS + ty; r0 < M; r0 += U * step) {
      float v[U][8];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long r = r0 + u * step;
        if (r < M) load8_f32(x + r * N + g * 8, v[u]);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long r = r0 + u * step;
        if (r < M) {
          if (drop.p > 0.f) {
            const uint32_t keep = dropout_keep8(dk, (unsigned long long)(r * N + g * 8) >> 3);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[u][j] = ((keep >> j) & 1u) ? v[u][j] * dk.inv : 0.f;
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            v[u][j] *= alpha;
            acc[j] += v[u][j];
          }
          store8_bf16(y + r * N + g * 8, v[u]);
        }
      }
    }
  }
  colsum_block_reduce(acc, red, tx, ty, N, out);
}

// out[n] += sum_m dy[m, n]   (dy bf16, row stride ld)
__global__ void __launch_bounds__(512, 2) colsum_kernel(const bf16* __restrict__ dy, int M, int N, long long ld,
                                                                 float* __restrict__ out) {
  b2_pdl_entry();  // programmatic dependent launch: see common.cuh
  __shared__ float red[CS_ROWS][CS_COLS + 1];
  const int tx = threadIdx.x & (CS_TX - 1), ty = threadIdx.x / CS_TX;
  const int g = blockIdx.x * CS_TX + tx;  // column group of 8
  const bool ok = g * 8 < N;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (ok) {
    const long long step = (long long)gridDim.y * CS_ROWS;
    for (long long r0 = (long long)blockIdx.y * CS_ROWS + ty; r0 < M; r0 += CS_U * step) {
      uint4 raw[CS_U];
#pragma unroll
      for (int u = 0; u < CS_U; ++u) {
        const long long r = r0 + u * step;
        raw[u] = r < M ? *reinterpret_cast<const uint4*>(dy + r * ld + g * 8) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int u = 0; u < CS_U; ++u) {
        const float2 a = unpack_bf16x2(raw[u].x), b = unpack_bf16x2(raw[u].y), c = unpack_bf16x2(raw[u].z),
                     d = unpack_bf16x2(raw[u].w);
        acc[0] += a.x; acc[1] += a.y; acc[2] += b.x; acc[3] += b.y; acc[4] += c.x; acc[5] += c.y; acc[6] += d.x; acc[7] += d.y;
      }
    }
  }
  colsum_block_reduce(acc, red, tx, ty, N, out);
}

// q_u = q + pos_bias_u, q_v = q + pos_bias_v  (q = first d columns of the fused qkv rows)
__global__ void __launch_bounds__(256) pos_bias_fwd_kernel(const bf16* __restrict__ q, long long ldq, const float* __restrict__ u,
                                                           const float* __restrict__ v, long long M, int d,
                                                           bf16* __restrict__ qu, bf16* __restrict__ qv) {
  b2_pdl_entry();  // programmatic dependent launch: see common.cuh
  const int gpr = d >> 3;
  const long long total = M * gpr;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / gpr;
    const int c = (int)(i - row * gpr) * 8;
    float x[8], bu[8], bv[8], o[8];
    load8_bf16(q + row * ldq + c, x);
    load8_f32(u + c, bu);
    load8_f32(v + c, bv);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = x[j] + bu[j];
    store8_bf16(qu + row * d + c, o);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = x[j] + bv[j];
    store8_bf16(qv + row * d + c, o);
  }
}

// dq = dq_u + dq_v (written into the fused dqkv rows); du += colsum(dq_u); dv += colsum(dq_v)
__global__ void __launch_bounds__(512, 1) pos_bias_bwd_kernel(const bf16* __restrict__ dqu, const bf16* __restrict__ dqv, int M,
                                                              int d, bf16* __restrict__ dq, long long lddq,
                                                              float* __restrict__ du, float* __restrict__ dv,
                                                              float* __restrict__ dbq) {
  b2_pdl_entry();  // programmatic dependent launch: see common.cuh
  __shared__ float red[2][CS_ROWS][CS_COLS + 1];
  const int tx = threadIdx.x & (CS_TX - 1), ty = threadIdx.x / CS_TX;
  const int g = blockIdx.x * CS_TX + tx;
  float au[8] = {0, 0, 0, 0, 0, 0, 0, 0}, av[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (g * 8 < d) {
    constexpr int U = 4;
    const long long step = (long long)gridDim.y * CS_ROWS;
    for (long long r0 = (long long)blockIdx.y * CS_ROWS + ty; r0 < M; r0 += U * step) {
      uint4 ra[U], rb[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long r = r0 + u * step;
        ra[u] = r < M ? *reinterpret_cast<const uint4*>(dqu + r * d + g * 8) : make_uint4(0u, 0u, 0u, 0u);
        rb[u] = r < M ? *reinterpret_cast<const uint4*>(dqv + r * d + g * 8) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long r = r0 + u * step;
        const uint32_t wa[4] = {ra[u].x, ra[u].y, ra[u].z, ra[u].w}, wb[4] = {rb[u].x, rb[u].y, rb[u].z, rb[u].w};
        float o[8];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float2 a = unpack_bf16x2(wa[q]), b = unpack_bf16x2(wb[q]);
          au[2 * q] += a.x; au[2 * q + 1] += a.y;
          av[2 * q] += b.x; av[2 * q + 1] += b.y;
          o[2 * q] = a.x + b.x; o[2 * q + 1] = a.y + b.y;
        }
        if (r < M) store8_bf16(dq + r * lddq + g * 8, o);
      }
    }
  }
  colsum_block_reduce(au, red[0], tx, ty, d, du);
  colsum_block_reduce(av, red[1], tx, ty, d, dv);
  if (dbq != nullptr) {  // colsum(dq) = colsum(dq_u) + colsum(dq_v): the q part of the fused projection's bias gradient
#pragma unroll
    for (int j = 0; j < 8; ++j) au[j] += av[j];
    __syncthreads();  // red[0] has been read out by the first reduction
    colsum_block_reduce(au, red[0], tx, ty, d, dbq);
  }
}

// x[r, :] = dropout(emb[tok[r], :] * scale + pe[r % L, :])   (fp32 residual stream)
__global__ void __launch_bounds__(256) embed_fwd_kernel(const float* __restrict__ emb, const int* __restrict__ tok,
                                                        const float* __restrict__ pe, long long R, int L, int d, float scale,
                                                        float* __restrict__ x, B2Dropout drop) {
  const int gpr = d >> 3;
  const long long total = R * gpr;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / gpr;
    const int c = (int)(i - row * gpr) * 8;
    float e[8], p[8];
    load8_f32(emb + (long long)tok[row] * d + c, e);
    load8_f32(pe + (row % L) * d + c, p);
#pragma unroll
    for (int j = 0; j < 8; ++j) e[j] = e[j] * scale + p[j];
    dropout8(drop, (unsigned long long)row * d + c, e);
    store8_f32(x + row * d + c, e);
  }
}

// demb[tok[r], :] += scale * dropout(dx[r, :])
__global__ void __launch_bounds__(256) embed_bwd_kernel(const float* __restrict__ dx, const int* __restrict__ tok, long long R,
                                                        int d, float scale, float* __restrict__ demb, B2Dropout drop) {
  const int gpr = d >> 3;
  const long long total = R * gpr;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / gpr;
    const int c = (int)(i - row * gpr) * 8;
    float g[8];
    load8_f32(dx + row * d + c, g);
    dropout8(drop, (unsigned long long)row * d + c, g);
    float* dst = demb + (long long)tok[row] * d + c;
#pragma unroll
    for (int j = 0; j < 8; ++j) atomicAdd(dst + j, g[j] * scale);
  }
}

// keep[e] = 1 iff element e (flat index) of dropout site (seed, site) is kept: the decisions every kernel of the
// library derives from dropout_keep8 / dropout_mul8 / dropout_lanes8 (parity tests replay them in the CPU oracle)
__global__ void __launch_bounds__(256) dropout_mask_kernel(uint8_t* __restrict__ keep, long long n, B2Dropout drop) {
  const long long ng = (n + 7) >> 3;
  const B2DropKey dk = dropout_key(drop);
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < ng; g += (long long)gridDim.x * blockDim.x) {
    const uint32_t m = dropout_keep8(dk, (unsigned long long)g);
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (g * 8 + j < n) keep[g * 8 + j] = (uint8_t)((m >> j) & 1u);
  }
}

int colsum_blocks_per_sm() {
  static const int v = getenv("B200SLU_COLSUM_BPS") ? atoi(getenv("B200SLU_COLSUM_BPS")) : 2;  // tuning knob
  return v > 0 ? v : 2;
}

}  // namespace

extern "C" int b200slu_cast_bf16(const float* x, void* y, long long n, float alpha, float drop_p, unsigned long long seed,
                                 unsigned int site, void* stream) {
  B2_CHECK_ARG(n > 0 && ((uintptr_t)x % 16) == 0 && ((uintptr_t)y % 16) == 0, "cast_bf16: n > 0 and 16-byte aligned pointers");
  B2_CHECK_ARG(!(drop_p > 0.f) || n % 8 == 0, "cast_bf16: dropout needs n %% 8 == 0");
  cudaStream_t st = (cudaStream_t)stream;
  B2_SCOPE(FAM_ELEMWISE, st, 0, 6.0 * n);
  B2_CUDA(b2_launch(cast_kernel, dim3(b2_grid_for(n >> 3, 256)), dim3(256), 0, st, x, (bf16*)y, n, alpha, b2_make_drop(drop_p, seed, site)));
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_dropout_mask(float drop_p, unsigned long long seed, unsigned int site, long long n, unsigned char* keep,
                                    void* stream) {
  B2_CHECK_ARG(n > 0 && keep != nullptr && drop_p >= 0.f && drop_p < 1.f, "dropout_mask: n > 0, 0 <= p < 1");
  cudaStream_t st = (cudaStream_t)stream;
  B2_SCOPE(FAM_ELEMWISE, st, 0, 1.0 * n);
  dropout_mask_kernel<<<b2_grid_for((n + 7) >> 3, 256), 256, 0, st>>>(keep, n, b2_make_drop(drop_p, seed, site));
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_cast_colsum_bf16(const float* x, void* y, int M, int N, float alpha, float drop_p,
                                        unsigned long long seed, unsigned int site, float* colsum_accum, void* stream) {
  B2_CHECK_ARG(M > 0 && N > 0 && N % 8 == 0 && ((uintptr_t)x % 16) == 0 && ((uintptr_t)y % 16) == 0,
               "cast_colsum: N %% 8 == 0 and 16-byte aligned pointers (M=%d N=%d)", M, N);
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid((N + CS_COLS - 1) / CS_COLS, 1);
  const int gy = (M + CS_ROWS * 3 - 1) / (CS_ROWS * 3);
  const int cap = (colsum_blocks_per_sm() * b2_num_sms() + grid.x - 1) / grid.x;
  grid.y = gy < cap ? gy : cap;
  B2_SCOPE(FAM_ELEMWISE, st, 0, 6.0 * M * N);
  B2_CUDA(b2_launch(cast_colsum_kernel, grid, dim3(512), 0, st, x, (bf16*)y, M, N, alpha, b2_make_drop(drop_p, seed, site), colsum_accum));
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_colsum_bf16(const void* dy, int M, int N, long long ld, float* out_accum, void* stream) {
  B2_CHECK_ARG(M > 0 && N > 0 && ld % 8 == 0 && ld >= ((N + 7) / 8) * 8, "colsum: bad shape M=%d N=%d ld=%lld", M, N, ld);
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid((N + CS_COLS - 1) / CS_COLS, 1);
  int gy = (M + CS_ROWS * CS_U - 1) / (CS_ROWS * CS_U);
  const int cap = (colsum_blocks_per_sm() * b2_num_sms() + grid.x - 1) / grid.x;
  grid.y = gy < cap ? gy : cap;
  B2_SCOPE(FAM_ELEMWISE, st, 0, 2.0 * M * N);
  B2_CUDA(b2_launch(colsum_kernel, grid, dim3(512), 0, st, (const bf16*)dy, M, N, ld, out_accum));
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_pos_bias_fwd(const void* q, long long ldq, const float* pos_bias_u, const float* pos_bias_v, int M,
                                    int d, void* q_u, void* q_v, void* stream) {
  B2_CHECK_ARG(M > 0 && d % 8 == 0 && ldq % 8 == 0, "pos_bias_fwd: d and ldq must be multiples of 8");
  cudaStream_t st = (cudaStream_t)stream;
  B2_SCOPE(FAM_ELEMWISE, st, 0, 6.0 * M * d);
  B2_CUDA(b2_launch(pos_bias_fwd_kernel, dim3(b2_grid_for((long long)M * (d >> 3), 256)), dim3(256), 0, st, (const bf16*)q, ldq,
                    pos_bias_u, pos_bias_v, M, d, (bf16*)q_u, (bf16*)q_v));
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_pos_bias_bwd_q(const void* dq_u, const void* dq_v, int M, int d, void* dq, long long lddq, float* du_accum,
                                      float* dv_accum, float* dq_colsum_accum, void* stream) {
  B2_CHECK_ARG(M > 0 && d % 8 == 0 && lddq % 8 == 0, "pos_bias_bwd: d and lddq must be multiples of 8");
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid((d + CS_COLS - 1) / CS_COLS, 1);
  int gy = (M + CS_ROWS * 4 - 1) / (CS_ROWS * 4);
  const int cap = (b2_num_sms() + grid.x - 1) / grid.x;  // one 512-thread block per SM
  grid.y = gy < cap ? gy : cap;
  B2_SCOPE(FAM_ELEMWISE, st, 0, 6.0 * M * d);
  B2_CUDA(b2_launch(pos_bias_bwd_kernel, grid, dim3(512), 0, st, (const bf16*)dq_u, (const bf16*)dq_v, M, d, (bf16*)dq, lddq, du_accum,
                    dv_accum, dq_colsum_accum));
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_pos_bias_bwd(const void* dq_u, const void* dq_v, int M, int d, void* dq, long long lddq, float* du_accum,
                                    float* dv_accum, void* stream) {
  return b200slu_pos_bias_bwd_q(dq_u, dq_v, M, d, dq, lddq, du_accum, dv_accum, nullptr, stream);
}

extern "C" int b200slu_embed_fwd(const float* emb, const int* tokens, const float* pe, int R, int L, int d, float scale, float* x,
                                 float drop_p, unsigned long long seed, unsigned int site, void* stream) {
  B2_CHECK_ARG(R > 0 && L > 0 && d % 8 == 0, "embed_fwd: d must be a multiple of 8");
  cudaStream_t st = (cudaStream_t)stream;
  B2_SCOPE(FAM_ELEMWISE, st, 0, 12.0 * R * d);
  embed_fwd_kernel<<<b2_grid_for((long long)R * (d >> 3), 256), 256, 0, st>>>(emb, tokens, pe, R, L, d, scale, x,
                                                                              b2_make_drop(drop_p, seed, site));
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_embed_bwd(const float* dx, const int* tokens, int R, int d, float scale, float* demb_accum, float drop_p,
                                 unsigned long long seed, unsigned int site, void* stream) {
  B2_CHECK_ARG(R > 0 && d % 8 == 0, "embed_bwd: d must be a multiple of 8");
  cudaStream_t st = (cudaStream_t)stream;
  B2_SCOPE(FAM_ELEMWISE, st, 0, 12.0 * R * d);
  embed_bwd_kernel<<<b2_grid_for((long long)R * (d >> 3), 256), 256, 0, st>>>(dx, tokens, R, d, scale, demb_accum,
                                                                              b2_make_drop(drop_p, seed, site));
  B2_LAUNCH_CHECK();
  return 0;
}
