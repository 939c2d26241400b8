// Common definitions for libb200slu (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/b200slu.h"

typedef __nv_bfloat16 bf16;

// ---------------------------------------------------------------------------
// error plumbing: never throw / abort across the C boundary
// ---------------------------------------------------------------------------
void b200slu_set_error(const char* fmt, ...);

#define B2_CHECK_ARG(cond, ...)                       \
  do {                                                \
    if (!(cond)) {                                    \
      b200slu_set_error(__VA_ARGS__);                 \
      return B200SLU_EINVAL;                          \
    }                                                 \
  } while (0)

#define B2_CUDA(call)                                                                  \
  do {                                                                                 \
    cudaError_t e_ = (call);                                                           \
    if (e_ != cudaSuccess) {                                                           \
      b200slu_set_error("%s:%d CUDA error: %s", __FILE__, __LINE__, cudaGetErrorString(e_)); \
      return B200SLU_ECUDA;                                                            \
    }                                                                                  \
  } while (0)

#define B2_LAUNCH_CHECK()                                                              \
  do {                                                                                 \
    cudaError_t e_ = cudaGetLastError();                                               \
    if (e_ != cudaSuccess) {                                                           \
      b200slu_set_error("%s:%d launch error: %s", __FILE__, __LINE__, cudaGetErrorString(e_)); \
      return B200SLU_ECUDA;                                                            \
    }                                                                                  \
  } while (0)

#define B2_TRY(expr)            \
  do {                          \
    int rc_ = (expr);           \
    if (rc_ != 0) return rc_;   \
  } while (0)

// ---------------------------------------------------------------------------
// per-family launch accounting + optional CUDA-event timing (bench.py roofline)
// ---------------------------------------------------------------------------
enum B2Family {
  FAM_GEMM = 0,
  FAM_ATTN,
  FAM_LN,
  FAM_CONVMOD,
  FAM_SUBSAMPLE,
  FAM_CTC,
  FAM_FRONTEND,
  FAM_LOSS,
  FAM_ELEMWISE,
  FAM_COUNT
};

struct B2ProfScope {
  int fam;
  cudaStream_t st;
  cudaEvent_t e0, e1;
  bool timed;
  B2ProfScope(int fam, cudaStream_t st, double flops, double bytes);
  ~B2ProfScope();
};
#define B2_SCOPE(fam, st, flops, bytes) B2ProfScope prof_scope_((fam), (st), (double)(flops), (double)(bytes))

int b2_num_sms();
// Per-device one-time set-up (cudaFuncSetAttribute, symbol addresses): `mask` holds one bit per device ordinal.  Two
// threads racing through the same set-up both run it (harmless); the bit is set only after the set-up succeeded.
int b2_current_device();
static inline bool b2_device_done(const unsigned long long& mask) { return (mask >> (b2_current_device() & 63)) & 1ull; }
static inline void b2_device_mark(unsigned long long& mask) { mask |= 1ull << (b2_current_device() & 63); }
// Debug scratch (tuning printouts only): a static device array, so that even the debug paths never allocate.
long long* b2_debug_scratch();  // 4096 long longs per device, or nullptr on error

// ---------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// sigmoid through ONE MUFU op (tanh.approx, rel. error 2^-11, far below the bf16 storage rounding of every consumer):
// the exp + reciprocal form made the Swish epilogues of the FFN GEMMs MUFU-bound.
__device__ __forceinline__ float sigmoidf_(float x) {
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.5f * x));
  return fmaf(0.5f, t, 0.5f);
}
__device__ __forceinline__ float swishf_(float x) { return x * sigmoidf_(x); }
// d/dx [x*sigmoid(x)] = s + x*s*(1-s)
__device__ __forceinline__ float swish_gradf_(float x) {
  float s = sigmoidf_(x);
  return s * (1.f + x * (1.f - s));
}

// two independent IEEE fp32 FMAs in one instruction (FFMA2): d = a * b + c per component
__device__ __forceinline__ float2 b2_fma2(float2 a, float2 b, float2 c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;"
      : "=l"(d)
      : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)),
        "l"(*reinterpret_cast<unsigned long long*>(&c)));
  return *reinterpret_cast<float2*>(&d);
}

// ... and the packed add / multiply (FADD2 / FMUL2): same rounding as the scalar instructions, half the issue slots
__device__ __forceinline__ float2 b2_add2(float2 a, float2 b) {
  unsigned long long d;
  asm("add.rn.f32x2 %0, %1, %2;"
      : "=l"(d)
      : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float2 b2_mul2(float2 a, float2 b) {
  unsigned long long d;
  asm("mul.rn.f32x2 %0, %1, %2;"
      : "=l"(d)
      : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
// sigmoid of a pair: 0.5 * tanh(0.5 x) + 0.5 per component (sigmoidf_ above), two MUFU.TANH and two packed instructions
__device__ __forceinline__ float2 b2_sigmoid2(float2 x) {
  const float2 half = make_float2(0.5f, 0.5f);
  const float2 h = b2_mul2(x, half);
  float2 t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t.x) : "f"(h.x));
  asm("tanh.approx.f32 %0, %1;" : "=f"(t.y) : "f"(h.y));
  return b2_fma2(t, half, half);
}

// d/dx [x * sigmoid(x)] of a pair: s * (1 + x * (1 - s))   (swish_gradf_ above)
__device__ __forceinline__ float2 b2_swish_grad2(float2 x) {
  const float2 one = make_float2(1.f, 1.f), mone = make_float2(-1.f, -1.f);
  const float2 s = b2_sigmoid2(x);
  return b2_mul2(s, b2_fma2(x, b2_fma2(s, mone, one), one));
}

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}

// ---------------------------------------------------------------------------
// rel_shift as an index map (attention.py:240-263 latest, :145-165 legacy): element (i, j) of the shifted T x T matrix
// is element (row, col) of the un-shifted matrix_bd; row < 0 = the literal zero the legacy variant inserts at j = i + 1.
// ONE definition for the host index entry (b200slu_rel_shift_index), softmax.cu (forward gather and the base column
// of the backward scatter) and attention.cu (base of the position window a query tile can reach).
// ---------------------------------------------------------------------------
struct B2ShiftSrc {
  int row, col;
};
__host__ __device__ __forceinline__ B2ShiftSrc b2_rel_shift_src(int i, int j, int T, int mode) {
  if (mode == B200SLU_RELPOS_LATEST) return B2ShiftSrc{i, T - 1 - i + j};
  if (j <= i) return B2ShiftSrc{i, T - 1 - i + j};  // legacy, lower triangle: same map as latest
  if (j == i + 1) return B2ShiftSrc{-1, 0};
  return B2ShiftSrc{i + 1, j - i - 2};
}

// ---------------------------------------------------------------------------
// counter-based dropout: splitmix64 keyed by (seed, site); two evaluations yield
// 128 random bits = 8 x 16-bit lanes -> keep decisions for 8 consecutive
// elements (element index = 8*group + lane).  Recomputed, never stored.
// ---------------------------------------------------------------------------
struct B2Dropout {
  unsigned long long seed;  // per training step
  unsigned int site;        // unique per dropout site
  float p;                  // drop probability; 0 => disabled
  const unsigned long long* dev_seed;  // optional device-resident per-step seed, XORed in (CUDA-graph replays)
};

// Device-resident per-step parameters registered with b200slu_set_step_params (runtime.cu): lets a captured CUDA
// graph of the whole step be replayed with a fresh dropout seed / learning rate by updating 24 bytes.
struct B2StepParams {
  unsigned long long seed;
  float lr, bc1, bc2, reserved;
};
const B2StepParams* b2_step_params();
static inline B2Dropout b2_make_drop(float p, unsigned long long seed, unsigned int site) {
  const B2StepParams* sp = b2_step_params();
  return B2Dropout{seed, site, p, sp ? &sp->seed : nullptr};
}

// splitmix64 finaliser: 64 well-mixed bits from a 64-bit counter (2 multiplies; ~3x cheaper than Philox4x32-10, which
// made the dropout epilogues of the FFN GEMMs ALU-bound -- profiles/r1_gemm_shapes_v0.txt)
__device__ __forceinline__ unsigned long long splitmix64(unsigned long long z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// Per-site dropout key, hoisted out of the element loops: one splitmix of (seed, site) per thread per kernel.
struct B2DropKey {
  unsigned long long key;
  uint32_t thr;  // drop when the 16-bit uniform lane < thr = p * 65536
  float inv;     // 1 / (1 - p)
};
__device__ __forceinline__ B2DropKey dropout_key(const B2Dropout& d) {
  B2DropKey k;
  const unsigned long long sd = d.dev_seed != nullptr ? (d.seed ^ *d.dev_seed) : d.seed;
  k.key = splitmix64(sd + 0x9E3779B97F4A7C15ull * (unsigned long long)(d.site + 1u));
  k.thr = (uint32_t)(d.p * 65536.f);
  k.inv = 1.f / (1.f - d.p);
  return k;
}
// 8 x 16-bit uniform lanes for elements [8*group, 8*group+8): ONE splitmix64 of the group counter plus one more
// multiply-xorshift round for the second 64 bits (two full splitmix64 per group were ~40 % of the instructions of the
// FFN w_1 forward epilogue, which is issue-bound).  Every kernel derives its decisions from these lanes, so forward and
// backward of a site agree whichever form (bit mask / multipliers) they use.
__device__ __forceinline__ void dropout_lanes8(const B2DropKey& k, unsigned long long group, uint32_t (&w)[4]) {
  const unsigned long long z0 = splitmix64(k.key + group * 0x9E3779B97F4A7C15ull);
  unsigned long long z1 = (z0 ^ (z0 >> 29)) * 0xD6E8FEB86659FD93ull;
  z1 ^= z1 >> 32;
  w[0] = (uint32_t)z0;
  w[1] = (uint32_t)(z0 >> 32);
  w[2] = (uint32_t)z1;
  w[3] = (uint32_t)(z1 >> 32);
}
// keep-mask (bit i set = keep): lane >= p * 65536
__device__ __forceinline__ uint32_t dropout_keep8(const B2DropKey& k, unsigned long long group) {
  uint32_t w[4];
  dropout_lanes8(k, group, w);
  const uint32_t thr = k.thr;
  uint32_t m = 0;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    m |= ((w[q] & 0xffffu) >= thr) << (2 * q);
    m |= ((w[q] >> 16) >= thr) << (2 * q + 1);
  }
  return m;
}
// the same decisions as multipliers: mul[i] = keep ? scale : 0
__device__ __forceinline__ void dropout_mul8(const B2DropKey& k, unsigned long long group, float scale, float (&mul)[8]) {
  uint32_t w[4];
  dropout_lanes8(k, group, w);
  const uint32_t thr = k.thr;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    mul[2 * q] = (w[q] & 0xffffu) >= thr ? scale : 0.f;
    mul[2 * q + 1] = (w[q] >> 16) >= thr ? scale : 0.f;
  }
}
__device__ __forceinline__ uint32_t dropout_keep8(const B2Dropout& d, unsigned long long group) {
  return dropout_keep8(dropout_key(d), group);
}

// ---------------------------------------------------------------------------
// 8-wide vector helpers (all memory-bound kernels work on groups of 8 contiguous elements:
// one uint4 of bf16 or two float4 of f32; one Philox call yields the 8 dropout decisions).
// ---------------------------------------------------------------------------
__device__ __forceinline__ void load8_f32(const float* p, float (&v)[8]) {
  float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void store8_f32(float* p, const float (&v)[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ void load8_bf16(const bf16* p, float (&v)[8]) {
  uint4 u = *reinterpret_cast<const uint4*>(p);
  float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
}
__device__ __forceinline__ void unpack8_bf16(const uint4& u, float (&v)[8]) {
  float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
}
__device__ __forceinline__ void store8_bf16(bf16* p, const float (&v)[8]) {
  *reinterpret_cast<uint4*>(p) = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                                            pack_bf16x2(v[6], v[7]));
}
// dropout on 8 consecutive elements whose first flat index is e0 (multiple of 8)
__device__ __forceinline__ void dropout8(const B2Dropout& d, unsigned long long e0, float (&v)[8]) {
  if (d.p > 0.f) {
    const uint32_t keep = dropout_keep8(d, e0 >> 3);
    const float inv = 1.f / (1.f - d.p);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = ((keep >> j) & 1u) ? v[j] * inv : 0.f;
  }
}
__device__ __forceinline__ float block_sum(float v, float* red /* >= 32 floats of smem */) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  v = (threadIdx.x < nw) ? red[threadIdx.x] : 0.f;
  if (w == 0) v = warp_sum(v);
  if (threadIdx.x == 0) red[0] = v;
  __syncthreads();
  return red[0];
}
// ---------------------------------------------------------------------------
// Programmatic dependent launch: a train step is ~1150 short kernels in one stream, so the launch-to-launch gap is a
// real share of the step.  Every kernel launched through b2_launch() (a) lets its successor start launching as soon as
// all of its own CTAs are running (griddepcontrol.launch_dependents at the top) and (b) does its local set-up, then
// blocks in griddepcontrol.wait until its predecessor has completed and flushed -- before its first global access.
// Ordering and memory visibility along the stream are unchanged; only launch latency and prologues overlap.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void b2_pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void b2_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void b2_pdl_entry() {
  b2_pdl_launch_dependents();
  b2_pdl_wait();
}
bool b2_pdl_enabled();  // runtime.cu: B200SLU_PDL=0 turns the launch attribute off (the device-side instructions are no-ops then)
template <typename... KArgs, typename... Args>
static inline cudaError_t b2_launch(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = b2_pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// ---------------------------------------------------------------------------
// Column-sum kernels (elementwise.cu, convmod.cu).  Same-address atomics from different SMs retire at one per ~70 ns (measured: +5.2 us for 74
// more blocks per column), so the number of blocks that contribute to one output column is what the kernel's tail
// costs.  A block is therefore a NARROW strip -- 8 column groups of 8 (64 columns: one 128-byte segment of a bf16
// row per 8 lanes, four rows per warp-wide load) x 64 row lanes -- which makes grid.x = N / 64 wide and grid.y
// (blocks per column) ~ SMs / grid.x small.  Every thread keeps CS_U independent 16-byte row loads in flight
// (512 threads x 8 x 16 B = 64 KB per block).  The block reduces its row lanes through shared memory and issues ONE
// atomic per column.
constexpr int CS_TX = 8;                       // column groups of 8 per block
constexpr int CS_ROWS = 512 / CS_TX;           // row lanes per block
constexpr int CS_COLS = CS_TX * 8;             // columns per block
constexpr int CS_U = 8;                        // rows in flight per thread

__device__ __forceinline__ void colsum_block_reduce(float (&acc)[8], float (*red)[CS_COLS + 1], int tx, int ty, int N,
                                                    float* __restrict__ out) {
#pragma unroll
  for (int j = 0; j < 8; ++j) red[ty][tx * 8 + j] = acc[j];
  __syncthreads();
  // 512 threads: 64 columns x 8 partial sums over 8 row lanes each, then a shuffle over the 8 partials
  const int col = threadIdx.x >> 3, part = threadIdx.x & 7;
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < CS_ROWS / 8; ++k) s += red[part * (CS_ROWS / 8) + k][col];
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  s += __shfl_xor_sync(0xffffffffu, s, 4);
  // a warp holds four consecutive columns (lanes 0, 8, 16, 24): one 16-byte vector atomic instead of four scalar ones
  const int lane = threadIdx.x & 31;
  const float s1 = __shfl_sync(0xffffffffu, s, 8), s2 = __shfl_sync(0xffffffffu, s, 16), s3 = __shfl_sync(0xffffffffu, s, 24);
  const int c = blockIdx.x * CS_COLS + col, c0 = c - (lane >> 3);  // c0: the warp's first column
  const bool vec = c0 + 3 < N && (((uintptr_t)(out + c0)) & 15) == 0;
  if (vec) {
    if (lane == 0) atomicAdd(reinterpret_cast<float4*>(out + c0), make_float4(s, s1, s2, s3));
  } else if (part == 0 && c < N) {
    atomicAdd(out + c, s);
  }
}

// gemm_sm100.cu: the GEMM behind b200slu_gemm_bf16_ex, for kernels families that contain one (subsample.cu)
int b2_gemm_ex(const b200slu_gemm_t* g, const b200slu_epilogue_t* e, cudaStream_t st);

static inline int b2_grid_for(long long work_items, int threads, int per_sm = 8) {
  long long blocks = (work_items + threads - 1) / threads;
  long long cap = (long long)b2_num_sms() * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}
