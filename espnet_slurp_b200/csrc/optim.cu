// Fused optimizer step over the flat parameter / gradient buffers: global L2 grad-norm, clip, non-finite
// skip, Adam (+ L2 weight decay as torch.optim.Adam applies it) -- SURVEY.md 8(f) row N1.
//
// Replaces espnet2/train/trainer.py:640-690: torch.nn.utils.clip_grad_norm_(max_norm=grad_clip, norm_type=2),
// "grad norm is not finite -> skip the update" (:651-667), optimizer.step() (torch.optim.Adam, registry
// espnet2/tasks/abs_task.py:856-880).  The learning rate for the step comes from the host (WarmupLR,
// espnet2/schedulers/warmup_lr.py:43-50).  Two launches: a grid-wide sum of squares into one double, then one
// pass that reads g, p, m, v and writes p, m, v: 28 bytes per parameter, memory-bound.
#include <math.h>

#include "common.cuh"

namespace {

__global__ void __launch_bounds__(256) sumsq_kernel(const float* __restrict__ g, long long n, float inv_scale,
                                                    double* __restrict__ out) {
  __shared__ float red[32];
  float s = 0.f;
  const long long n4 = n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = reinterpret_cast<const float4*>(g)[i];
    const float a = v.x * inv_scale, b = v.y * inv_scale, c = v.z * inv_scale, d = v.w * inv_scale;
    s += a * a + b * b + c * c + d * d;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const float a = g[(n4 << 2) + threadIdx.x] * inv_scale;
    s += a * a;
  }
  s = block_sum(s, red);
  if (threadIdx.x == 0) atomicAdd(out, (double)s);
}

// state[0] = sum of squares (in), state[1] = grad norm (out), state[2] = 1 if the update was applied (out)
__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                   float* __restrict__ v, long long n, float lr, float beta1, float beta2,
                                                   float eps, float weight_decay, float bc1, float bc2, float inv_scale,
                                                   float max_norm, double* __restrict__ state,
                                                   const B2StepParams* __restrict__ sp, float warmup, bf16* __restrict__ p16) {
  if (sp != nullptr) {  // per-step values live on the device (CUDA-graph replay): lr, 1 - beta1^t, 1 - beta2^t
    lr = sp->lr;
    bc1 = sp->bc1;
    bc2 = sp->bc2;
  }
  if (warmup > 0.f) {
    // scheduled form: t = applied updates so far + 1 lives on the device (state[3]); a skipped (non-finite) step advances
    // neither Adam's step count nor WarmupLR, like the reference (trainer.py:651-690, warmup_lr.py:43-50: lr is the
    // `lr` argument here = base_lr)
    const double t = state[3] + 1.0;
    lr = (float)((double)lr * sqrt((double)warmup) * fmin(1.0 / sqrt(t), t / ((double)warmup * sqrt((double)warmup))));
    bc1 = (float)(1.0 - pow((double)beta1, t));
    bc2 = (float)(1.0 - pow((double)beta2, t));
  }
  const double ss = state[0];
  const float norm = (float)sqrt(ss);
  const bool ok = isfinite(norm);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    state[1] = (double)norm;
    state[2] = ok ? 1.0 : 0.0;
  }
  if (!ok) return;
  float coef = inv_scale;
  if (max_norm > 0.f) coef *= fminf(1.f, max_norm / (norm + 1e-6f));
  const float step_size = lr / bc1;
  const float inv_sqrt_bc2 = rsqrtf(bc2);
  auto upd = [&](float gi, float& pi, float& mi, float& vi) {
    gi *= coef;
    if (weight_decay != 0.f) gi += weight_decay * pi;
    mi = beta1 * mi + (1.f - beta1) * gi;
    vi = beta2 * vi + (1.f - beta2) * gi * gi;
    pi = pi - step_size * mi / (sqrtf(vi) * inv_sqrt_bc2 + eps);
  };
  // 16-byte accesses over the aligned body (the flat buffers are; 28 bytes move per parameter and nothing else happens
  // here, so this kernel is as fast as its loads and stores are wide), scalar tail
  const bool vec = ((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0);
  const long long n4 = vec ? (n >> 2) : 0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 G = reinterpret_cast<const float4*>(g)[i];
    float4 P = reinterpret_cast<float4*>(p)[i], Mo = reinterpret_cast<float4*>(m)[i], V = reinterpret_cast<float4*>(v)[i];
    upd(G.x, P.x, Mo.x, V.x);
    upd(G.y, P.y, Mo.y, V.y);
    upd(G.z, P.z, Mo.z, V.z);
    upd(G.w, P.w, Mo.w, V.w);
    reinterpret_cast<float4*>(m)[i] = Mo;
    reinterpret_cast<float4*>(v)[i] = V;
    reinterpret_cast<float4*>(p)[i] = P;
    // the GEMM-operand copy of the new parameters, written here instead of by a cast pass over all of them next step
    if (p16 != nullptr) reinterpret_cast<uint2*>(p16)[i] = make_uint2(pack_bf16x2(P.x, P.y), pack_bf16x2(P.z, P.w));
  }
  for (long long i = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float pi = p[i], mi = m[i], vi = v[i];
    upd(g[i], pi, mi, vi);
    m[i] = mi;
    v[i] = vi;
    p[i] = pi;
    if (p16 != nullptr) p16[i] = __float2bfloat16(pi);
  }
}

// after the update: count it (state[3]) when it was applied
__global__ void adam_advance_kernel(double* __restrict__ state) {
  if (state[2] == 1.0) state[3] += 1.0;
}

// dst[i] += sum_k src[k * stride + i] (fp32, fixed order k = 0..nshards-1): the local reduction of the peer-memory
// gradient exchange (dist.PeerReducer) -- the shards the copy engines pulled from the other ranks are added onto this
// rank's own slice.  Streaming, 16-byte accesses; runs beside the backward on the communication stream, so the grid is
// kept to a fraction of the SMs.
__global__ void __launch_bounds__(256) add_shards_kernel(float* __restrict__ dst, const float* __restrict__ src, long long stride,
                                                         int nshards, long long n) {
  const long long n4 = n >> 2, step = (long long)gridDim.x * 256;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += step) {
    float4 a = reinterpret_cast<const float4*>(dst)[i];
    for (int k = 0; k < nshards; ++k) {
      const float4 b = __ldcs(reinterpret_cast<const float4*>(src + k * stride) + i);
      a.x += b.x;
      a.y += b.y;
      a.z += b.z;
      a.w += b.w;
    }
    reinterpret_cast<float4*>(dst)[i] = a;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {  // ragged tail
    const long long i = (n4 << 2) + threadIdx.x;
    float a = dst[i];
    for (int k = 0; k < nshards; ++k) a += src[k * stride + i];
    dst[i] = a;
  }
}

}  // namespace

// state: 3 doubles on the device (zeroed here).  grad_inv_scale multiplies every gradient first (e.g. 1/accum_grad).
// step >= 1: lr and the bias corrections come from the arguments; step == 0: from the registered device step params.
extern "C" int b200slu_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n, float lr,
                                 float beta1, float beta2, float eps, float weight_decay, int step, float grad_inv_scale,
                                 float max_norm, double* state3, void* stream) {
  B2_CHECK_ARG(n > 0 && step >= 0 && ((uintptr_t)grads % 16) == 0, "adam_step: n > 0, step >= 0, 16-byte aligned gradients");
  B2_CHECK_ARG(step >= 1 || b2_step_params() != nullptr, "adam_step: step == 0 needs b200slu_set_step_params");
  cudaStream_t st = (cudaStream_t)stream;
  B2_CUDA(cudaMemsetAsync(state3, 0, 3 * sizeof(double), st));
  B2_SCOPE(FAM_ELEMWISE, st, 0, 32.0 * n);
  sumsq_kernel<<<b2_grid_for(n >> 2, 256, 4), 256, 0, st>>>(grads, n, grad_inv_scale, state3);
  B2_LAUNCH_CHECK();
  const float bc1 = step >= 1 ? 1.f - powf(beta1, (float)step) : 1.f, bc2 = step >= 1 ? 1.f - powf(beta2, (float)step) : 1.f;
  adam_kernel<<<b2_grid_for((n + 3) >> 2, 256, 8), 256, 0, st>>>(params, grads, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, weight_decay,
                                                      bc1, bc2, grad_inv_scale, max_norm, state3,
                                                      step >= 1 ? nullptr : b2_step_params(), 0.f, nullptr);
  B2_LAUNCH_CHECK();
  return 0;
}

// Scheduled form: the step count lives on the device.  state4 = {sum of squares, grad norm, applied flag (as above; zeroed
// here), number of APPLIED updates so far (persistent: the caller zeroes it once)}.  lr_t = base_lr * warmup^0.5 *
// min(t^-0.5, t * warmup^-1.5) and the bias corrections use t = state4[3] + 1; a non-finite gradient norm skips the update
// AND leaves t unchanged.  Nothing depends on a host value, so the call is capturable in a CUDA graph as is.
static int adam_step_sched_impl(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n, float base_lr,
                                float warmup_steps, float beta1, float beta2, float eps, float weight_decay, float grad_inv_scale,
                                float max_norm, double* state4, void* params_bf16, void* stream) {
  B2_CHECK_ARG(n > 0 && warmup_steps > 0.f && ((uintptr_t)grads % 16) == 0, "adam_step_sched: n > 0, warmup_steps > 0, aligned grads");
  B2_CHECK_ARG(((uintptr_t)params_bf16 % 8) == 0, "adam_step_sched: the bf16 copy must be 8-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  B2_CUDA(cudaMemsetAsync(state4, 0, 3 * sizeof(double), st));
  B2_SCOPE(FAM_ELEMWISE, st, 0, 32.0 * n);
  sumsq_kernel<<<b2_grid_for(n >> 2, 256, 4), 256, 0, st>>>(grads, n, grad_inv_scale, state4);
  B2_LAUNCH_CHECK();
  adam_kernel<<<b2_grid_for((n + 3) >> 2, 256, 8), 256, 0, st>>>(params, grads, exp_avg, exp_avg_sq, n, base_lr, beta1, beta2, eps,
                                                      weight_decay, 1.f, 1.f, grad_inv_scale, max_norm, state4, nullptr, warmup_steps,
                                                      (bf16*)params_bf16);
  B2_LAUNCH_CHECK();
  adam_advance_kernel<<<1, 1, 0, st>>>(state4);
  B2_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200slu_adam_step_sched(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n,
                                       float base_lr, float warmup_steps, float beta1, float beta2, float eps, float weight_decay,
                                       float grad_inv_scale, float max_norm, double* state4, void* stream) {
  return adam_step_sched_impl(params, grads, exp_avg, exp_avg_sq, n, base_lr, warmup_steps, beta1, beta2, eps, weight_decay,
                              grad_inv_scale, max_norm, state4, nullptr, stream);
}

// ... and the bf16 copy of the updated parameters written in the same pass (what the next step's GEMMs read): saves the
// cast pass over all parameters at the start of every step.  A skipped update leaves both copies as they were.
extern "C" int b200slu_adam_step_sched_lp(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n,
                                          float base_lr, float warmup_steps, float beta1, float beta2, float eps,
                                          float weight_decay, float grad_inv_scale, float max_norm, double* state4,
                                          void* params_bf16, void* stream) {
  B2_CHECK_ARG(params_bf16 != nullptr, "adam_step_sched_lp: no bf16 destination");
  return adam_step_sched_impl(params, grads, exp_avg, exp_avg_sq, n, base_lr, warmup_steps, beta1, beta2, eps, weight_decay,
                              grad_inv_scale, max_norm, state4, params_bf16, stream);
}

extern "C" int b200slu_add_shards_f32(float* dst, const float* src, long long shard_stride, int nshards, long long n, void* stream) {
  B2_CHECK_ARG(n > 0 && nshards >= 1 && shard_stride >= n && shard_stride % 4 == 0 && ((uintptr_t)dst % 16) == 0 &&
                   ((uintptr_t)src % 16) == 0,
               "add_shards: n > 0, shards of stride >= n (multiple of 4 floats), 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  B2_SCOPE(FAM_ELEMWISE, st, 0, 4.0 * n * (nshards + 2));
  int grid = b2_grid_for(n >> 2, 256, 4);
  if (grid > 64) grid = 64;
  if (grid < 1) grid = 1;
  add_shards_kernel<<<grid, 256, 0, st>>>(dst, src, shard_stride, nshards, n);
  B2_LAUNCH_CHECK();
  return 0;
}
