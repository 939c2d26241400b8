// Runtime glue of libb200slu: error string, device properties, per-family launch accounting.
#include <stdarg.h>
#include <string.h>

#include <mutex>
#include <vector>

#include <stdlib.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

void b200slu_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int b2_current_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev;
}

// One registration slot per device ordinal: the pointer is device memory of THAT device (two devices driven from one
// process each keep their own; a kernel launched on device d only ever sees d's slot).
static const B2StepParams* g_step_params[64] = {};
const B2StepParams* b2_step_params() { return g_step_params[b2_current_device() & 63]; }
extern "C" int b200slu_set_step_params(const void* dev_ptr) {
  g_step_params[b2_current_device() & 63] = (const B2StepParams*)dev_ptr;
  return 0;
}

// SMs the persistent kernels size their grids for.  b200slu_set_sm_budget(n) leaves the others to whatever else runs on
// the device (the data-parallel trainer: NCCL's CTAs during the overlapped all-reduce -- a persistent grid of all 148 SMs
// would have its last CTAs wait for NCCL's to leave, and, being statically partitioned, end that much later).
static int g_sm_budget[64] = {};
int b2_num_sms() {
  static int n[64] = {};
  const int dev = b2_current_device();
  int& v = n[dev & 63];
  if (v == 0) {
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
  }
  const int lim = g_sm_budget[dev & 63];
  return (lim > 0 && lim < v) ? lim : v;
}

extern "C" int b200slu_set_sm_budget(int sms) {
  B2_CHECK_ARG(sms >= 0, "set_sm_budget: sms >= 0 (0 = all)");
  g_sm_budget[b2_current_device() & 63] = sms > 0 ? (sms < 8 ? 8 : (sms & ~1)) : 0;  // even: CTA pairs
  return 0;
}

static __device__ __align__(256) long long g_debug_scratch[4096];
long long* b2_debug_scratch() {
  static long long* ptrs[64] = {};
  long long*& p = ptrs[b2_current_device() & 63];
  if (p == nullptr && cudaGetSymbolAddress((void**)&p, g_debug_scratch) != cudaSuccess) p = nullptr;
  return p;
}

namespace {
struct FamStat {
  long long launches = 0;
  double flops = 0, bytes = 0, ms = 0;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pending;
};
FamStat g_fam[FAM_COUNT];
bool g_timing = false;
std::mutex g_mu;
const char* kFamNames[FAM_COUNT] = {"gemm_tcgen05", "attention", "layernorm", "convmodule", "subsample",
                                    "ctc", "frontend", "loss", "elementwise"};

void drain(FamStat& f) {
  for (auto& p : f.pending) {
    float ms = 0;
    cudaEventSynchronize(p.second);
    cudaEventElapsedTime(&ms, p.first, p.second);
    f.ms += ms;
    cudaEventDestroy(p.first);
    cudaEventDestroy(p.second);
  }
  f.pending.clear();
}
}  // namespace

B2ProfScope::B2ProfScope(int fam_, cudaStream_t st_, double flops, double bytes) : fam(fam_), st(st_), timed(false) {
  std::lock_guard<std::mutex> lk(g_mu);
  FamStat& f = g_fam[fam];
  f.launches++;
  f.flops += flops;
  f.bytes += bytes;
  if (g_timing) {
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(st, &cs);
    if (cs == cudaStreamCaptureStatusNone) {
      cudaEventCreate(&e0);
      cudaEventCreate(&e1);
      cudaEventRecord(e0, st);
      timed = true;
    }
  }
}
B2ProfScope::~B2ProfScope() {
  if (timed) {
    cudaEventRecord(e1, st);
    std::lock_guard<std::mutex> lk(g_mu);
    g_fam[fam].pending.emplace_back(e0, e1);
  }
}

extern "C" {
int b200slu_version(void) { return B200SLU_VERSION; }
const char* b200slu_last_error(void) { return g_err; }
void b200slu_prof_reset(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  for (auto& f : g_fam) {
    drain(f);
    f = FamStat();
  }
}
void b200slu_prof_enable_timing(int on) { g_timing = on != 0; }
int b200slu_prof_num_families(void) { return FAM_COUNT; }
const char* b200slu_prof_family_name(int fam) { return (fam >= 0 && fam < FAM_COUNT) ? kFamNames[fam] : ""; }
int b200slu_prof_get(int fam, double* out4) {
  if (fam < 0 || fam >= FAM_COUNT || out4 == nullptr) return B200SLU_EINVAL;
  std::lock_guard<std::mutex> lk(g_mu);
  FamStat& f = g_fam[fam];
  drain(f);
  out4[0] = (double)f.launches;
  out4[1] = f.ms;
  out4[2] = f.flops;
  out4[3] = f.bytes;
  return 0;
}
long long b200slu_prof_total_launches(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  long long n = 0;
  for (auto& f : g_fam) n += f.launches;
  return n;
}
}

// rel_shift index maps as pure integer functions (attention.py:240-263 latest, :145-165 legacy); the
// same arithmetic softmax.cu applies on the device.  Used by the bit-exact index tests.
extern "C" long long b200slu_rel_shift_index(int i, int j, int T, int mode) {
  if (i < 0 || j < 0 || i >= T || j >= T) return -1;
  if (mode != B200SLU_RELPOS_LATEST && mode != B200SLU_RELPOS_LEGACY) return -1;
  const B2ShiftSrc src = b2_rel_shift_src(i, j, T, mode);  // the function the kernels call (common.cuh)
  return src.row < 0 ? 0 : (long long)src.row * 65536 + src.col + 1;
}

bool b2_pdl_enabled() {
  static const bool on = !(getenv("B200SLU_PDL") != nullptr && atoi(getenv("B200SLU_PDL")) == 0);
  return on;
}
