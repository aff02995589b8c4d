// Library plumbing: error string, launch counter, index contract (host only).
#include <math.h>
#include <stdarg.h>

#include <mutex>
#include <utility>
#include <vector>

#include <atomic>

#include "common.cuh"

namespace tmil {
static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};   // process-wide: autograd runs the backward on its own thread
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// ---- per-stage CUDA-event profiler (bench.py's roofline numbers) -----------
static const char* kStageNames[ST_COUNT] = {"fc1_gemm0", "fc1_ln", "fc1_gemm1", "fc1_tail", "ln_pad", "qkv", "landmarks",
                                            "pinv", "attn3v", "zw", "nystrom_out", "to_out", "ppeg", "head", "prep", "x_landmarks",
                                            "x_attn3v", "x_conv_halo", "x_ppeg_halo", "x_logits", "end"};
static bool g_prof_on = false;
static std::vector<std::pair<int, cudaEvent_t>> g_marks;
static std::vector<cudaEvent_t> g_pool;
void prof_mark(int stage, cudaStream_t st) {
  if (!g_prof_on) return;
  cudaEvent_t ev;
  if (!g_pool.empty()) { ev = g_pool.back(); g_pool.pop_back(); }
  else if (cudaEventCreate(&ev) != cudaSuccess) return;
  cudaEventRecord(ev, st);
  g_marks.emplace_back(stage, ev);
}
bool prof_enabled() { return g_prof_on; }
bool pdl_enabled() {
  static const bool on = [] { const char* e = getenv("TM_PDL"); return !(e && e[0] == '0'); }();
  return on;
}

static thread_local bool g_pdl_plain = false;
static thread_local bool g_pdl_suspended = false;
void pdl_plain_next() { g_pdl_plain = true; }
void pdl_suspend(bool on) { g_pdl_suspended = on; }
bool pdl_take_plain() { const bool v = g_pdl_plain; g_pdl_plain = false; return v || g_pdl_suspended; }

static std::mutex g_side_mutex;
static std::vector<std::pair<int, SideStream>> g_side;
int side_stream(SideStream* out) {
  int dev = 0;
  TM_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(g_side_mutex);
  for (auto& e : g_side)
    if (e.first == dev) { *out = e.second; return TM_OK; }
  SideStream s;
  int lo = 0, hi = 0;
  TM_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  TM_CUDA(cudaStreamCreateWithPriority(&s.stream, cudaStreamNonBlocking, hi));
  TM_CUDA(cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming));
  TM_CUDA(cudaEventCreateWithFlags(&s.join, cudaEventDisableTiming));
  g_side.emplace_back(dev, s);
  *out = s;
  return TM_OK;
}
}  // namespace tmil

extern "C" {

const char* tm_version(void) { return "transmil_b200 0.1 (sm_100a)"; }
const char* tm_last_error(void) { return tmil::g_err; }
int64_t tm_launch_count(void) { return tmil::g_launches.load(std::memory_order_relaxed); }
void tm_reset_launch_count(void) { tmil::g_launches.store(0, std::memory_order_relaxed); }

int tm_profile_enable(int on) {
  tmil::g_prof_on = on != 0;
  return TM_OK;
}
int tm_profile_stage_count(void) { return tmil::ST_COUNT - 1; }
const char* tm_profile_stage_name(int i) { return (i >= 0 && i < tmil::ST_COUNT) ? tmil::kStageNames[i] : ""; }
int tm_profile_read(double* ms, int64_t* counts, int n) {
  // Synchronises the recorded events; accumulates the time from each mark to the next one into the mark's stage.
  using namespace tmil;
  for (int i = 0; i < n; ++i) { ms[i] = 0.0; counts[i] = 0; }
  for (size_t i = 0; i + 1 < g_marks.size(); ++i) {
    const int stg = g_marks[i].first;
    if (stg == ST_END) continue;
    float t = 0.f;
    cudaEventSynchronize(g_marks[i + 1].second);
    if (cudaEventElapsedTime(&t, g_marks[i].second, g_marks[i + 1].second) != cudaSuccess) continue;
    if (stg < n) { ms[stg] += t; counts[stg] += 1; }
  }
  for (auto& m : g_marks) g_pool.push_back(m.second);
  g_marks.clear();
  return TM_OK;
}

int tm_device_supported(void) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  int major = 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
  return major == 10 ? 1 : 0;
}

int tm_make_plan(int64_t n, tm_plan* out) {
  TM_CHECK_ARG(out != nullptr, TM_ERR_BAD_ARGUMENT, "tm_make_plan: out is NULL");
  TM_CHECK_ARG(n >= 1, TM_ERR_BAD_SHAPE, "tm_make_plan: bag must hold at least one patch (n=%lld)", (long long)n);
  // int(np.ceil(np.sqrt(H))) with a float64 sqrt  (TransMIL.py:178)
  int64_t G = (int64_t)ceil(sqrt((double)n));
  out->n = n;
  out->G = G;
  out->N = G * G;
  out->dup = out->N - n;
  out->S = out->N + 1;
  int64_t rem = out->S % TM_LANDMARKS;
  out->P = rem > 0 ? TM_LANDMARKS - rem : 0;
  out->Sp = out->S + out->P;
  out->l = (out->S + TM_LANDMARKS - 1) / TM_LANDMARKS;
  out->padding = rem > 0 ? 256 - rem : 0;  // TransMIL.py:190-193
  TM_CHECK_ARG(out->dup <= n, TM_ERR_BAD_SHAPE, "tm_make_plan: duplicate pad %lld exceeds bag %lld", (long long)out->dup, (long long)n);
  return TM_OK;
}
int64_t tm_dup_source(const tm_plan* p, int64_t t) { return t < p->n ? t : t - p->n; }
int64_t tm_landmark_group(const tm_plan* p, int64_t r) { return r / p->l; }
int64_t tm_padded_row(const tm_plan* p, int64_t s) { return p->P + s; }
void tm_ppeg_pixel(const tm_plan* p, int64_t t, int64_t* y, int64_t* x) {
  *y = t / p->G;
  *x = t % p->G;
}

}  // extern "C"
