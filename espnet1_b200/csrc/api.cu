// Library-level entry points: error string, launch counter, device info.
#include <stdarg.h>

#include <atomic>
#include <mutex>
#include <vector>

#include "common.cuh"

namespace b200ctc {

static thread_local char g_err[512] = "";
// process-wide: autograd runs backward launches on its own engine thread
static std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

void count_launch(int n) { g_launches += n; }

int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return B200CTC_ERR_DEVICE;
  }
  return B200CTC_OK;
}

// ---- optional per-phase CUDA-event profiling (bench.py roofline) -----------
static std::atomic<bool> g_prof{false};
static std::mutex g_prof_mu;
static std::vector<cudaEvent_t> g_ev[kProfPhases][2];

void prof_mark(int phase, int end, cudaStream_t st) {
  if (!g_prof || phase < 0 || phase >= kProfPhases) return;
  cudaEvent_t e;
  if (cudaEventCreate(&e) != cudaSuccess) return;
  cudaEventRecord(e, st);
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_ev[phase][end ? 1 : 0].push_back(e);
}

int sm_count() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, dev) != cudaSuccess) return 148;
    cached = p.multiProcessorCount;
    cached_dev = dev;
  }
  return cached;
}

}  // namespace b200ctc

extern "C" {

int b200ctc_abi_version(void) { return B200CTC_ABI_VERSION; }
const char* b200ctc_last_error(void) { return b200ctc::g_err; }
int64_t b200ctc_launch_count(void) { return b200ctc::g_launches.load(); }
void b200ctc_reset_launch_count(void) { b200ctc::g_launches = 0; }

void b200ctc_profile_enable(int on) { b200ctc::g_prof = on != 0; }

int b200ctc_profile_read(float* ms_total, int* counts, int n) {
  using namespace b200ctc;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (int p = 0; p < kProfPhases; ++p) {
    float tot = 0.f;
    int cnt = 0;
    auto& b = g_ev[p][0];
    auto& e = g_ev[p][1];
    const size_t m = b.size() < e.size() ? b.size() : e.size();
    for (size_t i = 0; i < m; ++i) {
      float ms = 0.f;
      if (cudaEventSynchronize(e[i]) == cudaSuccess &&
          cudaEventElapsedTime(&ms, b[i], e[i]) == cudaSuccess) { tot += ms; ++cnt; }
    }
    for (auto ev : b) cudaEventDestroy(ev);
    for (auto ev : e) cudaEventDestroy(ev);
    b.clear(); e.clear();
    if (p < n) { if (ms_total) ms_total[p] = tot; if (counts) counts[p] = cnt; }
  }
  return kProfPhases;
}

int b200ctc_device_info(int* sms, int* major, int* minor) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  cudaDeviceProp p;
  if (e == cudaSuccess) e = cudaGetDeviceProperties(&p, dev);
  if (e != cudaSuccess) {
    b200ctc::set_error("no CUDA device: %s (this library has no CPU fallback)",
                       cudaGetErrorString(e));
    return B200CTC_ERR_DEVICE;
  }
  if (sms) *sms = p.multiProcessorCount;
  if (major) *major = p.major;
  if (minor) *minor = p.minor;
  return B200CTC_OK;
}

}  // extern "C"
