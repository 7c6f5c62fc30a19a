// Library-level entry points: version, error string, device probe.
#include "hpc_common.cuh"

#include <stdlib.h>
#include <mutex>
#include <vector>

thread_local char g_hpc_err[512] = "";

unsigned long long g_hpc_launches = 0;
unsigned long long hpc_launch_count(void) { return __atomic_load_n(&g_hpc_launches, __ATOMIC_RELAXED); }

int hpc_version(void) { return HPC_VERSION; }
const char* hpc_last_error(void) { return g_hpc_err; }

int hpc_device_is_sm100(void) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  int major = 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return 0;
  return major == 10;
}

int hpc_num_sms() {
  thread_local int cached = 0;
  if (cached) return cached;
  int dev = 0, n = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  cached = n > 0 ? n : 148;
  return cached;
}

int hpc_pdl_enabled() {
  static int on = -1;
  if (on < 0) { const char* e = getenv("HPC_PDL"); on = (e && e[0] == '0') ? 0 : 1; }
  return on;
}

// ---------------------------------------------------------------------------------------
// kernel-class timing (bench.py roofline): event pairs on the launching stream
// ---------------------------------------------------------------------------------------
int g_hpc_prof_on = 0;
namespace {
struct ProfRec { cudaEvent_t a, b; int cls; double work; bool closed; };
std::vector<ProfRec> g_prof;
std::vector<cudaEvent_t> g_prof_pool;
std::mutex g_prof_mu;
cudaEvent_t prof_event() {
  if (!g_prof_pool.empty()) { cudaEvent_t e = g_prof_pool.back(); g_prof_pool.pop_back(); return e; }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}
}  // namespace

int hpc_prof_begin_(int cls, double work, cudaStream_t st) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  ProfRec r;
  r.a = prof_event(); r.b = prof_event(); r.cls = cls; r.work = work; r.closed = false;
  if (!r.a || !r.b) return -1;
  cudaEventRecord(r.a, st);
  g_prof.push_back(r);
  return (int)g_prof.size() - 1;
}

void hpc_prof_end_(int slot, cudaStream_t st) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (slot < 0 || slot >= (int)g_prof.size()) return;
  cudaEventRecord(g_prof[slot].b, st);
  g_prof[slot].closed = true;
}

int hpc_prof_enable(int on) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (on) {
    for (auto& r : g_prof) { g_prof_pool.push_back(r.a); g_prof_pool.push_back(r.b); }
    g_prof.clear();
  }
  g_hpc_prof_on = on ? 1 : 0;
  return HPC_OK;
}

int hpc_prof_pause(int paused) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_hpc_prof_on = paused ? 0 : 1;
  return HPC_OK;
}

int hpc_prof_read(int cls, double* ms_total, unsigned long long* launches, double* work_total) {
  HPC_REQUIRE(cls >= 0 && cls < HPC_PROF_NCLASSES, HPC_ERR_BAD_SHAPE, "hpc_prof_read: class %d", cls);
  std::lock_guard<std::mutex> lk(g_prof_mu);
  double ms = 0.0, work = 0.0;
  unsigned long long n = 0;
  for (auto& r : g_prof) {
    if (r.cls != cls || !r.closed) continue;
    HPC_CUDA(cudaEventSynchronize(r.b));
    float t = 0.f;
    HPC_CUDA(cudaEventElapsedTime(&t, r.a, r.b));
    ms += t; work += r.work; ++n;
  }
  if (ms_total) *ms_total = ms;
  if (launches) *launches = n;
  if (work_total) *work_total = work;
  return HPC_OK;
}

int hpc_prof_records(int* cls, float* ms, double* work, int max) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  int n = 0;
  for (auto& r : g_prof) {
    if (!r.closed) continue;
    if (n < max) {
      float t = 0.f;
      if (cudaEventSynchronize(r.b) == cudaSuccess) cudaEventElapsedTime(&t, r.a, r.b);
      if (cls) cls[n] = r.cls;
      if (ms) ms[n] = t;
      if (work) work[n] = r.work;
    }
    ++n;
  }
  return n;
}
