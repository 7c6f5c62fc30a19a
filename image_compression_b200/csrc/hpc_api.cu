// Library-level entry points: version, error string, device probe.
#include "hpc_common.cuh"

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
