// Shared device/host helpers for libhpcodec_sm100a.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <math.h>

#include "../../include/hpcodec.h"

// ---------------------------------------------------------------------------------------
// error convention: every export returns int, message via hpc_last_error()
// ---------------------------------------------------------------------------------------
extern thread_local char g_hpc_err[512];
extern unsigned long long g_hpc_launches;  // kernels launched by this library (bench.py's gpu_launches)

#define HPC_FAIL(code, ...)                                   \
  do {                                                        \
    snprintf(g_hpc_err, sizeof(g_hpc_err), __VA_ARGS__);      \
    return (code);                                            \
  } while (0)

#define HPC_REQUIRE(cond, code, ...) \
  do {                               \
    if (!(cond)) HPC_FAIL(code, __VA_ARGS__); \
  } while (0)

#define HPC_CUDA(expr)                                                                 \
  do {                                                                                 \
    cudaError_t _e = (expr);                                                           \
    if (_e != cudaSuccess)                                                             \
      HPC_FAIL(HPC_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),   \
               __FILE__, __LINE__);                                                    \
  } while (0)

#define HPC_LAUNCH_CHECK()                                                             \
  do {                                                                                 \
    __atomic_fetch_add(&g_hpc_launches, 1ull, __ATOMIC_RELAXED);                       \
    cudaError_t _e = cudaGetLastError();                                               \
    if (_e != cudaSuccess)                                                             \
      HPC_FAIL(HPC_ERR_CUDA, "kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), \
               __FILE__, __LINE__);                                                    \
  } while (0)

// kernel-class timing (hpc_prof_* in hpcodec.h).  Usage inside an export:
//   HpcProfScope prof(HPC_PROF_IGEMM_FWD, flops, stream);  ... launches ...   (the destructor records the end event)
extern int g_hpc_prof_on;
int hpc_prof_begin_(int cls, double work, cudaStream_t st);
void hpc_prof_end_(int slot, cudaStream_t st);
struct HpcProfScope {
  int slot;
  cudaStream_t st;
  HpcProfScope(int cls, double work, void* stream) : slot(-1), st((cudaStream_t)stream) {
    if (g_hpc_prof_on) slot = hpc_prof_begin_(cls, work, st);
  }
  ~HpcProfScope() {
    if (slot >= 0) hpc_prof_end_(slot, st);
  }
};

// ---------------------------------------------------------------------------------------
// Programmatic dependent launch (griddepcontrol): a kernel launched through hpc_launch_pdl may become resident
// while its predecessor in the stream is still draining -- its prologue (barrier init, TMEM allocation, tensor-map
// prefetch) then overlaps the predecessor's tail.  Contract for every kernel launched this way:
//   * pdl_wait() before the first global-memory access that depends on (or may race with) earlier work;
//   * pdl_launch_dependents() once the CTA no longer minds the next kernel's CTAs becoming resident beside it.
// Both are no-ops for a kernel launched the ordinary way.  HPC_PDL=0 in the environment disables the attribute.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
int hpc_pdl_enabled();

template <typename... KArgs, typename... Args>
static inline cudaError_t hpc_launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                         Args&&... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = hpc_pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

static inline bool hpc_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// number of SMs of the current device (cached per thread)
// K axis of the packed conv weights of the 3-channel layers (thin_umma.cu: kmaj): k' = (cs*5 + kh)*8 + kw.  Returns the
// index into the reference's (cs, kh*5 + kw) axis, or -1 where the packed operand holds a zero (kw >= 5, k' >= 120).
__host__ __device__ __forceinline__ int c3_kmaj_source(int k) {
  const int kw = k & 7, r = k >> 3;                                   // r = cs*5 + kh
  return (kw < 5 && r < 15) ? (r / 5) * 25 + (r % 5) * 5 + kw : -1;
}

int hpc_num_sms();

// One-time set-up that CUDA keeps PER DEVICE (cudaFuncSetAttribute: the opt-in dynamic shared memory size belongs to the
// function on one device).  A process-wide `static bool` would skip every device after the first when one process
// drives several GPUs (nn.DataParallel, a test that switches devices) and the launch there would fail with
// "invalid argument".  One bit per device ordinal; racing threads at worst repeat the idempotent set-up.
struct HpcOncePerDevice {
  unsigned long long bits[4] = {0, 0, 0, 0};          // ordinals 0..255
  bool need(int* dev) {
    if (cudaGetDevice(dev) != cudaSuccess || *dev < 0 || *dev >= 256) { *dev = -1; return true; }   // unknown: always set
    return !((__atomic_load_n(&bits[*dev >> 6], __ATOMIC_ACQUIRE) >> (*dev & 63)) & 1ull);
  }
  void mark(int dev) {
    if (dev >= 0) __atomic_fetch_or(&bits[dev >> 6], 1ull << (dev & 63), __ATOMIC_RELEASE);
  }
};

// ---------------------------------------------------------------------------------------
// low-precision element access (T = float or __nv_bfloat16)
// ---------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ float ld_as_float(const T* p, size_t i);
template <> __device__ __forceinline__ float ld_as_float<float>(const float* p, size_t i) { return p[i]; }
template <> __device__ __forceinline__ float ld_as_float<__nv_bfloat16>(const __nv_bfloat16* p, size_t i) {
  return __bfloat162float(p[i]);
}
template <typename T> __device__ __forceinline__ void st_from_float(T* p, size_t i, float v);
template <> __device__ __forceinline__ void st_from_float<float>(float* p, size_t i, float v) { p[i] = v; }
template <> __device__ __forceinline__ void st_from_float<__nv_bfloat16>(__nv_bfloat16* p, size_t i, float v) {
  p[i] = __float2bfloat16_rn(v);
}

// ---------------------------------------------------------------------------------------
// reductions
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Sum over the block (blockDim.x*blockDim.y*blockDim.z threads, multiple of 32, <= 1024).
// Result valid in thread 0.  `sm` must hold 32 floats.
__device__ __forceinline__ float block_sum(float v, float* sm) {
  const int tid = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
  const int nthreads = blockDim.x * blockDim.y * blockDim.z;
  v = warp_sum(v);
  __syncthreads();  // protect sm reuse across successive calls
  if ((tid & 31) == 0) sm[tid >> 5] = v;
  __syncthreads();
  if (tid < 32) {
    float t = (tid < (nthreads + 31) / 32) ? sm[tid] : 0.f;
    t = warp_sum(t);
    v = t;
  }
  return v;
}

// Deterministic cross-block sum: every block stores its partial; the last block to arrive
// (atomic ticket) adds the partials in index order in double and writes out[0].
// counter must be zero before the launch and is reset to zero by the last block.
__device__ __forceinline__ void finalize_partials(float partial, float* partials, int nblocks, int block_id,
                                                  unsigned int* counter, float* out, float scale) {
  __shared__ bool is_last;
  const int tid = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
  if (tid == 0) {
    partials[block_id] = partial;
    __threadfence();
    unsigned int t = atomicAdd(counter, 1u);
    is_last = (t == (unsigned int)nblocks - 1);
  }
  __syncthreads();
  if (is_last && tid < 32) {
    __threadfence();
    double s = 0.0;
    for (int i = tid; i < nblocks; i += 32) s += (double)((volatile float*)partials)[i];
    s = warp_sum_d(s);
    if (tid == 0) {
      out[0] = (float)(s * (double)scale);
      *counter = 0;
    }
  }
}

// ---------------------------------------------------------------------------------------
// Philox4x32-10 counter RNG (in-kernel uniform noise; key = seed, counter = (offset,idx))
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}
// U[0,1) with 24 bits, like torch.rand for fp32 (multiples of 2^-24, never 1.0)
__device__ __forceinline__ float u01_from_bits(uint32_t b) { return (float)(b >> 8) * (1.0f / 16777216.0f); }

// four uniforms for flat element indices 4*i4 .. 4*i4+3 (the same values philox_uniform returns one by one)
__device__ __forceinline__ void philox_uniform4(uint64_t seed, uint64_t offset, uint64_t i4, float* u) {
  uint4 ctr = make_uint4((uint32_t)i4, (uint32_t)(i4 >> 32), (uint32_t)offset, (uint32_t)(offset >> 32));
  uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
  const uint4 r = philox4x32_10(ctr, key);
  u[0] = u01_from_bits(r.x); u[1] = u01_from_bits(r.y); u[2] = u01_from_bits(r.z); u[3] = u01_from_bits(r.w);
}
// vector helpers for the element-wise kernels: 4 consecutive elements (index of the first = 4*i4)
template <typename T> __device__ __forceinline__ void ld4_as_float(const T* p, size_t i4, float* f);
template <> __device__ __forceinline__ void ld4_as_float<float>(const float* p, size_t i4, float* f) {
  const float4 v = reinterpret_cast<const float4*>(p)[i4];
  f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
}
template <> __device__ __forceinline__ void ld4_as_float<__nv_bfloat16>(const __nv_bfloat16* p, size_t i4, float* f) {
  const uint2 v = reinterpret_cast<const uint2*>(p)[i4];
  const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&v.x));
  const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&v.y));
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y;
}
template <typename T> __device__ __forceinline__ void st4_from_float(T* p, size_t i4, const float* f);
template <> __device__ __forceinline__ void st4_from_float<float>(float* p, size_t i4, const float* f) {
  reinterpret_cast<float4*>(p)[i4] = make_float4(f[0], f[1], f[2], f[3]);
}
template <> __device__ __forceinline__ void st4_from_float<__nv_bfloat16>(__nv_bfloat16* p, size_t i4, const float* f) {
  uint2 v;
  *reinterpret_cast<__nv_bfloat162*>(&v.x) = __floats2bfloat162_rn(f[0], f[1]);
  *reinterpret_cast<__nv_bfloat162*>(&v.y) = __floats2bfloat162_rn(f[2], f[3]);
  reinterpret_cast<uint2*>(p)[i4] = v;
}

// one uniform for flat element index i
__device__ __forceinline__ float philox_uniform(uint64_t seed, uint64_t offset, uint64_t i) {
  uint4 ctr = make_uint4((uint32_t)(i >> 2), (uint32_t)(i >> 34), (uint32_t)offset, (uint32_t)(offset >> 32));
  uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
  uint4 r = philox4x32_10(ctr, key);
  uint32_t w = (i & 3) == 0 ? r.x : (i & 3) == 1 ? r.y : (i & 3) == 2 ? r.z : r.w;
  return u01_from_bits(w);
}
