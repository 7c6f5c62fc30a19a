// Small elementwise kernels used by the composed (module-by-module) path.
#include "hpc_common.cuh"

static inline int grid_for(int64_t n, int threads, int max_blocks = 148 * 16) {
  int64_t b = (n + threads - 1) / threads;
  if (b < 1) b = 1;
  if (b > max_blocks) b = max_blocks;
  return (int)b;
}

__global__ void abs_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    y[i] = fabsf(x[i]);
}
__global__ void abs_bwd_kernel(const float* __restrict__ x, const float* __restrict__ g, float* __restrict__ gx,
                               int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float v = x[i];
    gx[i] = v > 0.f ? g[i] : (v < 0.f ? -g[i] : 0.f);  // torch: grad * sign(x)
  }
}
int hpc_abs_fwd(const float* x, float* y, int64_t n, hpc_stream_t stream) {
  if (n <= 0) return HPC_OK;
  abs_fwd_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(x, y, n);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}
int hpc_abs_bwd(const float* x, const float* g, float* gx, int64_t n, hpc_stream_t stream) {
  if (n <= 0) return HPC_OK;
  abs_bwd_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(x, g, gx, n);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}
