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

// ---------------------------------------------------------------------------------------
// hpc_pack_multi: every packed bf16 operand copy a training step needs (the [tap][row][col] copies of the
// 192 -> 192 conv / transposed-conv weights for the forward and the dgrad engines, the K-major / N-major copies of
// the 3-channel head and tail, the effective gamma / gamma^T / beta of every GDN layer) in ONE launch.
// A block owns one tile of one job; the job table travels by value in the kernel parameters.
// Same arithmetic as hpc_pack_weight / hpc_pack_c3_weight / hpc_gdn_prepare (pure data movement + the
// NonNegativeParam reparametrisation of gdn.py:59-62), so the engines see bit-identical operands.
// ---------------------------------------------------------------------------------------
namespace {
constexpr int PK_TR = 8, PK_TC = 32, PK_THREADS = 256;
constexpr int PK_FLAT = 2048;                 // dst elements per block of the flat job kinds

struct PackTable {
  hpc_pack_job job[HPC_PACK_MAX_JOBS];
  int block_begin[HPC_PACK_MAX_JOBS + 1];
  int njobs;
  int dtype;
};

template <typename T>
__device__ __forceinline__ void pack_taps_tile(const hpc_pack_job& J, int b, float* sm) {
  const int K = J.ntaps, rows = J.rows, cols = J.cols;
  const int tiles_c = (cols + PK_TC - 1) / PK_TC;
  const int tr0 = (b / tiles_c) * PK_TR, tc0 = (b % tiles_c) * PK_TC;
  const float* __restrict__ src = J.src;
  const int64_t sr = J.stride_r, sc = J.stride_c;
  const bool c_inner = sc <= sr;               // the (r | c) index whose K taps follow each other in memory goes second
  const int n = PK_TR * PK_TC * K;
  for (int i = threadIdx.x; i < n; i += PK_THREADS) {
    const int k = i % K, q = i / K;
    const int c = c_inner ? q % PK_TC : q / PK_TR;
    const int r = c_inner ? q / PK_TC : q % PK_TR;
    const bool in = tr0 + r < rows && tc0 + c < cols;
    sm[(r * PK_TC + c) * K + k] = in ? src[(tr0 + r) * sr + (tc0 + c) * sc + k] : 0.f;
  }
  __syncthreads();
  T* __restrict__ dst = (T*)J.dst;
  for (int i = threadIdx.x; i < n; i += PK_THREADS) {
    const int c = i % PK_TC, r = (i / PK_TC) % PK_TR, j = i / (PK_TC * PK_TR);
    if (tr0 + r < rows && tc0 + c < cols)
      st_from_float<T>(dst, ((size_t)j * rows + tr0 + r) * cols + tc0 + c, sm[(r * PK_TC + c) * K + J.tap_src[j]]);
  }
}

template <typename T>
__global__ void __launch_bounds__(PK_THREADS)
pack_multi_kernel(const __grid_constant__ PackTable t) {
  __shared__ float sm[PK_TR * PK_TC * HPC_MAX_TAPS];
  int lo = 0, hi = t.njobs;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if ((int)blockIdx.x >= t.block_begin[mid]) lo = mid; else hi = mid;
  }
  const hpc_pack_job& J = t.job[lo];
  const int b = (int)blockIdx.x - t.block_begin[lo];
  if (J.kind == HPC_PACK_TAPS) {
    pack_taps_tile<T>(J, b, sm);
  } else if (J.kind == HPC_PACK_C3_K || J.kind == HPC_PACK_C3_N) {
    // w[cw*stride_r + cs*stride_c + tap] -> kmaj[cw][128] (k' = (cs*5+kh)*8+kw, zeros elsewhere: c3_kmaj_source) or
    // nmaj[80][cw] (k = cs*25+tap); always bf16
    const int Cw = J.rows, total = J.kind == HPC_PACK_C3_K ? Cw * 128 : 80 * Cw;
    __nv_bfloat16* __restrict__ dst = (__nv_bfloat16*)J.dst;
    for (int i = b * PK_FLAT + threadIdx.x; i < min(total, (b + 1) * PK_FLAT); i += PK_THREADS) {
      const int cw = J.kind == HPC_PACK_C3_K ? i / 128 : i % Cw, k = J.kind == HPC_PACK_C3_K ? i % 128 : i / Cw;
      const int src = J.kind == HPC_PACK_C3_K ? c3_kmaj_source(k) : (k < 75 ? k : -1);
      dst[i] = __float2bfloat16_rn(src >= 0 ? J.src[cw * J.stride_r + (src / 25) * J.stride_c + (src % 25)] : 0.f);
    }
  } else {  // HPC_PACK_GDN: gdn.py:59-62 for gamma and beta + both bf16 operand copies of gamma (== gdn_prepare_kernel)
    const int C = J.rows, total = C * C;
    __nv_bfloat16* __restrict__ g = (__nv_bfloat16*)J.dst;
    __nv_bfloat16* __restrict__ gT = (__nv_bfloat16*)J.dst2;
    for (int i = b * PK_FLAT + threadIdx.x; i < min(total, (b + 1) * PK_FLAT); i += PK_THREADS) {
      const float v = fmaxf(J.src[i], J.f0);
      const __nv_bfloat16 h = __float2bfloat16_rn(v * v - J.f2);
      g[i] = h;
      gT[(i % C) * C + i / C] = h;
      if (i < C) {
        const float w = fmaxf(J.src2[i], J.f1);
        J.dst3[i] = w * w - J.f2;
      }
    }
  }
}
}  // namespace

int hpc_pack_multi(const hpc_pack_job* jobs, int njobs, int dtype, hpc_stream_t stream) {
  HPC_REQUIRE(jobs && njobs > 0, HPC_ERR_BAD_SHAPE, "hpc_pack_multi: no jobs");
  HPC_REQUIRE(dtype == HPC_BF16 || dtype == HPC_F32, HPC_ERR_UNSUPPORTED, "hpc_pack_multi: dtype %d", dtype);
  for (int base = 0; base < njobs; base += HPC_PACK_MAX_JOBS) {
    PackTable t;
    memset(&t, 0, sizeof(t));
    t.njobs = njobs - base < HPC_PACK_MAX_JOBS ? njobs - base : HPC_PACK_MAX_JOBS;
    t.dtype = dtype;
    int blocks = 0;
    for (int i = 0; i < t.njobs; ++i) {
      const hpc_pack_job& J = jobs[base + i];
      HPC_REQUIRE(J.src && J.dst && J.rows > 0, HPC_ERR_BAD_SHAPE, "hpc_pack_multi: job %d has a NULL pointer or no rows", base + i);
      t.job[i] = J;
      t.block_begin[i] = blocks;
      switch (J.kind) {
        case HPC_PACK_TAPS:
          HPC_REQUIRE(J.ntaps >= 1 && J.ntaps <= HPC_MAX_TAPS && J.cols > 0 && (J.stride_r == J.ntaps || J.stride_c == J.ntaps),
                      HPC_ERR_BAD_SHAPE, "hpc_pack_multi: job %d: taps=%d cols=%d strides (%lld,%lld): the taps of one "
                      "(row, col) must be contiguous", base + i, J.ntaps, J.cols, (long long)J.stride_r, (long long)J.stride_c);
          for (int k = 0; k < J.ntaps; ++k)
            HPC_REQUIRE(J.tap_src[k] < J.ntaps, HPC_ERR_BAD_SHAPE, "hpc_pack_multi: job %d: tap_src[%d] out of range", base + i, k);
          blocks += ((J.rows + PK_TR - 1) / PK_TR) * ((J.cols + PK_TC - 1) / PK_TC);
          break;
        case HPC_PACK_C3_K: blocks += (J.rows * 128 + PK_FLAT - 1) / PK_FLAT; break;
        case HPC_PACK_C3_N: blocks += (J.rows * 80 + PK_FLAT - 1) / PK_FLAT; break;
        case HPC_PACK_GDN:
          HPC_REQUIRE(J.src2 && J.dst2 && J.dst3 && J.rows <= PK_FLAT, HPC_ERR_BAD_SHAPE, "hpc_pack_multi: job %d: GDN job needs beta, "
                      "gamma^T and beta_eff pointers and C <= %d", base + i, PK_FLAT);
          blocks += (J.rows * J.rows + PK_FLAT - 1) / PK_FLAT;
          break;
        default: HPC_FAIL(HPC_ERR_UNSUPPORTED, "hpc_pack_multi: job %d: kind %d", base + i, J.kind);
      }
    }
    t.block_begin[t.njobs] = blocks;
    if (dtype == HPC_BF16)
      pack_multi_kernel<__nv_bfloat16><<<blocks, PK_THREADS, 0, (cudaStream_t)stream>>>(t);
    else
      pack_multi_kernel<float><<<blocks, PK_THREADS, 0, (cudaStream_t)stream>>>(t);
    HPC_LAUNCH_CHECK();
  }
  return HPC_OK;
}

// ---------------------------------------------------------------------------------------
// hpc_u8_crop_normalize: the reference's per-image input pipeline on the device:
//   RandomCrop window -> ToTensor (float) -> ChannelFirst -> Normalize = (v / 255 - mean) / std, stacked like Batch
//   (data_utils/transforms/transforms.py:23-35, 64-75, 84-127, 143-166; data_utils/batch.py:19-29).
// The host sends uint8 HWC pixels (4x fewer bytes than the fp32 NCHW batch the reference copies) and the crop offsets it
// drew; every arithmetic step is fp32 with IEEE division, like torch's, so the result is bit-identical.
// One thread per output pixel: a warp reads 32 consecutive pixels (96 B for RGB) and writes 32 consecutive floats to
// each channel plane.  A window that leaves the source image is never read (zeros are written): the host validates.
// ---------------------------------------------------------------------------------------
namespace {
struct CropNorm {
  const uint8_t* src; int64_t img_stride, row_stride;
  const int* offs;                 // device (N, 2) [y0, x0] or NULL
  const int* index;                // device (N) source image of output n, or NULL (= n)
  int n_src, src_h, src_w, N, C, oh, ow;
  float mean[4], std[4];
  float* out;
};
__global__ void u8_crop_normalize_kernel(const __grid_constant__ CropNorm p) {
  const int64_t plane = (int64_t)p.oh * p.ow, total = plane * p.N;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int n = (int)(t / plane);
    const int r = (int)(t % plane), i = r / p.ow, j = r % p.ow;
    const int y = i + (p.offs ? p.offs[2 * n] : 0), x = j + (p.offs ? p.offs[2 * n + 1] : 0);
    const int m = p.index ? p.index[n] : n;
    const bool in = m >= 0 && m < p.n_src && y >= 0 && y < p.src_h && x >= 0 && x < p.src_w;
    const uint8_t* s = p.src + m * p.img_stride + (int64_t)y * p.row_stride + (int64_t)x * p.C;
    float* o = p.out + (int64_t)n * p.C * plane + r;
#pragma unroll 4
    for (int c = 0; c < p.C; ++c) {
      const float v = in ? (float)s[c] : 0.f;
      o[c * plane] = (v / 255.0f - p.mean[c]) / p.std[c];
    }
  }
}
}  // namespace

int hpc_u8_crop_normalize(const unsigned char* src_u8_hwc, int64_t img_stride_bytes, int64_t row_stride_bytes, int n_src,
                          int src_h, int src_w, const int* src_index_dev, const int* offsets_yx_dev, int N, int C, int out_h, int out_w, const float* mean_host,
                          const float* std_host, float* out_nchw, hpc_stream_t stream) {
  HPC_REQUIRE(src_u8_hwc && out_nchw, HPC_ERR_BAD_SHAPE, "hpc_u8_crop_normalize: NULL pointer");
  HPC_REQUIRE(n_src > 0 && (src_index_dev || n_src >= N), HPC_ERR_BAD_SHAPE,
              "hpc_u8_crop_normalize: %d source images for %d outputs and no index", n_src, N);
  HPC_REQUIRE(N > 0 && C >= 1 && C <= 4 && out_h > 0 && out_w > 0 && src_h >= out_h && src_w >= out_w, HPC_ERR_BAD_SHAPE,
              "hpc_u8_crop_normalize: bad shape (N %d, C %d, source %dx%d, output %dx%d; the reference does not pad "
              "under-sized images either: transforms.py:158-159)", N, C, src_h, src_w, out_h, out_w);
  HPC_REQUIRE(row_stride_bytes >= (int64_t)src_w * C && img_stride_bytes >= 0, HPC_ERR_BAD_SHAPE,
              "hpc_u8_crop_normalize: row stride %lld smaller than a row (%d bytes)", (long long)row_stride_bytes, src_w * C);
  CropNorm p;
  memset(&p, 0, sizeof(p));
  p.src = src_u8_hwc; p.img_stride = img_stride_bytes; p.row_stride = row_stride_bytes; p.offs = offsets_yx_dev;
  p.index = src_index_dev; p.n_src = n_src;
  p.src_h = src_h; p.src_w = src_w; p.N = N; p.C = C; p.oh = out_h; p.ow = out_w; p.out = out_nchw;
  for (int c = 0; c < C; ++c) {
    p.mean[c] = mean_host ? mean_host[c] : 0.f;
    p.std[c] = std_host ? std_host[c] : 1.f;
    HPC_REQUIRE(p.std[c] != 0.f, HPC_ERR_BAD_SHAPE, "hpc_u8_crop_normalize: std[%d] is zero", c);
  }
  const int64_t total = (int64_t)N * out_h * out_w;
  u8_crop_normalize_kernel<<<grid_for(total, 256, 148 * 16), 256, 0, (cudaStream_t)stream>>>(p);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}
