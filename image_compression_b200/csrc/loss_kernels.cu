// Distortion path: clamp + MSE/PSNR (K6), (MS-)SSIM levels and pyramid (K7), RD combine.
// Images are NCHW float.  All kernels are HBM-bound streaming kernels.
#include "hpc_common.cuh"

static inline int grid_for(int64_t n, int threads, int max_blocks = 148 * 16) {
  int64_t b = (n + threads - 1) / threads;
  if (b < 1) b = 1;
  if (b > max_blocks) b = max_blocks;
  return (int)b;
}

// ---------------------------------------------------------------------------------------
// K6 forward: x~ = max(min(raw,1),0); per-image sum((scale*(x-x~))^2)
// grid = (blocks_per_image, N); deterministic two-level sum via a per-image ticket
// ---------------------------------------------------------------------------------------
struct SseScratch {
  unsigned int counters[1024];
  float partials[1];  // [N][blocks_per_image]
};

__global__ void __launch_bounds__(256)
clamp01_sse_kernel(const float* __restrict__ x, const float* __restrict__ raw, float* __restrict__ xt,
                   int64_t per_image, float scale, int do_clamp, float* __restrict__ sse, unsigned int* counters,
                   float* partials) {
  __shared__ float sm[32];
  __shared__ bool is_last;
  const int n = blockIdx.y;
  const int64_t base = (int64_t)n * per_image;
  float acc = 0.f;
  const int64_t nvec = per_image >> 2;
  const bool vec_ok = ((per_image & 3) == 0);
  if (vec_ok) {
    const float4* x4 = reinterpret_cast<const float4*>(x + base);
    const float4* r4 = reinterpret_cast<const float4*>(raw + base);
    float4* o4 = xt ? reinterpret_cast<float4*>(xt + base) : nullptr;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
      float4 a = x4[i], r = r4[i];
      if (do_clamp) {
        r.x = fmaxf(fminf(r.x, 1.f), 0.f); r.y = fmaxf(fminf(r.y, 1.f), 0.f);
        r.z = fmaxf(fminf(r.z, 1.f), 0.f); r.w = fmaxf(fminf(r.w, 1.f), 0.f);
      }
      if (o4) o4[i] = r;
      float d0 = scale * a.x - scale * r.x, d1 = scale * a.y - scale * r.y;
      float d2 = scale * a.z - scale * r.z, d3 = scale * a.w - scale * r.w;
      acc += d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
    }
  } else {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < per_image;
         i += (int64_t)gridDim.x * blockDim.x) {
      float a = x[base + i], r = raw[base + i];
      if (do_clamp) r = fmaxf(fminf(r, 1.f), 0.f);
      if (xt) xt[base + i] = r;
      float d = scale * a - scale * r;
      acc += d * d;
    }
  }
  float s = block_sum(acc, sm);
  if (threadIdx.x == 0) {
    partials[(int64_t)n * gridDim.x + blockIdx.x] = s;
    __threadfence();
    unsigned int t = atomicAdd(&counters[n], 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last && threadIdx.x < 32) {
    __threadfence();
    double tot = 0.0;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += 32)
      tot += (double)((volatile float*)partials)[(int64_t)n * gridDim.x + i];
    tot = warp_sum_d(tot);
    if (threadIdx.x == 0) {
      sse[n] = (float)tot;
      counters[n] = 0;
    }
  }
}

// scratch layout: 1024 counters then partials; hpc_reduce_scratch_bytes() is too small for large N,
// so the Python side allocates 4096 + 4*N*kSseBlocks bytes (see _lib.sse_scratch_bytes).
static constexpr int kSseBlocksMax = 64;

int hpc_clamp01_sse_fwd(const float* x, const float* raw, float* xt, int N, int64_t per_image, float scale,
                        int do_clamp, float* sse_per_image, void* scratch, hpc_stream_t stream) {
  HPC_REQUIRE(N > 0 && per_image > 0, HPC_ERR_BAD_SHAPE, "hpc_clamp01_sse_fwd: empty input");
  HPC_REQUIRE(N <= 1024, HPC_ERR_UNSUPPORTED, "hpc_clamp01_sse_fwd: N=%d > 1024 images per call", N);
  HPC_REQUIRE(scratch, HPC_ERR_BAD_SHAPE, "hpc_clamp01_sse_fwd: scratch is NULL");
  HPC_REQUIRE((per_image & 3) != 0 || (hpc_aligned16(x) && hpc_aligned16(raw) && (!xt || hpc_aligned16(xt))),
              HPC_ERR_BAD_ALIGNMENT, "hpc_clamp01_sse_fwd: pointers must be 16-byte aligned");
  int bpi = (int)((per_image / 4 + 255) / 256);
  int want = (148 * 8 + N - 1) / N;
  if (bpi > want) bpi = want;
  if (bpi > kSseBlocksMax) bpi = kSseBlocksMax;
  if (bpi < 1) bpi = 1;
  unsigned int* counters = (unsigned int*)scratch;
  float* partials = (float*)((char*)scratch + 4096);
  HpcProfScope prof(HPC_PROF_DISTORTION, (double)N * per_image * (8 + (xt ? 4 : 0)), stream);
  clamp01_sse_kernel<<<dim3(bpi, N), 256, 0, (cudaStream_t)stream>>>(x, raw, xt, per_image, scale, do_clamp,
                                                                   sse_per_image, counters, partials);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

__global__ void clamp01_bwd_kernel(const float* __restrict__ raw, const float* __restrict__ x,
                                   const float* __restrict__ g_xt, int64_t n, float coef,
                                   const float* __restrict__ coef_dev, int do_clamp, float* __restrict__ g_raw) {
  const float c2 = coef * (coef_dev ? coef_dev[0] : 1.f);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float r = raw[i];
    float u = do_clamp ? fminf(r, 1.f) : r;   // UpperBound(1)
    float c = do_clamp ? fmaxf(u, 0.f) : r;   // LowerBound(0)
    float g = g_xt ? g_xt[i] : 0.f;
    if (x) g += c2 * 2.f * (c - x[i]);
    if (!do_clamp) { g_raw[i] = g; continue; }
    // LowerBound backward (input u, bound 0): pass if u >= 0 or g < 0
    float gu = (u >= 0.f || g < 0.f) ? g : 0.f;
    // UpperBound backward (input raw, bound 1): pass if raw <= 1 or g > 0
    g_raw[i] = (r <= 1.f || gu > 0.f) ? gu : 0.f;
  }
}
int hpc_clamp01_bwd(const float* raw, const float* x, const float* g_xt, int64_t n, float mse_coef,
                    const float* mse_coef_dev, int do_clamp, float* g_raw, hpc_stream_t stream) {
  if (n <= 0) return HPC_OK;
  HpcProfScope prof(HPC_PROF_DISTORTION, (double)n * (8 + (x ? 4 : 0) + (g_xt ? 4 : 0)), stream);
  clamp01_bwd_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(raw, x, g_xt, n, mse_coef, mse_coef_dev,
                                                                       do_clamp, g_raw);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

__global__ void psnr_kernel(const float* __restrict__ sse, int N, float inv_count, float c0, float* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < N) {
    float mse = sse[i] * inv_count;
    out[i] = 10.f * (c0 - logf(mse) / 2.302585092994046f);
  }
}
int hpc_psnr_from_sse(const float* sse_per_image, int N, int64_t per_image, float max_val, float* psnr,
                      hpc_stream_t stream) {
  HPC_REQUIRE(N > 0 && per_image > 0, HPC_ERR_BAD_SHAPE, "hpc_psnr_from_sse: empty input");
  psnr_kernel<<<(N + 127) / 128, 128, 0, (cudaStream_t)stream>>>(sse_per_image, N, (float)(1.0 / (double)per_image),
                                                                (float)(2.0 * log10((double)max_val)), psnr);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

// ---------------------------------------------------------------------------------------
// K7: one SSIM level.  Block = 32x8 threads, output tile 32x16, separable fs-tap window.
// ---------------------------------------------------------------------------------------
constexpr int SS_TW = 32, SS_TH = 16, SS_MAXF = 15;
constexpr int SS_PW = SS_TW + SS_MAXF - 1, SS_PH = SS_TH + SS_MAXF - 1;

struct SsimTerms {
  float ssim, cs;
  float dA_s, dB_s, dC_s;  // d ssim / d(mu2, E22, E12)
  float dA_c, dB_c, dC_c;  // d cs   / d(mu2, E22, E12)
};
__device__ __forceinline__ SsimTerms ssim_terms(float mu1, float mu2, float e11, float e22, float e12, float c1,
                                                float c2, bool want_grad) {
  SsimTerms t;
  float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
  float s1 = e11 - mu1_sq, s2 = e22 - mu2_sq, s12 = e12 - mu12;
  float dcs = s1 + s2 + c2;
  float cs = (2.f * s12 + c2) / dcs;
  float dl = mu1_sq + mu2_sq + c1;
  float lum = (2.f * mu12 + c1) / dl;
  t.cs = cs;
  t.ssim = cs * lum;
  if (want_grad) {
    // one reciprocal per denominator for the four gradient quotients (the loss VALUES above keep the true division)
    const float inv_dcs = 1.f / dcs, inv_dl = 1.f / dl;
    float dcs_dmu2 = (-2.f * mu1 + 2.f * mu2 * cs) * inv_dcs;
    float dcs_dE22 = -cs * inv_dcs;
    float dcs_dE12 = 2.f * inv_dcs;
    float dlum_dmu2 = 2.f * (mu1 - lum * mu2) * inv_dl;
    t.dA_c = dcs_dmu2; t.dB_c = dcs_dE22; t.dC_c = dcs_dE12;
    t.dA_s = lum * dcs_dmu2 + cs * dlum_dmu2;
    t.dB_s = lum * dcs_dE22;
    t.dC_s = lum * dcs_dE12;
  }
  return t;
}

// MODE 0: forward sums.  MODE 1: backward maps A,B,C.
template <int MODE>
__global__ void __launch_bounds__(256)
ssim_level_kernel(const float* __restrict__ img1, const float* __restrict__ img2, int C, int H, int W,
                  const float* __restrict__ filt, int fs, float in_scale, float c1, float c2,
                  unsigned long long* __restrict__ fx_sums, const float* __restrict__ g_ssim,
                  const float* __restrict__ g_cs, float* __restrict__ maps) {
  __shared__ float p1[SS_PH][SS_PW + 1], p2[SS_PH][SS_PW + 1];
  __shared__ float hbuf[5][SS_PH][SS_TW + 1];
  __shared__ float f[SS_MAXF];
  __shared__ float red[32];
  const int plane = blockIdx.z;  // n*C + c
  const int n = plane / C;
  const int OH = H - fs + 1, OW = W - fs + 1;
  const int ox0 = blockIdx.x * SS_TW, oy0 = blockIdx.y * SS_TH;
  const int tid = threadIdx.y * 32 + threadIdx.x;
  if (tid < fs) f[tid] = filt[tid];
  const float* a = img1 + (int64_t)plane * H * W;
  const float* b = img2 + (int64_t)plane * H * W;
  const int ph = SS_TH + fs - 1, pw = SS_TW + fs - 1;
  for (int i = tid; i < ph * pw; i += 256) {
    int r = i / pw, c = i % pw;
    int gy = oy0 + r, gx = ox0 + c;
    float v1 = 0.f, v2 = 0.f;
    if (gy < H && gx < W) {
      v1 = a[(int64_t)gy * W + gx] * in_scale;
      v2 = b[(int64_t)gy * W + gx] * in_scale;
    }
    p1[r][c] = v1;
    p2[r][c] = v2;
  }
  __syncthreads();
  // horizontal pass: 5 quantities for ph rows x 32 cols
  for (int i = tid; i < ph * SS_TW; i += 256) {
    int r = i / SS_TW, c = i % SS_TW;
    float m1 = 0.f, m2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
    for (int k = 0; k < fs; ++k) {
      float w = f[k], u = p1[r][c + k], v = p2[r][c + k];
      m1 += w * u; m2 += w * v; e11 += w * u * u; e22 += w * v * v; e12 += w * u * v;
    }
    hbuf[0][r][c] = m1; hbuf[1][r][c] = m2; hbuf[2][r][c] = e11; hbuf[3][r][c] = e22; hbuf[4][r][c] = e12;
  }
  __syncthreads();
  float acc_s = 0.f, acc_c = 0.f;
  float gs = 0.f, gc = 0.f;
  if (MODE == 1) { gs = g_ssim[n]; gc = g_cs[n]; }
#pragma unroll
  for (int rep = 0; rep < 2; ++rep) {
    int ty = threadIdx.y + rep * 8, tx = threadIdx.x;
    int oy = oy0 + ty, ox = ox0 + tx;
    if (oy < OH && ox < OW) {
      float m1 = 0.f, m2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
      for (int k = 0; k < fs; ++k) {
        float w = f[k];
        m1 += w * hbuf[0][ty + k][tx]; m2 += w * hbuf[1][ty + k][tx];
        e11 += w * hbuf[2][ty + k][tx]; e22 += w * hbuf[3][ty + k][tx]; e12 += w * hbuf[4][ty + k][tx];
      }
      SsimTerms t = ssim_terms(m1, m2, e11, e22, e12, c1, c2, MODE == 1);
      if (MODE == 0) {
        acc_s += t.ssim;
        acc_c += t.cs;
      } else {
        const int64_t msz = (int64_t)gridDim.z * OH * OW;
        const int64_t o = (int64_t)plane * OH * OW + (int64_t)oy * OW + ox;
        maps[o] = gs * t.dA_s + gc * t.dA_c;
        maps[msz + o] = gs * t.dB_s + gc * t.dB_c;
        maps[2 * msz + o] = gs * t.dC_s + gc * t.dC_c;
      }
    }
  }
  if (MODE == 0) {
    float s = block_sum(acc_s, red);
    float c = block_sum(acc_c, red);
    // block partials are accumulated in 32.32 fixed point: integer addition is associative, so the per-image
    // sums do not depend on the order in which the blocks retire (|partial| <= 512, <= 2^22 blocks per image)
    if (tid == 0) {
      atomicAdd(&fx_sums[2 * n], (unsigned long long)__double2ll_rn((double)s * 4294967296.0));
      atomicAdd(&fx_sums[2 * n + 1], (unsigned long long)__double2ll_rn((double)c * 4294967296.0));
    }
  }
}

__global__ void ssim_sums_finish_kernel(const unsigned long long* __restrict__ fx, int N, float* __restrict__ ssim_sum,
                                        float* __restrict__ cs_sum) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  ssim_sum[n] = (float)((double)(long long)fx[2 * n] * (1.0 / 4294967296.0));
  cs_sum[n] = (float)((double)(long long)fx[2 * n + 1] * (1.0 / 4294967296.0));
}

// backward filter: g_img2(p) = in_scale * ( FA + 2*x2*FB + x1*FC ),  F* = full correlation with g (x) g
__global__ void __launch_bounds__(256)
ssim_bwd_filter_kernel(const float* __restrict__ img1, const float* __restrict__ img2, int H, int W,
                       const float* __restrict__ filt, int fs, float in_scale, const float* __restrict__ maps,
                       int64_t msz, float* __restrict__ g_img2, int accumulate) {
  __shared__ float pm[3][SS_PH][SS_PW + 1];
  __shared__ float hb[3][SS_PH][SS_TW + 1];
  __shared__ float f[SS_MAXF];
  const int plane = blockIdx.z;
  const int OH = H - fs + 1, OW = W - fs + 1;
  const int x0 = blockIdx.x * SS_TW, y0 = blockIdx.y * SS_TH;
  const int tid = threadIdx.y * 32 + threadIdx.x;
  if (tid < fs) f[tid] = filt[tid];
  const int ph = SS_TH + fs - 1, pw = SS_TW + fs - 1;
  // patch covers map coords [y0-(fs-1), y0+TH) x [x0-(fs-1), x0+TW)
  for (int i = tid; i < ph * pw; i += 256) {
    int r = i / pw, c = i % pw;
    int qy = y0 - (fs - 1) + r, qx = x0 - (fs - 1) + c;
    bool ok = (qy >= 0 && qy < OH && qx >= 0 && qx < OW);
    int64_t o = (int64_t)plane * OH * OW + (int64_t)qy * OW + qx;
    pm[0][r][c] = ok ? maps[o] : 0.f;
    pm[1][r][c] = ok ? maps[msz + o] : 0.f;
    pm[2][r][c] = ok ? maps[2 * msz + o] : 0.f;
  }
  __syncthreads();
  // horizontal: out col c (pixel x0+c) = sum_d f[d] * M(x0+c-d) ; patch col of (x0+c-d) = c + fs-1 - d
  for (int i = tid; i < ph * SS_TW; i += 256) {
    int r = i / SS_TW, c = i % SS_TW;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f;
    for (int d = 0; d < fs; ++d) {
      float w = f[d];
      int cc = c + fs - 1 - d;
      s0 += w * pm[0][r][cc]; s1 += w * pm[1][r][cc]; s2 += w * pm[2][r][cc];
    }
    hb[0][r][c] = s0; hb[1][r][c] = s1; hb[2][r][c] = s2;
  }
  __syncthreads();
  const float* a = img1 + (int64_t)plane * H * W;
  const float* b = img2 + (int64_t)plane * H * W;
  float* g = g_img2 + (int64_t)plane * H * W;
#pragma unroll
  for (int rep = 0; rep < 2; ++rep) {
    int ty = threadIdx.y + rep * 8, tx = threadIdx.x;
    int y = y0 + ty, x = x0 + tx;
    if (y < H && x < W) {
      float s0 = 0.f, s1 = 0.f, s2 = 0.f;
      for (int d = 0; d < fs; ++d) {
        float w = f[d];
        int rr = ty + fs - 1 - d;
        s0 += w * hb[0][rr][tx]; s1 += w * hb[1][rr][tx]; s2 += w * hb[2][rr][tx];
      }
      int64_t o = (int64_t)y * W + x;
      float x1 = a[o] * in_scale, x2 = b[o] * in_scale;
      float v = in_scale * (s0 + 2.f * x2 * s1 + x1 * s2);
      g[o] = accumulate ? (g[o] + v) : v;
    }
  }
}

// ---------------------------------------------------------------------------------------
// Register-blocked SSIM level for the shipped 11-tap window (loss.py:57-66 / metric.py:47-56).  The kernels above
// issue one shared-memory load per multiply-add (~120 LDS per output pixel: bound by shared-memory bandwidth at 5% of
// HBM speed).  Here a block owns a 64x32 output tile (halo overhead 1.5x instead of 2.1x) and every thread produces
// 8 outputs along the filtered direction from 18 loaded values:
//   vertical pass   thread = (column, strip of 8 rows): 18 x 2 loads, the five products formed once per pixel;
//   horizontal pass thread = (row, group of 8 columns): 18 x 5 loads (lanes differ in the ROW: the odd row pitch
//                   makes the accesses conflict free), then the SSIM terms of its 8 pixels.
// ~16 LDS and ~120 FMA per output pixel.  The filter taps sit in registers.  Same outputs as MODE 0 / 1 above.
// ---------------------------------------------------------------------------------------
constexpr int S2_TW = 64, S2_TH = 32, S2_R = 8, S2_FS = 11;
constexpr int S2_PW = S2_TW + S2_FS - 1, S2_PH = S2_TH + S2_FS - 1, S2_PP = S2_PW | 1;     // 74, 42, pitch 75
constexpr int S2_SMEM_FWD = (2 * S2_PH + 5 * S2_TH) * S2_PP * (int)sizeof(float);
constexpr int S2_SMEM_BWD = (3 * S2_PH + 3 * S2_TH) * S2_PP * (int)sizeof(float);

template <int MODE>
__global__ void __launch_bounds__(256)
ssim_level_rb_kernel(const float* __restrict__ img1, const float* __restrict__ img2, int C, int H, int W,
                     const float* __restrict__ filt, float in_scale, float c1, float c2,
                     unsigned long long* __restrict__ fx_sums, const float* __restrict__ g_ssim,
                     const float* __restrict__ g_cs, float* __restrict__ maps) {
  extern __shared__ float s2m[];
  float* p1 = s2m;                          // [PH][PP]
  float* p2 = p1 + S2_PH * S2_PP;           // [PH][PP]
  float* vb = p2 + S2_PH * S2_PP;           // [5][TH][PP]   vertically filtered moments
  __shared__ float red[32];
  constexpr int FS = S2_FS, R = S2_R;
  const int plane = blockIdx.z, n = plane / C;
  const int OH = H - FS + 1, OW = W - FS + 1;
  const int ox0 = blockIdx.x * S2_TW, oy0 = blockIdx.y * S2_TH;
  const int tid = threadIdx.x;
  float f[FS];
#pragma unroll
  for (int k = 0; k < FS; ++k) f[k] = __ldg(filt + k);
  const float* a = img1 + (int64_t)plane * H * W;
  const float* b = img2 + (int64_t)plane * H * W;
  // patch load: a warp per row, lanes along x (no per-element division, one row pointer per row)
  for (int r = tid >> 5; r < S2_PH; r += 8) {
    const int gy = oy0 + r;
    const float* ar = a + (int64_t)gy * W + ox0;
    const float* br = b + (int64_t)gy * W + ox0;
#pragma unroll
    for (int c = tid & 31; c < S2_PW; c += 32) {
      const bool ok = gy < H && ox0 + c < W;
      p1[r * S2_PP + c] = ok ? ar[c] * in_scale : 0.f;
      p2[r * S2_PP + c] = ok ? br[c] * in_scale : 0.f;
    }
  }
  __syncthreads();
  // vertical pass
  for (int i = tid; i < S2_PW * (S2_TH / R); i += 256) {
    const int c = i % S2_PW, r0 = (i / S2_PW) * R;
    float acc[5][R];
#pragma unroll
    for (int m = 0; m < 5; ++m)
#pragma unroll
      for (int j = 0; j < R; ++j) acc[m][j] = 0.f;
#pragma unroll
    for (int rr = 0; rr < R + FS - 1; ++rr) {
      const float u = p1[(r0 + rr) * S2_PP + c], v = p2[(r0 + rr) * S2_PP + c];
      const float q[5] = {u, v, u * u, v * v, u * v};
#pragma unroll
      for (int j = 0; j < R; ++j) {
        const int k = rr - j;                       // compile-time after unrolling
        if (k >= 0 && k < FS) {
#pragma unroll
          for (int m = 0; m < 5; ++m) acc[m][j] += f[k] * q[m];
        }
      }
    }
#pragma unroll
    for (int m = 0; m < 5; ++m)
#pragma unroll
      for (int j = 0; j < R; ++j) vb[(m * S2_TH + r0 + j) * S2_PP + c] = acc[m][j];
  }
  __syncthreads();
  // horizontal pass + SSIM terms: S2_TH rows x (S2_TW / R) groups = 256 tasks
  float acc_s = 0.f, acc_c = 0.f, gs = 0.f, gc = 0.f;
  if (MODE == 1) { gs = g_ssim[n]; gc = g_cs[n]; }
  {
    const int row = tid % S2_TH, x0 = (tid / S2_TH) * R;
    float o[5][R];
#pragma unroll
    for (int m = 0; m < 5; ++m) {
      float w[R + FS - 1];
#pragma unroll
      for (int j = 0; j < R + FS - 1; ++j) w[j] = vb[(m * S2_TH + row) * S2_PP + x0 + j];
#pragma unroll
      for (int j = 0; j < R; ++j) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < FS; ++k) t += f[k] * w[j + k];
        o[m][j] = t;
      }
    }
    const int oy = oy0 + row;
    float mo[3][R];
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const int ox = ox0 + x0 + j;
      mo[0][j] = mo[1][j] = mo[2][j] = 0.f;
      if (oy < OH && ox < OW) {
        const SsimTerms t = ssim_terms(o[0][j], o[1][j], o[2][j], o[3][j], o[4][j], c1, c2, MODE == 1);
        if (MODE == 0) {
          acc_s += t.ssim;
          acc_c += t.cs;
        } else {
          mo[0][j] = gs * t.dA_s + gc * t.dA_c;
          mo[1][j] = gs * t.dB_s + gc * t.dB_c;
          mo[2][j] = gs * t.dC_s + gc * t.dC_c;
        }
      }
    }
    if (MODE == 1) {
      // lanes of the pass above differ in the ROW: stage the three maps in shared memory (over the moments, once
      // every thread has read them) and store them with the lanes along x
      __syncthreads();
#pragma unroll
      for (int m = 0; m < 3; ++m)
#pragma unroll
        for (int j = 0; j < R; ++j) vb[(m * S2_TH + row) * S2_PP + x0 + j] = mo[m][j];
      __syncthreads();
      const int64_t msz = (int64_t)gridDim.z * OH * OW;
      for (int r = tid >> 5; r < S2_TH; r += 8) {
        const int yy = oy0 + r;
        if (yy >= OH) break;
        float* mr = maps + (int64_t)plane * OH * OW + (int64_t)yy * OW + ox0;
#pragma unroll
        for (int x = tid & 31; x < S2_TW; x += 32)
          if (ox0 + x < OW) {
#pragma unroll
            for (int m = 0; m < 3; ++m) mr[m * msz + x] = vb[(m * S2_TH + r) * S2_PP + x];
          }
      }
    }
  }
  if (MODE == 0) {
    const float ss = block_sum(acc_s, red);
    const float cc = block_sum(acc_c, red);
    if (tid == 0) {        // 32.32 fixed point (|partial| <= 2048): order-independent integer accumulation
      atomicAdd(&fx_sums[2 * n], (unsigned long long)__double2ll_rn((double)ss * 4294967296.0));
      atomicAdd(&fx_sums[2 * n + 1], (unsigned long long)__double2ll_rn((double)cc * 4294967296.0));
    }
  }
}

// backward filter, register blocked the same way: g_img2(p) = in_scale * (FA + 2 x2 FB + x1 FC), F* = full
// correlation of the three maps with g (x) g
__global__ void __launch_bounds__(256)
ssim_bwd_filter_rb_kernel(const float* __restrict__ img1, const float* __restrict__ img2, int H, int W,
                          const float* __restrict__ filt, float in_scale, const float* __restrict__ maps, int64_t msz,
                          float* __restrict__ g_img2, int accumulate) {
  extern __shared__ float s2m[];
  float* pm = s2m;                          // [3][PH][PP]
  float* vb = pm + 3 * S2_PH * S2_PP;       // [3][TH][PP]
  constexpr int FS = S2_FS, R = S2_R;
  const int plane = blockIdx.z;
  const int OH = H - FS + 1, OW = W - FS + 1;
  const int x0t = blockIdx.x * S2_TW, y0t = blockIdx.y * S2_TH;
  const int tid = threadIdx.x;
  float f[FS];
#pragma unroll
  for (int k = 0; k < FS; ++k) f[k] = __ldg(filt + k);
  // patch covers map coords [y0-(FS-1), y0+TH) x [x0-(FS-1), x0+TW)
  for (int r = tid >> 5; r < S2_PH; r += 8) {
    const int qy = y0t - (FS - 1) + r;
    const bool oky = qy >= 0 && qy < OH;
    const float* m0 = maps + (int64_t)plane * OH * OW + (int64_t)qy * OW + (x0t - (FS - 1));
#pragma unroll
    for (int c = tid & 31; c < S2_PW; c += 32) {
      const int qx = x0t - (FS - 1) + c;
      const bool ok = oky && qx >= 0 && qx < OW;
      pm[(0 * S2_PH + r) * S2_PP + c] = ok ? m0[c] : 0.f;
      pm[(1 * S2_PH + r) * S2_PP + c] = ok ? m0[msz + c] : 0.f;
      pm[(2 * S2_PH + r) * S2_PP + c] = ok ? m0[2 * msz + c] : 0.f;
    }
  }
  __syncthreads();
  // vertical: out row j (pixel y0+j) = sum_d f[d] M(y0+j-d): patch row of (y0+j-d) = j + FS-1 - d
  for (int i = tid; i < S2_PW * (S2_TH / R); i += 256) {
    const int c = i % S2_PW, r0 = (i / S2_PW) * R;
    float acc[3][R];
#pragma unroll
    for (int m = 0; m < 3; ++m)
#pragma unroll
      for (int j = 0; j < R; ++j) acc[m][j] = 0.f;
#pragma unroll
    for (int rr = 0; rr < R + FS - 1; ++rr) {
      float q[3];
#pragma unroll
      for (int m = 0; m < 3; ++m) q[m] = pm[(m * S2_PH + r0 + rr) * S2_PP + c];
#pragma unroll
      for (int j = 0; j < R; ++j) {
        const int d = j + FS - 1 - rr;
        if (d >= 0 && d < FS) {
#pragma unroll
          for (int m = 0; m < 3; ++m) acc[m][j] += f[d] * q[m];
        }
      }
    }
#pragma unroll
    for (int m = 0; m < 3; ++m)
#pragma unroll
      for (int j = 0; j < R; ++j) vb[(m * S2_TH + r0 + j) * S2_PP + c] = acc[m][j];
  }
  __syncthreads();
  {
    const int row = tid % S2_TH, xg = (tid / S2_TH) * R;
    float o[3][R];
#pragma unroll
    for (int m = 0; m < 3; ++m) {
      float w[R + FS - 1];
#pragma unroll
      for (int j = 0; j < R + FS - 1; ++j) w[j] = vb[(m * S2_TH + row) * S2_PP + xg + j];
#pragma unroll
      for (int j = 0; j < R; ++j) {
        float t = 0.f;
#pragma unroll
        for (int d = 0; d < FS; ++d) t += f[d] * w[j + FS - 1 - d];
        o[m][j] = t;
      }
    }
    // stage the filtered maps (lanes differ in the row) and finish with the lanes along x: coalesced image loads
    // and gradient stores
    __syncthreads();
#pragma unroll
    for (int m = 0; m < 3; ++m)
#pragma unroll
      for (int j = 0; j < R; ++j) vb[(m * S2_TH + row) * S2_PP + xg + j] = o[m][j];
    __syncthreads();
  }
  for (int r = tid >> 5; r < S2_TH; r += 8) {
    const int y = y0t + r;
    if (y >= H) break;
    const int64_t ro = (int64_t)plane * H * W + (int64_t)y * W + x0t;
#pragma unroll
    for (int xl = tid & 31; xl < S2_TW; xl += 32)
      if (x0t + xl < W) {
        const float x1 = img1[ro + xl] * in_scale, x2 = img2[ro + xl] * in_scale;
        const float v = in_scale * (vb[(0 * S2_TH + r) * S2_PP + xl] + 2.f * x2 * vb[(1 * S2_TH + r) * S2_PP + xl] +
                                    x1 * vb[(2 * S2_TH + r) * S2_PP + xl]);
        g_img2[ro + xl] = accumulate ? (g_img2[ro + xl] + v) : v;
      }
  }
}

static int ssim_rb_attrs() {
  static HpcOncePerDevice once;
  int dev;
  if (once.need(&dev)) {
    HPC_CUDA(cudaFuncSetAttribute(ssim_level_rb_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, S2_SMEM_FWD));
    HPC_CUDA(cudaFuncSetAttribute(ssim_level_rb_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, S2_SMEM_FWD));
    HPC_CUDA(cudaFuncSetAttribute(ssim_bwd_filter_rb_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, S2_SMEM_BWD));
    once.mark(dev);
  }
  return HPC_OK;
}
// the register-blocked kernels serve the 11-tap window on planes that give their 64x32 tiles something to do
static bool ssim_use_rb(int H, int W, int fs) {
  static int off = -1;
  if (off < 0) { const char* e = getenv("HPC_SSIM_RB"); off = (e && e[0] == '0') ? 1 : 0; }
  return !off && fs == S2_FS && H >= 24 && W >= 24;
}

static int ssim_check(int N, int C, int H, int W, int fs) {
  HPC_REQUIRE(N > 0 && C > 0, HPC_ERR_BAD_SHAPE, "ssim: empty batch");
  HPC_REQUIRE(fs >= 1 && fs <= SS_MAXF, HPC_ERR_UNSUPPORTED, "ssim: filter size %d not in [1,%d]", fs, SS_MAXF);
  HPC_REQUIRE(H >= fs && W >= fs, HPC_ERR_BAD_SHAPE, "ssim: image %dx%d smaller than the %d-tap window", H, W, fs);
  HPC_REQUIRE((int64_t)N * C <= 65535, HPC_ERR_UNSUPPORTED, "ssim: N*C=%lld > 65535", (long long)N * C);
  return HPC_OK;
}

int hpc_ssim_level_fwd(const float* img1, const float* img2, int N, int C, int H, int W, const float* filt1d,
                       int fs, float in_scale, float c1, float c2, float* ssim_sum, float* cs_sum,
                       void* scratch, hpc_stream_t stream) {
  int rc = ssim_check(N, C, H, W, fs);
  if (rc) return rc;
  HPC_REQUIRE(scratch && ((uintptr_t)scratch & 7) == 0, HPC_ERR_BAD_ALIGNMENT,
              "hpc_ssim_level_fwd: scratch (16*N bytes, 8B aligned) is required");
  auto st = (cudaStream_t)stream;
  auto* fx = (unsigned long long*)scratch;
  HPC_CUDA(cudaMemsetAsync(fx, 0, (size_t)N * 16, st));
  const int OH = H - fs + 1, OW = W - fs + 1;
  if (ssim_use_rb(H, W, fs)) {
    if ((rc = ssim_rb_attrs())) return rc;
    dim3 grid((OW + S2_TW - 1) / S2_TW, (OH + S2_TH - 1) / S2_TH, N * C);
    ssim_level_rb_kernel<0><<<grid, 256, S2_SMEM_FWD, st>>>(img1, img2, C, H, W, filt1d, in_scale, c1, c2, fx, nullptr,
                                                           nullptr, nullptr);
  } else {
    dim3 grid((OW + SS_TW - 1) / SS_TW, (OH + SS_TH - 1) / SS_TH, N * C);
    ssim_level_kernel<0><<<grid, dim3(32, 8), 0, st>>>(img1, img2, C, H, W, filt1d, fs, in_scale, c1, c2, fx, nullptr,
                                                      nullptr, nullptr);
  }
  HPC_LAUNCH_CHECK();
  ssim_sums_finish_kernel<<<(N + 127) / 128, 128, 0, st>>>(fx, N, ssim_sum, cs_sum);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

int hpc_ssim_level_bwd(const float* img1, const float* img2, int N, int C, int H, int W, const float* filt1d,
                       int fs, float in_scale, float c1, float c2, const float* g_ssim, const float* g_cs,
                       float* maps, float* g_img2, int accumulate, hpc_stream_t stream) {
  int rc = ssim_check(N, C, H, W, fs);
  if (rc) return rc;
  auto st = (cudaStream_t)stream;
  const int OH = H - fs + 1, OW = W - fs + 1;
  if (ssim_use_rb(H, W, fs)) {
    if ((rc = ssim_rb_attrs())) return rc;
    dim3 grid((OW + S2_TW - 1) / S2_TW, (OH + S2_TH - 1) / S2_TH, N * C);
    ssim_level_rb_kernel<1><<<grid, 256, S2_SMEM_FWD, st>>>(img1, img2, C, H, W, filt1d, in_scale, c1, c2, nullptr,
                                                           g_ssim, g_cs, maps);
    HPC_LAUNCH_CHECK();
    dim3 grid2((W + S2_TW - 1) / S2_TW, (H + S2_TH - 1) / S2_TH, N * C);
    ssim_bwd_filter_rb_kernel<<<grid2, 256, S2_SMEM_BWD, st>>>(img1, img2, H, W, filt1d, in_scale, maps,
                                                               (int64_t)N * C * OH * OW, g_img2, accumulate);
    HPC_LAUNCH_CHECK();
    return HPC_OK;
  }
  dim3 grid((OW + SS_TW - 1) / SS_TW, (OH + SS_TH - 1) / SS_TH, N * C);
  ssim_level_kernel<1><<<grid, dim3(32, 8), 0, st>>>(img1, img2, C, H, W, filt1d, fs, in_scale, c1, c2, nullptr,
                                                    g_ssim, g_cs, maps);
  HPC_LAUNCH_CHECK();
  dim3 grid2((W + SS_TW - 1) / SS_TW, (H + SS_TH - 1) / SS_TH, N * C);
  ssim_bwd_filter_kernel<<<grid2, dim3(32, 8), 0, st>>>(img1, img2, H, W, filt1d, fs, in_scale, maps,
                                                        (int64_t)N * C * OH * OW, g_img2, accumulate);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

// ---------------------------------------------------------------------------------------
// 2x2 average pool with reflect padding of an odd last row/column
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ int reflect_idx(int i, int n) { return i < n ? i : 2 * n - 2 - i; }

// grid = (ceil(OW / 256), OH, planes): no 64-bit index arithmetic per element
__global__ void avgpool2_fwd_kernel(const float* __restrict__ in, float* __restrict__ out, int H, int W, int OH, int OW,
                                    float scale) {
  const int ox = blockIdx.x * blockDim.x + threadIdx.x, oy = blockIdx.y;
  if (ox >= OW) return;
  const float* p = in + (int64_t)blockIdx.z * H * W;
  const int y0 = 2 * oy, y1 = reflect_idx(2 * oy + 1, H), x0 = 2 * ox, x1 = reflect_idx(2 * ox + 1, W);
  // F.avg_pool2d sums the window then divides
  const float s = p[(int64_t)y0 * W + x0] + p[(int64_t)y0 * W + x1] + p[(int64_t)y1 * W + x0] + p[(int64_t)y1 * W + x1];
  out[((int64_t)blockIdx.z * OH + oy) * OW + ox] = scale * (s * 0.25f);
}
// gather form of the backward: each input pixel collects from the (<=2x2) outputs it fed
__global__ void avgpool2_bwd_kernel(const float* __restrict__ gout, float* __restrict__ gin, int H, int W, int OH,
                                    int OW, float scale, int accumulate) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  const float* g = gout + (int64_t)blockIdx.z * OH * OW;
  const int oy = y >> 1, ox = x >> 1;
  float acc = g[(int64_t)oy * OW + ox];
  // rows of the padded image that map to y: y itself, plus the reflected row H (== H-2) when H is odd
  const bool ry = (H & 1) && (y == H - 2);  // padded row H reads row H-2, belongs to output row (H)/2 = OH-1
  const bool rx = (W & 1) && (x == W - 2);
  if (ry) acc += g[(int64_t)(OH - 1) * OW + ox];
  if (rx) acc += g[(int64_t)oy * OW + (OW - 1)];
  if (ry && rx) acc += g[(int64_t)(OH - 1) * OW + (OW - 1)];
  const float v = scale * 0.25f * acc;
  const int64_t i = ((int64_t)blockIdx.z * H + y) * W + x;
  gin[i] = accumulate ? gin[i] + v : v;
}
int hpc_avgpool2_fwd(const float* in, float* out, int NC, int H, int W, float scale, hpc_stream_t stream) {
  HPC_REQUIRE(NC > 0 && H >= 2 && W >= 2, HPC_ERR_BAD_SHAPE, "hpc_avgpool2_fwd: bad shape NC=%d H=%d W=%d", NC, H, W);
  int OH = (H + 1) / 2, OW = (W + 1) / 2;
  HPC_REQUIRE(NC <= 65535 && OH <= 65535, HPC_ERR_UNSUPPORTED, "hpc_avgpool2_fwd: NC=%d or OH=%d > 65535", NC, OH);
  const int th = OW >= 256 ? 256 : (OW >= 128 ? 128 : 64);
  avgpool2_fwd_kernel<<<dim3((OW + th - 1) / th, OH, NC), th, 0, (cudaStream_t)stream>>>(in, out, H, W, OH, OW, scale);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}
int hpc_avgpool2_bwd(const float* gout, float* gin, int NC, int H, int W, float scale, int accumulate,
                     hpc_stream_t stream) {
  HPC_REQUIRE(NC > 0 && H >= 2 && W >= 2, HPC_ERR_BAD_SHAPE, "hpc_avgpool2_bwd: bad shape");
  int OH = (H + 1) / 2, OW = (W + 1) / 2;
  HPC_REQUIRE(NC <= 65535 && H <= 65535, HPC_ERR_UNSUPPORTED, "hpc_avgpool2_bwd: NC=%d or H=%d > 65535", NC, H);
  const int th = W >= 256 ? 256 : (W >= 128 ? 128 : 64);
  avgpool2_bwd_kernel<<<dim3((W + th - 1) / th, H, NC), th, 0, (cudaStream_t)stream>>>(gout, gin, H, W, OH, OW, scale,
                                                                                    accumulate);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

// ---------------------------------------------------------------------------------------
// per-image level means -> loss / metric (tiny; one thread per image, fixed-order mean)
// ---------------------------------------------------------------------------------------
struct CombineArgs {
  int L, N, kind, flags;
  float eps;
  float inv_count[8];
  float w[8];
};

__global__ void msssim_combine_fwd_kernel(const float* __restrict__ sums, CombineArgs a, float* __restrict__ out) {
  __shared__ float red[32];
  float my = 0.f;
  const int L = a.L, N = a.N;
  for (int n = threadIdx.x; n < N; n += blockDim.x) {
    float r;
    if (a.kind == HPC_MSSSIM_LOSS_LOG) {
      r = 0.f;
      for (int l = 0; l < L; ++l) {
        float v = sums[(l * 2 + (l == L - 1 ? 0 : 1)) * N + n] * a.inv_count[l];
        v = fmaxf(v, a.eps);  // LowerBound(eps), loss.py:93-98
        r = r + logf(v) * a.w[l];
      }
      my += r;
    } else if (a.kind == HPC_MSSSIM_LOSS_PLAIN || a.kind == HPC_MSSSIM_METRIC) {
      r = 1.f;
      for (int l = 0; l < L; ++l) {
        float v = sums[(l * 2 + (l == L - 1 ? 0 : 1)) * N + n] * a.inv_count[l];
        v = fmaxf(v, 0.f);  // LowerBound(0) (loss) / clamp_min_(0) (metric)
        r = r * powf(v, a.w[l]);
      }
      if (a.kind == HPC_MSSSIM_METRIC) {
        out[n] = (a.flags & 1) ? (-10.f * logf(1.f - r) / 2.302585092994046f) : r;
      } else {
        my += r;
      }
    } else if (a.kind == HPC_SSIM_LOSS) {
      float v = sums[n] * a.inv_count[0];
      if (a.flags & 1) {
        out[n] = -logf(fmaxf(v, a.eps));  // loss.py:72 returns -ssim (log) per image
      } else {
        my += fmaxf(v, 0.f);
      }
    } else {  // HPC_SSIM_METRIC
      float v = sums[n] * a.inv_count[0];
      if (a.flags & 2) v = fmaxf(v, 0.f);
      out[n] = (a.flags & 1) ? (-10.f * logf(1.f - v) / 2.302585092994046f) : v;
    }
  }
  bool scalar = (a.kind == HPC_MSSSIM_LOSS_LOG) || (a.kind == HPC_MSSSIM_LOSS_PLAIN) ||
                (a.kind == HPC_SSIM_LOSS && !(a.flags & 1));
  if (scalar) {
    float s = block_sum(my, red);
    if (threadIdx.x == 0) {
      float mean = s / (float)N;
      out[0] = (a.kind == HPC_MSSSIM_LOSS_LOG) ? -mean : (1.f - mean);
    }
  }
}

__global__ void msssim_combine_bwd_kernel(const float* __restrict__ sums, CombineArgs a, const float* __restrict__ g_out,
                                          float g_scale, float* __restrict__ g_sums) {
  const int L = a.L, N = a.N;
  for (int n = threadIdx.x; n < N; n += blockDim.x) {
    for (int l = 0; l < L; ++l) { g_sums[(l * 2) * N + n] = 0.f; g_sums[(l * 2 + 1) * N + n] = 0.f; }
    if (a.kind == HPC_MSSSIM_LOSS_LOG) {
      // loss = -(1/N) sum_n sum_l w_l log(LB(v_ln, eps)); LowerBound passes grad if v>=eps or grad<0
      float go = g_out[0] * g_scale;
      for (int l = 0; l < L; ++l) {
        int idx = (l * 2 + (l == L - 1 ? 0 : 1)) * N + n;
        float v = sums[idx] * a.inv_count[l];
        float vb = fmaxf(v, a.eps);
        float g = go * (-1.f / (float)N) * a.w[l] / vb;  // dL/d(bounded mean)
        bool pass = (v >= a.eps) || (g < 0.f);
        g_sums[idx] = pass ? g * a.inv_count[l] : 0.f;
      }
    } else if (a.kind == HPC_MSSSIM_LOSS_PLAIN) {
      float go = g_out[0] * g_scale;
      float vb[8], r = 1.f;
      for (int l = 0; l < L; ++l) {
        int idx = (l * 2 + (l == L - 1 ? 0 : 1)) * N + n;
        vb[l] = fmaxf(sums[idx] * a.inv_count[l], 0.f);
        r *= powf(vb[l], a.w[l]);
      }
      for (int l = 0; l < L; ++l) {
        int idx = (l * 2 + (l == L - 1 ? 0 : 1)) * N + n;
        float v = sums[idx] * a.inv_count[l];
        // d r / d vb_l = w_l * r / vb_l   (computed as product of the other factors for vb=0 safety)
        float others = 1.f;
        for (int k = 0; k < L; ++k) if (k != l) others *= powf(vb[k], a.w[k]);
        float dr = others * a.w[l] * powf(vb[l], a.w[l] - 1.f);
        float g = go * (-1.f / (float)N) * dr;
        bool pass = (v >= 0.f) || (g < 0.f);
        g_sums[idx] = pass ? g * a.inv_count[l] : 0.f;
      }
      (void)r;
    } else if (a.kind == HPC_SSIM_LOSS) {
      float v = sums[n] * a.inv_count[0];
      if (a.flags & 1) {
        float vb = fmaxf(v, a.eps);
        float g = g_out[n] * g_scale * (-1.f / vb);
        bool pass = (v >= a.eps) || (g < 0.f);
        g_sums[n] = pass ? g * a.inv_count[0] : 0.f;
      } else {
        float g = g_out[0] * g_scale * (-1.f / (float)N);
        bool pass = (v >= 0.f) || (g < 0.f);
        g_sums[n] = pass ? g * a.inv_count[0] : 0.f;
      }
    }
  }
}

static int fill_combine(CombineArgs* a, const double* counts_host, int L, int N, const float* weights_host, int kind,
                        int flags, float eps) {
  HPC_REQUIRE(L >= 1 && L <= 8, HPC_ERR_UNSUPPORTED, "msssim_combine: %d levels (max 8)", L);
  HPC_REQUIRE(N >= 1, HPC_ERR_BAD_SHAPE, "msssim_combine: empty batch");
  a->L = L; a->N = N; a->kind = kind; a->flags = flags; a->eps = eps;
  for (int l = 0; l < L; ++l) {
    a->inv_count[l] = (float)(1.0 / counts_host[l]);
    a->w[l] = weights_host ? weights_host[l] : 1.f;
  }
  return HPC_OK;
}
int hpc_msssim_combine_fwd(const float* sums, const double* counts_host, int L, int N, const float* weights_host,
                           int kind, int flags, float eps, float* out, hpc_stream_t stream) {
  CombineArgs a;
  int rc = fill_combine(&a, counts_host, L, N, weights_host, kind, flags, eps);
  if (rc) return rc;
  msssim_combine_fwd_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(sums, a, out);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}
int hpc_msssim_combine_bwd(const float* sums, const double* counts_host, int L, int N, const float* weights_host,
                           int kind, int flags, float eps, const float* g_out, float g_scale, float* g_sums,
                           hpc_stream_t stream) {
  CombineArgs a;
  int rc = fill_combine(&a, counts_host, L, N, weights_host, kind, flags, eps);
  if (rc) return rc;
  HPC_REQUIRE(kind == HPC_MSSSIM_LOSS_LOG || kind == HPC_MSSSIM_LOSS_PLAIN || kind == HPC_SSIM_LOSS,
              HPC_ERR_UNSUPPORTED, "msssim_combine_bwd: metrics have no backward");
  msssim_combine_bwd_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(sums, a, g_out, g_scale, g_sums);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

// ---------------------------------------------------------------------------------------
// rate-distortion combine (bmshl2018.py:88-96)
// ---------------------------------------------------------------------------------------
__global__ void rd_combine_kernel(const float* z_bits, const float* y_bits, hpc_dist_terms t, float lambda,
                                  float inv_np, float* out) {
  float zb = z_bits[0], yb = y_bits[0];
  float total_d = 0.f;
  for (int k = 0; k < t.n_terms; ++k) {
    float d = 0.f;
    for (int i = 0; i < t.n[k]; ++i) d += t.dist[k][i];
    d *= t.scale[k];
    out[4 + k] = d;
    total_d += d;
  }
  float ent = (zb + yb) * inv_np;  // (z_ce + y_ce) / num_pixels
  out[0] = zb * inv_np;
  out[1] = yb * inv_np;
  out[2] = ent;
  out[3] = lambda * total_d + ent;
}
int hpc_rd_combine(const float* z_bits, const float* y_bits, const hpc_dist_terms* terms, float lambda,
                   float inv_num_pixels, float* out8, hpc_stream_t stream) {
  HPC_REQUIRE(terms && terms->n_terms >= 0 && terms->n_terms <= HPC_MAX_DIST_TERMS, HPC_ERR_BAD_SHAPE,
              "hpc_rd_combine: bad number of distortion terms");
  rd_combine_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(z_bits, y_bits, *terms, lambda, inv_num_pixels, out8);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}
