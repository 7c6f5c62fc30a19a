// Thin-channel layers: the Cin=3 head of g_a (analysis.py:55, K=75) and the Cout=3 tail of g_s
// (synthesis.py:55-57).  Both are HBM-bound: the wide 192-channel tensor is streamed
// once; weights (14 400 floats) sit in registers / shared memory.
#include "hpc_common.cuh"

constexpr int C3_T = 8;                 // 8x8 pixel tile on the wide (low-res) grid
constexpr int C3_P = 2 * C3_T + 3;      // 19x19 thin patch
constexpr int C3_MAXS = 4;              // thin channels supported (3 in the model)

// ---------------------------------------------------------------------------------------
// Conv2d(Cs->Cw, 5, s2, p2): out[n,i,j,cw] = b[cw] + sum_{cs,kh,kw} x[n,cs,2i-2+kh,2j-2+kw]*w[cw,cs,kh,kw]
// block = Cw-tile of threads (one cw each, 25*Cs weights in registers), loops over tiles
// ---------------------------------------------------------------------------------------
template <typename LP, int CS>
__global__ void __launch_bounds__(256)
conv_c3_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, int64_t w_sr, int64_t w_sc,
                   const float* __restrict__ bias, int N, int H, int W, int Cw, float* __restrict__ out_f32,
                   LP* __restrict__ out_lp, LP* __restrict__ out_sq) {
  __shared__ float patch[CS][C3_P][C3_P + 1];
  const int OH = H / 2, OW = W / 2;
  const int cw = blockIdx.y * blockDim.x + threadIdx.x;
  float wr[CS * 25];
  float b = 0.f;
  if (cw < Cw) {
#pragma unroll
    for (int cs = 0; cs < CS; ++cs)
#pragma unroll
      for (int k = 0; k < 25; ++k) wr[cs * 25 + k] = w[cw * w_sr + cs * w_sc + k];
    if (bias) b = bias[cw];
  }
  const int tiles_x = (OW + C3_T - 1) / C3_T, tiles_y = (OH + C3_T - 1) / C3_T;
  const int64_t ntiles = (int64_t)N * tiles_x * tiles_y;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int tx = (int)(tile % tiles_x), ty = (int)((tile / tiles_x) % tiles_y), n = (int)(tile / ((int64_t)tiles_x * tiles_y));
    const int i0 = ty * C3_T, j0 = tx * C3_T;
    __syncthreads();
    for (int e = threadIdx.x; e < CS * C3_P * C3_P; e += blockDim.x) {
      int c = e % C3_P, r = (e / C3_P) % C3_P, cs = e / (C3_P * C3_P);
      int gy = 2 * i0 - 2 + r, gx = 2 * j0 - 2 + c;
      patch[cs][r][c] = (gy >= 0 && gy < H && gx >= 0 && gx < W) ? x[(((int64_t)n * CS + cs) * H + gy) * W + gx] : 0.f;
    }
    __syncthreads();
    if (cw < Cw) {
      for (int pi = 0; pi < C3_T; ++pi) {
        const int i = i0 + pi;
        if (i >= OH) break;
        for (int pj = 0; pj < C3_T; ++pj) {
          const int j = j0 + pj;
          if (j >= OW) break;
          float acc = 0.f;
#pragma unroll
          for (int cs = 0; cs < CS; ++cs)
#pragma unroll
            for (int kh = 0; kh < 5; ++kh)
#pragma unroll
              for (int kw = 0; kw < 5; ++kw)
                acc = fmaf(patch[cs][2 * pi + kh][2 * pj + kw], wr[cs * 25 + kh * 5 + kw], acc);
          acc += b;
          const int64_t o = (((int64_t)n * OH + i) * OW + j) * Cw + cw;
          if (out_f32) out_f32[o] = acc;
          if (out_lp) st_from_float<LP>(out_lp, o, acc);
          if (out_sq) st_from_float<LP>(out_sq, o, acc * acc);
        }
      }
    }
  }
}

static thread_local double g_thin_work = 0.0;

int hpc_conv_c3_fwd(const float* x_nchw, const float* w, int64_t w_sr, int64_t w_sc, const float* bias, int N,
                    int Cs, int H, int W, int Cw, float* out_f32, void* out_lp, void* out_sq_lp, int lp_dtype,
                    hpc_stream_t stream) {
  HPC_REQUIRE(N > 0 && H > 0 && W > 0 && Cw > 0, HPC_ERR_BAD_SHAPE, "hpc_conv_c3_fwd: empty tensor");
  HPC_REQUIRE(Cs >= 1 && Cs <= C3_MAXS, HPC_ERR_UNSUPPORTED, "hpc_conv_c3_fwd: thin channels %d not in [1,%d]", Cs,
              C3_MAXS);
  HPC_REQUIRE(H % 2 == 0 && W % 2 == 0, HPC_ERR_BAD_SHAPE, "hpc_conv_c3_fwd: H, W must be even (got %dx%d)", H, W);
  int threads = Cw >= 192 ? 192 : ((Cw + 31) / 32) * 32;
  if (Cw > 192 && Cw % 256 == 0) threads = 256;
  int gy = (Cw + threads - 1) / threads;
  const int64_t ntiles = (int64_t)N * ((H / 2 + C3_T - 1) / C3_T) * ((W / 2 + C3_T - 1) / C3_T);
  int gx = (int)(ntiles < 148 * 6 ? ntiles : 148 * 6);
  dim3 grid(gx, gy);
  auto st = (cudaStream_t)stream;
  {
    const int lpb = lp_dtype == HPC_BF16 ? 2 : 4;
    const double opix = (double)N * (H / 2) * (W / 2) * Cw;
    g_thin_work = (double)N * Cs * H * W * 4 + opix * ((out_f32 ? 4 : 0) + (out_lp ? lpb : 0) + (out_sq_lp ? lpb : 0));
  }
  HpcProfScope prof(HPC_PROF_THIN, g_thin_work, stream);
#define LAUNCH(T, CS_)                                                                                            \
  conv_c3_fwd_kernel<T, CS_><<<grid, threads, 0, st>>>(x_nchw, w, w_sr, w_sc, bias, N, H, W, Cw, out_f32,         \
                                                        (T*)out_lp, (T*)out_sq_lp)
#define LAUNCH_T(CS_) do { if (lp_dtype == HPC_BF16) LAUNCH(__nv_bfloat16, CS_); else LAUNCH(float, CS_); } while (0)
  switch (Cs) {
    case 1: LAUNCH_T(1); break;
    case 2: LAUNCH_T(2); break;
    case 3: LAUNCH_T(3); break;
    default: LAUNCH_T(4); break;
  }
#undef LAUNCH_T
#undef LAUNCH
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

// ---------------------------------------------------------------------------------------
// ConvTranspose2d(Cw->Cs, 5, s2, p2, op1): out[n,cs,2i+a,2j+b] = bias[cs] +
//     sum_{kh=a(2),kw=b(2)} sum_cw x[n, i+(a+2-kh)/2, j+(b+2-kw)/2, cw] * w[cw,cs,kh,kw]
// One warp per INPUT-grid pixel (i,j): produces the 2x2 output pixels; lanes split channels.
// weights staged in shared memory as [k][cs][cw] (conflict-free across lanes).
// ---------------------------------------------------------------------------------------
template <typename LP, int CS>
__global__ void __launch_bounds__(256)
deconv_c3_fwd_kernel(const LP* __restrict__ x, const float* __restrict__ w, int64_t w_sr, int64_t w_sc,
                     const float* __restrict__ bias, int N, int Cw, int H, int W, float* __restrict__ out) {
  extern __shared__ float wsm[];  // [25][CS][Cw]
  for (int e = threadIdx.x; e < 25 * CS * Cw; e += blockDim.x) {
    int cw = e % Cw, cs = (e / Cw) % CS, k = e / (Cw * CS);
    wsm[e] = w[cw * w_sr + cs * w_sc + k];
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int64_t npix = (int64_t)N * H * W;
  const int OH = 2 * H, OW = 2 * W;
  for (int64_t pix = (int64_t)blockIdx.x * nwarps + warp; pix < npix; pix += (int64_t)gridDim.x * nwarps) {
    const int j = (int)(pix % W), i = (int)((pix / W) % H), n = (int)(pix / ((int64_t)W * H));
    float acc[2][2][CS];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int cs = 0; cs < CS; ++cs) acc[a][b][cs] = 0.f;
    // input neighbourhood rows i-1..i+1, cols j-1..j+1
#pragma unroll
    for (int di = -1; di <= 1; ++di) {
      const int ii = i + di;
      if (ii < 0 || ii >= H) continue;
#pragma unroll
      for (int dj = -1; dj <= 1; ++dj) {
        const int jj = j + dj;
        if (jj < 0 || jj >= W) continue;
        const LP* xr = x + (((int64_t)n * H + ii) * W + jj) * Cw;
        for (int cw = lane; cw < Cw; cw += 32) {
          const float xv = ld_as_float<LP>(xr, cw);
          // output row parity a uses kh = a + 2 - 2*di  (dh = di = (a+2-kh)/2)
#pragma unroll
          for (int a = 0; a < 2; ++a) {
            const int kh = a + 2 - 2 * di;
            if (kh < 0 || kh > 4) continue;
#pragma unroll
            for (int b = 0; b < 2; ++b) {
              const int kw = b + 2 - 2 * dj;
              if (kw < 0 || kw > 4) continue;
              const float* wp = wsm + ((kh * 5 + kw) * CS) * Cw + cw;
#pragma unroll
              for (int cs = 0; cs < CS; ++cs) acc[a][b][cs] = fmaf(xv, wp[cs * Cw], acc[a][b][cs]);
            }
          }
        }
      }
    }
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int cs = 0; cs < CS; ++cs) {
          float v = warp_sum(acc[a][b][cs]);
          if (lane == 0) {
            v += bias ? bias[cs] : 0.f;
            out[(((int64_t)n * CS + cs) * OH + (2 * i + a)) * OW + (2 * j + b)] = v;
          }
        }
  }
}

int hpc_deconv_c3_fwd(const void* x_nhwc, int lp_dtype, const float* w, int64_t w_sr, int64_t w_sc,
                      const float* bias, int N, int Cw, int H, int W, int Cs, float* out_nchw,
                      hpc_stream_t stream) {
  HPC_REQUIRE(N > 0 && H > 0 && W > 0 && Cw > 0, HPC_ERR_BAD_SHAPE, "hpc_deconv_c3_fwd: empty tensor");
  HPC_REQUIRE(Cs >= 1 && Cs <= C3_MAXS, HPC_ERR_UNSUPPORTED, "hpc_deconv_c3_fwd: thin channels %d not in [1,%d]", Cs,
              C3_MAXS);
  size_t smem = (size_t)25 * Cs * Cw * sizeof(float);
  HPC_REQUIRE(smem <= 200 * 1024, HPC_ERR_UNSUPPORTED, "hpc_deconv_c3_fwd: weights (%zu B) exceed shared memory", smem);
  const int64_t npix = (int64_t)N * H * W;
  int grid = (int)((npix + 7) / 8 < 148 * 4 ? (npix + 7) / 8 : 148 * 4);
  auto st = (cudaStream_t)stream;
  HpcProfScope prof(HPC_PROF_THIN,
                    (double)npix * Cw * (lp_dtype == HPC_BF16 ? 2 : 4) + (double)npix * 4 * Cs * 4, stream);
#define LAUNCH(T, CS_)                                                                                             \
  do {                                                                                                             \
    HPC_CUDA(cudaFuncSetAttribute(deconv_c3_fwd_kernel<T, CS_>, cudaFuncAttributeMaxDynamicSharedMemorySize,       \
                                  (int)smem));                                                                     \
    deconv_c3_fwd_kernel<T, CS_><<<grid, 256, smem, st>>>((const T*)x_nhwc, w, w_sr, w_sc, bias, N, Cw, H, W,      \
                                                          out_nchw);                                               \
  } while (0)
#define LAUNCH_T(CS_) do { if (lp_dtype == HPC_BF16) LAUNCH(__nv_bfloat16, CS_); else LAUNCH(float, CS_); } while (0)
  switch (Cs) {
    case 1: LAUNCH_T(1); break;
    case 2: LAUNCH_T(2); break;
    case 3: LAUNCH_T(3); break;
    default: LAUNCH_T(4); break;
  }
#undef LAUNCH_T
#undef LAUNCH
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

// ---------------------------------------------------------------------------------------
// weight gradient of both thin layers:
//   gw[cw,cs,kh,kw] = sum_{n,i,j} wide[n,i,j,cw] * thin[n,cs,2i-2+kh,2j-2+kw]
// stage 1: per-block register partials (75 per thread) over a grid-stride set of tiles
// stage 2: deterministic sum over blocks
// ---------------------------------------------------------------------------------------
template <typename LP, int CS>
__global__ void __launch_bounds__(256)
wgrad_c3_stage1(const LP* __restrict__ wide, const float* __restrict__ thin, int N, int Cw, int H, int W,
                float* __restrict__ part) {
  __shared__ float patch[CS][C3_P][C3_P + 1];
  const int TH = 2 * H, TW = 2 * W;
  const int cw = blockIdx.y * blockDim.x + threadIdx.x;
  float acc[CS * 25];
#pragma unroll
  for (int k = 0; k < CS * 25; ++k) acc[k] = 0.f;
  const int tiles_x = (W + C3_T - 1) / C3_T, tiles_y = (H + C3_T - 1) / C3_T;
  const int64_t ntiles = (int64_t)N * tiles_x * tiles_y;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int tx = (int)(tile % tiles_x), ty = (int)((tile / tiles_x) % tiles_y), n = (int)(tile / ((int64_t)tiles_x * tiles_y));
    const int i0 = ty * C3_T, j0 = tx * C3_T;
    __syncthreads();
    for (int e = threadIdx.x; e < CS * C3_P * C3_P; e += blockDim.x) {
      int c = e % C3_P, r = (e / C3_P) % C3_P, cs = e / (C3_P * C3_P);
      int gy = 2 * i0 - 2 + r, gx = 2 * j0 - 2 + c;
      patch[cs][r][c] = (gy >= 0 && gy < TH && gx >= 0 && gx < TW) ? thin[(((int64_t)n * CS + cs) * TH + gy) * TW + gx] : 0.f;
    }
    __syncthreads();
    if (cw < Cw) {
      for (int pi = 0; pi < C3_T; ++pi) {
        const int i = i0 + pi;
        if (i >= H) break;
        for (int pj = 0; pj < C3_T; ++pj) {
          const int j = j0 + pj;
          if (j >= W) break;
          const float v = ld_as_float<LP>(wide, (((int64_t)n * H + i) * W + j) * Cw + cw);
#pragma unroll
          for (int cs = 0; cs < CS; ++cs)
#pragma unroll
            for (int kh = 0; kh < 5; ++kh)
#pragma unroll
              for (int kw = 0; kw < 5; ++kw)
                acc[cs * 25 + kh * 5 + kw] = fmaf(v, patch[cs][2 * pi + kh][2 * pj + kw], acc[cs * 25 + kh * 5 + kw]);
        }
      }
    }
  }
  if (cw < Cw) {
    float* dst = part + ((int64_t)blockIdx.x * Cw + cw) * (CS * 25);
#pragma unroll
    for (int k = 0; k < CS * 25; ++k) dst[k] = acc[k];
  }
}
__global__ void wgrad_c3_stage2(const float* __restrict__ part, int nblocks, int Cw, int CS, float* __restrict__ gw,
                                int64_t w_sr, int64_t w_sc) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const int K = CS * 25;
  if (e >= Cw * K) return;
  const int cw = e / K, k = e % K;
  double s = 0.0;
  for (int b = 0; b < nblocks; ++b) s += (double)part[((int64_t)b * Cw + cw) * K + k];
  gw[cw * w_sr + (k / 25) * w_sc + (k % 25)] = (float)s;
}

static int wgrad_c3_blocks(int N, int H, int W) {
  const int64_t ntiles = (int64_t)N * ((H + C3_T - 1) / C3_T) * ((W + C3_T - 1) / C3_T);
  return (int)(ntiles < 148 * 2 ? ntiles : 148 * 2);
}
size_t hpc_wgrad_c3_scratch_bytes(int N, int Cw, int Cs, int H, int W) {
  return (size_t)wgrad_c3_blocks(N, H, W) * Cw * Cs * 25 * sizeof(float);
}
int hpc_wgrad_c3(const void* wide_nhwc, int lp_dtype, const float* thin_nchw, int N, int Cw, int Cs, int H, int W,
                 float* gw, int64_t w_sr, int64_t w_sc, void* scratch, size_t scratch_bytes,
                 hpc_stream_t stream) {
  HPC_REQUIRE(N > 0 && H > 0 && W > 0 && Cw > 0, HPC_ERR_BAD_SHAPE, "hpc_wgrad_c3: empty tensor");
  HPC_REQUIRE(Cs >= 1 && Cs <= C3_MAXS, HPC_ERR_UNSUPPORTED, "hpc_wgrad_c3: thin channels %d not in [1,%d]", Cs, C3_MAXS);
  HPC_REQUIRE(scratch && scratch_bytes >= hpc_wgrad_c3_scratch_bytes(N, Cw, Cs, H, W), HPC_ERR_BAD_SHAPE,
              "hpc_wgrad_c3: scratch too small");
  int threads = Cw >= 192 ? 192 : ((Cw + 31) / 32) * 32;
  if (Cw > 192 && Cw % 256 == 0) threads = 256;
  int gy = (Cw + threads - 1) / threads;
  int nblocks = wgrad_c3_blocks(N, H, W);
  dim3 grid(nblocks, gy);
  auto st = (cudaStream_t)stream;
  HpcProfScope prof(HPC_PROF_THIN,
                    (double)N * H * W * Cw * (lp_dtype == HPC_BF16 ? 2 : 4) + (double)N * Cs * 4 * H * W * 4, stream);
#define LAUNCH(T, CS_) \
  wgrad_c3_stage1<T, CS_><<<grid, threads, 0, st>>>((const T*)wide_nhwc, thin_nchw, N, Cw, H, W, (float*)scratch)
#define LAUNCH_T(CS_) do { if (lp_dtype == HPC_BF16) LAUNCH(__nv_bfloat16, CS_); else LAUNCH(float, CS_); } while (0)
  switch (Cs) {
    case 1: LAUNCH_T(1); break;
    case 2: LAUNCH_T(2); break;
    case 3: LAUNCH_T(3); break;
    default: LAUNCH_T(4); break;
  }
#undef LAUNCH_T
#undef LAUNCH
  HPC_LAUNCH_CHECK();
  wgrad_c3_stage2<<<(Cw * Cs * 25 + 255) / 256, 256, 0, st>>>((const float*)scratch, nblocks, Cw, Cs, gw, w_sr, w_sc);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}
