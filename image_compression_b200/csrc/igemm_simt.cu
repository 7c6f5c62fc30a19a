// CUDA-core (SIMT) engine of the implicit-GEMM convolution family plus the small helper
// kernels around it (weight packing, column sums, layout conversion, thin-channel layers).
// The SIMT engine accumulates in fp32 over fp32 or bf16 operands for ANY channel count; it is
// the exact-parity engine (dtype fp32) and the on-device cross-check of the tcgen05 engine.
#include "igemm_epilogue.cuh"

int hpc_igemm_fwd_umma(const hpc_igemm_fwd_params* p, hpc_stream_t stream);      // igemm_umma.cu
size_t hpc_igemm_fwd_umma_workspace_bytes(const hpc_igemm_fwd_params* p);
int hpc_igemm_wgrad_umma(const hpc_igemm_wgrad_params* p, hpc_stream_t stream);  // igemm_umma.cu
size_t hpc_igemm_wgrad_umma_workspace_bytes(const hpc_igemm_wgrad_params* p);

static inline int grid_for(int64_t n, int threads, int64_t max_blocks = 1 << 30) {
  int64_t b = (n + threads - 1) / threads;
  if (b < 1) b = 1;
  if (b > max_blocks) b = max_blocks;
  return (int)b;
}

// ---------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------
struct FwdGeom {
  int N, H, W, Ci, GH, GW, is, os, OH, OW, Co, nphases;
  hpc_igemm_phase phase[HPC_MAX_PHASES];
  int dh[HPC_MAX_TAPS], dw[HPC_MAX_TAPS];
};

template <typename T>
__global__ void __launch_bounds__(256)
igemm_fwd_simt_kernel(const T* __restrict__ x, const T* __restrict__ w, FwdGeom g, EpiArgs e, int64_t total) {
  int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int co = (int)(idx % g.Co);
  int64_t m = idx / g.Co;
  const int j = (int)(m % g.GW); m /= g.GW;
  const int i = (int)(m % g.GH); m /= g.GH;
  const int n = (int)(m % g.N);
  const int ph = (int)(m / g.N);
  const hpc_igemm_phase P = g.phase[ph];
  const int oh = i * g.os + P.oh_off, ow = j * g.os + P.ow_off;
  if (oh >= g.OH || ow >= g.OW) return;
  float acc = 0.f;
  for (int t = P.tap_begin; t < P.tap_begin + P.ntaps; ++t) {
    const int ih = i * g.is + g.dh[t], iw = j * g.is + g.dw[t];
    if (ih < 0 || ih >= g.H || iw < 0 || iw >= g.W) continue;
    const T* xr = x + (((int64_t)n * g.H + ih) * g.W + iw) * g.Ci;
    const T* wr = w + ((int64_t)t * g.Co + co) * g.Ci;
    float a = 0.f;
    for (int ci = 0; ci < g.Ci; ++ci) a = fmaf(ld_as_float<T>(xr, ci), ld_as_float<T>(wr, ci), a);
    acc += a;
  }
  const int64_t o = (((int64_t)n * g.OH + oh) * g.OW + ow) * g.Co + co;
  epi_store_scalar<T>(e, acc, o, co);
}

static int check_fwd(const hpc_igemm_fwd_params* p) {
  HPC_REQUIRE(p, HPC_ERR_BAD_SHAPE, "hpc_igemm_fwd: NULL params");
  HPC_REQUIRE(p->N > 0 && p->H > 0 && p->W > 0 && p->Ci > 0 && p->Co > 0 && p->GH > 0 && p->GW > 0 && p->OH > 0 &&
                  p->OW > 0,
              HPC_ERR_BAD_SHAPE, "hpc_igemm_fwd: empty tensor (N=%d H=%d W=%d Ci=%d Co=%d)", p->N, p->H, p->W, p->Ci,
              p->Co);
  HPC_REQUIRE(p->nphases >= 1 && p->nphases <= HPC_MAX_PHASES, HPC_ERR_BAD_SHAPE, "hpc_igemm_fwd: nphases=%d",
              p->nphases);
  HPC_REQUIRE(p->is >= 1 && p->os >= 1, HPC_ERR_BAD_SHAPE, "hpc_igemm_fwd: strides must be >= 1");
  for (int i = 0; i < p->nphases; ++i)
    HPC_REQUIRE(p->phase[i].ntaps >= 1 && p->phase[i].tap_begin >= 0 &&
                    p->phase[i].tap_begin + p->phase[i].ntaps <= HPC_MAX_TAPS,
                HPC_ERR_BAD_SHAPE, "hpc_igemm_fwd: phase %d tap range invalid", i);
  HPC_REQUIRE(p->dtype == HPC_F32 || p->dtype == HPC_BF16, HPC_ERR_UNSUPPORTED, "hpc_igemm_fwd: dtype %d", p->dtype);
  HPC_REQUIRE(p->x && p->w, HPC_ERR_BAD_SHAPE, "hpc_igemm_fwd: NULL operand");
  HPC_REQUIRE(p->out_f32 || p->out_lp || p->out_sq_lp || p->out_abs_lp, HPC_ERR_BAD_SHAPE,
              "hpc_igemm_fwd: no output requested");
  switch (p->epi) {
    case HPC_EPI_GDN: case HPC_EPI_IGDN: case HPC_EPI_RELU_MASK:
      HPC_REQUIRE(p->aux0, HPC_ERR_BAD_SHAPE, "hpc_igemm_fwd: epilogue %d needs aux0", p->epi); break;
    case HPC_EPI_GDN_BWD: case HPC_EPI_IGDN_BWD:
      HPC_REQUIRE(p->aux0 && p->aux1 && p->aux2, HPC_ERR_BAD_SHAPE, "hpc_igemm_fwd: GDN_BWD needs aux0..2"); break;
    case HPC_EPI_SIGN_ADD:
      HPC_REQUIRE(p->aux0 && p->aux1, HPC_ERR_BAD_SHAPE, "hpc_igemm_fwd: SIGN_ADD needs aux0, aux1"); break;
    case HPC_EPI_LINEAR: case HPC_EPI_RELU: case HPC_EPI_EXP_CLAMP: break;
    default: HPC_FAIL(HPC_ERR_UNSUPPORTED, "hpc_igemm_fwd: unknown epilogue %d", p->epi);
  }
  return HPC_OK;
}

int hpc_igemm_fwd(const hpc_igemm_fwd_params* p, hpc_stream_t stream) {
  int rc = check_fwd(p);
  if (rc) return rc;
  if (p->engine == HPC_ENGINE_UMMA) return hpc_igemm_fwd_umma(p, stream);
  HPC_REQUIRE(p->engine == HPC_ENGINE_SIMT, HPC_ERR_UNSUPPORTED, "hpc_igemm_fwd: unknown engine %d", p->engine);
  FwdGeom g;
  g.N = p->N; g.H = p->H; g.W = p->W; g.Ci = p->Ci; g.GH = p->GH; g.GW = p->GW; g.is = p->is; g.os = p->os;
  g.OH = p->OH; g.OW = p->OW; g.Co = p->Co; g.nphases = p->nphases;
  memcpy(g.phase, p->phase, sizeof(g.phase));
  memcpy(g.dh, p->dh, sizeof(g.dh));
  memcpy(g.dw, p->dw, sizeof(g.dw));
  EpiArgs e = make_epi_args(p);
  const int64_t total = (int64_t)p->nphases * p->N * p->GH * p->GW * p->Co;
  if (p->dtype == HPC_BF16)
    igemm_fwd_simt_kernel<__nv_bfloat16><<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)p->x, (const __nv_bfloat16*)p->w, g, e, total);
  else
    igemm_fwd_simt_kernel<float><<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(
        (const float*)p->x, (const float*)p->w, g, e, total);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

// ---------------------------------------------------------------------------------------
// weight gradient
// ---------------------------------------------------------------------------------------
struct WgradGeom {
  int N, GH, GW, R, HH, HW, Cc, hs, nphases;
  hpc_igemm_phase phase[HPC_MAX_PHASES];
  int dh[HPC_MAX_TAPS], dw[HPC_MAX_TAPS];
  int tap_phase[HPC_MAX_TAPS];
  int ntaps;
};

// block = 32x8 threads: 32 consecutive c, 8 consecutive r of one tap; pixels are split over
// blockIdx.y and reduced with atomics (fp32; the SIMT engine is a checker, not the product path)
template <typename T>
__global__ void __launch_bounds__(256)
igemm_wgrad_simt_kernel(const T* __restrict__ lo, const T* __restrict__ hi, WgradGeom g, float scale,
                        float* __restrict__ gw, int pix_per_block) {
  const int ctiles = (g.Cc + 31) / 32, rtiles = (g.R + 7) / 8;
  int b = blockIdx.x;
  const int ct = b % ctiles; b /= ctiles;
  const int rt = b % rtiles; b /= rtiles;
  const int t = b;
  const int c = ct * 32 + threadIdx.x, r = rt * 8 + threadIdx.y;
  const hpc_igemm_phase P = g.phase[g.tap_phase[t]];
  const int64_t npix = (int64_t)g.N * g.GH * g.GW;
  int64_t p0 = (int64_t)blockIdx.y * pix_per_block, p1 = p0 + pix_per_block;
  if (p1 > npix) p1 = npix;
  float acc = 0.f;
  if (c < g.Cc && r < g.R) {
    for (int64_t pix = p0; pix < p1; ++pix) {
      const int j = (int)(pix % g.GW);
      const int i = (int)((pix / g.GW) % g.GH);
      const int n = (int)(pix / ((int64_t)g.GW * g.GH));
      const int li = i + g.dh[t], lj = j + g.dw[t];
      const int hh = i * g.hs + P.oh_off, hw = j * g.hs + P.ow_off;
      if (li < 0 || li >= g.GH || lj < 0 || lj >= g.GW || hh >= g.HH || hw >= g.HW) continue;
      float a = ld_as_float<T>(lo, (((int64_t)n * g.GH + li) * g.GW + lj) * g.R + r);
      float bb = ld_as_float<T>(hi, (((int64_t)n * g.HH + hh) * g.HW + hw) * g.Cc + c);
      acc = fmaf(a, bb, acc);
    }
    atomicAdd(&gw[((int64_t)t * g.R + r) * g.Cc + c], acc * scale);
  }
}

static int check_wgrad(const hpc_igemm_wgrad_params* p) {
  HPC_REQUIRE(p, HPC_ERR_BAD_SHAPE, "hpc_igemm_wgrad: NULL params");
  HPC_REQUIRE(p->N > 0 && p->GH > 0 && p->GW > 0 && p->R > 0 && p->HH > 0 && p->HW > 0 && p->Cc > 0,
              HPC_ERR_BAD_SHAPE, "hpc_igemm_wgrad: empty tensor");
  HPC_REQUIRE(p->nphases >= 1 && p->nphases <= HPC_MAX_PHASES, HPC_ERR_BAD_SHAPE, "hpc_igemm_wgrad: nphases=%d",
              p->nphases);
  HPC_REQUIRE(p->hs >= 1, HPC_ERR_BAD_SHAPE, "hpc_igemm_wgrad: hs must be >= 1");
  HPC_REQUIRE(p->dtype == HPC_F32 || p->dtype == HPC_BF16, HPC_ERR_UNSUPPORTED, "hpc_igemm_wgrad: dtype %d", p->dtype);
  HPC_REQUIRE(p->lo && p->hi && p->gw_f32, HPC_ERR_BAD_SHAPE, "hpc_igemm_wgrad: NULL operand");
  int nt = 0;
  for (int i = 0; i < p->nphases; ++i) {
    HPC_REQUIRE(p->phase[i].ntaps >= 1 && p->phase[i].tap_begin == nt, HPC_ERR_BAD_SHAPE,
                "hpc_igemm_wgrad: phases must tile the tap list contiguously");
    nt += p->phase[i].ntaps;
  }
  HPC_REQUIRE(nt <= HPC_MAX_TAPS, HPC_ERR_BAD_SHAPE, "hpc_igemm_wgrad: %d taps > %d", nt, HPC_MAX_TAPS);
  return HPC_OK;
}

size_t hpc_igemm_fwd_workspace_bytes(const hpc_igemm_fwd_params* p) {
  if (p && p->engine == HPC_ENGINE_UMMA && p->dtype == HPC_BF16 && p->Ci % 64 == 0 && p->Ci > 0 && p->N > 0 &&
      p->GH > 0 && p->GW > 0 && p->nphases >= 1 && p->nphases <= HPC_MAX_PHASES)
    return hpc_igemm_fwd_umma_workspace_bytes(p);
  return 0;
}

size_t hpc_igemm_wgrad_workspace_bytes(const hpc_igemm_wgrad_params* p) {
  if (p && p->engine == HPC_ENGINE_UMMA) return hpc_igemm_wgrad_umma_workspace_bytes(p);
  return 0;
}

int hpc_igemm_wgrad(const hpc_igemm_wgrad_params* p, hpc_stream_t stream) {
  int rc = check_wgrad(p);
  if (rc) return rc;
  if (p->engine == HPC_ENGINE_UMMA) return hpc_igemm_wgrad_umma(p, stream);
  HPC_REQUIRE(p->engine == HPC_ENGINE_SIMT, HPC_ERR_UNSUPPORTED, "hpc_igemm_wgrad: unknown engine %d", p->engine);
  HPC_REQUIRE(p->out_mode == 0 && p->out_accumulate == 0, HPC_ERR_UNSUPPORTED,
              "hpc_igemm_wgrad: the SIMT engine writes the packed layout only (use hpc_unpack_wgrad)");
  WgradGeom g;
  g.N = p->N; g.GH = p->GH; g.GW = p->GW; g.R = p->R; g.HH = p->HH; g.HW = p->HW; g.Cc = p->Cc; g.hs = p->hs;
  g.nphases = p->nphases;
  memcpy(g.phase, p->phase, sizeof(g.phase));
  memcpy(g.dh, p->dh, sizeof(g.dh));
  memcpy(g.dw, p->dw, sizeof(g.dw));
  int nt = 0;
  for (int i = 0; i < p->nphases; ++i)
    for (int k = 0; k < p->phase[i].ntaps; ++k) g.tap_phase[nt++] = i;
  g.ntaps = nt;
  auto st = (cudaStream_t)stream;
  HPC_CUDA(cudaMemsetAsync(p->gw_f32, 0, (size_t)nt * p->R * p->Cc * sizeof(float), st));
  const int64_t npix = (int64_t)p->N * p->GH * p->GW;
  int ysplit = (int)((npix + 2047) / 2048);
  if (ysplit > 1024) ysplit = 1024;
  int ppb = (int)((npix + ysplit - 1) / ysplit);
  dim3 grid(nt * ((p->R + 7) / 8) * ((p->Cc + 31) / 32), ysplit);
  if (p->dtype == HPC_BF16)
    igemm_wgrad_simt_kernel<__nv_bfloat16><<<grid, dim3(32, 8), 0, st>>>(
        (const __nv_bfloat16*)p->lo, (const __nv_bfloat16*)p->hi, g, p->scale, p->gw_f32, ppb);
  else
    igemm_wgrad_simt_kernel<float><<<grid, dim3(32, 8), 0, st>>>((const float*)p->lo, (const float*)p->hi, g, p->scale,
                                                                p->gw_f32, ppb);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

// ---------------------------------------------------------------------------------------
// weight packing
// ---------------------------------------------------------------------------------------
struct TapMap { int src[HPC_MAX_TAPS]; };

template <typename T>
__global__ void pack_weight_kernel(const float* __restrict__ src, T* __restrict__ dst, int ntaps, int rows, int cols,
                                   int64_t sr, int64_t sc, TapMap tm) {
  const int64_t total = (int64_t)ntaps * rows * cols;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int c = (int)(i % cols);
    int r = (int)((i / cols) % rows);
    int t = (int)(i / ((int64_t)cols * rows));
    st_from_float<T>(dst, i, src[r * sr + c * sc + tm.src[t]]);
  }
}
__global__ void unpack_wgrad_kernel(const float* __restrict__ src, float* __restrict__ dst, int ntaps, int rows,
                                    int cols, int64_t sr, int64_t sc, TapMap tm, int accumulate) {
  const int64_t total = (int64_t)ntaps * rows * cols;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int c = (int)(i % cols);
    int r = (int)((i / cols) % rows);
    int t = (int)(i / ((int64_t)cols * rows));
    float* d = dst + r * sr + c * sc + tm.src[t];
    *d = accumulate ? (*d + src[i]) : src[i];
  }
}
int hpc_pack_weight(const float* src, void* dst, int dtype, int ntaps, int rows, int cols, int64_t stride_r,
                    int64_t stride_c, const int* tap_src_host, hpc_stream_t stream) {
  HPC_REQUIRE(ntaps >= 1 && ntaps <= HPC_MAX_TAPS && rows > 0 && cols > 0, HPC_ERR_BAD_SHAPE,
              "hpc_pack_weight: bad shape taps=%d rows=%d cols=%d", ntaps, rows, cols);
  TapMap tm;
  for (int i = 0; i < ntaps; ++i) tm.src[i] = tap_src_host[i];
  const int64_t total = (int64_t)ntaps * rows * cols;
  int grid = grid_for(total, 256, 148 * 8);
  if (dtype == HPC_BF16)
    pack_weight_kernel<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst, ntaps, rows,
                                                                             cols, stride_r, stride_c, tm);
  else if (dtype == HPC_F32)
    pack_weight_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(src, (float*)dst, ntaps, rows, cols, stride_r,
                                                                     stride_c, tm);
  else
    HPC_FAIL(HPC_ERR_UNSUPPORTED, "hpc_pack_weight: dtype %d", dtype);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}
int hpc_unpack_wgrad(const float* src, float* dst, int ntaps, int rows, int cols, int64_t stride_r,
                     int64_t stride_c, const int* tap_src_host, int accumulate, hpc_stream_t stream) {
  HPC_REQUIRE(ntaps >= 1 && ntaps <= HPC_MAX_TAPS && rows > 0 && cols > 0, HPC_ERR_BAD_SHAPE,
              "hpc_unpack_wgrad: bad shape");
  TapMap tm;
  for (int i = 0; i < ntaps; ++i) tm.src[i] = tap_src_host[i];
  const int64_t total = (int64_t)ntaps * rows * cols;
  unpack_wgrad_kernel<<<grid_for(total, 256, 148 * 8), 256, 0, (cudaStream_t)stream>>>(src, dst, ntaps, rows, cols,
                                                                                     stride_r, stride_c, tm,
                                                                                     accumulate);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

// ---------------------------------------------------------------------------------------
// column sums: out[c] = scale * sum_p x[p][c]   (deterministic two-stage)
// ---------------------------------------------------------------------------------------
static constexpr int kColsumParts = 148 * 4;  // row blocks in stage 1

// generic (any C): one thread per column, rows split over blockIdx.y
template <typename T>
__global__ void colsum_stage1(const T* __restrict__ x, int64_t P, int C, float* __restrict__ part) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const int64_t rows_per = (P + gridDim.y - 1) / gridDim.y;
  int64_t r0 = blockIdx.y * rows_per, r1 = r0 + rows_per;
  if (r1 > P) r1 = P;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  int64_t r = r0;
  for (; r + 3 < r1; r += 4) {
    a0 += ld_as_float<T>(x, r * C + c);
    a1 += ld_as_float<T>(x, (r + 1) * C + c);
    a2 += ld_as_float<T>(x, (r + 2) * C + c);
    a3 += ld_as_float<T>(x, (r + 3) * C + c);
  }
  for (; r < r1; ++r) a0 += ld_as_float<T>(x, r * C + c);
  part[(int64_t)blockIdx.y * C + c] = (a0 + a1) + (a2 + a3);
}
// vectorised bf16 (C % 8 == 0, C <= 2048): a thread owns one 16-byte unit column and every RL-th row of the
// block's row range; 256 threads = (C/8 units) x RL row lanes; lanes are combined through shared memory
__global__ void __launch_bounds__(256)
colsum_stage1_bf16x8(const __nv_bfloat16* __restrict__ x, int64_t P, int C, float* __restrict__ part) {
  extern __shared__ float csm[];                // [RL][C]
  const int units = C >> 3;
  const int RL = 256 / units;
  const int u = threadIdx.x % units, rl = threadIdx.x / units;
  const int64_t rows_per = (P + gridDim.x - 1) / gridDim.x;
  int64_t r0 = blockIdx.x * rows_per, r1 = r0 + rows_per;
  if (r1 > P) r1 = P;
  float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  pdl_wait();
  pdl_launch_dependents();
  if (rl < RL) {
    for (int64_t r = r0 + rl; r < r1; r += RL) {
      const uint4 v = *reinterpret_cast<const uint4*>(x + r * C + u * 8);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 t = __bfloat1622float2(h[i]);
        a[2 * i] += t.x; a[2 * i + 1] += t.y;
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) csm[rl * C + u * 8 + i] = a[i];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += 256) {
    float s = 0.f;
    for (int l = 0; l < RL; ++l) s += csm[l * C + c];
    part[(int64_t)blockIdx.x * C + c] = s;
  }
}
// 32 lanes per column: fixed-order fp64 partial sums + shuffle tree (deterministic)
__global__ void colsum_stage2(const float* __restrict__ part, int nparts, int C, float scale, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  pdl_wait();
  pdl_launch_dependents();
  if (c >= C) return;
  double s = 0.0;
  for (int i = lane; i < nparts; i += 32) s += (double)part[(int64_t)i * C + c];
  s = warp_sum_d(s);
  if (lane == 0) out[c] = (float)(s * (double)scale);
}
size_t hpc_colsum_scratch_bytes(int64_t P, int C) {
  (void)P;
  return (size_t)kColsumParts * (size_t)C * sizeof(float);
}
int hpc_colsum(const void* x, int dtype, int64_t P, int C, float scale, float* out, void* scratch,
               size_t scratch_bytes, hpc_stream_t stream) {
  HPC_REQUIRE(P > 0 && C > 0, HPC_ERR_BAD_SHAPE, "hpc_colsum: empty input");
  int nparts = kColsumParts;
  if ((int64_t)nparts > P) nparts = (int)P;
  HPC_REQUIRE(scratch && scratch_bytes >= (size_t)nparts * C * sizeof(float), HPC_ERR_BAD_SHAPE,
              "hpc_colsum: scratch too small (%zu < %zu)", scratch_bytes, (size_t)nparts * C * sizeof(float));
  auto st = (cudaStream_t)stream;
  if (dtype == HPC_BF16 && (C & 7) == 0 && C <= 2048 && hpc_aligned16(x)) {
    const int RL = 256 / (C >> 3);
    HPC_CUDA(hpc_launch_pdl(colsum_stage1_bf16x8, dim3(nparts), dim3(256), (size_t)RL * C * sizeof(float), st,
                            (const __nv_bfloat16*)x, P, C, (float*)scratch));
  } else {
    int threads = C >= 128 ? 128 : ((C + 31) / 32) * 32;
    dim3 grid((C + threads - 1) / threads, nparts);
    if (dtype == HPC_BF16)
      colsum_stage1<__nv_bfloat16><<<grid, threads, 0, st>>>((const __nv_bfloat16*)x, P, C, (float*)scratch);
    else
      colsum_stage1<float><<<grid, threads, 0, st>>>((const float*)x, P, C, (float*)scratch);
  }
  HPC_LAUNCH_CHECK();
  HPC_CUDA(hpc_launch_pdl(colsum_stage2, dim3((C + 7) / 8), dim3(256), 0, st, (const float*)scratch, nparts, C, scale, out));
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

// ---------------------------------------------------------------------------------------
// hpc_colsum_static: hpc_colsum for callers that cannot allocate device memory (code under CUDA-graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The stage-1 partial sums live in a file-scope device array instead of a caller buffer:
//   * size: kColsumParts x HPC_COLSUM_STATIC_MAX_C floats = 592 x 512 x 4 B = 1.16 MiB, held by EVERY GPU that loads
//     this library until the process exits; C beyond the maximum is rejected;
//   * in host code `g_colsum_static` is only a stand-in for the symbol, not a device address: its address on the
//     current device comes from cudaGetSymbolAddress, once per device, with that device current;
//   * nothing has to start at zero: stage 1 overwrites every partial (nparts x C) that stage 2 of the same call reads,
//     so what an earlier call left behind is never seen;
//   * every call on a device shares the one array: two calls whose kernels overlap on the GPU (different streams
//     without an ordering between them) race.  Calls on one stream, or ordered by events, are safe.
// The entry point launches kernels only: no cudaMalloc / cudaFree / allocating library routine.
// ---------------------------------------------------------------------------------------
static __device__ __align__(256) float g_colsum_static[(size_t)kColsumParts * HPC_COLSUM_STATIC_MAX_C];

static int colsum_static_address(float** out) {
  constexpr int kMaxDevices = 64;
  static float* addr[kMaxDevices] = {nullptr};     // per device; a benign race: every thread resolves the same address
  int dev = -1;
  HPC_CUDA(cudaGetDevice(&dev));
  HPC_REQUIRE(dev >= 0 && dev < kMaxDevices, HPC_ERR_UNSUPPORTED, "hpc_colsum_static: device ordinal %d not in [0,%d)", dev,
              kMaxDevices);
  float* p = __atomic_load_n(&addr[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    HPC_CUDA(cudaGetSymbolAddress((void**)&p, g_colsum_static));      // address on the CURRENT device (= dev)
    HPC_REQUIRE(p != nullptr, HPC_ERR_CUDA, "hpc_colsum_static: cudaGetSymbolAddress returned NULL on device %d", dev);
    __atomic_store_n(&addr[dev], p, __ATOMIC_RELEASE);
  }
  *out = p;
  return HPC_OK;
}

int hpc_colsum_static(const void* x, int dtype, int64_t P, int C, float scale, float* out, hpc_stream_t stream) {
  HPC_REQUIRE(x && out, HPC_ERR_BAD_SHAPE, "hpc_colsum_static: NULL pointer");
  HPC_REQUIRE(P > 0 && C > 0, HPC_ERR_BAD_SHAPE, "hpc_colsum_static: empty input");
  HPC_REQUIRE(C <= HPC_COLSUM_STATIC_MAX_C, HPC_ERR_BAD_SHAPE,
              "hpc_colsum_static: C = %d exceeds the built-in scratch (at most %d columns); use hpc_colsum with a caller buffer",
              C, HPC_COLSUM_STATIC_MAX_C);
  float* scratch = nullptr;
  int rc = colsum_static_address(&scratch);
  if (rc) return rc;
  return hpc_colsum(x, dtype, P, C, scale, out, scratch, sizeof(float) * (size_t)kColsumParts * HPC_COLSUM_STATIC_MAX_C, stream);
}

// ---------------------------------------------------------------------------------------
// GDN elementwise helpers
// ---------------------------------------------------------------------------------------
template <typename T>
__global__ void square_kernel(const T* __restrict__ x, T* __restrict__ out, int64_t n, int relu, T* __restrict__ ro) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float v = ld_as_float<T>(x, i);
    if (relu) v = fmaxf(v, 0.f);
    if (ro) st_from_float<T>(ro, i, v);
    st_from_float<T>(out, i, v * v);
  }
}
template <typename T>
__global__ void gdn_bwd_t_kernel(const T* __restrict__ g, const T* __restrict__ x, const T* __restrict__ nrm,
                                 T* __restrict__ t, int64_t n, int inverse) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float r = ld_as_float<T>(nrm, i);
    float v = ld_as_float<T>(g, i) * ld_as_float<T>(x, i) * (inverse ? r : r * r * r);
    st_from_float<T>(t, i, v);
  }
}
int hpc_square(const void* x, void* out, int dtype, int64_t n, int relu, void* relu_out, hpc_stream_t stream) {
  if (n <= 0) return HPC_OK;
  int grid = grid_for(n, 256, 148 * 16);
  if (dtype == HPC_BF16)
    square_kernel<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)out,
                                                                        n, relu, (__nv_bfloat16*)relu_out);
  else
    square_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)x, (float*)out, n, relu,
                                                                (float*)relu_out);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}
int hpc_gdn_bwd_t(const void* g, const void* x, const void* nrm, void* t, int dtype, int64_t n, int inverse,
                  hpc_stream_t stream) {
  if (n <= 0) return HPC_OK;
  int grid = grid_for(n, 256, 148 * 16);
  if (dtype == HPC_BF16)
    gdn_bwd_t_kernel<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)g, (const __nv_bfloat16*)x, (const __nv_bfloat16*)nrm, (__nv_bfloat16*)t, n, inverse);
  else
    gdn_bwd_t_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)g, (const float*)x,
                                                                   (const float*)nrm, (float*)t, n, inverse);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

template <typename T>
__global__ void relu_mask_kernel(const T* __restrict__ g, const T* __restrict__ x, T* __restrict__ out, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    st_from_float<T>(out, i, ld_as_float<T>(x, i) > 0.f ? ld_as_float<T>(g, i) : 0.f);
}
int hpc_relu_mask(const void* g, const void* x, void* out, int dtype, int64_t n, hpc_stream_t stream) {
  if (n <= 0) return HPC_OK;
  int grid = grid_for(n, 256, 148 * 16);
  if (dtype == HPC_BF16)
    relu_mask_kernel<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)g,
                                                                           (const __nv_bfloat16*)x, (__nv_bfloat16*)out, n);
  else
    relu_mask_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)g, (const float*)x, (float*)out, n);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

// ---------------------------------------------------------------------------------------
// layout / dtype conversion
// ---------------------------------------------------------------------------------------
template <typename T>
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ src, T* __restrict__ dst, int C, int64_t HW) {
  // tile transpose 32(c) x 32(hw) through shared memory
  __shared__ float tile[32][33];
  const int n = blockIdx.z;
  const int64_t hw0 = (int64_t)blockIdx.x * 32;
  const int c0 = blockIdx.y * 32;
  for (int k = threadIdx.y; k < 32; k += blockDim.y) {
    int c = c0 + k;
    int64_t hw = hw0 + threadIdx.x;
    tile[k][threadIdx.x] = (c < C && hw < HW) ? src[((int64_t)n * C + c) * HW + hw] : 0.f;
  }
  __syncthreads();
  for (int k = threadIdx.y; k < 32; k += blockDim.y) {
    int64_t hw = hw0 + k;
    int c = c0 + threadIdx.x;
    if (c < C && hw < HW) st_from_float<T>(dst, ((int64_t)n * HW + hw) * C + c, tile[threadIdx.x][k]);
  }
}
template <typename T>
__global__ void nhwc_to_nchw_kernel(const T* __restrict__ src, float* __restrict__ dst, int C, int64_t HW) {
  __shared__ float tile[32][33];
  const int n = blockIdx.z;
  const int64_t hw0 = (int64_t)blockIdx.x * 32;
  const int c0 = blockIdx.y * 32;
  for (int k = threadIdx.y; k < 32; k += blockDim.y) {
    int64_t hw = hw0 + k;
    int c = c0 + threadIdx.x;
    tile[k][threadIdx.x] = (c < C && hw < HW) ? ld_as_float<T>(src, ((int64_t)n * HW + hw) * C + c) : 0.f;
  }
  __syncthreads();
  for (int k = threadIdx.y; k < 32; k += blockDim.y) {
    int c = c0 + k;
    int64_t hw = hw0 + threadIdx.x;
    if (c < C && hw < HW) dst[((int64_t)n * C + c) * HW + hw] = tile[threadIdx.x][k];
  }
}
int hpc_nchw_to_nhwc(const float* src, void* dst, int dtype, int N, int C, int H, int W, hpc_stream_t stream) {
  HPC_REQUIRE(N > 0 && C > 0 && H > 0 && W > 0, HPC_ERR_BAD_SHAPE, "hpc_nchw_to_nhwc: empty tensor");
  HPC_REQUIRE(N <= 65535, HPC_ERR_UNSUPPORTED, "hpc_nchw_to_nhwc: N > 65535");
  const int64_t HW = (int64_t)H * W;
  dim3 grid((unsigned)((HW + 31) / 32), (C + 31) / 32, N);
  if (dtype == HPC_BF16)
    nchw_to_nhwc_kernel<__nv_bfloat16><<<grid, dim3(32, 8), 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst, C, HW);
  else
    nchw_to_nhwc_kernel<float><<<grid, dim3(32, 8), 0, (cudaStream_t)stream>>>(src, (float*)dst, C, HW);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}
int hpc_nhwc_to_nchw(const void* src, int dtype, float* dst, int N, int C, int H, int W, hpc_stream_t stream) {
  HPC_REQUIRE(N > 0 && C > 0 && H > 0 && W > 0, HPC_ERR_BAD_SHAPE, "hpc_nhwc_to_nchw: empty tensor");
  HPC_REQUIRE(N <= 65535, HPC_ERR_UNSUPPORTED, "hpc_nhwc_to_nchw: N > 65535");
  const int64_t HW = (int64_t)H * W;
  dim3 grid((unsigned)((HW + 31) / 32), (C + 31) / 32, N);
  if (dtype == HPC_BF16)
    nhwc_to_nchw_kernel<__nv_bfloat16><<<grid, dim3(32, 8), 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)src, dst,
                                                                                      C, HW);
  else
    nhwc_to_nchw_kernel<float><<<grid, dim3(32, 8), 0, (cudaStream_t)stream>>>((const float*)src, dst, C, HW);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

template <typename S, typename D>
__global__ void cast_kernel(const S* __restrict__ s, D* __restrict__ d, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    st_from_float<D>(d, i, ld_as_float<S>(s, i));
}
int hpc_cast(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t n, hpc_stream_t stream) {
  if (n <= 0) return HPC_OK;
  int grid = grid_for(n, 256, 148 * 16);
  auto st = (cudaStream_t)stream;
  if (src_dtype == HPC_F32 && dst_dtype == HPC_BF16)
    cast_kernel<float, __nv_bfloat16><<<grid, 256, 0, st>>>((const float*)src, (__nv_bfloat16*)dst, n);
  else if (src_dtype == HPC_BF16 && dst_dtype == HPC_F32)
    cast_kernel<__nv_bfloat16, float><<<grid, 256, 0, st>>>((const __nv_bfloat16*)src, (float*)dst, n);
  else if (src_dtype == HPC_F32 && dst_dtype == HPC_F32)
    cast_kernel<float, float><<<grid, 256, 0, st>>>((const float*)src, (float*)dst, n);
  else if (src_dtype == HPC_BF16 && dst_dtype == HPC_BF16)
    cast_kernel<__nv_bfloat16, __nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)src, (__nv_bfloat16*)dst, n);
  else
    HPC_FAIL(HPC_ERR_UNSUPPORTED, "hpc_cast: dtypes %d -> %d", src_dtype, dst_dtype);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

// per-channel sum of an NCHW tensor: one block per (n, c) plane, then a fixed-order fp64 sum over n
__global__ void nchw_plane_sum_kernel(const float* __restrict__ x, int64_t HW, float* __restrict__ part) {
  __shared__ float sm[32];
  const float* p = x + (int64_t)blockIdx.x * HW;
  float a = 0.f;
  if ((HW & 3) == 0 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
    const float4* p4 = reinterpret_cast<const float4*>(p);
    for (int64_t i = threadIdx.x; i < (HW >> 2); i += blockDim.x) {
      const float4 v = p4[i];
      a += (v.x + v.y) + (v.z + v.w);
    }
  } else {
    for (int64_t i = threadIdx.x; i < HW; i += blockDim.x) a += p[i];
  }
  const float s = block_sum(a, sm);
  if (threadIdx.x == 0) part[blockIdx.x] = s;
}
__global__ void nchw_plane_sum_final(const float* __restrict__ part, int N, int C, float* __restrict__ out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double tot = 0.0;
  for (int n = 0; n < N; ++n) tot += (double)part[(int64_t)n * C + c];
  out[c] = (float)tot;
}
int hpc_nchw_channel_sum(const float* x, int N, int C, int64_t HW, float* out, float* scratch, hpc_stream_t stream) {
  HPC_REQUIRE(N > 0 && C > 0 && HW > 0, HPC_ERR_BAD_SHAPE, "hpc_nchw_channel_sum: empty tensor");
  HPC_REQUIRE(scratch, HPC_ERR_BAD_SHAPE, "hpc_nchw_channel_sum: scratch (N*C floats) is NULL");
  nchw_plane_sum_kernel<<<N * C, 256, 0, (cudaStream_t)stream>>>(x, HW, scratch);
  HPC_LAUNCH_CHECK();
  nchw_plane_sum_final<<<(C + 63) / 64, 64, 0, (cudaStream_t)stream>>>(scratch, N, C, out);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}
