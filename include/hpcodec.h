/*
 * hpcodec.h -- C ABI of libhpcodec_sm100a.so: the B200-native (sm_100a) kernels behind the
 * scale-hyperprior codec hot path of hieu1999210/image_compression.
 *
 * The reference has no FFI of its own: every arithmetic op on the path is a stock torch call
 * inside its nn.Modules.  This header is the native seam one level below
 * that module surface; each entry cites the reference code (path:line under the reference
 * repository root) whose arithmetic it replaces.  The Python modules in
 * image_compression_b200/modelling bind these through ctypes.
 *
 * Conventions
 *  - plain pointers and sizes only; all pointers are DEVICE pointers unless stated otherwise
 *  - every call enqueues work on `stream` (a cudaStream_t) and returns without synchronising
 *  - return value: HPC_OK or an HPC_ERR_* code; text via hpc_last_error() (thread local)
 *  - activations are NHWC ("channels last"): element (n,h,w,c) at ((n*H+h)*W+w)*C+c
 *  - `lp` ("low precision") tensors have element type `dtype`: HPC_F32 or HPC_BF16; tensors
 *    named *_f32 are always float
 *  - the library never allocates device memory; scratch is passed in by the caller
 */
#ifndef HPCODEC_H_
#define HPCODEC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HPC_VERSION 100

#define HPC_OK 0
#define HPC_ERR_BAD_SHAPE 1
#define HPC_ERR_BAD_ALIGNMENT 2
#define HPC_ERR_UNSUPPORTED 3
#define HPC_ERR_CUDA 4

#define HPC_F32 0
#define HPC_BF16 1

typedef void* hpc_stream_t; /* cudaStream_t */

int hpc_version(void);
const char* hpc_last_error(void);
/* number of kernels this library has launched so far in this process (monotonic) */
unsigned long long hpc_launch_count(void);
/* 1 when the current device is compute capability 10.x (tcgen05/TMEM/TMA kernels usable) */
int hpc_device_is_sm100(void);

/* Kernel-class timing for bench.py's roofline (no reference counterpart: the reference has no
 * profiler below cProfile, main.py:76-86).  While enabled, every launch of a kernel of one of the
 * classes below is bracketed by a cudaEvent pair on ITS OWN stream and tagged with its algorithmic
 * work (flops for the GEMM classes, bytes for the streaming classes).  hpc_prof_read synchronises on
 * the recorded events and returns the summed device time, launch count and work of one class. */
#define HPC_PROF_IGEMM_FWD 0   /* igemm_fwd_umma_kernel: conv / deconv / dgrad / GDN contraction  (flops) */
#define HPC_PROF_IGEMM_WGRAD 1 /* igemm_wgrad_umma_kernel + its split-K reduce                    (flops) */
#define HPC_PROF_CONDITIONAL 2 /* conditional likelihood fwd + bwd (K5)                           (bytes) */
#define HPC_PROF_FACTORIZED 3  /* factorised prior fwd + bwd (K4)                                 (bytes) */
#define HPC_PROF_THIN 4        /* conv_c3 / deconv_c3 / wgrad_c3 (3-channel head and tail)        (bytes) */
#define HPC_PROF_DISTORTION 5  /* clamp+SSE fwd/bwd, SSIM levels (K6/K7)                          (bytes) */
#define HPC_PROF_GDN 6        /* fused GDN forward / backward (gdn_umma.cu)                      (bytes) */
#define HPC_PROF_OPTIM 7      /* fused clip + AdamW step (optim_kernels.cu)                      (bytes) */
#define HPC_PROF_NCLASSES 8
int hpc_prof_enable(int on); /* on!=0: clear the records and start; 0: stop recording (records kept) */
int hpc_prof_pause(int paused); /* suspend / resume recording, records kept (needed around CUDA-graph capture) */
int hpc_prof_read(int cls, double* ms_total, unsigned long long* launches, double* work_total);
/* the individual records in call order: class, milliseconds and work of record i (returns the record count; fills at
 * most `max` entries; any output pointer may be NULL) */
int hpc_prof_records(int* cls, float* ms, double* work, int max);

/* ------------------------------------------------------------------------------------------
 * Bounds and the GDN non-negative reparameterisation
 * ---------------------------------------------------------------------------------------- */

/* LowerBound / UpperBound forward: y = max(x,bound) or min(x,bound).
 * Replaces modelling/layers/bound.py:30-33 and :47-50. */
int hpc_bound_fwd(const float* x, float* y, int64_t n, float bound, int is_upper, hpc_stream_t stream);
/* backward: gx = g where (x>=bound or g<0) [lower] / (x<=bound or g>0) [upper], else 0.
 * Replaces modelling/layers/bound.py:36-42 and :53-59. */
int hpc_bound_bwd(const float* x, const float* gy, float* gx, int64_t n, float bound, int is_upper,
                  hpc_stream_t stream);

/* NonNegativeParam.forward: out = max(param,bound)^2 - pedestal.
 * Replaces modelling/layers/gdn.py:59-62 (bound/pedestal as built at :51-53). */
int hpc_nonneg_fwd(const float* param, float* out, int64_t n, float bound, float pedestal, hpc_stream_t stream);
/* backward: gparam = m * 2*max(param,bound) * gout, m = (param>=bound) or (2*val*gout < 0).
 * `scale` multiplies gout first (GDN kernels hand over -1/2-free sums; see hpc_igemm_wgrad). */
int hpc_nonneg_bwd(const float* param, const float* gout, float* gparam, int64_t n, float bound, float scale,
                   hpc_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Factorised prior (EntropyModel + CDFEstimator, DIMS=[3,3,3]) -- kernel K4
 * ---------------------------------------------------------------------------------------- */

/* The 11 parameter tensors of the per-channel cumulative MLP 1->3->3->3->1, in the reference's
 * own memory layout (entropy_model.py:54-65): weight (1,C,out,in), bias (1,C,out,1),
 * factor (1,C,out,1).  The same struct carries the gradient pointers in the backward call. */
typedef struct {
  float* w[4]; /* (C,3,1) (C,3,3) (C,3,3) (C,1,3) */
  float* b[4]; /* (C,3) (C,3) (C,3) (C,1) */
  float* f[3]; /* (C,3) x3 */
} hpc_cdf_params;

/* CDFLayer.forward stand-alone (one layer of the MLP): x (N,C,din,1) -> out (N,C,dout,1),
 *   pre = softplus(w[c]) x + b[c];  out = f ? pre + tanh(pre)*tanh(f[c]) : pre   (f == NULL: the last layer, act=False)
 * w (C,dout,din), b (C,dout), f (C,dout) in the reference's layouts; din, dout <= 4.
 * Replaces modelling/blocks/entropy_model.py:67-78 (module-surface parity; the hot path evaluates the whole MLP in
 * hpc_factorized_fwd / hpc_cdf_logits_fwd).  bwd: gx may be NULL; gw, gb, gf are overwritten (fixed-order sums). */
int hpc_cdf_layer_fwd(const float* x, const float* w, const float* b, const float* f, int64_t N, int C, int din,
                      int dout, float* out, hpc_stream_t stream);
int hpc_cdf_layer_bwd(const float* x, const float* w, const float* b, const float* f, int64_t N, int C, int din,
                      int dout, const float* g, float* gx, float* gw, float* gb, float* gf, hpc_stream_t stream);

/* EntropyModel.forward: quantise, pmf = s*(sigmoid(s*u)-sigmoid(s*l)), bits.
 * Replaces modelling/blocks/entropy_model.py:204-214 (_quantize :216-236, _prob_mass :259-269,
 * CDFLayer.forward :67-78, CDFEstimator.forward :101-114, _ce_loss :185).
 *   z            (P,C) NHWC rows, float
 *   mode         0 = "quantize" (round half to even), 1 = "noise"
 *   noise        optional (P,C) U[0,1) tensor; when NULL in mode 1 Philox(seed,offset) is used
 *   z_tilde_f32  (P,C) out;  z_tilde_lp optional copy in `lp_dtype` (feeds h_s)
 *   probs        optional (P,C) out (NHWC; the reference's (N,W,C,H) result is a view of it)
 *   bits_out     1 float: sum of clamp(-log2(p+1e-10),0,50); deterministic reduction
 *   scratch      >= hpc_reduce_scratch_bytes() bytes, zero-initialised once by the caller; required: besides the
 *                reduction state it holds the transformed-parameter table (67 x C floats, C <= 1024) that a small
 *                pre-kernel of this call fills (calls sharing one scratch buffer must be stream ordered) */
int hpc_factorized_fwd(const float* z, const float* noise, const hpc_cdf_params* params, int64_t P, int C,
                       int mode, float bin, uint64_t seed, uint64_t offset, float* z_tilde_f32,
                       void* z_tilde_lp, int lp_dtype, float* probs, float* bits_out, void* scratch,
                       hpc_stream_t stream);
/* backward of the above (autograd in the reference).  g_bits: device scalar dL/dbits (may be
 * NULL), g_probs: optional (P,C) dL/dprobs, g_zt: optional (P,C) dL/dz_tilde (float).
 * gz (P,C) out; gparams accumulated into caller-zeroed buffers.  g_scale
 * multiplies *g_bits (e.g. 1/num_pixels). */
int hpc_factorized_bwd(const float* z_tilde, const hpc_cdf_params* params, int64_t P, int C, int mode,
                       float bin, const float* g_bits, float g_scale, const float* g_probs, const float* g_zt,
                       float* gz, const hpc_cdf_params* gparams, void* scratch, size_t scratch_bytes,
                       hpc_stream_t stream);
/* scratch (required, >= hpc_factorized_bwd_scratch_bytes(P, C)): per-row-block partials of the 43 x C parameter
 * gradients, summed in a fixed order by a finishing kernel (bit-reproducible), and the transformed-parameter table. */
size_t hpc_factorized_bwd_scratch_bytes(int64_t P, int C);

/* CDFEstimator.forward stand-alone: logits of the per-channel MLP at x (P,C) (entropy_model.py:101-114)
 * and its backward (g (P,C) upstream; gx may be NULL; gparams accumulated into zeroed buffers). */
int hpc_cdf_logits_fwd(const float* x, const hpc_cdf_params* params, int64_t P, int C, float* out,
                       hpc_stream_t stream);
int hpc_cdf_logits_bwd(const float* x, const hpc_cdf_params* params, int64_t P, int C, const float* g, float* gx,
                       const hpc_cdf_params* gparams, hpc_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Conditional (scale-only) Laplace / Gauss likelihood -- kernel K5
 * ---------------------------------------------------------------------------------------- */
#define HPC_DIST_LAPLACE 0
#define HPC_DIST_GAUSS 1

/* SymmetricConditionalModel.forward: p = cdf((bin/2-|y~-mean|)/s) - cdf((-bin/2-|y~-mean|)/s).
 * Replaces modelling/blocks/entropy_model.py:280-295, :319-339, :341-352, :362 / :375 and the
 * bit count of :185 (called from modelling/meta_arch/bmshl2018.py:76-77).
 * n elements, any layout (pure pointwise).  Outputs as in hpc_factorized_fwd. */
int hpc_conditional_fwd(const float* y, const float* scale, const float* noise, int64_t n, int dist, int mode,
                        float bin, float mean, uint64_t seed, uint64_t offset, float* y_tilde_f32,
                        void* y_tilde_lp, int lp_dtype, float* probs, float* bits_out, void* scratch,
                        hpc_stream_t stream);
/* backward: gy = gp*dp/dy~ + g_yt ; gscale = gp*dp/dscale, gp = *g_bits*g_scale*dbits/dp + g_probs.
 * g_yt optional, dtype g_yt_dtype (the dgrad of g_s arrives in bf16).  If fuse_exp_clamp != 0
 * the scale gradient is chained through sigma = clamp(exp(pre),1e-10,1e10)
 * (modelling/blocks/prior_synthesis.py:72): gscale_out = gscale*sigma*[1e-10<=sigma<=1e10].
 * gscale_f32 and/or gscale_lp may be NULL. */
int hpc_conditional_bwd(const float* y_tilde, const float* scale, int64_t n, int dist, int mode, float bin,
                        float mean, const float* g_bits, float g_scale, const float* g_probs, const void* g_yt,
                        int g_yt_dtype, int fuse_exp_clamp, float* gy, float* gscale_f32, void* gscale_lp,
                        int lp_dtype, hpc_stream_t stream);

/* backward of sigma = clamp(exp(pre),1e-10,1e10) (modelling/blocks/prior_synthesis.py:72) for the
 * stand-alone HyperpriorSynthesisTransform: g_pre = g_sigma*sigma*[1e-10<=sigma<=1e10] (lp out). */
int hpc_exp_clamp_bwd(const float* sigma, const float* g_sigma, void* g_pre_lp, int lp_dtype, int64_t n,
                      hpc_stream_t stream);

/* BaseEntropyModel._ce_loss standalone: bits = sum clamp(-log(p+1e-10)/ln2,0,50).
 * Replaces modelling/blocks/entropy_model.py:171-185. */
int hpc_bits_fwd(const float* probs, int64_t n, float* bits_out, void* scratch, hpc_stream_t stream);
int hpc_bits_bwd(const float* probs, const float* g_bits, int64_t n, float* g_probs, hpc_stream_t stream);

size_t hpc_reduce_scratch_bytes(void);

/* ------------------------------------------------------------------------------------------
 * Clamp + MSE / PSNR -- kernel K6.  Images are NCHW float (the user-facing layout).
 * ---------------------------------------------------------------------------------------- */

/* x~ = LowerBound(UpperBound(raw,1),0) and per-image sum of squared error against x.
 * Replaces modelling/meta_arch/bmshl2018.py:80-81 + modelling/loss.py:25 (nn.MSELoss) and, with
 * scale=255, utils/metric.py:26-36 (psnr).  raw/x/xt: (N, per_image) floats; xt may alias raw or
 * be NULL; sse_per_image (N) out = sum((scale*(x - x~))^2).  grid-deterministic. */
int hpc_clamp01_sse_fwd(const float* x, const float* raw, float* xt, int N, int64_t per_image, float scale,
                        int do_clamp, float* sse_per_image, void* scratch, hpc_stream_t stream);
/* g_raw = bounds_bwd( g_xt + mse_coef * 2*(x~ - x) ), masks per bound.py:36-42,:53-59.
 * g_xt may be NULL (no extra upstream gradient), x may be NULL (no MSE term);
 * mse_coef_dev: optional device scalar multiplied into mse_coef (upstream dL/dloss).
 * do_clamp=0 skips the bounds (plain nn.MSELoss backward wrt its target, modelling/loss.py:25). */
int hpc_clamp01_bwd(const float* raw, const float* x, const float* g_xt, int64_t n, float mse_coef,
                    const float* mse_coef_dev, int do_clamp, float* g_raw, hpc_stream_t stream);
/* psnr_n = 10*(2*log10(max_val) - ln(sse_n/per_image)/ln10); utils/metric.py:26-36 */
int hpc_psnr_from_sse(const float* sse_per_image, int N, int64_t per_image, float max_val, float* psnr,
                      hpc_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * (MS-)SSIM -- kernel K7.  Planes are NCHW float.
 * ---------------------------------------------------------------------------------------- */

/* One SSIM level: valid 11x11 (size `fs`<=15) Gaussian moments, cs and ssim maps, per-image sums.
 * Replaces modelling/loss.py:75-91,103-121 and utils/metric.py:56-72,79-97.
 *   img1,img2 (N,C,H,W); in_scale multiplies both on load (the x255 of loss.py:67-68)
 *   filt1d    fs floats on the device (the 2-D softmax window is exactly g (x) g)
 *   ssim_sum, cs_sum (N) out: sums over (C,H',W'), H'=H-fs+1 (caller divides)
 *   scratch   16*N bytes, 8-byte aligned: 32.32 fixed-point accumulators, so the sums are bit-reproducible
 *             from run to run (no floating-point atomics) */
int hpc_ssim_level_fwd(const float* img1, const float* img2, int N, int C, int H, int W, const float* filt1d,
                       int fs, float in_scale, float c1, float c2, float* ssim_sum, float* cs_sum,
                       void* scratch, hpc_stream_t stream);
/* backward wrt img2: g_ssim, g_cs (N) are dL/d(sum) per image (already divided by the map size).
 * maps: scratch of 3*N*C*H'*W' floats.  g_img2 (N,C,H,W): if accumulate!=0 adds to the
 * existing content (the pyramid's upsampled gradient), else overwrites.  The result is
 * multiplied by in_scale (chain rule of the load scaling). */
int hpc_ssim_level_bwd(const float* img1, const float* img2, int N, int C, int H, int W, const float* filt1d,
                       int fs, float in_scale, float c1, float c2, const float* g_ssim, const float* g_cs,
                       float* maps, float* g_img2, int accumulate, hpc_stream_t stream);
/* 2x2 average pool with reflect padding of odd edges; out = scale*avg.
 * Replaces modelling/loss.py:176-188 and utils/metric.py:126-138. */
int hpc_avgpool2_fwd(const float* in, float* out, int NC, int H, int W, float scale, hpc_stream_t stream);
/* gin (+)= scale * pooled-gradient scatter (accumulate flag as above). */
int hpc_avgpool2_bwd(const float* gout, float* gin, int NC, int H, int W, float scale, int accumulate,
                     hpc_stream_t stream);

#define HPC_MSSSIM_LOSS_LOG 0   /* MS_SSIMLoss._log_forward  modelling/loss.py:141-156 */
#define HPC_MSSSIM_LOSS_PLAIN 1 /* MS_SSIMLoss._forward      modelling/loss.py:158-174 */
#define HPC_MSSSIM_METRIC 2     /* utils/metric.py:110-124 (clamp_min 0, optional dB)   */
#define HPC_SSIM_LOSS 3         /* SSIMLoss.forward          modelling/loss.py:65-73     */
#define HPC_SSIM_METRIC 4       /* SSIM.__call__             utils/metric.py:52-54       */
/* Per-image level means -> loss / metric.  sums: (L,2,N) floats = [level][0:ssim,1:cs][n] sums,
 * counts (L) map sizes C*H'*W' (host array).  out: 1 float for losses, N floats for metrics
 * (SSIM_LOSS with log_scale returns N values: loss.py:72 returns -ssim unreduced).
 * flags: bit0 = log_scale (SSIM_LOSS), bit0 = in_dB (metrics), bit1 = non_negative (SSIM metric) */
int hpc_msssim_combine_fwd(const float* sums, const double* counts_host, int L, int N, const float* weights_host,
                           int kind, int flags, float eps, float* out, hpc_stream_t stream);
/* backward of the loss kinds: g_out device scalar (or N values for SSIM_LOSS log), multiplied by
 * the host scalar g_scale (lambda); g_sums (L,2,N) out = dL/d(sum) per level/image. */
int hpc_msssim_combine_bwd(const float* sums, const double* counts_host, int L, int N, const float* weights_host,
                           int kind, int flags, float eps, const float* g_out, float g_scale, float* g_sums,
                           hpc_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Convolution family as implicit GEMM -- kernels K1/K2/K3/K8
 * ---------------------------------------------------------------------------------------- */
#define HPC_MAX_TAPS 25
#define HPC_MAX_PHASES 4

#define HPC_ENGINE_SIMT 0 /* CUDA-core fp32-accumulate reference engine (any channel count) */
#define HPC_ENGINE_UMMA 1 /* tcgen05/TMEM/TMA engine (bf16 operands, Ci%64==0, Co%16==0, Co<=256) */

#define HPC_EPI_LINEAR 0     /* out = acc + bias */
#define HPC_EPI_RELU 1       /* out = max(acc+bias,0)                         prior_analysis.py:66, prior_synthesis.py:64 */
#define HPC_EPI_EXP_CLAMP 2  /* out = clamp(exp(acc+bias),1e-10,1e10)         prior_synthesis.py:72 */
#define HPC_EPI_GDN 3        /* n=acc+bias; out = aux0*rsqrt(n); nrm=rsqrt(n) gdn.py:84-86 (aux0 = x, A = x^2) */
#define HPC_EPI_IGDN 4       /* out = aux0*sqrt(n); nrm=rsqrt(n)              gdn.py:86 inverse */
#define HPC_EPI_GDN_BWD 5    /* out = aux0*aux2 - aux1*acc   (g*nrm - x*u)    SURVEY A.4 */
#define HPC_EPI_IGDN_BWD 6   /* out = aux0/aux2 + aux1*acc                    SURVEY A.4 */
#define HPC_EPI_RELU_MASK 7  /* out = acc * [aux0 > 0]                        autograd of ReLU */
#define HPC_EPI_SIGN_ADD 8   /* out = acc*sign(aux0_f32) + aux1_f32           autograd of abs, bmshl2018.py:72 */

/* One dense stride-1 sub-problem of the implicit GEMM (a stride-2 transposed convolution is four
 * of them, one per output-pixel parity). */
typedef struct {
  int ntaps;     /* taps of this phase */
  int tap_begin; /* first tap: index into dh/dw and into the packed weights [tap][Co][Ci] */
  int oh_off, ow_off; /* output pixel of grid point (i,j) = (i*os+oh_off, j*os+ow_off) */
} hpc_igemm_phase;

/* out[n, i*os+oh_off, j*os+ow_off, co] = epi( sum_{t in phase} sum_ci
 *        x[n, i*is+dh[t], j*is+dw[t], ci] * w[t][co][ci] )          (x = 0 outside the image)
 * Conv2d(k,s=is,p):   one phase, taps (kh-p, kw-p), os=1, grid = output size
 * ConvTranspose2d(5,2,2,1): four phases over the INPUT grid, is=1, os=2
 * Replaces the nn.Conv2d / nn.ConvTranspose2d calls at modelling/blocks/analysis.py:55,
 * synthesis.py:55-57, prior_analysis.py:54-56, prior_synthesis.py:54-56, the 1x1 gamma
 * contraction of modelling/layers/gdn.py:84, and (with transposed packing) their autograd dgrads. */
typedef struct {
  int engine;   /* HPC_ENGINE_* */
  int dtype;    /* element type of x, w and the lp outputs / lp aux inputs */
  const void* x; /* (N,H,W,Ci) */
  const void* w; /* packed (ntaps_total, Co, Ci) */
  int N, H, W, Ci;
  int GH, GW;   /* iteration grid per image */
  int is, os;
  int OH, OW, Co;
  int nphases;
  hpc_igemm_phase phase[HPC_MAX_PHASES];
  int dh[HPC_MAX_TAPS], dw[HPC_MAX_TAPS];
  /* epilogue */
  int epi;
  const float* bias;  /* (Co) or NULL */
  const void* aux0;   /* see HPC_EPI_*; lp (N,OH,OW,Co) unless stated float */
  const void* aux1;
  const void* aux2;
  float* out_f32;     /* optional (N,OH,OW,Co) */
  void* out_lp;       /* optional */
  void* out_sq_lp;    /* optional out^2 (the A operand of the following GDN) */
  void* out_abs_lp;   /* optional |out| (input of h_a, bmshl2018.py:72) */
  void* out_nrm_lp;   /* GDN/IGDN: rsqrt(n), saved for backward */
  /* Optional scratch for the tcgen05 engine's split-K path (layers with so few 128-pixel tiles that most SMs would
   * idle: the K loop is split over CTAs, fp32 partials go here and a second kernel sums them in a fixed order and
   * applies the epilogue).  >= hpc_igemm_fwd_workspace_bytes(p) (0 = the launch does not split); NULL = never split. */
  void* workspace;
  size_t workspace_bytes;
} hpc_igemm_fwd_params;

int hpc_igemm_fwd(const hpc_igemm_fwd_params* p, hpc_stream_t stream);
size_t hpc_igemm_fwd_workspace_bytes(const hpc_igemm_fwd_params* p);
/* Host-side introspection of the tcgen05 engine's plan for `p` (no GPU work; the CPU tests of the host logic use it):
 * out = {CTA pairs, M tiles per CTA, shared activation boxes, split-K slices, phase interleave, staged epilogue,
 * tw, th, nb (tile = nb x th x tw = 128 pixels), tiles per phase, tiles, grid}.  Returns 12, or -HPC_ERR_*. */
int hpc_igemm_fwd_plan_describe(const hpc_igemm_fwd_params* p, int* out, int cap);

/* Weight gradient as implicit GEMM over pixels:
 * gw[t][r][c] = scale * sum_{n,i,j} lo[n, i+dh[t], j+dw[t], r] * hi[n, i*hs+oh_off, j*hs+ow_off, c]
 * for every tap t of every phase ("lo" = tensor on the iteration grid, "hi" = strided tensor).
 *   Conv2d(k,s=2) wgrad:            lo = dy (grid GH x GW), hi = x, phases = input-pixel parities
 *   ConvTranspose2d(5,2) wgrad:     lo = x, hi = dy
 *   stride-1 conv / GDN gamma grad: one phase, hs = 1
 * Output gw_f32 (ntaps_total, R, Cc) float in the packed tap order (hpc_unpack_wgrad maps it to
 * the reference layout).  Replaces autograd's convolution_backward weight path. */
typedef struct {
  int engine, dtype;
  const void* lo; /* (N,GH,GW,R) */
  const void* hi; /* (N,HH,HW,Cc) */
  int N, GH, GW, R;
  int HH, HW, Cc;
  int hs;
  int nphases;
  hpc_igemm_phase phase[HPC_MAX_PHASES];
  int dh[HPC_MAX_TAPS], dw[HPC_MAX_TAPS];
  float scale;
  float* gw_f32;
  void* workspace;        /* UMMA engine split-K partials; >= hpc_igemm_wgrad_workspace_bytes() */
  size_t workspace_bytes;
  /* Output layout.  out_mode 0: gw_f32 is the packed [tap][R][Cc] array.  out_mode 1 (UMMA engine only):
   * gw_f32 is the reference's weight-gradient tensor and element (t, r, c) goes to
   * gw_f32[r*out_stride_r + c*out_stride_c + out_tap_src[t]] -- nn.Conv2d's (Co,Ci,kh,kw) or
   * nn.ConvTranspose2d's (Ci,Co,kh,kw) -- so no separate hpc_unpack_wgrad pass is needed.
   * out_accumulate != 0 adds to what gw_f32 holds (.grad accumulation) instead of overwriting. */
  int out_mode, out_accumulate;
  int64_t out_stride_r, out_stride_c;
  int out_tap_src[HPC_MAX_TAPS];
} hpc_igemm_wgrad_params;

int hpc_igemm_wgrad(const hpc_igemm_wgrad_params* p, hpc_stream_t stream);
size_t hpc_igemm_wgrad_workspace_bytes(const hpc_igemm_wgrad_params* p);
/* Host-side introspection of the tcgen05 engine's work-unit plan for `p` (no GPU work; the CPU tests of the host logic
 * use it): out = {nunits, nsplits, lo_area_bytes, kblocks} then per unit 35 ints: phase, natoms, nboxes,
 * 4 x (tap, r-chunk, offset KB) for the atoms, 4 x (r-chunk, dw, dh, rows, offset KB) for the `lo` TMA boxes (-1 = unused).
 * Returns the number of ints written, or -HPC_ERR_*. */
int hpc_igemm_wgrad_plan_describe(const hpc_igemm_wgrad_params* p, int* out, int cap);

/* Weight (un)packing between the reference layouts and the packed (tap, row, col) layout.
 * src element (r, c, k) of a weight with K taps (k = kh*KW+kw) sits at r*stride_r + c*stride_c + k:
 *   Conv2d weight (Co,Ci,kh,kw), rows=Co, cols=Ci:          stride_r = Ci*K, stride_c = K
 *   ConvTranspose2d weight (Ci,Co,kh,kw), rows=Co, cols=Ci: stride_r = K,    stride_c = Co*K
 * dst[t][r][c] = src(r, c, tap_src[t]) converted to `dtype`.  tap_src: host array (ntaps). */
int hpc_pack_weight(const float* src, void* dst, int dtype, int ntaps, int rows, int cols, int64_t stride_r,
                    int64_t stride_c, const int* tap_src_host, hpc_stream_t stream);
/* inverse for gradients: dst(r,c,tap_src[t]) (+)= src[t][r][c]; float both sides. */
int hpc_unpack_wgrad(const float* src, float* dst, int ntaps, int rows, int cols, int64_t stride_r,
                     int64_t stride_c, const int* tap_src_host, int accumulate, hpc_stream_t stream);

/* The reference's per-image input pipeline on the device: crop window -> float -> channels first
 * -> (v / 255 - mean[c]) / std[c], stacked into the (N,C,out_h,out_w) fp32 batch.  Bit-identical to the reference
 * (fp32 steps, IEEE division).  Replaces data_utils/transforms/transforms.py:23-35 (RandomCrop :143-166 with the
 * offsets the HOST drew, ToTensor :64-75, ChannelFirst :84-100, Normalize :103-127) and the stack + fp32 host-to-device
 * copy of data_utils/batch.py:19-29: the host sends uint8 pixels, 4x fewer bytes.
 *   src_u8_hwc      device uint8, n_src images; pixel (m,y,x,c) at m*img_stride_bytes + y*row_stride_bytes + x*C + c
 *                   (C <= 4)
 *   src_index_dev   device int32 (N): output n is cut from source image src_index[n] (a uint8 image set kept resident
 *                   in HBM, sampled by index); NULL = image n
 *   offsets_yx_dev  device int32 (N,2) = (y0, x0) per image, or NULL for (0,0); a window or index that leaves the source is
 *                   not read (zeros are written): the host validates its offsets (the reference asserts, :158-159)
 *   mean_host, std_host   HOST arrays of C floats (cfg.DATA.MEAN / STD), NULL = 0 / 1 */
int hpc_u8_crop_normalize(const unsigned char* src_u8_hwc, int64_t img_stride_bytes, int64_t row_stride_bytes, int n_src,
                          int src_h, int src_w, const int* src_index_dev, const int* offsets_yx_dev, int N, int C, int out_h, int out_w, const float* mean_host,
                          const float* std_host, float* out_nchw, hpc_stream_t stream);

/* All packed operand copies of a model in ONE launch (a training step changes every weight, so every copy is
 * rebuilt once per step: 27 pack + 6 gdn_prepare launches otherwise).  Job kinds:
 *   HPC_PACK_TAPS  dst[t][r][c] = src(r, c, tap_src[t]) as hpc_pack_weight (`dtype` of the call); the taps of one
 *                  (r, c) pair must be contiguous in src (stride_r == ntaps or stride_c == ntaps)
 *   HPC_PACK_C3_K  as hpc_pack_c3_weight's kmaj (rows = Cw; bf16)    HPC_PACK_C3_N  its nmaj (rows = Cw; bf16)
 *   HPC_PACK_GDN   as hpc_gdn_prepare: src = gamma_param, src2 = beta_param, rows = C, f0 = gamma_bound,
 *                  f1 = beta_bound, f2 = pedestal; dst = gamma_bf16, dst2 = gammaT_bf16, dst3 = beta_eff
 * Replaces the same reference code as those three entries (weights are read by F.conv2d / conv_transpose2d in
 * place: analysis.py:55, synthesis.py:55-57, prior_*.py:54-56; gdn.py:59-62). */
#define HPC_PACK_TAPS 0
#define HPC_PACK_C3_K 1
#define HPC_PACK_C3_N 2
#define HPC_PACK_GDN 3
#define HPC_PACK_MAX_JOBS 48
typedef struct {
  int kind, ntaps, rows, cols;
  int64_t stride_r, stride_c;
  const float* src; const float* src2;
  void* dst; void* dst2; float* dst3;
  float f0, f1, f2;
  unsigned char tap_src[28];
} hpc_pack_job;
int hpc_pack_multi(const hpc_pack_job* jobs_host, int njobs, int dtype, hpc_stream_t stream);

/* out[c] = scale * sum_p x[p][c]  (bias gradients; GDN beta gradient).  x (P,C) lp. */
int hpc_colsum(const void* x, int dtype, int64_t P, int C, float scale, float* out, void* scratch,
               size_t scratch_bytes, hpc_stream_t stream);
size_t hpc_colsum_scratch_bytes(int64_t P, int C);
/* hpc_colsum without a caller scratch buffer, for callers that cannot allocate device memory (CUDA-graph capture,
 * serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * The partial sums use a static device array inside the library (592 x HPC_COLSUM_STATIC_MAX_C floats =
 * 1.16 MiB on every GPU that loads the library, until exit); C > HPC_COLSUM_STATIC_MAX_C returns HPC_ERR_BAD_SHAPE.
 * The call allocates nothing (no cudaMalloc / cudaFree).  All calls on a device share the array: calls whose kernels
 * may overlap on the GPU (unordered streams) race; calls on one stream are safe.  Same result as hpc_colsum, bit for bit. */
#define HPC_COLSUM_STATIC_MAX_C 512
int hpc_colsum_static(const void* x, int dtype, int64_t P, int C, float scale, float* out, hpc_stream_t stream);

/* GDN helpers (elementwise, NHWC lp tensors of n elements).
 * hpc_square: out = x^2 (optionally relu first: gdn.py:80-81); used when the producer did not
 * emit x^2 itself.  hpc_gdn_bwd_t: t = g*x*nrm^3 (GDN) or g*x*nrm (IGDN), SURVEY A.4. */
int hpc_square(const void* x, void* out, int dtype, int64_t n, int relu, void* relu_out, hpc_stream_t stream);
int hpc_gdn_bwd_t(const void* g, const void* x, const void* nrm, void* t, int dtype, int64_t n, int inverse,
                  hpc_stream_t stream);

/* |x| and its autograd (torch.abs at modelling/meta_arch/bmshl2018.py:72), float, n elements */
int hpc_abs_fwd(const float* x, float* y, int64_t n, hpc_stream_t stream);
int hpc_abs_bwd(const float* x, const float* g, float* gx, int64_t n, hpc_stream_t stream);

/* out = g * [x > 0]  (autograd of F.relu in the stand-alone relu-GDN, gdn.py:80-81) */
int hpc_relu_mask(const void* g, const void* x, void* out, int dtype, int64_t n, hpc_stream_t stream);

/* Layout / dtype conversion: NCHW float <-> NHWC lp (module boundaries). */
int hpc_nchw_to_nhwc(const float* src, void* dst, int dtype, int N, int C, int H, int W, hpc_stream_t stream);
int hpc_nhwc_to_nchw(const void* src, int dtype, float* dst, int N, int C, int H, int W, hpc_stream_t stream);
/* elementwise cast between float and lp (n elements) */
int hpc_cast(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t n, hpc_stream_t stream);

/* Fused clip_grad_value_ + AdamW step over every parameter tensor in one launch.
 * Replaces engine/trainer.py:189-192 (clip_grad_value_(params, GRAD_CLIP); optimizer.step()) for the optimizer
 * solver/optim.py:27-39 builds: torch.optim.AdamW over one parameter group per tensor, each with its own lr and
 * weight decay.  One segment per tensor; the host passes the scalars torch derives in double per group and step t:
 *   decay = 1 - lr*weight_decay,  neg_step_size = -lr / (1 - beta1^t),  bc2_sqrt = sqrt(1 - beta2^t).
 * beta1, beta2, eps, clip are doubles (torch forms 1 - beta in double before the fp32 cast).  clip <= 0: no clipping.  write_clipped_grad != 0: the clamped gradient is also written back (what
 * clip_grad_value_ leaves in .grad).  p, g, m (exp_avg), v (exp_avg_sq): fp32 device pointers, n elements each. */
#define HPC_ADAMW_MAX_SEGS 56
typedef struct {
  float* p; float* g; float* m; float* v;
  int64_t n;
  float decay, neg_step_size, bc2_sqrt, reserved;
} hpc_adamw_seg;
int hpc_adamw_step(const hpc_adamw_seg* segs, int nseg, double beta1, double beta2, double eps, double clip,
                   int write_clipped_grad, hpc_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Entropy coding of the quantised latents: what the reference's `compress` / `decompress` /
 * "symbol" stubs stand for (modelling/meta_arch/bmshl2018.py:106-110, modelling/blocks/entropy_model.py:157-169,
 * 235-251).  Parity unpinned by the reference; the format is stated in csrc/rans_kernels.cu and restated in
 * oracle/rans_oracle.py.  rANS, 32-bit state, 16-bit words, 16-bit probabilities; one stream per image and group of
 * `cpg` consecutive channels; one thread codes one stream.
 *   symbols   (n_img*npos, C) float, NHWC rows, integer valued (the rounded latent)
 *   level     optional (n_img*npos, C) table index per symbol (conditional model: quantised scale); NULL: table = channel
 *   cum       (T, A+1) uint32 cumulative frequencies, cum[t][0] = 0, cum[t][A] = 65536, strictly increasing
 *   encode -> stream_buf (n_img*C/cpg, capacity) uint16, the stream's words in its LAST lengths[s] slots;
 *   compact -> words[offsets[s] .. + lengths[s]);  decode reads that layout.  err_flag: device int, set non-zero on a
 *   symbol outside the alphabet (encode) or a malformed stream (decode); the caller zeroes and checks it. */
int hpc_rans_stream_capacity(int npos, int cpg);
int hpc_rans_encode(const float* symbols, const unsigned char* level, const uint32_t* cum, int n_img, int npos, int C,
                    int cpg, int A, int vmin, unsigned short* stream_buf, int* lengths, int* err_flag,
                    hpc_stream_t stream);
int hpc_rans_compact(const unsigned short* stream_buf, int cap, const int* lengths, const int64_t* offsets, int nstreams,
                     unsigned short* words, hpc_stream_t stream);
int hpc_rans_decode(const unsigned short* words, const int64_t* offsets, const int* lengths, const unsigned char* level,
                    const uint32_t* cum, int n_img, int npos, int C, int cpg, int A, int vmin, float* symbols,
                    void* symbols_lp, int lp_dtype, int* err_flag, hpc_stream_t stream);
/* level[i] = smallest l in [0, nlevels) with sigma[i] <= scale_min * (scale_max/scale_min)^(l/(nlevels-1)) (the scale
 * table of the conditional model's coder; the likelihood itself is entropy_model.py:341-352 at those scales) */
int hpc_scale_levels(const float* sigma, int64_t n, float scale_min, float scale_max, int nlevels, unsigned char* level,
                     hpc_stream_t stream);
/* out2[0] = min, out2[1] = max over rint(x[i]) (int32); the caller initialises out2 to {INT_MAX, INT_MIN} */
int hpc_minmax_rint(const float* x, int64_t n, int* out2, hpc_stream_t stream);

/* Thin-channel layers (Cin=3 head of g_a, Cout=3 tail of g_s): K=75 is memory-bound.
 * conv_c3_fwd:   Conv2d(Cs->Cw,5,s2,p2): x NCHW float (N,Cs,H,W), w float (Cw,Cs,5,5) with
 *                strides (w_sr, w_sc) for (Cw,Cs) so the same kernel serves the dgrad of the tail
 *                deconv (weight (Cw,Cs,5,5) read transposed).  out NHWC (N,H/2,W/2,Cw).
 * deconv_c3_fwd: ConvTranspose2d(Cw->Cs,5,s2,p2,op1): x NHWC lp (N,H,W,Cw) -> out NCHW float
 *                (N,Cs,2H,2W) (+bias).
 * wgrad_c3:      gw(Cw,Cs,5,5)-strided float = sum_pixels wide[n,i,j,cw]*thin[n,cs,2i-2+kh,2j-2+kw]
 *                (serves Conv2d(3->C) weight grads and ConvTranspose2d(C->3) weight grads). */
int hpc_conv_c3_fwd(const float* x_nchw, const float* w, int64_t w_sr, int64_t w_sc, const float* bias, int N,
                    int Cs, int H, int W, int Cw, float* out_f32, void* out_lp, void* out_sq_lp, int lp_dtype,
                    hpc_stream_t stream);
int hpc_deconv_c3_fwd(const void* x_nhwc, int lp_dtype, const float* w, int64_t w_sr, int64_t w_sc,
                      const float* bias, int N, int Cw, int H, int W, int Cs, float* out_nchw,
                      hpc_stream_t stream);
int hpc_wgrad_c3(const void* wide_nhwc, int lp_dtype, const float* thin_nchw, int N, int Cw, int Cs, int H, int W,
                 float* gw, int64_t w_sr, int64_t w_sc, void* scratch, size_t scratch_bytes,
                 hpc_stream_t stream);
size_t hpc_wgrad_c3_scratch_bytes(int N, int Cw, int Cs, int H, int W);
/* Fused GDN / IGDN on tcgen05 (sm_100a, bf16 activations, fp32 accumulate) -- modelling/layers/gdn.py:79-88.
 * The activation is a [P pixels][C] row-major bf16 matrix (NHWC flattened), C in {64,128,192}.
 *  gdn_fwd_umma: n = beta + x^2 . gamma^T ; rnorm = rsqrt(n) ; y = x*rnorm (inverse: x*n*rnorm).
 *                gamma_bf16 = effective gamma (gdn.py:59-62 applied) as [C][C] bf16, row i = output channel.
 *  gdn_bwd_umma: SURVEY A.4.  t = g x rnorm^3 (inverse: g x rnorm); dx = g rnorm - x (t . gamma) (inverse:
 *                g / rnorm + x (t . gamma)); also writes t and x^2 (operands of the d(gamma) contraction through
 *                hpc_igemm_wgrad) and per-CTA column sums of t: colsum_part[parts][C], parts =
 *                hpc_gdn_bwd_umma_parts(P); d(beta) = -+1/2 * sum over parts.  gammaT_bf16[k][i] = gamma[i][k]. */
/*  gdn_prepare: NonNegativeParam.forward (gdn.py:59-62) of gamma and beta and both bf16 operand copies of gamma in
 *               one launch: beta_eff[C] fp32, gamma_bf16[i][k], gammaT_bf16[k][i]. */
int hpc_gdn_prepare(const float* gamma_param, const float* beta_param, int C, float gamma_bound, float beta_bound,
                    float pedestal, float* beta_eff, void* gamma_bf16, void* gammaT_bf16, hpc_stream_t stream);
/*  rnorm_bf16 may be NULL (evaluation under no_grad: r = n^-1/2 is only needed by the backward): the kernel then
 *  makes 2 HBM passes (read x, write y) instead of 3. */
int hpc_gdn_fwd_umma(const void* x_bf16, const void* gamma_bf16, const float* beta, int64_t P, int C, int inverse,
                     void* y_bf16, void* rnorm_bf16, hpc_stream_t stream);
int hpc_gdn_bwd_umma_parts(int64_t P);
/*  gdn_bwd_fused_umma (C in {128,192}): the whole backward of a GDN layer in two launches.  As gdn_bwd_umma, but the
 *                d(gamma) contraction sum_p t[p][i] x[p][j]^2 runs inside the kernel on the same shared-memory tiles
 *                (t and x^2 are never written to HBM), accumulated per CTA in TMEM; a finishing kernel reduces the
 *                per-CTA partials of d(gamma), of the column sums of t and of the column sums of dx in a fixed order
 *                (through `workspace`, >= hpc_gdn_bwd_fused_umma_workspace_bytes()) and writes
 *                  dgamma[C][C] = R(-+1/2 sum_p t x^2),  dbeta[C] = R(-+1/2 sum_p t),  dbias[C] = sum_p dx (nullable)
 *                where R is the backward of the non-negative reparametrisation (gdn.py:42-62 with bound.py:36-46)
 *                if the raw parameter (gamma_param / beta_param) is given, the identity if it is NULL.  dbias is
 *                the bias gradient of the conv / transposed conv whose output feeds the GDN (analysis.py:55-58). */
size_t hpc_gdn_bwd_fused_umma_workspace_bytes(int64_t P, int C);
int hpc_gdn_bwd_fused_umma(const void* g_bf16, const void* x_bf16, const void* rnorm_bf16, const void* gammaT_bf16,
                           int64_t P, int C, int inverse, void* dx_bf16, const float* gamma_param, float gamma_bound,
                           const float* beta_param, float beta_bound, float* dgamma, float* dbeta, float* dbias,
                           void* workspace, size_t workspace_bytes, hpc_stream_t stream);
/*  gdn_bwd_fused_rc_umma: the same without the saved r = n^-1/2: the kernel rebuilds n = beta_eff + x^2 . gamma^T on the
 *                tensor cores (the resident gamma^T tile read through an MN-major descriptor) and uses the forward's
 *                fp32 r instead of its bf16 rounding; the forward then need not write r at all (hpc_gdn_fwd_umma with
 *                rnorm_bf16 == NULL): 5 HBM passes per layer and step instead of 7.  beta_eff: C floats (hpc_gdn_prepare).
 *                Same reference code: modelling/layers/gdn.py:79-88 and its autograd. */
int hpc_gdn_bwd_fused_rc_umma(const void* g_bf16, const void* x_bf16, const float* beta_eff, const void* gammaT_bf16,
                           int64_t P, int C, int inverse, void* dx_bf16, const float* gamma_param, float gamma_bound,
                           const float* beta_param, float beta_bound, float* dgamma, float* dbeta, float* dbias,
                           void* workspace, size_t workspace_bytes, hpc_stream_t stream);
int hpc_gdn_bwd_umma(const void* g_bf16, const void* x_bf16, const void* rnorm_bf16, const void* gammaT_bf16, int64_t P,
                     int C, int inverse, void* dx_bf16, void* t_bf16, void* xsq_bf16, float* colsum_part,
                     hpc_stream_t stream);

/* tcgen05 (sm_100a, bf16 operands, fp32 accumulate) versions of the three thin-channel kernels above; same
 * arithmetic, same reference call sites (analysis.py:55, synthesis.py:55-57 and their autograd).
 *   pack_c3_weight:  w[cw*w_sr + cs*w_sc + kh*5+kw] -> kmaj [Cw][128] bf16 (k' = (cs*5+kh)*8 + kw: one 16-byte unit of 8
 *                    taps per window row, zero for kw >= 5 and k' >= 120; operand of conv_c3_fwd_umma, an opaque
 *                    format) and/or nmaj [80][Cw] bf16 (k = cs*25+kh*5+kw; operand of deconv_c3_fwd_umma)
 *   conv_c3_fwd_umma:   x NCHW float (N,3,H,W) -> out NHWC bf16 (N,H/2,W/2,Cw) (+ squares, may be NULL)
 *   deconv_c3_fwd_umma: x NHWC bf16 (N,H,W,Cw) -> out NCHW float (N,3,2H,2W)
 *   wgrad_c3_umma:      as wgrad_c3 with wide bf16; workspace from hpc_wgrad_c3_umma_workspace_bytes */
int hpc_pack_c3_weight(const float* w, int64_t w_sr, int64_t w_sc, int Cw, void* kmaj_bf16, void* nmaj_bf16,
                       hpc_stream_t stream);
/* index into the reference's (cs, kh*5+kw) axis held at position k' of a kmaj row, or -1 where it holds a zero */
int hpc_c3_kmaj_source(int k);
int hpc_conv_c3_fwd_umma(const float* x_nchw, const void* w_kmaj_bf16, const float* bias, int N, int H, int W, int Cw,
                         void* out_lp_bf16, void* out_sq_bf16, hpc_stream_t stream);
int hpc_deconv_c3_fwd_umma(const void* x_nhwc_bf16, const void* w_nmaj_bf16, const float* bias, int N, int Cw, int H,
                           int W, float* out_nchw, hpc_stream_t stream);
int hpc_wgrad_c3_umma(const void* wide_nhwc_bf16, const float* thin_nchw, int N, int Cw, int H, int W, float* gw,
                      int64_t w_sr, int64_t w_sc, void* workspace, size_t workspace_bytes, hpc_stream_t stream);
size_t hpc_wgrad_c3_umma_workspace_bytes(int N, int Cw, int H, int W);
/* per-channel sum of an NCHW float tensor (bias grad of the tail deconv); scratch: N*C floats */
int hpc_nchw_channel_sum(const float* x, int N, int C, int64_t HW, float* out, float* scratch, hpc_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Rate-distortion combine (bmshl2018.py:84-96): device scalars in, device scalars out, no sync
 *   out[0]=z_entropy out[1]=y_entropy out[2]=bpp out[3]=total_loss out[4+i]=distortion term i
 * term i = scale[i] * sum(dist[i][0..n[i]))   (scale = 1/(N*3*H*W) turns per-image SSEs into the
 * MSE of modelling/loss.py:25); total = lambda * sum_i term_i + bpp. */
#define HPC_MAX_DIST_TERMS 4
typedef struct {
  const float* dist[HPC_MAX_DIST_TERMS];
  int n[HPC_MAX_DIST_TERMS];
  float scale[HPC_MAX_DIST_TERMS];
  int n_terms;
} hpc_dist_terms;
int hpc_rd_combine(const float* z_bits, const float* y_bits, const hpc_dist_terms* terms, float lambda,
                   float inv_num_pixels, float* out8, hpc_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* HPCODEC_H_ */
