// Range-ANS entropy coder for the quantised latents -- the arithmetic behind `compress` / `decompress` / "symbol"
// mode, which the reference leaves as stubs (modelling/meta_arch/bmshl2018.py:106-110,
// modelling/blocks/entropy_model.py:157-169,235-251).  Parity is UNPINNED by the reference (nothing there implements or
// tests a coder); the checker is oracle/rans_oracle.py, a numpy restatement of this very format.
//
// Format (rANS, Duda 2013; 32-bit state, 16-bit renormalisation words, 16-bit probabilities):
//   * a latent of one image, (pos, C) rounded values in NHWC order, is cut into S = C / cpg streams of `cpg` consecutive
//     channels; stream s holds the symbols (pos, c) for pos = 0..npos-1, c = s*cpg .. s*cpg+cpg-1, in that order;
//   * symbol v is coded with table t = c (factorised prior: one table per channel) or t = level[pos][c] (conditional
//     model: one table per quantised scale), as index v - vmin into cum[t][0..A] (cum[t][A] = 65536, every freq >= 1);
//   * encoder: x = 65536; for the symbols in REVERSE order { if x >= freq << 16: emit x & 0xffff, x >>= 16;
//     x = (x / freq << 16) + x % freq + start };  finally emit x & 0xffff, then x >> 16.  Words are stored so that the
//     decoder reads them forwards: state high, state low, then the renormalisation words in decoding order;
//   * decoder: x = w0 << 16 | w1; per symbol { slot = x & 0xffff; i with cum[i] <= slot < cum[i+1];
//     x = freq * (x >> 16) + slot - cum[i]; if x < 65536: x = x << 16 | next word }.
// One thread codes one stream (the recurrence is serial); a batch of 64 crops gives 64 * 24 * 2 independent streams.
#include "hpc_common.cuh"

namespace {
constexpr uint32_t RANS_L = 1u << 16;
constexpr int RANS_THREADS = 64;

struct RansGeom {
  int n_img, npos, C, cpg, S;      // S = C / cpg streams per image
  int A, vmin;                     // alphabet size, value of index 0
  int cap;                         // words per stream buffer (encode) = npos * cpg + 2
};

__global__ void __launch_bounds__(RANS_THREADS)
rans_encode_kernel(const float* __restrict__ sym, const uint8_t* __restrict__ level, const uint32_t* __restrict__ cum,
                   RansGeom g, uint16_t* __restrict__ buf, int* __restrict__ lengths, int* __restrict__ err) {
  const int sid = blockIdx.x * blockDim.x + threadIdx.x;
  if (sid >= g.n_img * g.S) return;
  const int n = sid / g.S, s = sid % g.S;
  uint16_t* out = buf + (int64_t)sid * g.cap;
  int w = g.cap;                                   // words are written backwards from the end of the stream buffer
  uint32_t x = RANS_L;
  for (int pos = g.npos - 1; pos >= 0; --pos) {
    const int64_t base = ((int64_t)n * g.npos + pos) * g.C + s * g.cpg;
    for (int j = g.cpg - 1; j >= 0; --j) {
      const int c = s * g.cpg + j;
      const int v = (int)rintf(sym[base + j]) - g.vmin;
      if (v < 0 || v >= g.A) { atomicExch(err, 1); continue; }
      const int t = level ? (int)level[base + j] : c;
      const uint32_t* ct = cum + (int64_t)t * (g.A + 1);
      const uint32_t start = ct[v], freq = ct[v + 1] - start;
      if ((uint64_t)x >= ((uint64_t)freq << 16)) { out[--w] = (uint16_t)(x & 0xffffu); x >>= 16; }
      x = ((x / freq) << 16) + (x % freq) + start;
    }
  }
  out[--w] = (uint16_t)(x & 0xffffu);
  out[--w] = (uint16_t)(x >> 16);
  lengths[sid] = g.cap - w;
}

// stream sid: the last lengths[sid] words of its buffer -> words[offsets[sid] ...]
__global__ void rans_compact_kernel(const uint16_t* __restrict__ buf, int cap, const int* __restrict__ lengths,
                                    const int64_t* __restrict__ offsets, uint16_t* __restrict__ words) {
  const int sid = blockIdx.x;
  const int L = lengths[sid];
  const uint16_t* src = buf + (int64_t)sid * cap + (cap - L);
  uint16_t* dst = words + offsets[sid];
  for (int i = threadIdx.x; i < L; i += blockDim.x) dst[i] = src[i];
}

template <typename LP>
__global__ void __launch_bounds__(RANS_THREADS)
rans_decode_kernel(const uint16_t* __restrict__ words, const int64_t* __restrict__ offsets, const int* __restrict__ lengths,
                   const uint8_t* __restrict__ level, const uint32_t* __restrict__ cum, RansGeom g,
                   float* __restrict__ sym, LP* __restrict__ sym_lp, int* __restrict__ err) {
  const int sid = blockIdx.x * blockDim.x + threadIdx.x;
  if (sid >= g.n_img * g.S) return;
  const int n = sid / g.S, s = sid % g.S;
  const uint16_t* in = words + offsets[sid];
  const int L = lengths[sid];
  if (L < 2) { atomicExch(err, 2); return; }
  uint32_t x = ((uint32_t)in[0] << 16) | in[1];
  int r = 2;
  for (int pos = 0; pos < g.npos; ++pos) {
    const int64_t base = ((int64_t)n * g.npos + pos) * g.C + s * g.cpg;
    for (int j = 0; j < g.cpg; ++j) {
      const int t = level ? (int)level[base + j] : s * g.cpg + j;
      const uint32_t* ct = cum + (int64_t)t * (g.A + 1);
      const uint32_t slot = x & 0xffffu;
      int lo = 0, hi = g.A;                        // largest i with ct[i] <= slot
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (ct[mid] <= slot) lo = mid; else hi = mid;
      }
      const uint32_t start = ct[lo], freq = ct[lo + 1] - start;
      x = freq * (x >> 16) + slot - start;
      if (x < RANS_L) {
        if (r >= L) { atomicExch(err, 3); return; }
        x = (x << 16) | in[r++];
      }
      const float v = (float)(lo + g.vmin);
      sym[base + j] = v;
      if (sym_lp) st_from_float<LP>(sym_lp, base + j, v);
    }
  }
  if (r != L || x != RANS_L) atomicExch(err, 4);    // a well-formed stream ends exactly here, back in the initial state
}

// level[i] = smallest l with sigma[i] <= smin * exp(l * step), clamped to nlevels-1 (tables are built for those scales)
__global__ void scale_levels_kernel(const float* __restrict__ sigma, int64_t n, float log_smin, float inv_step, int nlevels,
                                    uint8_t* __restrict__ level) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float l = ceilf((logf(sigma[i]) - log_smin) * inv_step);
    level[i] = (uint8_t)(l < 0.f ? 0.f : (l > (float)(nlevels - 1) ? (float)(nlevels - 1) : l));
  }
}

// out[0] = min, out[1] = max of rint(x[i]) as int32 (out initialised by the caller to INT_MAX / INT_MIN)
__global__ void minmax_rint_kernel(const float* __restrict__ x, int64_t n, int* __restrict__ out) {
  int lo = 2147483647, hi = -2147483647 - 1;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int v = (int)rintf(x[i]);
    lo = v < lo ? v : lo; hi = v > hi ? v : hi;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const int a = __shfl_xor_sync(0xffffffffu, lo, o), b = __shfl_xor_sync(0xffffffffu, hi, o);
    lo = a < lo ? a : lo; hi = b > hi ? b : hi;
  }
  if ((threadIdx.x & 31) == 0) { atomicMin(out, lo); atomicMax(out + 1, hi); }
}

int rans_geom(RansGeom* g, int n_img, int npos, int C, int cpg, int A, int vmin, const char* who) {
  HPC_REQUIRE(n_img > 0 && npos > 0 && C > 0 && cpg > 0 && C % cpg == 0, HPC_ERR_BAD_SHAPE,
              "%s: bad geometry (images %d, positions %d, C %d, channels per stream %d)", who, n_img, npos, C, cpg);
  HPC_REQUIRE(A >= 1 && A <= 65536 - 1, HPC_ERR_BAD_SHAPE, "%s: alphabet size %d", who, A);
  g->n_img = n_img; g->npos = npos; g->C = C; g->cpg = cpg; g->S = C / cpg; g->A = A; g->vmin = vmin;
  g->cap = npos * cpg + 2;
  return HPC_OK;
}
}  // namespace

int hpc_rans_stream_capacity(int npos, int cpg) { return npos * cpg + 2; }

int hpc_rans_encode(const float* symbols, const unsigned char* level, const uint32_t* cum, int n_img, int npos, int C,
                    int cpg, int A, int vmin, unsigned short* stream_buf, int* lengths, int* err_flag,
                    hpc_stream_t stream) {
  RansGeom g;
  int rc = rans_geom(&g, n_img, npos, C, cpg, A, vmin, "hpc_rans_encode");
  if (rc) return rc;
  HPC_REQUIRE(symbols && cum && stream_buf && lengths && err_flag, HPC_ERR_BAD_SHAPE, "hpc_rans_encode: NULL pointer");
  const int ns = n_img * g.S;
  rans_encode_kernel<<<(ns + RANS_THREADS - 1) / RANS_THREADS, RANS_THREADS, 0, (cudaStream_t)stream>>>(
      symbols, level, cum, g, stream_buf, lengths, err_flag);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

int hpc_rans_compact(const unsigned short* stream_buf, int cap, const int* lengths, const int64_t* offsets, int nstreams,
                     unsigned short* words, hpc_stream_t stream) {
  HPC_REQUIRE(stream_buf && lengths && offsets && words && nstreams > 0 && cap > 0, HPC_ERR_BAD_SHAPE,
              "hpc_rans_compact: bad arguments");
  rans_compact_kernel<<<nstreams, 128, 0, (cudaStream_t)stream>>>(stream_buf, cap, lengths, offsets, words);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

int hpc_rans_decode(const unsigned short* words, const int64_t* offsets, const int* lengths, const unsigned char* level,
                    const uint32_t* cum, int n_img, int npos, int C, int cpg, int A, int vmin, float* symbols,
                    void* symbols_lp, int lp_dtype, int* err_flag, hpc_stream_t stream) {
  RansGeom g;
  int rc = rans_geom(&g, n_img, npos, C, cpg, A, vmin, "hpc_rans_decode");
  if (rc) return rc;
  HPC_REQUIRE(words && offsets && lengths && cum && symbols && err_flag, HPC_ERR_BAD_SHAPE, "hpc_rans_decode: NULL pointer");
  const int ns = n_img * g.S;
  const int grid = (ns + RANS_THREADS - 1) / RANS_THREADS;
  if (symbols_lp && lp_dtype == HPC_BF16)
    rans_decode_kernel<__nv_bfloat16><<<grid, RANS_THREADS, 0, (cudaStream_t)stream>>>(
        words, offsets, lengths, level, cum, g, symbols, (__nv_bfloat16*)symbols_lp, err_flag);
  else
    rans_decode_kernel<float><<<grid, RANS_THREADS, 0, (cudaStream_t)stream>>>(
        words, offsets, lengths, level, cum, g, symbols, (float*)symbols_lp, err_flag);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

int hpc_scale_levels(const float* sigma, int64_t n, float scale_min, float scale_max, int nlevels, unsigned char* level,
                     hpc_stream_t stream) {
  HPC_REQUIRE(sigma && level && n > 0 && nlevels >= 2 && nlevels <= 256 && scale_min > 0.f && scale_max > scale_min,
              HPC_ERR_BAD_SHAPE, "hpc_scale_levels: bad arguments");
  const float step = (logf(scale_max) - logf(scale_min)) / (float)(nlevels - 1);
  int64_t b = (n + 255) / 256;
  if (b > 148 * 8) b = 148 * 8;
  scale_levels_kernel<<<(int)b, 256, 0, (cudaStream_t)stream>>>(sigma, n, logf(scale_min), 1.f / step, nlevels, level);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

int hpc_minmax_rint(const float* x, int64_t n, int* out2, hpc_stream_t stream) {
  HPC_REQUIRE(x && out2 && n > 0, HPC_ERR_BAD_SHAPE, "hpc_minmax_rint: bad arguments");
  int64_t b = (n + 255) / 256;
  if (b > 148 * 8) b = 148 * 8;
  minmax_rint_kernel<<<(int)b, 256, 0, (cudaStream_t)stream>>>(x, n, out2);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}
