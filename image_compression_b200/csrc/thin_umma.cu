// tcgen05 versions of the thin-channel layers (sm_100a): the Cin=3 head of g_a (analysis.py:55), the
// Cout=3 tail of g_s (synthesis.py:55-57) and their weight gradients.  All three are HBM-bound
// (SURVEY 8d): the 192-channel tensor is streamed exactly once; the 75-long (cs,kh,kw) axis is padded to
// 80 and becomes the K (conv), N (transposed conv) or M (wgrad) dimension of a dense UMMA.
//
//   conv    D[128 px x Cw]   = im2col(x)[128 x 80] . W[Cw x 80]^T        im2col rows built in smem by 4 warps
//   deconv  P[128 px x 80]   = X[128 x Cw] . W'[80 x Cw]^T  then col2im gather of the 2x2 output phases in smem
//   wgrad   G[80 x Cw]      += im2col(thin)[px x 80]^T . wide[px x Cw]   split-K over pixel blocks, 2-pass reduce
//
// k = cs*25 + kh*5 + kw everywhere (the reference's own weight memory order for both (Cw,Cs,5,5) layouts).
#include <cuda.h>

#include "tmap.cuh"
#include "umma_ptx.cuh"

using namespace umma;

namespace {

constexpr int KPAD = 80;            // 75 taps*channels padded to a multiple of 16
constexpr int TCS = 3;              // thin channels (RGB)
constexpr int ATOM_BYTES = 8192;    // 64 rows x 128 B

__device__ __forceinline__ void named_bar(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__device__ __forceinline__ uint32_t pack2_bf16(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

}  // namespace

// ---------------------------------------------------------------------------------------
// weight packing: w[cw*w_sr + cs*w_sc + t] (t = kh*5+kw) ->
//   kmaj [Cw][128] bf16, k' = (cs*5 + kh)*8 + kw: every (cs, kh) row of the 5x5 window is one 16-byte unit of 8 taps,
//                  kw = 5..7 and k' >= 120 zero       : B operand of the conv (its im2col rows copy 8 consecutive
//                  pixels per unit and let the zero weights mask the 3 extra ones)
//   nmaj [80][Cw]  bf16 (k = cs*25 + t, cw contiguous, rows >= 75 zero): B operand of the transposed conv
// ---------------------------------------------------------------------------------------
// (c3_kmaj_source() in hpc_common.cuh maps k' back to the (cs, t) axis)
__global__ void pack_c3_weight_kernel(const float* __restrict__ w, int64_t w_sr, int64_t w_sc, int Cw,
                                      __nv_bfloat16* __restrict__ kmaj, __nv_bfloat16* __restrict__ nmaj) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (kmaj && i < Cw * 128) {
    const int cw = i / 128, k = i % 128;
    const int src = c3_kmaj_source(k);
    kmaj[i] = __float2bfloat16_rn(src >= 0 ? w[cw * w_sr + (src / 25) * w_sc + (src % 25)] : 0.f);
  }
  if (nmaj && i < KPAD * Cw) {
    const int k = i / Cw, cw = i % Cw;
    nmaj[i] = __float2bfloat16_rn(k < 75 ? w[cw * w_sr + (k / 25) * w_sc + (k % 25)] : 0.f);
  }
}

int hpc_c3_kmaj_source(int k) { return c3_kmaj_source(k); }

int hpc_pack_c3_weight(const float* w, int64_t w_sr, int64_t w_sc, int Cw, void* kmaj_bf16, void* nmaj_bf16,
                       hpc_stream_t stream) {
  HPC_REQUIRE(w && Cw > 0 && (kmaj_bf16 || nmaj_bf16), HPC_ERR_BAD_SHAPE, "hpc_pack_c3_weight: bad arguments");
  const int n = Cw * 128;
  pack_c3_weight_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(w, w_sr, w_sc, Cw, (__nv_bfloat16*)kmaj_bf16,
                                                                         (__nv_bfloat16*)nmaj_bf16);
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

// ---------------------------------------------------------------------------------------
// T1: Conv2d(3 -> Cw, 5, s2, p2) : x NCHW fp32 -> out NHWC bf16 (+ squares)
//   warp 0: weights via TMA, MMA issue, TMEM owner | warps 1-4: im2col builders | warps 5-12: epilogue (the kernel is
//   bound by the epilogue: K = 80 makes the MMA negligible, so two warps share each TMEM lane quarter)
// ---------------------------------------------------------------------------------------
namespace {
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* m, const void* src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void bulk_commit_group() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_group_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void bulk_wait_group_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

constexpr int T1_TH = 8, T1_TW = 16;                      // 128 output pixels per tile
constexpr int T1_PH = 2 * T1_TH + 3, T1_PWV = 2 * T1_TW + 3;   // 19 x 35 input pixels per channel
constexpr int T1_PWF = 40;     // floats per patch row as the TMA box lands it.  The box starts FOUR pixels left of the tile's
                               // first tap (TMA wants the innermost start coordinate on a 16-byte boundary: x0 = 2 j0 - 4;
                               // with 2 j0 - 2 the load faults with "illegal instruction"), so pixel (jl, kw) is column
                               // 2 jl + 2 + kw and a unit's 8 values end at column 39
constexpr int T1_PX0 = 4;      // image columns left of 2 j0 in the box
constexpr int T1_NPATCH = 2;   // patch buffers: the box of tile it+2 is in flight while tile it+1 is gathered
constexpr int T1_PATCH_BYTES = ((TCS * T1_PH * T1_PWF * 4 + 127) / 128) * 128;   // 9216
constexpr int T1_UNITS = TCS * 5;                         // 16-byte units per im2col row: one per (cs, kh); unit 15 stays zero
constexpr int T1_THREADS = 416;
constexpr int T1_EPI = 256;                               // epilogue threads
constexpr int T1_A_BYTES = 2 * 16384;                     // two 64-wide K chunks of [128 x 128 B]
constexpr int T1_ST_BYTES = 128 * 128;                    // staging tile of one tensor, one 64-channel chunk: [128 pixels x 128 B],
                                                          // SWIZZLE_128B (the layout a TMA tensor store reads)
constexpr int T1_NSTAGE = 3;                              // staging tiles in flight (see the epilogue)

struct T1Params {
  const float* x; const float* bias;
  __nv_bfloat16* out_lp; __nv_bfloat16* out_sq;
  int N, H, W, OH, OW, Cw;
  int tiles_w, tiles_h, total_tiles;
};

template <int CW>
struct T1Cfg {
  static constexpr int B_BYTES = 2 * CW * 128;
  static constexpr int ACC_STRIDE = CW <= 64 ? 64 : (CW <= 128 ? 128 : 256);
  static constexpr int TMEM_COLS = 2 * ACC_STRIDE;
  static constexpr int PATCH_BYTES = T1_NPATCH * T1_PATCH_BYTES;
  // no alignment slack (CW = 192 fills the 227 KB exactly; the kernel traps if the dynamic window is not 1 KB aligned)
  static constexpr int SMEM = 2 * T1_A_BYTES + B_BYTES + 2 * T1_NSTAGE * T1_ST_BYTES + PATCH_BYTES + CW * 4 /*bias*/ + 256;
  static_assert(SMEM <= 232448, "shared memory budget");
};

template <int CW>
__global__ void __launch_bounds__(T1_THREADS, 1)
conv_c3_umma_kernel(const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmOut,
                    const __grid_constant__ CUtensorMap tmSq, const __grid_constant__ CUtensorMap tmX,
                    const __grid_constant__ T1Params p) {
  using Cfg = T1Cfg<CW>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if (smem_u32(smem_raw) & 1023u) {                         // no slack budgeted (T1Cfg::SMEM)
    if (threadIdx.x == 0) printf("hpcodec: conv_c3_umma: dynamic shared memory window not 1 KB aligned\n");
    __trap();
  }
  uint8_t* smem_a = smem;                                   // 2 x 32 KB
  uint8_t* smem_b = smem_a + 2 * T1_A_BYTES;                // 2 chunks x [CW x 128 B]
  uint8_t* stage = smem_b + Cfg::B_BYTES;                    // [T1_NSTAGE][2 tensors][128 x 128 B], 1 KB aligned
  uint8_t* patch = stage + 2 * T1_NSTAGE * T1_ST_BYTES;      // [T1_NPATCH][TCS][T1_PH][T1_PWF] fp32, written by the TMA
  float* bias_s = reinterpret_cast<float*>(patch + Cfg::PATCH_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(bias_s + CW);
  uint64_t* full_a = bars;          // [2] builders -> MMA   (128 arrivals)
  uint64_t* empty_a = bars + 2;     // [2] MMA -> builders   (tcgen05.commit)
  uint64_t* tfull = bars + 4;       // [2] MMA -> epilogue
  uint64_t* tempty = bars + 6;      // [2] epilogue -> MMA   (4 arrivals)
  uint64_t* b_full = bars + 8;
  uint64_t* p_full = bars + 9;      // [T1_NPATCH] TMA -> builders: the input patch of a tile has landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9 + T1_NPATCH);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // zero the A buffers once: the last unit (k' >= 120) is never written again
  for (int i = threadIdx.x; i < 2 * T1_A_BYTES / 16; i += T1_THREADS) reinterpret_cast<uint4*>(smem_a)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x == 0) {
    prefetch_tmap(&tmB); prefetch_tmap(&tmOut); prefetch_tmap(&tmSq); prefetch_tmap(&tmX);
    for (int s = 0; s < T1_NPATCH; ++s) mbar_init(&p_full[s], 1);
    for (int s = 0; s < 2; ++s) { mbar_init(&full_a[s], 128); mbar_init(&empty_a[s], 1); mbar_init(&tfull[s], 1); mbar_init(&tempty[s], T1_EPI / 32); }
    mbar_init(b_full, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
  pdl_wait();                      // the prologue above overlaps the previous kernel's tail
  for (int i = threadIdx.x; i < CW; i += T1_THREADS) bias_s[i] = p.bias ? p.bias[i] : 0.f;
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  const int tiles_per_img = p.tiles_w * p.tiles_h;

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(b_full, Cfg::B_BYTES);
      tma_load_2d(smem_b, &tmB, b_full, 0, 0);
      tma_load_2d(smem_b + CW * 128, &tmB, b_full, 64, 0);
      mbar_wait(b_full, 0);
      constexpr uint32_t idesc = make_idesc_bf16(128, CW, 0, 0);
      int it = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
        const int buf = it & 1;
        const uint32_t ph = (it >> 1) & 1;
        mbar_wait(&tempty[buf], ph ^ 1);
        mbar_wait(&full_a[buf], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem_a + buf * T1_A_BYTES), b_addr = smem_u32(smem_b);
        const uint32_t d = tmem_base + buf * Cfg::ACC_STRIDE;
#pragma unroll
        for (int k = 0; k < 128 / 16; ++k) {                 // K = 128: 15 units of 8 taps (5 live each) + one zero unit
          const uint32_t ao = (k >> 2) * 16384 + (k & 3) * 32, bo = (k >> 2) * (CW * 128) + (k & 3) * 32;
          mma_bf16_ss(d, make_smem_desc_sw128(a_addr + ao, 16, 1024), make_smem_desc_sw128(b_addr + bo, 16, 1024), idesc,
                      k != 0);
        }
        mma_commit(&empty_a[buf]);
        mma_commit(&tfull[buf]);
      }
    }
  } else if (warp <= 4) {
    // ============================ im2col builders (128 threads) ============================
    // The input patch of a tile ((2 x 8 + 3) x 40 pixels of the 3 channels, the convolution's zero padding included) is
    // ONE TMA box of the NCHW fp32 image (out-of-bounds = zero fill), two tiles ahead.  (Fetching it with 16 predicated
    // loads per thread was 2/3 of the builders' instructions and a third of their time was the wait for those loads:
    // ncu r02u -- and the builders, not the epilogue, bound the kernel.)
    const int t = threadIdx.x - 32;
    const int il = t >> 4, jl = t & 15;
    auto issue_patch = [&](int tile, int pb) {
      const int n = tile / tiles_per_img, r = tile % tiles_per_img;
      mbar_arrive_expect_tx(&p_full[pb], TCS * T1_PH * T1_PWF * 4);
      tma_load_3d(patch + pb * T1_PATCH_BYTES, &tmX, &p_full[pb], 2 * (r % p.tiles_w) * T1_TW - T1_PX0, 2 * (r / p.tiles_w) * T1_TH - 2,
                  n * TCS);
    };
    if (t == 0)
      for (int k = 0; k < T1_NPATCH; ++k)
        if ((int)blockIdx.x + k * (int)gridDim.x < p.total_tiles) issue_patch(blockIdx.x + k * gridDim.x, k);
    int it = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
      const int buf = it & 1, pb = it % T1_NPATCH;
      const uint32_t ph = (it >> 1) & 1;
      mbar_wait(&p_full[pb], (it / T1_NPATCH) & 1);
      mbar_wait(&empty_a[buf], ph ^ 1);
      uint8_t* arow = smem_a + buf * T1_A_BYTES + t * 128;
      // unit u = (cs, kh): the 8 patch values at row 2 il + kh from column 2 jl + 2 (= image column 2 (j0 + jl) - 2; 8-byte
      // aligned: four 64-bit loads, a half-warp reads 32 consecutive words) -- taps kw = 5..7 meet zero weights (kmaj
      // layout).  (A non-finite pixel would therefore also reach the outputs up to 1.5 pixels to its left; the
      // reference's inputs are images.)
      const float2* prow = reinterpret_cast<const float2*>(patch + pb * T1_PATCH_BYTES) + il * T1_PWF + jl + (T1_PX0 - 2) / 2;
#pragma unroll
      for (int u = 0; u < T1_UNITS; ++u) {
        const float2* src = prow + ((u / 5) * T1_PH + (u % 5)) * (T1_PWF / 2);
        const float2 a = src[0], b = src[1], c = src[2], d = src[3];
        *reinterpret_cast<uint4*>(arow + (u >> 3) * 16384 + (((u & 7) ^ (t & 7)) << 4)) =
            make_uint4(pack2_bf16(a.x, a.y), pack2_bf16(b.x, b.y), pack2_bf16(c.x, c.y), pack2_bf16(d.x, d.y));
      }
      fence_proxy_async();
      mbar_arrive(&full_a[buf]);
      named_bar(1, 128);                                   // every builder has finished reading patch[pb]
      const int nxt = tile + T1_NPATCH * gridDim.x;
      if (t == 0 && nxt < p.total_tiles) issue_patch(nxt, pb);
    }
  } else {
    // ============================ epilogue (256 threads) ===================================
    const int t = threadIdx.x - 160;
    const int q = warp & 3, half = (warp - 5) >> 2;         // TMEM lane quarter, 32-column half of each 64-channel chunk
    const int row = q * 32 + lane;                          // TMEM lane == tile pixel
    int it = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
      const int buf = it & 1;
      const uint32_t ph = (it >> 1) & 1;
      const int n = tile / tiles_per_img, r = tile % tiles_per_img;
      const int i0 = (r / p.tiles_w) * T1_TH, j0 = (r % p.tiles_w) * T1_TW;
      mbar_wait(&tfull[buf], ph);
      tc_fence_after();
      const uint32_t taddr = tmem_base + buf * Cfg::ACC_STRIDE + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
      for (int cc = 0; cc < CW / 64; ++cc) {
        // A thread owns one pixel (= TMEM lane) and 32 channels of the chunk: it writes them as four 16-byte units of its
        // 128-byte row of a SWIZZLE_128B staging tile (unit u of row r at (u ^ (r & 7)) * 16: the 8 lanes of a store
        // wavefront hit 8 different bank groups), and ONE thread then hands the whole [8 x 16 pixels][64 channels] tile to
        // the TMA (one tensor store per tile and chunk; rows past the image edge are clipped by the tensor map).
        // Three staging tiles rotate on the running chunk count: the issuing thread waits, after committing the store of
        // chunk j, until the store of chunk j-1 has been read out of shared memory, and only then joins the barrier of
        // chunk j+1 -- whose passing is what lets the writers of chunk j+2 start on the tile chunk j-1 used.
        const int j = it * (CW / 64) + cc;
        uint8_t* st_lp = stage + (j % T1_NSTAGE) * 2 * T1_ST_BYTES;
        uint8_t* st_sq = st_lp + T1_ST_BYTES;
        float v2[2][16];
#pragma unroll
        for (int c16 = 0; c16 < 2; ++c16) tmem_ld16(taddr + cc * 64 + half * 32 + c16 * 16, v2[c16]);
        tmem_ld_wait();
#pragma unroll
        for (int c16 = 0; c16 < 2; ++c16) {
          const float* v = v2[c16];
          const int cb = half * 32 + c16 * 16;              // first channel of this group within the chunk
          uint32_t lo[8], sq[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const float a = v[2 * e] + bias_s[cc * 64 + cb + 2 * e];
            const float b = v[2 * e + 1] + bias_s[cc * 64 + cb + 2 * e + 1];
            lo[e] = pack2_bf16(a, b);
            if (p.out_sq) sq[e] = pack2_bf16(a * a, b * b);      // (uniform branch: the fused GDN path never asks for x^2)
          }
          const int u0 = cb >> 3;                            // 16-byte unit of the first 8 channels
          const uint32_t o0 = row * 128 + (((u0) ^ (row & 7)) << 4), o1 = row * 128 + (((u0 + 1) ^ (row & 7)) << 4);
          *reinterpret_cast<uint4*>(st_lp + o0) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
          *reinterpret_cast<uint4*>(st_lp + o1) = make_uint4(lo[4], lo[5], lo[6], lo[7]);
          if (p.out_sq) {
            *reinterpret_cast<uint4*>(st_sq + o0) = make_uint4(sq[0], sq[1], sq[2], sq[3]);
            *reinterpret_cast<uint4*>(st_sq + o1) = make_uint4(sq[4], sq[5], sq[6], sq[7]);
          }
        }
        if (cc == CW / 64 - 1) {                           // accumulator fully read: hand it back early
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty[buf]);
        }
        fence_proxy_async();                               // the staged rows must be visible to the TMA (async proxy)
        named_bar(2, T1_EPI);
        if (t == 0) {
          tma_store_4d(&tmOut, st_lp, cc * 64, j0, i0, n);
          if (p.out_sq) tma_store_4d(&tmSq, st_sq, cc * 64, j0, i0, n);
          bulk_commit_group();
          bulk_wait_group_read<1>();
        }
      }
    }
    if (t == 0) bulk_wait_group_all();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

template <int CW>
int launch_t1(const CUtensorMap& tmB, const CUtensorMap& tmOut, const CUtensorMap& tmSq, const CUtensorMap& tmX,
              const T1Params& kp, cudaStream_t st) {
  using Cfg = T1Cfg<CW>;
  static HpcOncePerDevice attr_once;
  int attr_dev;
  if (attr_once.need(&attr_dev)) {
    HPC_CUDA(cudaFuncSetAttribute(conv_c3_umma_kernel<CW>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
    attr_once.mark(attr_dev);
  }
  const int grid = kp.total_tiles < hpc_num_sms() ? kp.total_tiles : hpc_num_sms();
  HPC_CUDA(hpc_launch_pdl(conv_c3_umma_kernel<CW>, dim3(grid), dim3(T1_THREADS), Cfg::SMEM, st, tmB, tmOut, tmSq, tmX, kp));
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}
}  // namespace

int hpc_conv_c3_fwd_umma(const float* x_nchw, const void* w_kmaj_bf16, const float* bias, int N, int H, int W, int Cw,
                         void* out_lp_bf16, void* out_sq_bf16, hpc_stream_t stream) {
  HPC_REQUIRE(N > 0 && H > 0 && W > 0, HPC_ERR_BAD_SHAPE, "hpc_conv_c3_fwd_umma: empty tensor");
  HPC_REQUIRE(H % 2 == 0 && W % 2 == 0, HPC_ERR_BAD_SHAPE, "hpc_conv_c3_fwd_umma: H, W must be even (got %dx%d)", H, W);
  HPC_REQUIRE(Cw == 64 || Cw == 128 || Cw == 192, HPC_ERR_UNSUPPORTED,
              "hpc_conv_c3_fwd_umma: Cw=%d not in {64,128,192} (256 channels do not fit the 227 KB of shared memory beside the "
              "staging tiles: use hpc_conv_c3_fwd)", Cw);
  HPC_REQUIRE(x_nchw && w_kmaj_bf16 && out_lp_bf16, HPC_ERR_BAD_SHAPE, "hpc_conv_c3_fwd_umma: NULL pointer");
  HPC_REQUIRE(W % 4 == 0 && hpc_aligned16(x_nchw), HPC_ERR_BAD_ALIGNMENT,
              "hpc_conv_c3_fwd_umma: the image rows are fetched by TMA: W %% 4 == 0 and a 16-byte aligned x (W = %d); use "
              "hpc_conv_c3_fwd otherwise", W);
  HPC_REQUIRE(hpc_aligned16(w_kmaj_bf16) && hpc_aligned16(out_lp_bf16) && hpc_aligned16(out_sq_bf16), HPC_ERR_BAD_ALIGNMENT,
              "hpc_conv_c3_fwd_umma: pointers must be 16B aligned");
  T1Params kp;
  kp.x = x_nchw; kp.bias = bias; kp.out_lp = (__nv_bfloat16*)out_lp_bf16; kp.out_sq = (__nv_bfloat16*)out_sq_bf16;
  kp.N = N; kp.H = H; kp.W = W; kp.OH = H / 2; kp.OW = W / 2; kp.Cw = Cw;
  kp.tiles_w = (kp.OW + T1_TW - 1) / T1_TW; kp.tiles_h = (kp.OH + T1_TH - 1) / T1_TH;
  kp.total_tiles = N * kp.tiles_w * kp.tiles_h;
  CUtensorMap tmB;
  uint64_t dims[2] = {128, (uint64_t)Cw};
  uint64_t strides[1] = {256};
  uint32_t box[2] = {64, (uint32_t)Cw};
  uint32_t estr[2] = {1, 1};
  int rc = make_tmap(&tmB, w_kmaj_bf16, 2, dims, strides, box, estr);
  if (rc) return rc;
  // output tiles leave through TMA tensor stores: NHWC as a 4-D map (C, OW, OH, N), box = 64 channels x 16 x 8 pixels
  CUtensorMap tmOut, tmSq;
  uint64_t odims[4] = {(uint64_t)Cw, (uint64_t)kp.OW, (uint64_t)kp.OH, (uint64_t)N};
  uint64_t ostr[3] = {(uint64_t)Cw * 2, (uint64_t)kp.OW * Cw * 2, (uint64_t)kp.OH * kp.OW * Cw * 2};
  uint32_t obox[4] = {64, T1_TW, T1_TH, 1};
  uint32_t oestr[4] = {1, 1, 1, 1};
  if ((rc = make_tmap(&tmOut, out_lp_bf16, 4, odims, ostr, obox, oestr))) return rc;
  if ((rc = make_tmap(&tmSq, out_sq_bf16 ? out_sq_bf16 : out_lp_bf16, 4, odims, ostr, obox, oestr))) return rc;
  // the input as a 3-D fp32 map (W, H, N*3): one box = the patch of a tile, T1_PWF columns x T1_PH rows x 3 channels
  CUtensorMap tmX;
  uint64_t xdims[3] = {(uint64_t)W, (uint64_t)H, (uint64_t)N * TCS};
  uint64_t xstr[2] = {(uint64_t)W * 4, (uint64_t)H * W * 4};
  uint32_t xbox[3] = {T1_PWF, T1_PH, TCS};
  if ((rc = make_tmap_f32_plain(&tmX, x_nchw, 3, xdims, xstr, xbox))) return rc;
  const double opix = (double)N * kp.OH * kp.OW * Cw;
  HpcProfScope prof(HPC_PROF_THIN, (double)N * TCS * H * W * 4 + opix * (2 + (out_sq_bf16 ? 2 : 0)), stream);
  auto st = (cudaStream_t)stream;
  switch (Cw) {
    case 64: return launch_t1<64>(tmB, tmOut, tmSq, tmX, kp, st);
    case 128: return launch_t1<128>(tmB, tmOut, tmSq, tmX, kp, st);
    default: return launch_t1<192>(tmB, tmOut, tmSq, tmX, kp, st);
  }
}

// ---------------------------------------------------------------------------------------
// T2: ConvTranspose2d(Cw -> 3, 5, s2, p2, op1): x NHWC bf16 -> out NCHW fp32
//   P[pixel][k] = sum_cw x[pixel][cw] * w[cw][k]   (one UMMA chain per 8x16 input patch, halo included)
//   out[n][cs][2i+a][2j+b] = bias[cs] + sum_{kh = a (2), kw = b (2)} P[i + (a+2-kh)/2][j + (b+2-kw)/2][cs,kh,kw]
//   The patch carries a 1-pixel halo, so every output of its 6x14 interior is complete inside the CTA:
//   no atomics, deterministic.  warp 0: TMA | warp 1: MMA, TMEM | warps 2-5: epilogue (TMEM -> smem -> gather)
// ---------------------------------------------------------------------------------------
namespace {
constexpr int T2_TH = 8, T2_TW = 16, T2_IH = T2_TH - 2, T2_IW = T2_TW - 2;
constexpr int T2_THREADS = 320;                          // TMA, MMA, 8 epilogue warps
constexpr int T2_STAGES = 3;
constexpr int T2_PROW = 77;                                // floats per pixel row of P in smem (75 + pad, odd)

struct T2Params {
  const float* bias; float* out;
  int N, H, W, Cw;                                         // input grid
  int tiles_w, tiles_h, total_tiles;
};

template <int CW>
struct T2Cfg {
  static constexpr int A_BYTES = (CW / 64) * 16384;       // whole K of one patch
  static constexpr int B_BYTES = (CW / 64) * KPAD * 128;
  static constexpr int P_BYTES = 128 * T2_PROW * 4;
  static constexpr int SMEM = 1024 + T2_STAGES * A_BYTES + ((B_BYTES + 1023) / 1024) * 1024 + P_BYTES + 256;
};

template <int CW>
__global__ void __launch_bounds__(T2_THREADS, 1)
deconv_c3_umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const __grid_constant__ T2Params p) {
  using Cfg = T2Cfg<CW>;
  constexpr int KC = CW / 64;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1 KB aligned; pointer arithmetic keeps the
                                                                                  // shared address space (LDS/STS, not generic LD/ST)
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + T2_STAGES * Cfg::A_BYTES;
  float* Ps = reinterpret_cast<float*>(smem_b + ((Cfg::B_BYTES + 1023) / 1024) * 1024);
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(Ps) + Cfg::P_BYTES);
  uint64_t* full_bar = bars;                    // [STAGES]
  uint64_t* empty_bar = bars + T2_STAGES;       // [STAGES]
  uint64_t* tfull = bars + 2 * T2_STAGES;       // [2]
  uint64_t* tempty = tfull + 2;                 // [2]
  uint64_t* b_full = tempty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(b_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int s = 0; s < T2_STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&tfull[s], 1); mbar_init(&tempty[s], 8); }
    mbar_init(b_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                      // the prologue above overlaps the previous kernel's tail
  pdl_launch_dependents();
  const int tiles_per_img = p.tiles_w * p.tiles_h;

  if (warp == 0) {
    if (lane < KC) {                    // lane c issues channel chunk c (TMA instructions of a stage in parallel)
      constexpr uint32_t lanes = (1u << KC) - 1;
      if (lane == 0) mbar_arrive_expect_tx(b_full, Cfg::B_BYTES);
      __syncwarp(lanes);
      tma_load_2d(smem_b + lane * KPAD * 128, &tmB, b_full, lane * 64, 0);
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        const int n = tile / tiles_per_img, r = tile % tiles_per_img;
        const int i0 = (r / p.tiles_w) * T2_IH - 1, j0 = (r % p.tiles_w) * T2_IW - 1;   // patch origin (halo)
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (lane == 0) mbar_arrive_expect_tx(&full_bar[stage], Cfg::A_BYTES);
        __syncwarp(lanes);
        tma_load_4d(smem_a + stage * Cfg::A_BYTES + lane * 16384, &tmA, &full_bar[stage], lane * 64, j0, i0, n);
        if (++stage == T2_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, KPAD, 0, 0);
      mbar_wait(b_full, 0);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
        const int acc = it & 1;
        const uint32_t aph = (it >> 1) & 1;
        mbar_wait(&tempty[acc], aph ^ 1);
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem_a + stage * Cfg::A_BYTES), b_addr = smem_u32(smem_b);
        const uint32_t d = tmem_base + acc * 128;
#pragma unroll
        for (int k = 0; k < CW / 16; ++k) {
          const uint32_t ao = (k >> 2) * 16384 + (k & 3) * 32, bo = (k >> 2) * (KPAD * 128) + (k & 3) * 32;
          mma_bf16_ss(d, make_smem_desc_sw128(a_addr + ao, 16, 1024), make_smem_desc_sw128(b_addr + bo, 16, 1024), idesc,
                      k != 0);
        }
        mma_commit(&empty_bar[stage]);
        mma_commit(&tfull[acc]);
        if (++stage == T2_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    const int t = threadIdx.x - 64;                         // 0..255
    const int q = warp & 3, half = (warp - 2) >> 2;         // warps w and w+4 share a TMEM lane quarter
    const int row = q * 32 + lane;                          // patch pixel index = il*16 + jl
    const int OH = 2 * p.H, OW = 2 * p.W;
    // gather work item: one interior input pixel x one output channel -> its 2x2 outputs (252 of 256 threads)
    const int g_cs = t / (T2_IH * T2_IW), g_px = t % (T2_IH * T2_IW);
    const int g_i = g_px / T2_IW, g_j = g_px % T2_IW;
    const bool g_on = t < TCS * T2_IH * T2_IW;
    const float bs = (g_on && p.bias) ? p.bias[g_cs] : 0.f;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
      const int acc = it & 1;
      const uint32_t aph = (it >> 1) & 1;
      const int n = tile / tiles_per_img, r = tile % tiles_per_img;
      const int ii0 = (r / p.tiles_w) * T2_IH, jj0 = (r % p.tiles_w) * T2_IW;            // interior origin
      mbar_wait(&tfull[acc], aph);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * 128 + ((uint32_t)(q * 32) << 16);
      named_bar(3, 256);                                    // previous tile's gather is done with Ps
      if (half == 0) {
        float v[3][16];
#pragma unroll
        for (int c = 0; c < 3; ++c) tmem_ld16(taddr + c * 16, v[c]);
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < 3; ++c)
#pragma unroll
          for (int e = 0; e < 16; ++e) Ps[row * T2_PROW + c * 16 + e] = v[c][e];
      } else {
        float v[2][16];
#pragma unroll
        for (int c = 0; c < 2; ++c) tmem_ld16(taddr + 48 + c * 16, v[c]);
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
          for (int e = 0; e < 16; ++e)
            if (48 + c * 16 + e < 75) Ps[row * T2_PROW + 48 + c * 16 + e] = v[c][e];
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);
      named_bar(3, 256);
      const int gi = ii0 + g_i, gj = jj0 + g_j;
      if (g_on && gi < p.H && gj < p.W) {
        const float* P0 = Ps + ((g_i + 1) * T2_TW + (g_j + 1)) * T2_PROW + g_cs * 25;
        float* o = p.out + (((int64_t)n * TCS + g_cs) * OH + 2 * gi) * OW + 2 * gj;
#pragma unroll
        for (int a = 0; a < 2; ++a) {
          float s0 = bs, s1 = bs;
#pragma unroll
          for (int kh = a; kh < 5; kh += 2) {
            const int dh = (a + 2 - kh) / 2;
#pragma unroll
            for (int kw = 0; kw < 5; ++kw) {
              const int b = kw & 1, dw = (b + 2 - kw) / 2;
              const float pv = P0[(dh * T2_TW + dw) * T2_PROW + kh * 5 + kw];
              if (b == 0) s0 += pv; else s1 += pv;
            }
          }
          *reinterpret_cast<float2*>(o + (int64_t)a * OW) = make_float2(s0, s1);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 256);
}

template <int CW>
int launch_t2(const CUtensorMap& tmA, const CUtensorMap& tmB, const T2Params& kp, cudaStream_t st) {
  using Cfg = T2Cfg<CW>;
  static HpcOncePerDevice attr_once;
  int attr_dev;
  if (attr_once.need(&attr_dev)) {
    HPC_CUDA(cudaFuncSetAttribute(deconv_c3_umma_kernel<CW>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
    attr_once.mark(attr_dev);
  }
  const int grid = kp.total_tiles < hpc_num_sms() ? kp.total_tiles : hpc_num_sms();
  HPC_CUDA(hpc_launch_pdl(deconv_c3_umma_kernel<CW>, dim3(grid), dim3(T2_THREADS), Cfg::SMEM, st, tmA, tmB, kp));
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}
}  // namespace

int hpc_deconv_c3_fwd_umma(const void* x_nhwc_bf16, const void* w_nmaj_bf16, const float* bias, int N, int Cw, int H,
                           int W, float* out_nchw, hpc_stream_t stream) {
  HPC_REQUIRE(N > 0 && H > 0 && W > 0, HPC_ERR_BAD_SHAPE, "hpc_deconv_c3_fwd_umma: empty tensor");
  HPC_REQUIRE(Cw == 64 || Cw == 128 || Cw == 192, HPC_ERR_UNSUPPORTED, "hpc_deconv_c3_fwd_umma: Cw=%d not in {64,128,192}", Cw);
  HPC_REQUIRE(x_nhwc_bf16 && w_nmaj_bf16 && out_nchw, HPC_ERR_BAD_SHAPE, "hpc_deconv_c3_fwd_umma: NULL pointer");
  HPC_REQUIRE(hpc_aligned16(x_nhwc_bf16) && hpc_aligned16(w_nmaj_bf16), HPC_ERR_BAD_ALIGNMENT,
              "hpc_deconv_c3_fwd_umma: pointers must be 16B aligned");
  T2Params kp;
  kp.bias = bias; kp.out = out_nchw; kp.N = N; kp.H = H; kp.W = W; kp.Cw = Cw;
  kp.tiles_w = (W + T2_IW - 1) / T2_IW; kp.tiles_h = (H + T2_IH - 1) / T2_IH;
  kp.total_tiles = N * kp.tiles_w * kp.tiles_h;
  CUtensorMap tmA, tmB;
  {
    uint64_t dims[4] = {(uint64_t)Cw, (uint64_t)W, (uint64_t)H, (uint64_t)N};
    uint64_t strides[3] = {(uint64_t)Cw * 2, (uint64_t)W * Cw * 2, (uint64_t)H * W * Cw * 2};
    uint32_t box[4] = {64, T2_TW, T2_TH, 1};
    uint32_t estr[4] = {1, 1, 1, 1};
    int rc = make_tmap(&tmA, x_nhwc_bf16, 4, dims, strides, box, estr);
    if (rc) return rc;
  }
  {
    uint64_t dims[2] = {(uint64_t)Cw, KPAD};
    uint64_t strides[1] = {(uint64_t)Cw * 2};
    uint32_t box[2] = {64, KPAD};
    uint32_t estr[2] = {1, 1};
    int rc = make_tmap(&tmB, w_nmaj_bf16, 2, dims, strides, box, estr);
    if (rc) return rc;
  }
  const double npix = (double)N * H * W;
  HpcProfScope prof(HPC_PROF_THIN, npix * Cw * 2 + npix * 4 * TCS * 4, stream);
  auto st = (cudaStream_t)stream;
  switch (Cw) {
    case 64: return launch_t2<64>(tmA, tmB, kp, st);
    case 128: return launch_t2<128>(tmA, tmB, kp, st);
    default: return launch_t2<192>(tmA, tmB, kp, st);
  }
}

// ---------------------------------------------------------------------------------------
// T3: weight gradient of both thin layers:
//   gw[cw][cs][kh][kw] = sum_{n,i,j} wide[n,i,j,cw] * thin[n,cs,2i-2+kh,2j-2+kw]
//   D[M = 128 (k, 80 live)][N = Cw] += A^T B over K = pixels.  Both operands are MN-major SWIZZLE_128B atoms
//   ([64 pixels] x [64 values]): `wide` straight from NHWC by TMA, the im2col of `thin` written by 4 warps.
//   Split-K over 64-pixel blocks across the grid; partials go to the workspace, a second kernel reduces them
//   in a fixed order (deterministic) into the reference weight layout.
//   warp 0: TMA | warp 1: MMA, TMEM | warps 2-9: two im2col builder groups; group 0 then runs the TMEM -> workspace epilogue
// ---------------------------------------------------------------------------------------
namespace {
constexpr int T3_TH = 4, T3_TW = 16;                      // 64 pixels per K block
constexpr int T3_PH = 2 * T3_TH + 3;                      // 11 input rows per K block
constexpr int T3_PWF = T1_PWF, T3_PX0 = T1_PX0;           // the patch box of the conv kernel (T1), 11 rows high
constexpr int T3_PATCH_BYTES = ((TCS * T3_PH * T3_PWF * 4 + 127) / 128) * 128;       // 5376
constexpr int T3_KROWS = 120;                             // live rows of the result in the k' = (cs*5+kh)*8 + kw order
constexpr int T3_GROUPS = 2;                              // independent im2col builder groups (128 threads each): the
                                                          // kernel is bound by the builders, not by TMA or the MMAs
constexpr int T3_THREADS = 64 + 128 * T3_GROUPS;
constexpr int T3_STAGES = 4;

struct T3Params {
  const float* thin; float* ws;
  int N, H, W, Cw;                                         // wide grid (thin is 2H x 2W)
  int tiles_w, tiles_h, kblocks, nsplits;
};

template <int CW>
struct T3Cfg {
  static constexpr int STAGE_BYTES = (2 + CW / 64) * ATOM_BYTES;
  static constexpr int PATCH_BYTES = 2 * T3_PATCH_BYTES;                 // per builder group: two boxes in flight
  static constexpr int SMEM = 1024 + T3_STAGES * STAGE_BYTES + T3_GROUPS * PATCH_BYTES + 256;
};

template <int CW>
__global__ void __launch_bounds__(T3_THREADS, 1)
wgrad_c3_umma_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX,
                     const __grid_constant__ T3Params p) {
  using Cfg = T3Cfg<CW>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1 KB aligned; pointer arithmetic keeps the
                                                                                  // shared address space (LDS/STS, not generic LD/ST)
  uint8_t* patch = smem + T3_STAGES * Cfg::STAGE_BYTES;         // [T3_GROUPS][2][TCS][T3_PH][T3_PWF] fp32, written by the TMA
  uint64_t* bars = reinterpret_cast<uint64_t*>(patch + T3_GROUPS * Cfg::PATCH_BYTES);
  uint64_t* full_bar = bars;                    // [STAGES] 1 TMA arrival (+tx) + 128 builder arrivals
  uint64_t* empty_bar = bars + T3_STAGES;       // [STAGES] tcgen05.commit
  uint64_t* done_bar = bars + 2 * T3_STAGES;
  uint64_t* p_full = done_bar + 1;              // [T3_GROUPS][2] TMA -> builders: an input patch has landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(p_full + 2 * T3_GROUPS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int split = blockIdx.x;
  const int kb0 = (int)(((int64_t)split * p.kblocks) / p.nsplits);
  const int kb1 = (int)(((int64_t)(split + 1) * p.kblocks) / p.nsplits);
  const int tiles_per_img = p.tiles_w * p.tiles_h;

  // zero all im2col atoms once: the last unit (k' >= 120) of the second atom is never written again
  for (int s = 0; s < T3_STAGES; ++s)
    for (int i = threadIdx.x; i < 2 * ATOM_BYTES / 16; i += T3_THREADS)
      reinterpret_cast<uint4*>(smem + s * Cfg::STAGE_BYTES)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x == 0) {
    prefetch_tmap(&tmW); prefetch_tmap(&tmX);
    for (int s = 0; s < T3_STAGES; ++s) { mbar_init(&full_bar[s], 129); mbar_init(&empty_bar[s], 1); }
    for (int s = 0; s < 2 * T3_GROUPS; ++s) mbar_init(&p_full[s], 1);
    mbar_init(done_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 256);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                      // the prologue above overlaps the previous kernel's tail
  pdl_launch_dependents();

  if (warp == 0) {
    if (lane < CW / 64) {               // lane c issues channel chunk c
      constexpr uint32_t lanes = (1u << (CW / 64)) - 1;
      int stage = 0; uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        const int n = kb / tiles_per_img, r = kb % tiles_per_img;
        const int i0 = (r / p.tiles_w) * T3_TH, j0 = (r % p.tiles_w) * T3_TW;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (lane == 0) mbar_arrive_expect_tx(&full_bar[stage], (CW / 64) * ATOM_BYTES);
        __syncwarp(lanes);
        uint8_t* sb = smem + stage * Cfg::STAGE_BYTES + 2 * ATOM_BYTES;
        tma_load_4d(sb + lane * ATOM_BYTES, &tmW, &full_bar[stage], lane * 64, j0, i0, n);
        if (++stage == T3_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, CW, 1, 1);
      int stage = 0; uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + stage * Cfg::STAGE_BYTES);
        const uint32_t b_addr = a_addr + 2 * ATOM_BYTES;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)
          mma_bf16_ss(tmem_base, make_smem_desc_sw128(a_addr + ks * 2048, ATOM_BYTES, 1024),
                      make_smem_desc_sw128(b_addr + ks * 2048, ATOM_BYTES, 1024), idesc, (kb != kb0) || ks != 0);
        mma_commit(&empty_bar[stage]);
        if (++stage == T3_STAGES) { stage = 0; phase ^= 1; }
      }
      mma_commit(done_bar);
    }
  } else {
    // Builder group g (128 threads, its own patch buffer and named barrier) takes the K blocks kb0 + g, kb0 + g + 2, ...;
    // the pipeline stage of a block follows from its index, so TMA, MMA and both groups agree without talking.
    const int grp = (threadIdx.x - 64) >> 7;
    const int t = (threadIdx.x - 64) & 127;
    const int px = t & 63, half = t >> 6;                  // pixel of the K block, which atom (8 units) of its row
    const int il = px >> 4, jl = px & 15;
    uint8_t* mypatch = patch + grp * Cfg::PATCH_BYTES;
    // the input patch of a K block is ONE TMA box, fetched two of the group's blocks ahead (see the conv kernel, T1)
    auto issue_patch = [&](int kb, int pb) {
      const int n = kb / tiles_per_img, r = kb % tiles_per_img;
      mbar_arrive_expect_tx(&p_full[grp * 2 + pb], TCS * T3_PH * T3_PWF * 4);
      tma_load_3d(mypatch + pb * T3_PATCH_BYTES, &tmX, &p_full[grp * 2 + pb], 2 * (r % p.tiles_w) * T3_TW - T3_PX0,
                  2 * (r / p.tiles_w) * T3_TH - 2, n * TCS);
    };
    if (t == 0)
      for (int k = 0; k < 2; ++k)
        if (kb0 + grp + k * T3_GROUPS < kb1) issue_patch(kb0 + grp + k * T3_GROUPS, k);
    int m = 0;
    for (int kb = kb0 + grp; kb < kb1; kb += T3_GROUPS, ++m) {
      const int seq = kb - kb0, stage = seq % T3_STAGES, pb = m & 1;
      const uint32_t phase = (uint32_t)(seq / T3_STAGES) & 1;
      mbar_wait(&p_full[grp * 2 + pb], (m >> 1) & 1);
      mbar_wait(&empty_bar[stage], phase ^ 1);
      // unit u = (cs, kh): 8 consecutive pixels of patch row 2 il + kh from column 2 jl + 2 (the conv kernel's im2col;
      // the result rows of taps kw = 5..7 are never read).  Thread `half` writes atom `half` of the pixel's row: units
      // 0..7 | 8..14 (unit 15 stays zero).
      uint8_t* arow = smem + stage * Cfg::STAGE_BYTES + half * ATOM_BYTES + px * 128;
      const float2* prow = reinterpret_cast<const float2*>(mypatch + pb * T3_PATCH_BYTES) + il * T3_PWF + jl + (T3_PX0 - 2) / 2;
#pragma unroll
      for (int uu = 0; uu < 8; ++uu) {
        const int u = half * 8 + uu;
        if (u < TCS * 5) {
          const float2* src = prow + ((u / 5) * T3_PH + (u % 5)) * (T3_PWF / 2);
          const float2 a = src[0], b = src[1], c = src[2], d = src[3];
          *reinterpret_cast<uint4*>(arow + ((uu ^ (px & 7)) << 4)) =
              make_uint4(pack2_bf16(a.x, a.y), pack2_bf16(b.x, b.y), pack2_bf16(c.x, c.y), pack2_bf16(d.x, d.y));
        }
      }
      fence_proxy_async();
      mbar_arrive(&full_bar[stage]);
      named_bar(1 + grp, 128);                             // the group has finished reading patch buffer pb
      if (t == 0 && kb + 2 * T3_GROUPS < kb1) issue_patch(kb + 2 * T3_GROUPS, pb);
    }
    if (grp == 0) {
    // epilogue: row = k' (TMEM lane); rows >= 120 are padding
    const int q = warp & 3;
    const int row = q * 32 + lane;
    mbar_wait(done_bar, 0);
    tc_fence_after();
    if (q * 32 < T3_KROWS) {
      float* dst = p.ws + ((int64_t)split * T3_KROWS + row) * CW;
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
      for (int c0 = 0; c0 < CW; c0 += 16) {
        float v[16];
        if (kb1 > kb0) {
          tmem_ld16(taddr + c0, v);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int e = 0; e < 16; ++e) v[e] = 0.f;
        }
        if (row < T3_KROWS) {
#pragma unroll
          for (int e = 0; e < 16; e += 4) *reinterpret_cast<float4*>(dst + c0 + e) = make_float4(v[e], v[e + 1], v[e + 2], v[e + 3]);
        }
      }
    }
    }   // grp == 0
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 256);
}

// gw[cw*w_sr + cs*w_sc + kh*5 + kw] = sum_s ws[s][(cs*5+kh)*8 + kw][cw]
__global__ void wgrad_c3_umma_reduce(const float* __restrict__ ws, int nsplits, int Cw, float* __restrict__ gw, int64_t w_sr,
                                     int64_t w_sc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  pdl_wait();
  pdl_launch_dependents();
  if (i >= 75 * Cw) return;
  const int k = i / Cw, cw = i % Cw;
  const int kp = ((k / 25) * 5 + (k % 25) / 5) * 8 + k % 5;
  float a = 0.f;
  for (int s = 0; s < nsplits; ++s) a += ws[((int64_t)s * T3_KROWS + kp) * Cw + cw];
  gw[cw * w_sr + (k / 25) * w_sc + (k % 25)] = a;
}

int t3_splits(int kblocks) {
  int ns = hpc_num_sms();
  return ns < kblocks ? ns : kblocks;
}
}  // namespace

size_t hpc_wgrad_c3_umma_workspace_bytes(int N, int Cw, int H, int W) {
  const int kblocks = N * ((H + T3_TH - 1) / T3_TH) * ((W + T3_TW - 1) / T3_TW);
  return (size_t)t3_splits(kblocks) * T3_KROWS * Cw * sizeof(float);
}

int hpc_wgrad_c3_umma(const void* wide_nhwc_bf16, const float* thin_nchw, int N, int Cw, int H, int W, float* gw,
                      int64_t w_sr, int64_t w_sc, void* workspace, size_t workspace_bytes, hpc_stream_t stream) {
  HPC_REQUIRE(N > 0 && H > 0 && W > 0, HPC_ERR_BAD_SHAPE, "hpc_wgrad_c3_umma: empty tensor");
  HPC_REQUIRE(Cw == 64 || Cw == 128 || Cw == 192 || Cw == 256, HPC_ERR_UNSUPPORTED,
              "hpc_wgrad_c3_umma: Cw=%d not in {64,128,192,256}", Cw);
  HPC_REQUIRE(wide_nhwc_bf16 && thin_nchw && gw, HPC_ERR_BAD_SHAPE, "hpc_wgrad_c3_umma: NULL pointer");
  HPC_REQUIRE(hpc_aligned16(wide_nhwc_bf16) && hpc_aligned16(workspace), HPC_ERR_BAD_ALIGNMENT,
              "hpc_wgrad_c3_umma: pointers must be 16B aligned");
  T3Params kp;
  kp.thin = thin_nchw; kp.ws = (float*)workspace; kp.N = N; kp.H = H; kp.W = W; kp.Cw = Cw;
  kp.tiles_w = (W + T3_TW - 1) / T3_TW; kp.tiles_h = (H + T3_TH - 1) / T3_TH;
  kp.kblocks = N * kp.tiles_w * kp.tiles_h;
  kp.nsplits = t3_splits(kp.kblocks);
  const size_t need = (size_t)kp.nsplits * T3_KROWS * Cw * sizeof(float);
  HPC_REQUIRE(workspace && workspace_bytes >= need, HPC_ERR_BAD_SHAPE, "hpc_wgrad_c3_umma: workspace too small (%zu < %zu)",
              workspace_bytes, need);
  HPC_REQUIRE(W % 2 == 0 && hpc_aligned16(thin_nchw), HPC_ERR_BAD_ALIGNMENT,
              "hpc_wgrad_c3_umma: the image rows are fetched by TMA: an even W (rows of 2W floats in whole 16-byte units) and "
              "a 16-byte aligned image (W = %d); use hpc_wgrad_c3 otherwise", W);
  CUtensorMap tmX;      // the thin tensor as a 3-D fp32 map (2W, 2H, N*3): one box = the patch of a K block
  {
    uint64_t xdims[3] = {(uint64_t)2 * W, (uint64_t)2 * H, (uint64_t)N * TCS};
    uint64_t xstr[2] = {(uint64_t)2 * W * 4, (uint64_t)4 * H * W * 4};
    uint32_t xbox[3] = {T3_PWF, T3_PH, TCS};
    int rcx = make_tmap_f32_plain(&tmX, thin_nchw, 3, xdims, xstr, xbox);
    if (rcx) return rcx;
  }
  CUtensorMap tmW;
  uint64_t dims[4] = {(uint64_t)Cw, (uint64_t)W, (uint64_t)H, (uint64_t)N};
  uint64_t strides[3] = {(uint64_t)Cw * 2, (uint64_t)W * Cw * 2, (uint64_t)H * W * Cw * 2};
  uint32_t box[4] = {64, T3_TW, T3_TH, 1};
  uint32_t estr[4] = {1, 1, 1, 1};
  int rc = make_tmap(&tmW, wide_nhwc_bf16, 4, dims, strides, box, estr);
  if (rc) return rc;
  HpcProfScope prof(HPC_PROF_THIN, (double)N * H * W * Cw * 2 + (double)N * TCS * 4 * H * W * 4, stream);
  auto st = (cudaStream_t)stream;
#define LAUNCH_T3(CW_)                                                                                               \
  do {                                                                                                               \
    static HpcOncePerDevice attr_once;                                                                                    \
    int attr_dev;                                                                                    \
    if (attr_once.need(&attr_dev)) {                                                                                                 \
      HPC_CUDA(cudaFuncSetAttribute(wgrad_c3_umma_kernel<CW_>, cudaFuncAttributeMaxDynamicSharedMemorySize,          \
                                    T3Cfg<CW_>::SMEM));                                                              \
      attr_once.mark(attr_dev);                                                                                               \
    }                                                                                                                \
    HPC_CUDA(hpc_launch_pdl(wgrad_c3_umma_kernel<CW_>, dim3(kp.nsplits), dim3(T3_THREADS), T3Cfg<CW_>::SMEM, st, tmW, tmX, kp)); \
  } while (0)
  switch (Cw) {
    case 64: LAUNCH_T3(64); break;
    case 128: LAUNCH_T3(128); break;
    case 192: LAUNCH_T3(192); break;
    default: LAUNCH_T3(256); break;
  }
#undef LAUNCH_T3
  HPC_LAUNCH_CHECK();
  HPC_CUDA(hpc_launch_pdl(wgrad_c3_umma_reduce, dim3((75 * Cw + 255) / 256), dim3(256), 0, st, (const float*)kp.ws, kp.nsplits,
                          Cw, gw, w_sr, w_sc));
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}
