// Fused GDN / IGDN kernels on tcgen05 (sm_100a) -- modelling/layers/gdn.py:79-88 and its autograd.
//
// GDN is a 1x1 "convolution": the activation is a [P pixels][C] bf16 matrix, so a tile is 128 consecutive
// rows (C/64 TMA boxes of [128 x 64] in the canonical K-major SWIZZLE_128B layout).  gamma (C x C bf16,
// 72 KB for C = 192) stays resident in shared memory for the whole persistent CTA.
//
//   forward : x tile --TMA--> smem;  the workers square it straight into TMEM (tcgen05.st), where it is the A operand
//             of the UMMA n = x^2 . gamma^T (A-from-TMEM form) -- x^2 exists neither in global nor in shared memory,
//             and the squares of tile i+1 overlap the MMA of tile i (two A buffers);  epilogue r = rsqrt(beta + n),
//             y = x r written in place over x, r into a staging tile;  TMA stores of y and r.
//             HBM traffic: read x, write y, write r (3 passes).
//   backward: g, x, r tiles --TMA--> smem;  t = g x r^3 (in place over r), u = g r (in place over g);
//             UMMA a = t . gamma;  dx = u - x a (over u), x^2 (over x);  TMA stores of dx, t, x^2;  column sums
//             of t (d beta) accumulated per CTA.  6 passes instead of 10 for the unfused sequence.
//
// warp 0: TMA loads | warp 1: MMA issue + TMEM owner | warp 2: TMA stores | warps 3-18: 512 workers, four
// threads per tile row (each owns a quarter of the channels; warps w, w+4, w+8, w+12 share a TMEM lane quarter).
#include <cuda.h>

#include <cstdlib>
#include <initializer_list>

#include "tmap.cuh"
#include "umma_ptx.cuh"

using namespace umma;

namespace {

constexpr int G_WORKERS = 512;                 // 16 worker warps: the elementwise phases are latency bound, warps hide it
constexpr int G_THREADS = 96 + G_WORKERS;
constexpr int G_CHUNK = 16384;                 // [128 rows x 128 B]

__device__ __forceinline__ void named_bar(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, const void* src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(smem_u32(src)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ void unpack8(const uint4& u, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x; f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return u;
}
// rsqrtf() without its denormal-input path (an FSETP, two scaling multiplies and selects per element, a fifth of the
// forward kernel's instructions): n = beta + gamma . x^2 >= beta >= 1e-6 is never denormal, and for normal inputs this
// is the same MUFU.RSQ result, bit for bit.
__device__ __forceinline__ float rsqrt_normal(float x) {
  float y;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {      // one MUFU.RCP (1 ulp); the results it feeds are rounded to bf16
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// acc[0..7] += the sums over 8 tile rows of one 16-byte unit (8 channels): the 8 rows are added as packed bf16 pairs in a
// 3-level tree (28 HADD2 instead of 64 unpacks + 64 FADD: the column sums were 22 % of the backward kernel's
// instructions, ncu r02u) and only the 8 results are widened.  Each level rounds to bf16: the error of a group sum is
// ~2^-8 of its magnitude, unbiased (round to nearest), against the 2^-9 every summand already carries.
__device__ __forceinline__ void colsum8_tree(const uint4 (&r)[8], float* acc) {
  const __nv_bfloat162* h[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) h[i] = reinterpret_cast<const __nv_bfloat162*>(&r[i]);
#pragma unroll
  for (int w = 0; w < 4; ++w) {
    const __nv_bfloat162 s01 = __hadd2(h[0][w], h[1][w]), s23 = __hadd2(h[2][w], h[3][w]);
    const __nv_bfloat162 s45 = __hadd2(h[4][w], h[5][w]), s67 = __hadd2(h[6][w], h[7][w]);
    const float2 f = __bfloat1622float2(__hadd2(__hadd2(s01, s23), __hadd2(s45, s67)));
    acc[2 * w] += f.x; acc[2 * w + 1] += f.y;
  }
}
// byte offset of 16-byte unit U (8 channels) of tile row r in the [chunk][128 rows][128 B] swizzled tile
__device__ __forceinline__ uint32_t unit_off(int r, int U) {
  return (uint32_t)((U >> 3) * G_CHUNK + r * 128 + (((U & 7) ^ (r & 7)) << 4));
}

struct GdnParams {
  const float* beta;       // forward only
  float* colsum_part;      // backward only: [gridDim.x][C]
  int64_t P;
  int tiles;
  int inverse;
  int store_r;             // forward only: 0 = r is not wanted (no-grad evaluation): 2 HBM passes instead of 3
};

template <int C>
struct GCfg {
  static constexpr int KC = C / 64;
  static constexpr int TILE = KC * G_CHUNK;
  static constexpr int GAMMA = C * C * 2;
  static constexpr int TMEM_COLS = C <= 64 ? 64 : (C <= 128 ? 128 : 256);
  // forward: accumulator (C columns) + two x^2 A-operand buffers of C/2 columns each
  static constexpr int FWD_A_COL = C;
  static constexpr int FWD_TMEM_COLS = C <= 64 ? 128 : (C <= 128 ? 256 : 512);
  static constexpr int SMEM = 1024 + GAMMA + 3 * TILE + 1024 /*beta*/ + 256;
};

// ---------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------
template <int C>
__global__ void __launch_bounds__(G_THREADS, 1)
gdn_fwd_umma_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmY,
                    const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmG,
                    const __grid_constant__ GdnParams p) {
  using Cfg = GCfg<C>;
  constexpr int KC = Cfg::KC;
  constexpr int A_COL = Cfg::FWD_A_COL, A_COLS = C / 2;     // x^2 as the TMEM-resident A operand, double buffered
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1 KB aligned; pointer arithmetic keeps the
                                                                                  // shared address space (LDS/STS, not generic LD/ST)
  uint8_t* gam = smem;                           // KC chunks of [C rows x 128 B]
  uint8_t* xs = gam + Cfg::GAMMA;                // [2][TILE]   x, later y
  uint8_t* rs = xs + 2 * Cfg::TILE;              // [TILE]      r staging for the TMA store
  float* beta_s = reinterpret_cast<float*>(rs + Cfg::TILE);
  uint64_t* bars = reinterpret_cast<uint64_t*>(beta_s + 256);
  uint64_t* g_full = bars;
  uint64_t* xs_full = bars + 1;     // [2]
  uint64_t* xs_empty = bars + 3;    // [2]  store warp, after the y store has finished READING smem
  uint64_t* a_full = bars + 5;      // workers: x^2 of the tile is in TMEM (and the accumulator has been drained)
  uint64_t* rs_empty = bars + 6;    // store warp, after the r store has finished reading smem
  uint64_t* tfull = bars + 7;       // [2] tcgen05.commit of the two halves of the output channels (see the MMA warp)
  uint64_t* out_full = bars + 9;    // workers: y and r written
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    prefetch_tmap(&tmX); prefetch_tmap(&tmY); prefetch_tmap(&tmR); prefetch_tmap(&tmG);
    mbar_init(g_full, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(&xs_full[s], 1); mbar_init(&xs_empty[s], 1); }
    mbar_init(a_full, G_WORKERS); mbar_init(rs_empty, 1); mbar_init(&tfull[0], 1); mbar_init(&tfull[1], 1);
    mbar_init(out_full, G_WORKERS);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::FWD_TMEM_COLS);
  pdl_wait();                      // the prologue above overlaps the previous kernel's tail (beta comes from gdn_prepare)
  for (int i = threadIdx.x; i < 256; i += G_THREADS) beta_s[i] = i < C ? p.beta[i] : 1.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();

  if (warp == 0) {
    // lane c issues chunk c: the TMA instructions of a tile go out in parallel (each costs its thread ~100+ cycles)
    if (lane < KC) {
      constexpr uint32_t lanes = (1u << KC) - 1;
      if (lane == 0) mbar_arrive_expect_tx(g_full, Cfg::GAMMA);
      __syncwarp(lanes);
      tma_load_2d(gam + lane * (C * 128), &tmG, g_full, lane * 64, 0);
      int it = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
        const int b = it & 1;
        mbar_wait(&xs_empty[b], ((it >> 1) & 1) ^ 1);
        if (lane == 0) mbar_arrive_expect_tx(&xs_full[b], Cfg::TILE);
        __syncwarp(lanes);
        tma_load_2d(xs + b * Cfg::TILE + lane * G_CHUNK, &tmX, &xs_full[b], lane * 64, tile * 128);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // The accumulator is single-buffered (TMEM: 192 columns of it + 2 x 96 of x^2), so the workers wait out every
      // MMA.  It is issued as two halves of the output channels (N = C/2 each, rows C/2.. of gamma^T start 12 KB into
      // every chunk) with a commit each: the warps of the first two channel quarters start their epilogue after half
      // of the MMA time, while the second half runs.
      constexpr uint32_t idesc = make_idesc_bf16(128, C / 2, 0, 0);
      static_assert((C / 2) % 16 == 0 && ((C / 2) * 128) % 1024 == 0, "N halves");
      mbar_wait(g_full, 0);
      const uint32_t b_addr = smem_u32(gam);
      int it = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
        mbar_wait(a_full, it & 1);
        tc_fence_after();
        const uint32_t a_tmem = tmem_base + A_COL + (it & 1) * A_COLS;
#pragma unroll
        for (int nh = 0; nh < 2; ++nh) {
#pragma unroll
          for (int k = 0; k < C / 16; ++k)
            mma_bf16_ts(tmem_base + nh * (C / 2), a_tmem + k * 8,
                        make_smem_desc_sw128(b_addr + (k >> 2) * (C * 128) + nh * (C / 2) * 128 + (k & 3) * 32, 16, 1024),
                        idesc, k != 0);
          mma_commit(&tfull[nh]);
        }
      }
    }
  } else if (warp == 2) {
    // lanes 0..KC-1 store the r chunks, lanes KC..2KC-1 the y chunks (bulk groups are per thread); each group frees
    // its own buffer as soon as ITS stores have been read out of shared memory
    if (lane < 2 * KC && (p.store_r || lane >= KC)) {
      const int c = lane % KC;
      int it = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
        const int b = it & 1;
        mbar_wait(out_full, it & 1);
        if (lane < KC) tma_store_2d(&tmR, rs + c * G_CHUNK, c * 64, tile * 128);
        else tma_store_2d(&tmY, xs + b * Cfg::TILE + c * G_CHUNK, c * 64, tile * 128);
        bulk_commit();
        bulk_wait_read0();
        if (lane < KC) {
          __syncwarp((1u << KC) - 1);
          if (lane == 0) mbar_arrive(rs_empty);
        } else {
          __syncwarp(((1u << KC) - 1) << KC);
          if (lane == KC) mbar_arrive(&xs_empty[b]);
        }
      }
      bulk_wait0();
    }
  } else {
    const int w = warp - 3;                       // 0..15
    const int q = warp & 3, half = w >> 2;        // `half` = which quarter of the channels (0..3)
    const int row = q * 32 + lane;
    constexpr int UH = C / 32;                    // 16-byte units (8 channels) per quarter row
    static_assert(UH % 2 == 0, "two units per tcgen05.st");
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;

    // x^2 of tile `it2` -> TMEM A buffer it2 & 1 (bf16 pairs, K order): 4 columns per 8-channel unit
    auto square_to_tmem = [&](int it2) {
      const uint8_t* xt = xs + (it2 & 1) * Cfg::TILE;
      mbar_wait(&xs_full[it2 & 1], (it2 >> 1) & 1);
      const uint32_t abase = tmem_base + lane_off + A_COL + (it2 & 1) * A_COLS + half * (UH * 4);
#pragma unroll 1
      for (int uu = 0; uu < UH; uu += 2) {
        const uint4 raw0 = *reinterpret_cast<const uint4*>(xt + unit_off(row, half * UH + uu));
        const uint4 raw1 = *reinterpret_cast<const uint4*>(xt + unit_off(row, half * UH + uu + 1));
        // squares on the packed pairs: the product of two bf16 values is exact in fp32, so one HMUL2 per pair gives the
        // bits of two fp32 multiplies and a pack
        uint32_t v[8] = {raw0.x, raw0.y, raw0.z, raw0.w, raw1.x, raw1.y, raw1.z, raw1.w};
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          __nv_bfloat162 h = *reinterpret_cast<__nv_bfloat162*>(&v[e]);
          h = __hmul2(h, h);
          v[e] = *reinterpret_cast<uint32_t*>(&h);
        }
        tmem_st8(abase + uu * 4, v);
      }
      tmem_st_wait();
    };

    // software pipeline with ONE copy of each phase: iteration `it` publishes x^2(it) (computed at the end of the
    // previous iteration, after x(it) had a whole tile time to land), then runs the epilogue of tile it; the squares
    // of tile it+1 follow it, so the MMA of tile it+1 never waits for a staging buffer to drain
    const int ntile = ((int)blockIdx.x < p.tiles) ? (p.tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
    for (int it = -1; it < ntile; ++it) {
      if (it >= 0) {
        // x^2(it) is in TMEM and (program order) this thread has finished reading the accumulator of tile it-1
        tc_fence_before();
        mbar_arrive(a_full);
      }
      if (it < 0) {
        if (ntile > 0) square_to_tmem(0);
        continue;
      }
      const int tile = blockIdx.x + it * gridDim.x;
      (void)tile;
      const int b = it & 1;
      uint8_t* xt = xs + b * Cfg::TILE;
      mbar_wait(&tfull[half >> 1], it & 1);             // this warp's channel quarter lies in half `half >> 1`
      tc_fence_after();
      if (p.store_r) mbar_wait(rs_empty, (it & 1) ^ 1);   // r of the previous tile has left the staging buffer
      const uint32_t taddr = tmem_base + lane_off + half * (C / 4);
      constexpr int GB = (C / 64) % 3 == 0 ? 3 : ((C / 64) % 2 == 0 ? 2 : 1);      // ld16 groups per batch
#pragma unroll 1
      for (int gb = 0; gb < C / 64; gb += GB) {
        float v[GB][16];
        uint4 xr[GB][2];
#pragma unroll
        for (int j = 0; j < GB; ++j) tmem_ld16(taddr + (gb + j) * 16, v[j]);
#pragma unroll
        for (int j = 0; j < GB; ++j)
#pragma unroll
          for (int hh = 0; hh < 2; ++hh)
            xr[j][hh] = *reinterpret_cast<const uint4*>(xt + unit_off(row, ((half * (C / 4) + (gb + j) * 16) >> 3) + hh));
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < GB; ++j) {
          const int c0 = half * (C / 4) + (gb + j) * 16;
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            const uint32_t o = unit_off(row, (c0 >> 3) + hh);
            float x[8], y[8], r[8];
            unpack8(xr[j][hh], x);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const float n = v[j][hh * 8 + e] + beta_s[c0 + hh * 8 + e];
              r[e] = rsqrt_normal(n);
              y[e] = p.inverse ? x[e] * (n * r[e]) : x[e] * r[e];
            }
            *reinterpret_cast<uint4*>(xt + o) = pack8(y);
            if (p.store_r) *reinterpret_cast<uint4*>(rs + o) = pack8(r);
          }
        }
      }
      fence_proxy_async();
      mbar_arrive(out_full);
      if (it + 1 < ntile) square_to_tmem(it + 1);   // other A buffer: its last reader (MMA it-1) retired before tfull(it-1)
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::FWD_TMEM_COLS);
}

// ---------------------------------------------------------------------------------------
// backward:  t = g x r^3 | g x r (inverse);  a = t . gamma;  dx = g r - x a | g / r + x a
// ---------------------------------------------------------------------------------------
template <int C>
__global__ void __launch_bounds__(G_THREADS, 1)
gdn_bwd_umma_kernel(const __grid_constant__ CUtensorMap tmGr, const __grid_constant__ CUtensorMap tmX,
                    const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmDx,
                    const __grid_constant__ CUtensorMap tmT, const __grid_constant__ CUtensorMap tmSq,
                    const __grid_constant__ CUtensorMap tmG, const __grid_constant__ GdnParams p) {
  using Cfg = GCfg<C>;
  constexpr int KC = Cfg::KC;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1 KB aligned; pointer arithmetic keeps the
                                                                                  // shared address space (LDS/STS, not generic LD/ST)
  uint8_t* gam = smem;                           // gamma^T as the K-major B operand
  uint8_t* bg = gam + Cfg::GAMMA;                // g   -> u = g r -> dx
  uint8_t* bx = bg + Cfg::TILE;                  // x   -> x^2
  uint8_t* bn = bx + Cfg::TILE;                  // r   -> t
  uint64_t* bars = reinterpret_cast<uint64_t*>(bn + Cfg::TILE + 1024);
  uint64_t* g_full = bars;
  uint64_t* in_full = bars + 1;     // TMA: g, x, r tiles landed
  uint64_t* in_empty = bars + 2;    // store warp: all three buffers have been read by the TMA stores
  uint64_t* a_full = bars + 3;      // 256 workers: t (and u) written
  uint64_t* tfull = bars + 4;
  uint64_t* out_full = bars + 5;    // 256 workers: dx, x^2 written
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 6);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    prefetch_tmap(&tmGr); prefetch_tmap(&tmX); prefetch_tmap(&tmR); prefetch_tmap(&tmDx); prefetch_tmap(&tmT);
    prefetch_tmap(&tmSq); prefetch_tmap(&tmG);
    mbar_init(g_full, 1); mbar_init(in_full, 1); mbar_init(in_empty, 1);
    mbar_init(a_full, G_WORKERS); mbar_init(tfull, 1); mbar_init(out_full, G_WORKERS);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                      // the prologue above overlaps the previous kernel's tail
  pdl_launch_dependents();

  if (warp == 0) {
    // lane l issues chunk l % KC of tensor l / KC (g, x, r): 3 KC TMA instructions per tile, issued in parallel
    if (lane < 3 * KC) {
      constexpr uint32_t lanes = (1u << (3 * KC)) - 1;
      const int c = lane % KC, which = lane / KC;
      if (lane == 0) mbar_arrive_expect_tx(g_full, Cfg::GAMMA);
      __syncwarp(lanes);
      if (lane < KC) tma_load_2d(gam + lane * (C * 128), &tmG, g_full, lane * 64, 0);
      uint8_t* dst = (which == 0 ? bg : (which == 1 ? bx : bn)) + c * G_CHUNK;
      const CUtensorMap* tm = which == 0 ? &tmGr : (which == 1 ? &tmX : &tmR);
      int it = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
        mbar_wait(in_empty, (it & 1) ^ 1);
        if (lane == 0) mbar_arrive_expect_tx(in_full, 3 * Cfg::TILE);
        __syncwarp(lanes);
        tma_load_2d(dst, tm, in_full, c * 64, tile * 128);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, C, 0, 0);
      mbar_wait(g_full, 0);
      int it = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
        mbar_wait(a_full, it & 1);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(bn), b_addr = smem_u32(gam);
#pragma unroll
        for (int k = 0; k < C / 16; ++k)
          mma_bf16_ss(tmem_base, make_smem_desc_sw128(a_addr + (k >> 2) * G_CHUNK + (k & 3) * 32, 16, 1024),
                      make_smem_desc_sw128(b_addr + (k >> 2) * (C * 128) + (k & 3) * 32, 16, 1024), idesc, k != 0);
        mma_commit(tfull);
      }
    }
  } else if (warp == 2) {
    if (lane < 3 * KC) {
      constexpr uint32_t lanes = (1u << (3 * KC)) - 1;
      const int c = lane % KC, which = lane / KC;
      const uint8_t* src = (which == 0 ? bg : (which == 1 ? bn : bx)) + c * G_CHUNK;
      const CUtensorMap* tm = which == 0 ? &tmDx : (which == 1 ? &tmT : &tmSq);
      int it = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
        mbar_wait(out_full, it & 1);
        tma_store_2d(tm, src, c * 64, tile * 128);
        bulk_commit();
        bulk_wait_read0();
        __syncwarp(lanes);
        if (lane == 0) mbar_arrive(in_empty);
      }
      bulk_wait0();
    }
  } else {
    const int w = warp - 3;
    const int q = warp & 3, half = w >> 2;
    const int row = q * 32 + lane;
    const int wt = threadIdx.x - 96;              // 0..511: column owner for the d(beta) sums
    constexpr int UH = C / 32;
    float bsum = 0.f;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
      mbar_wait(in_full, it & 1);
      constexpr int UB = UH % 4 == 0 ? 4 : (UH % 3 == 0 ? 3 : (UH % 2 == 0 ? 2 : 1));
#pragma unroll 1
      for (int ub = 0; ub < UH; ub += UB) {
        uint4 rg[UB], rx[UB], rr[UB];
#pragma unroll
        for (int j = 0; j < UB; ++j) {
          const uint32_t o = unit_off(row, half * UH + ub + j);
          rg[j] = *reinterpret_cast<const uint4*>(bg + o);
          rx[j] = *reinterpret_cast<const uint4*>(bx + o);
          rr[j] = *reinterpret_cast<const uint4*>(bn + o);
        }
#pragma unroll
        for (int j = 0; j < UB; ++j) {
          const uint32_t o = unit_off(row, half * UH + ub + j);
          float g[8], x[8], r[8], t[8], u[8];
          unpack8(rg[j], g);
          unpack8(rx[j], x);
          unpack8(rr[j], r);
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            t[e] = g[e] * x[e] * (p.inverse ? r[e] : r[e] * r[e] * r[e]);
            u[e] = p.inverse ? g[e] / r[e] : g[e] * r[e];
          }
          *reinterpret_cast<uint4*>(bn + o) = pack8(t);
          *reinterpret_cast<uint4*>(bg + o) = pack8(u);
        }
      }
      fence_proxy_async();
      mbar_arrive(a_full);
      // d(beta): column sums of the bf16 t tile, while the MMA runs (rows past P are zero: TMA zero fill)
      named_bar(4, G_WORKERS);
      if (wt < C) {
        const int U = wt >> 3, e = wt & 7;
        float s = 0.f;
#pragma unroll 8
        for (int r2 = 0; r2 < 128; ++r2)
          s += __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(bn + unit_off(r2, U) + e * 2));
        bsum += s;
      }
      mbar_wait(tfull, it & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + half * (C / 4);
      constexpr int GB = 1;   // 96-register budget at 608 threads: the 16 warps hide latency instead
#pragma unroll 1
      for (int gb = 0; gb < C / 64; gb += GB) {
        float v[GB][16];
        uint4 ur[GB][2], xr[GB][2];
#pragma unroll
        for (int j = 0; j < GB; ++j) tmem_ld16(taddr + (gb + j) * 16, v[j]);
#pragma unroll
        for (int j = 0; j < GB; ++j)
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            const uint32_t o = unit_off(row, ((half * (C / 4) + (gb + j) * 16) >> 3) + hh);
            ur[j][hh] = *reinterpret_cast<const uint4*>(bg + o);
            xr[j][hh] = *reinterpret_cast<const uint4*>(bx + o);
          }
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < GB; ++j) {
          const int c0 = half * (C / 4) + (gb + j) * 16;
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            const uint32_t o = unit_off(row, (c0 >> 3) + hh);
            float u[8], x[8], d[8];
            unpack8(ur[j][hh], u);
            unpack8(xr[j][hh], x);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              d[e] = p.inverse ? u[e] + x[e] * v[j][hh * 8 + e] : u[e] - x[e] * v[j][hh * 8 + e];
              x[e] *= x[e];
            }
            *reinterpret_cast<uint4*>(bg + o) = pack8(d);
            *reinterpret_cast<uint4*>(bx + o) = pack8(x);
          }
        }
      }
      tc_fence_before();
      fence_proxy_async();
      mbar_arrive(out_full);
    }
    if (wt < C) p.colsum_part[(int64_t)blockIdx.x * C + wt] = bsum;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

// ---------------------------------------------------------------------------------------
// backward with d(gamma) and the bias gradient of the preceding layer folded in (C = 128, 192):
//   d(gamma)[i][j] = sum_p t[p][i] x^2[p][j] is a second UMMA on the SAME shared-memory tiles -- a K-major
//   [128 rows][64 ch] SWIZZLE_128B tile read as an MN-major operand has M/N = channel and K = row -- so t and
//   x^2 never go to HBM (4 passes instead of 6, and no separate one-tap wgrad over them).  The C x C fp32
//   accumulator stays in TMEM for the whole persistent CTA: rows i < 128 as one M=128 x N=C accumulator; for
//   C = 192 the remaining 64 rows are held transposed (x^2^T . t[:,128:192], two M=128 x N=64 accumulators over
//   j = 0..127 and j = 64..191), 192 + 192 + 64 + 64 = 512 columns with the `a` accumulator.  Each CTA writes its
//   partial to dgamma_part[blockIdx]; a second kernel reduces the partials in a fixed order (deterministic).
//   Column sums of dx (= d bias of the conv / deconv that feeds this GDN) are accumulated like those of t.
// ---------------------------------------------------------------------------------------
struct GdnFusedParams {
  float* colsum_part;      // [gridDim.x][C]      column sums of t  (d beta)
  float* dxsum_part;       // [gridDim.x][C]      column sums of dx (d bias of the producer layer)
  float* dgamma_part;      // [gridDim.x][C][C]   sum_p t[p][i] x^2[p][j]
  int64_t P;
  int tiles;
  int inverse;
  int prefetch;            // dual kernel: L2 prefetch of the next half-tile of each set
  int packed;              // dual kernel: the t / u phase in packed bf16 arithmetic (see the kernel)
  const float* beta;       // recompute form (RC): effective beta[C]; r = rsqrt(beta + x^2 . gamma^T) is rebuilt in the kernel
};

template <int C>
__global__ void __launch_bounds__(G_THREADS, 1)
gdn_bwd_fused_kernel(const __grid_constant__ CUtensorMap tmGr, const __grid_constant__ CUtensorMap tmX,
                     const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmDx,
                     const __grid_constant__ CUtensorMap tmG, const __grid_constant__ GdnFusedParams p) {
  static_assert(C == 128 || C == 192, "fused d(gamma): C in {128, 192}");
  using Cfg = GCfg<C>;
  constexpr int KC = Cfg::KC;
  constexpr int REM = C - 128;                    // 0 or 64 rows of d(gamma) held transposed
  constexpr int COL_DGA = C, COL_DGB = 2 * C;     // TMEM columns: a | dgamma[0:128][:] | transposed remainder x2
  constexpr int TCOLS = C == 128 ? 256 : 512;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1 KB aligned; pointer arithmetic keeps the
                                                                                  // shared address space (LDS/STS, not generic LD/ST)
  uint8_t* gam = smem;                           // gamma^T as the K-major B operand
  uint8_t* bg = gam + Cfg::GAMMA;                // g   -> u = g r -> dx
  uint8_t* bx = bg + Cfg::TILE;                  // x   -> x^2
  uint8_t* bn = bx + Cfg::TILE;                  // r   -> t
  uint64_t* bars = reinterpret_cast<uint64_t*>(bn + Cfg::TILE + 1024);
  uint64_t* g_full = bars;
  uint64_t* in_full = bars + 1;     // TMA: g, x, r tiles landed
  uint64_t* in_empty = bars + 2;    // store warp (dx read out) + d(gamma) MMA commit (t, x^2 read) + 16 worker warps (sums)
  uint64_t* a_full = bars + 3;      // workers: t (and u) written
  uint64_t* tfull = bars + 4;
  uint64_t* out_full = bars + 5;    // workers: dx, x^2 written
  uint64_t* dg_done = bars + 6;     // last d(gamma) MMA retired
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 7);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    prefetch_tmap(&tmGr); prefetch_tmap(&tmX); prefetch_tmap(&tmR); prefetch_tmap(&tmDx); prefetch_tmap(&tmG);
    mbar_init(g_full, 1); mbar_init(in_full, 1); mbar_init(in_empty, 2 + G_WORKERS / 32);
    mbar_init(a_full, G_WORKERS); mbar_init(tfull, 1); mbar_init(out_full, G_WORKERS); mbar_init(dg_done, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TCOLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                      // the prologue above overlaps the previous kernel's tail
  pdl_launch_dependents();

  if (warp == 0) {
    // lane l issues chunk l % KC of tensor l / KC (g, x, r): 3 KC TMA instructions per tile, issued in parallel
    if (lane < 3 * KC) {
      constexpr uint32_t lanes = (1u << (3 * KC)) - 1;
      const int c = lane % KC, which = lane / KC;
      if (lane == 0) mbar_arrive_expect_tx(g_full, Cfg::GAMMA);
      __syncwarp(lanes);
      if (lane < KC) tma_load_2d(gam + lane * (C * 128), &tmG, g_full, lane * 64, 0);
      uint8_t* dst = (which == 0 ? bg : (which == 1 ? bx : bn)) + c * G_CHUNK;
      const CUtensorMap* tm = which == 0 ? &tmGr : (which == 1 ? &tmX : &tmR);
      int it = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
        mbar_wait(in_empty, (it & 1) ^ 1);
        if (lane == 0) mbar_arrive_expect_tx(in_full, 3 * Cfg::TILE);
        __syncwarp(lanes);
        tma_load_2d(dst, tm, in_full, c * 64, tile * 128);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, C, 0, 0);
      constexpr uint32_t idesc_dga = make_idesc_bf16(128, C, 1, 1);     // both operands MN-major
      constexpr uint32_t idesc_dgb = make_idesc_bf16(128, 64, 1, 1);
      mbar_wait(g_full, 0);
      const uint32_t t_addr = smem_u32(bn), sq_addr = smem_u32(bx), b_addr = smem_u32(gam);
      int it = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
        mbar_wait(a_full, it & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < C / 16; ++k)
          mma_bf16_ss(tmem_base, make_smem_desc_sw128(t_addr + (k >> 2) * G_CHUNK + (k & 3) * 32, 16, 1024),
                      make_smem_desc_sw128(b_addr + (k >> 2) * (C * 128) + (k & 3) * 32, 16, 1024), idesc, k != 0);
        mma_commit(tfull);
        mbar_wait(out_full, it & 1);
        tc_fence_after();
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {            // K = 128 tile rows, 16 per MMA (16 rows x 128 B = 2 KB)
          const uint32_t acc = (it | ks) != 0;
          const uint64_t td = make_smem_desc_sw128(t_addr + ks * 2048, G_CHUNK, 1024);
          const uint64_t sd = make_smem_desc_sw128(sq_addr + ks * 2048, G_CHUNK, 1024);
          mma_bf16_ss(tmem_base + COL_DGA, td, sd, idesc_dga, acc);
          if (REM) {
            const uint64_t tr = make_smem_desc_sw128(t_addr + 2 * G_CHUNK + ks * 2048, G_CHUNK, 1024);
            const uint64_t s1 = make_smem_desc_sw128(sq_addr + G_CHUNK + ks * 2048, G_CHUNK, 1024);
            mma_bf16_ss(tmem_base + COL_DGB, sd, tr, idesc_dgb, acc);         // j = 0..127   x  i = 128..191
            mma_bf16_ss(tmem_base + COL_DGB + 64, s1, tr, idesc_dgb, acc);    // j = 64..191  x  i = 128..191
          }
        }
        mma_commit(in_empty);                       // arrives once the MMAs above have read t and x^2
      }
      mma_commit(dg_done);
    }
  } else if (warp == 2) {
    if (lane < KC) {
      constexpr uint32_t lanes = (1u << KC) - 1;
      int it = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
        mbar_wait(out_full, it & 1);
        tma_store_2d(&tmDx, bg + lane * G_CHUNK, lane * 64, tile * 128);
        bulk_commit();
        bulk_wait_read0();
        __syncwarp(lanes);
        if (lane == 0) mbar_arrive(in_empty);
      }
      bulk_wait0();
    }
  } else {
    const int w = warp - 3;
    const int q = warp & 3, half = w >> 2;
    const int row = q * 32 + lane;
    const int wt = threadIdx.x - 96;              // 0..511: column owner for the column sums
    constexpr int UH = C / 32;
    // column sums: warp w owns tile rows 8w..8w+7, lane l < C/8 the 8 channels of unit l (16-byte loads); the
    // per-thread partials live in registers for the whole kernel and are combined across warps once at the end
    float tsum[8], dsum[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) tsum[e] = dsum[e] = 0.f;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
      mbar_wait(in_full, it & 1);
      constexpr int UB = UH % 4 == 0 ? 4 : (UH % 3 == 0 ? 3 : (UH % 2 == 0 ? 2 : 1));
#pragma unroll 1
      for (int ub = 0; ub < UH; ub += UB) {
        uint4 rg[UB], rx[UB], rr[UB];
#pragma unroll
        for (int j = 0; j < UB; ++j) {
          const uint32_t o = unit_off(row, half * UH + ub + j);
          rg[j] = *reinterpret_cast<const uint4*>(bg + o);
          rx[j] = *reinterpret_cast<const uint4*>(bx + o);
          rr[j] = *reinterpret_cast<const uint4*>(bn + o);
        }
#pragma unroll
        for (int j = 0; j < UB; ++j) {
          const uint32_t o = unit_off(row, half * UH + ub + j);
          float g[8], x[8], r[8], t[8], u[8];
          unpack8(rg[j], g);
          unpack8(rx[j], x);
          unpack8(rr[j], r);
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            t[e] = g[e] * x[e] * (p.inverse ? r[e] : r[e] * r[e] * r[e]);
            u[e] = p.inverse ? g[e] / r[e] : g[e] * r[e];
          }
          *reinterpret_cast<uint4*>(bn + o) = pack8(t);
          *reinterpret_cast<uint4*>(bg + o) = pack8(u);
        }
      }
      fence_proxy_async();
      mbar_arrive(a_full);
      // d(beta): column sums of the bf16 t tile, while the MMA runs (rows past P are zero: TMA zero fill)
      named_bar(4, G_WORKERS);
      if (lane < C / 8) {
#pragma unroll
        for (int rr = 0; rr < 8; ++rr) {
          float f[8];
          unpack8(*reinterpret_cast<const uint4*>(bn + unit_off(w * 8 + rr, lane)), f);
#pragma unroll
          for (int e = 0; e < 8; ++e) tsum[e] += f[e];
        }
      }
      mbar_wait(tfull, it & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + half * (C / 4);
#pragma unroll 1
      for (int gb = 0; gb < C / 64; ++gb) {
        float v[16];
        uint4 ur[2], xr[2];
        tmem_ld16(taddr + gb * 16, v);
        const int c0 = half * (C / 4) + gb * 16;
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          const uint32_t o = unit_off(row, (c0 >> 3) + hh);
          ur[hh] = *reinterpret_cast<const uint4*>(bg + o);
          xr[hh] = *reinterpret_cast<const uint4*>(bx + o);
        }
        tmem_ld_wait();
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          const uint32_t o = unit_off(row, (c0 >> 3) + hh);
          float u[8], x[8], d[8];
          unpack8(ur[hh], u);
          unpack8(xr[hh], x);
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            d[e] = p.inverse ? u[e] + x[e] * v[hh * 8 + e] : u[e] - x[e] * v[hh * 8 + e];
            x[e] *= x[e];
          }
          *reinterpret_cast<uint4*>(bg + o) = pack8(d);
          *reinterpret_cast<uint4*>(bx + o) = pack8(x);
        }
      }
      tc_fence_before();
      fence_proxy_async();
      mbar_arrive(out_full);
      // d(bias) of the producer layer: column sums of the bf16 dx tile, while it is being stored.  Rows past P
      // are skipped: the inverse form divides the zero-filled g by the zero-filled r there.
      named_bar(4, G_WORKERS);
      if (lane < C / 8) {
        const int64_t left = p.P - (int64_t)tile * 128;
        const int nrows = left < 128 ? (int)left : 128;
#pragma unroll
        for (int rr = 0; rr < 8; ++rr) {
          if (w * 8 + rr < nrows) {
            float f[8];
            unpack8(*reinterpret_cast<const uint4*>(bg + unit_off(w * 8 + rr, lane)), f);
#pragma unroll
            for (int e = 0; e < 8; ++e) dsum[e] += f[e];
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(in_empty);
    }
    // ---- d(gamma) partial of this CTA: TMEM -> dgamma_part[blockIdx] ----
    mbar_wait(dg_done, 0);
    tc_fence_after();
    {   // cross-warp combination of the column sums through the (now idle) t buffer, in a fixed order
      float* sc_t = reinterpret_cast<float*>(bn);
      float* sc_d = sc_t + (G_WORKERS / 32) * C;
      if (lane < C / 8) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          sc_t[w * C + lane * 8 + e] = tsum[e];
          sc_d[w * C + lane * 8 + e] = dsum[e];
        }
      }
      named_bar(4, G_WORKERS);
      if (wt < C) {
        float a = 0.f, b = 0.f;
#pragma unroll
        for (int k = 0; k < G_WORKERS / 32; ++k) { a += sc_t[k * C + wt]; b += sc_d[k * C + wt]; }
        p.colsum_part[(int64_t)blockIdx.x * C + wt] = a;
        p.dxsum_part[(int64_t)blockIdx.x * C + wt] = b;
      }
    }
    float* dst = p.dgamma_part + (int64_t)blockIdx.x * C * C;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
#pragma unroll 1
    for (int c0 = 0; c0 < C / 4; c0 += 16) {       // rows i = row (0..127), this thread's quarter of the columns
      float v[16];
      tmem_ld16(tmem_base + lane_off + COL_DGA + half * (C / 4) + c0, v);
      tmem_ld_wait();
      float4* d4 = reinterpret_cast<float4*>(dst + (int64_t)row * C + half * (C / 4) + c0);
#pragma unroll
      for (int k = 0; k < 4; ++k) d4[k] = make_float4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
    }
    if (REM) {                                     // rows i = 128 + i', held as [j][i']: lanes run along j (coalesced)
      float v[16];
      tmem_ld16(tmem_base + lane_off + COL_DGB + half * 16, v);
      tmem_ld_wait();
#pragma unroll
      for (int k = 0; k < 16; ++k) dst[(int64_t)(128 + half * 16 + k) * C + row] = v[k];
      if (q >= 2) {                                // second accumulator: j = 64 + row, only j >= 128 is new
        tmem_ld16(tmem_base + lane_off + COL_DGB + 64 + half * 16, v);
        tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < 16; ++k) dst[(int64_t)(128 + half * 16 + k) * C + 64 + row] = v[k];
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TCOLS);
}

// ---------------------------------------------------------------------------------------
// Dual-pipeline form of the fused backward.  With gamma^T resident there is shared memory for ONE set of 128-row
// tile buffers, so in gdn_bwd_fused_kernel the loads of tile i+1 wait for the tail of tile i (ncu: a third of the
// worker time is spent on the load barrier).  Here the same three 48 KB buffers are split by ROWS into two
// independent 64-row pipelines (set 0 = rows 0..63, set 1 = rows 64..127 of every [128 x 64] operand chunk), each with
// its own barriers, its own half of the worker warps (set s = the warps of TMEM lane quarters 2s, 2s+1) and its own
// half-tiles.  While one set waits for its loads the other computes.  The a = t . gamma UMMA still has the M = 128
// footprint over both sets (the rows of the other set are whatever is there) but is issued with the
// disable-output-lane mask of the other half, so the two sets share one accumulator; the d(gamma) UMMAs contract over
// the 64 rows of their own set only (K = 64).  TMEM budget unchanged (512 columns for C = 192).
// ---------------------------------------------------------------------------------------
//
// RC = true ("recompute"): r is NOT read from HBM (and the forward never writes it: 7 -> 5 HBM passes per layer and
// step).  The workers first write x^2 into the r buffer; an extra UMMA n = x^2 . gamma^T reads it as its A operand and
// the RESIDENT gamma^T tile through an MN-major descriptor as its B operand (element (N = i, K = k) of gamma is
// gammaT[k][i]: i contiguous), accumulating into the `a` columns, which are time-shared: n, read out as
// r = rsqrt(beta + n), then a = t . gamma.  TMEM stays at 512 columns; r is the forward's fp32 value instead of its
// bf16 rounding.
template <int C, bool RC>
__global__ void __launch_bounds__(G_THREADS, 1)
gdn_bwd_dual_kernel(const __grid_constant__ CUtensorMap tmGr, const __grid_constant__ CUtensorMap tmX,
                    const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmDx,
                    const __grid_constant__ CUtensorMap tmG, const __grid_constant__ GdnFusedParams p) {
  static_assert(C == 128 || C == 192, "fused d(gamma): C in {128, 192}");
  constexpr int NT = RC ? 2 : 3;                            // tensors loaded per half-tile: g, x (, r)
  using Cfg = GCfg<C>;
  constexpr int KC = Cfg::KC;
  constexpr int REM = C - 128;
  constexpr int COL_DGA = C, COL_DGB = 2 * C;
  constexpr int TCOLS = C == 128 ? 256 : 512;
  constexpr int HROWS = 64, HBYTES = HROWS * 128;          // rows per half-tile, bytes of a half chunk
  constexpr int GW = G_WORKERS / 2;                        // worker threads per set
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* gam = smem;                           // gamma^T as the K-major B operand
  uint8_t* bg = gam + Cfg::GAMMA;                // g   -> u = g r -> dx
  uint8_t* bx = bg + Cfg::TILE;                  // x   -> x^2
  uint8_t* bn = bx + Cfg::TILE;                  // r   -> t          (RC: x^2 -> t)
  float* beta_s = reinterpret_cast<float*>(bn + Cfg::TILE);              // RC: 256 floats
  uint64_t* bars = reinterpret_cast<uint64_t*>(bn + Cfg::TILE + 1024);
  uint64_t* g_full = bars;
  uint64_t* in_full = bars + 1;     // [2] TMA: g, x (, r) half-tiles landed
  uint64_t* in_empty = bars + 3;    // [2] KC store lanes + d(gamma) MMA commit + 8 worker warps
  uint64_t* a_full = bars + 5;      // [2] workers of the set: t (and u) written
  uint64_t* tfull = bars + 7;       // [2]
  uint64_t* out_full = bars + 9;    // [2] workers of the set: dx, x^2 written
  uint64_t* dg_done = bars + 11;
  uint64_t* sq_full = bars + 12;    // [2] RC: workers of the set: x^2 written
  uint64_t* nfull = bars + 14;      // [2] RC: n = x^2 . gamma^T is in TMEM
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nh = (int)((p.P + HROWS - 1) / HROWS);                              // half-tiles in the tensor
  const int mine = (int)blockIdx.x < nh ? (nh - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;   // ... of this CTA
  // half-tile k of this CTA (rows (blockIdx.x + k gridDim.x) * 64 ...) belongs to set k & 1
  if (threadIdx.x == 0) {
    prefetch_tmap(&tmGr); prefetch_tmap(&tmX); prefetch_tmap(&tmR); prefetch_tmap(&tmDx); prefetch_tmap(&tmG);
    mbar_init(g_full, 1); mbar_init(dg_done, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&in_full[s], 1); mbar_init(&in_empty[s], KC + 1 + GW / 32);
      mbar_init(&a_full[s], GW); mbar_init(&tfull[s], 1); mbar_init(&out_full[s], GW);
      mbar_init(&sq_full[s], GW); mbar_init(&nfull[s], 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TCOLS);
  if constexpr (RC) {
    pdl_wait();                    // beta was written by an earlier kernel of the stream
    for (int i = threadIdx.x; i < 256; i += G_THREADS) beta_s[i] = i < C ? p.beta[i] : 1.f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if constexpr (!RC) pdl_wait();   // the prologue above overlaps the previous kernel's tail
  pdl_launch_dependents();

  if (warp == 0) {
    // The whole warp walks the half-tiles of this CTA in order (k = 0, 1, 2, ...: the sets alternate) and waits
    // CONVERGED for the buffers of half-tile k; lane l then issues tensor l / KC, chunk l % KC of it.  (Independent
    // per-lane loops made 18 diverged threads spin on their own: ncu r02p, 1.6 M try_wait executions in this warp.)
    if (lane == 0) {
      mbar_arrive_expect_tx(g_full, Cfg::GAMMA);
      for (int c = 0; c < KC; ++c) tma_load_2d(gam + c * (C * 128), &tmG, g_full, c * 64, 0);
    }
    __syncwarp();
    const bool active = lane < NT * KC;
    const int which = active ? lane / KC : 0, c = lane % KC;
    uint8_t* dst0 = (which == 0 ? bg : (which == 1 ? bx : bn)) + c * G_CHUNK;
    const CUtensorMap* tm = which == 0 ? &tmGr : (which == 1 ? &tmX : &tmR);
    // L2 prefetch two half-tiles ahead: the load that follows the buffer release then pays L2 latency instead of HBM
    // latency (the buffers, not the bandwidth, bound the bytes in flight: HPC_GDN_PREFETCH=0 disables)
    if (p.prefetch && active)
      for (int k = 0; k < 2 && k < mine; ++k) tma_prefetch_2d(tm, c * 64, (blockIdx.x + k * gridDim.x) * HROWS);
    for (int k = 0; k < mine; ++k) {
      const int s = k & 1, j = k >> 1;
      const int h = blockIdx.x + k * gridDim.x;
      if (p.prefetch && active && k + 2 < mine) tma_prefetch_2d(tm, c * 64, (h + 2 * (int)gridDim.x) * HROWS);
      mbar_wait(&in_empty[s], (j & 1) ^ 1);
      if (lane == 0) mbar_arrive_expect_tx(&in_full[s], NT * Cfg::TILE / 2);
      __syncwarp();
      if (active) tma_load_2d(dst0 + s * HBYTES, tm, &in_full[s], c * 64, h * HROWS);
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, C, 0, 0);
      constexpr uint32_t idesc_dga = make_idesc_bf16(128, C, 1, 1);     // both operands MN-major
      constexpr uint32_t idesc_dgb = make_idesc_bf16(128, 64, 1, 1);
      mbar_wait(g_full, 0);
      const uint32_t t_addr = smem_u32(bn), sq_addr = smem_u32(bx), b_addr = smem_u32(gam);
      constexpr uint32_t idesc_n = make_idesc_bf16(128, C, 0, 1);       // RC: A = x^2 K-major, B = gamma^T read MN-major
      int nset[2] = {(mine + 1) / 2, mine / 2};
      // per set, the MMAs of half-tile j go out in the order N (RC only), A, D; st = which one is due
      int j[2] = {0, 0}, st[2] = {RC ? 0 : 1, RC ? 0 : 1};
      uint32_t dg_started = 0, spins = 0;
      int nextD = 0;      // d(gamma) accumulates the half-tiles in index order (k = 2 j + s): bit-reproducible sums
      while (j[0] < nset[0] || j[1] < nset[1]) {
        bool progressed = false;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          if (j[s] >= nset[s]) continue;
          const uint32_t keep = s ? 0u : 0xFFFFFFFFu;       // disable the OTHER set's lanes
          if (RC && st[s] == 0) {
            if (mbar_try_wait(&sq_full[s], j[s] & 1)) {
              tc_fence_after();
              // n[p][i] = sum_k x^2[p][k] gamma[i][k]: K = k runs over the ROWS of the resident gamma^T tile (16 rows =
              // 2 KB per step), N = i over its KC chunks (leading byte offset = chunk stride)
#pragma unroll
              for (int k = 0; k < C / 16; ++k)
                mma_bf16_ss_masked(tmem_base, make_smem_desc_sw128(t_addr + (k >> 2) * G_CHUNK + (k & 3) * 32, 16, 1024),
                                   make_smem_desc_sw128(b_addr + k * 2048, C * 128, 1024), idesc_n, k != 0, ~keep, ~keep,
                                   keep, keep);
              mma_commit(&nfull[s]);
              st[s] = 1;
              progressed = true;
            }
          } else if (st[s] == 1) {
            if (mbar_try_wait(&a_full[s], j[s] & 1)) {
              tc_fence_after();
#pragma unroll
              for (int k = 0; k < C / 16; ++k)
                mma_bf16_ss_masked(tmem_base, make_smem_desc_sw128(t_addr + (k >> 2) * G_CHUNK + (k & 3) * 32, 16, 1024),
                                   make_smem_desc_sw128(b_addr + (k >> 2) * (C * 128) + (k & 3) * 32, 16, 1024), idesc,
                                   k != 0, ~keep, ~keep, keep, keep);
              mma_commit(&tfull[s]);
              st[s] = 2;
              progressed = true;
            }
          } else if (st[s] == 2 && 2 * j[s] + s == nextD) {
            if (mbar_try_wait(&out_full[s], j[s] & 1)) {
              tc_fence_after();
#pragma unroll
              for (int ks = 0; ks < HROWS / 16; ++ks) {     // K = the 64 rows of this set
                const uint32_t acc = dg_started | (uint32_t)ks;
                const uint32_t ro = s * HBYTES + ks * 2048;
                const uint64_t td = make_smem_desc_sw128(t_addr + ro, G_CHUNK, 1024);
                const uint64_t sd = make_smem_desc_sw128(sq_addr + ro, G_CHUNK, 1024);
                mma_bf16_ss(tmem_base + COL_DGA, td, sd, idesc_dga, acc);
                if (REM) {
                  const uint64_t tr = make_smem_desc_sw128(t_addr + 2 * G_CHUNK + ro, G_CHUNK, 1024);
                  const uint64_t s1 = make_smem_desc_sw128(sq_addr + G_CHUNK + ro, G_CHUNK, 1024);
                  mma_bf16_ss(tmem_base + COL_DGB, sd, tr, idesc_dgb, acc);
                  mma_bf16_ss(tmem_base + COL_DGB + 64, s1, tr, idesc_dgb, acc);
                }
              }
              dg_started = 1;
              mma_commit(&in_empty[s]);
              st[s] = RC ? 0 : 1;
              ++j[s];
              ++nextD;
              progressed = true;
            }
          }
        }
        if (progressed) spins = 0;
        else if (++spins > (1u << 26)) { printf("hpcodec: gdn_bwd_dual MMA loop timed out (block %d)\n", blockIdx.x); __trap(); }
      }
      mma_commit(dg_done);
    }
  } else if (warp == 2) {
    // the warp walks the half-tiles in order, converged (see the producer); lane c < KC stores chunk c and arrives for
    // itself once its store has finished reading shared memory
    for (int k = 0; k < mine; ++k) {
      const int s = k & 1, j = k >> 1;
      const int h = blockIdx.x + k * gridDim.x;
      mbar_wait(&out_full[s], j & 1);
      if (lane < KC) {
        tma_store_2d(&tmDx, bg + lane * G_CHUNK + s * HBYTES, lane * 64, h * HROWS);
        bulk_commit();
        bulk_wait_read0();
        mbar_arrive(&in_empty[s]);
      }
      __syncwarp();
    }
    if (lane < KC) bulk_wait0();
  } else {
    const int w = warp - 3;
    const int q = warp & 3, half = w >> 2;        // TMEM lane quarter, channel quarter
    const int s = q >> 1;                         // set of this warp
    const int row = q * 32 + lane;                // row of the 128-row operand layout == TMEM lane
    const int gidx = (q & 1) * 4 + half;          // warp index within the set (0..7): owns rows s*64 + 8 gidx .. +7 of the sums
    const int wt = threadIdx.x - 96;
    constexpr int UH = C / 32;
    float tsum[8], dsum[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) tsum[e] = dsum[e] = 0.f;
    for (int k = s, j = 0; k < mine; k += 2, ++j) {
      const int h = blockIdx.x + k * gridDim.x;
      mbar_wait(&in_full[s], j & 1);
      constexpr int UB = UH % 4 == 0 ? 4 : (UH % 3 == 0 ? 3 : (UH % 2 == 0 ? 2 : 1));
      if constexpr (RC) {
        // phase 1: x^2 -> the r buffer (A operand of n = x^2 . gamma^T)
#pragma unroll 1
        for (int ub = 0; ub < UH; ub += UB) {
          uint4 rx[UB];
#pragma unroll
          for (int jj = 0; jj < UB; ++jj) rx[jj] = *reinterpret_cast<const uint4*>(bx + unit_off(row, half * UH + ub + jj));
#pragma unroll
          for (int jj = 0; jj < UB; ++jj) {
            float x[8];
            unpack8(rx[jj], x);
#pragma unroll
            for (int e = 0; e < 8; ++e) x[e] *= x[e];
            *reinterpret_cast<uint4*>(bn + unit_off(row, half * UH + ub + jj)) = pack8(x);
          }
        }
        fence_proxy_async();
        tc_fence_before();          // (the accumulator reads of the previous half-tile precede this arrival)
        mbar_arrive(&sq_full[s]);
        mbar_wait(&nfull[s], j & 1);
        tc_fence_after();
        // phase 2: r = rsqrt(beta + n) from TMEM;  t = g x r^3 (over x^2),  u = g r (over g)
        const uint32_t naddr = tmem_base + ((uint32_t)(q * 32) << 16) + half * (C / 4);
#pragma unroll 1
        for (int gb = 0; gb < C / 64; ++gb) {
          float v[16];
          uint4 gr[2], xr[2];
          tmem_ld16(naddr + gb * 16, v);
          const int c0 = half * (C / 4) + gb * 16;
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            const uint32_t o = unit_off(row, (c0 >> 3) + hh);
            gr[hh] = *reinterpret_cast<const uint4*>(bg + o);
            xr[hh] = *reinterpret_cast<const uint4*>(bx + o);
          }
          tmem_ld_wait();
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            const uint32_t o = unit_off(row, (c0 >> 3) + hh);
            float g[8], x[8], t[8], u[8];
            unpack8(gr[hh], g);
            unpack8(xr[hh], x);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const float n = v[hh * 8 + e] + beta_s[c0 + hh * 8 + e];
              const float r = rsqrt_normal(n);
              t[e] = g[e] * x[e] * (p.inverse ? r : r * r * r);
              u[e] = p.inverse ? g[e] * (n * r) : g[e] * r;        // g / r = g sqrt(n) = g n r
            }
            *reinterpret_cast<uint4*>(bn + o) = pack8(t);
            *reinterpret_cast<uint4*>(bg + o) = pack8(u);
          }
        }
      } else {
#pragma unroll 1
      for (int ub = 0; ub < UH; ub += UB) {
        uint4 rg[UB], rx[UB], rr[UB];
#pragma unroll
        for (int jj = 0; jj < UB; ++jj) {
          const uint32_t o = unit_off(row, half * UH + ub + jj);
          rg[jj] = *reinterpret_cast<const uint4*>(bg + o);
          rx[jj] = *reinterpret_cast<const uint4*>(bx + o);
          rr[jj] = *reinterpret_cast<const uint4*>(bn + o);
        }
#pragma unroll
        for (int jj = 0; jj < UB; ++jj) {
          const uint32_t o = unit_off(row, half * UH + ub + jj);
          if (p.packed) {
            // Packed bf16 arithmetic: g, x, r ARE bf16 values and t, u are stored as bf16, so the phase can run on
            // pairs without widening: 4 HMUL2 per pair instead of 3 unpacks, 4 multiplies and a pack per element
            // (this phase was 41 % of the kernel's instructions, ncu r02u).  Price: t carries three bf16 roundings
            // (~2^-8 relative) instead of one; u one, as before.  t only enters sums (a = t . gamma over 192 channels,
            // d(gamma) and d(beta) over all pixels), where the extra rounding noise averages out.
            const __nv_bfloat162* g2 = reinterpret_cast<const __nv_bfloat162*>(&rg[jj]);
            const __nv_bfloat162* x2 = reinterpret_cast<const __nv_bfloat162*>(&rx[jj]);
            const __nv_bfloat162* r2 = reinterpret_cast<const __nv_bfloat162*>(&rr[jj]);
            uint4 tq, uq;
            __nv_bfloat162* t2 = reinterpret_cast<__nv_bfloat162*>(&tq);
            __nv_bfloat162* u2 = reinterpret_cast<__nv_bfloat162*>(&uq);
            if (p.inverse) {            // t = g x r, u = g / r
#pragma unroll
              for (int w2 = 0; w2 < 4; ++w2) {
                t2[w2] = __hmul2(__hmul2(g2[w2], x2[w2]), r2[w2]);
                u2[w2] = __hmul2(g2[w2], h2rcp(r2[w2]));
              }
            } else {                    // u = g r, t = u (x r^2)
#pragma unroll
              for (int w2 = 0; w2 < 4; ++w2) {
                u2[w2] = __hmul2(g2[w2], r2[w2]);
                t2[w2] = __hmul2(u2[w2], __hmul2(x2[w2], __hmul2(r2[w2], r2[w2])));
              }
            }
            *reinterpret_cast<uint4*>(bn + o) = tq;
            *reinterpret_cast<uint4*>(bg + o) = uq;
            continue;
          }
          float g[8], x[8], r[8], t[8], u[8];
          unpack8(rg[jj], g);
          unpack8(rx[jj], x);
          unpack8(rr[jj], r);
          if (p.inverse) {              // t = g x r, u = g / r (4 instructions per element; r > 0 is a bf16 value)
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              t[e] = (g[e] * x[e]) * r[e];
              u[e] = g[e] * rcp_approx(r[e]);
            }
          } else {                      // u = g r, t = g x r^3 = u (x r^2)
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              u[e] = g[e] * r[e];
              t[e] = u[e] * (x[e] * (r[e] * r[e]));
            }
          }
          *reinterpret_cast<uint4*>(bn + o) = pack8(t);
          *reinterpret_cast<uint4*>(bg + o) = pack8(u);
        }
      }
      }
      fence_proxy_async();
      tc_fence_before();            // (the accumulator reads of the previous half-tile precede this arrival)
      mbar_arrive(&a_full[s]);
      // d(beta): column sums of the bf16 t half-tile, while the MMA runs (rows past P are zero: TMA zero fill)
      named_bar(4 + s, GW);
      if (lane < C / 8) {
        uint4 rows8[8];
#pragma unroll
        for (int rr2 = 0; rr2 < 8; ++rr2)
          rows8[rr2] = *reinterpret_cast<const uint4*>(bn + unit_off(s * HROWS + gidx * 8 + rr2, lane));
        colsum8_tree(rows8, tsum);
      }
      mbar_wait(&tfull[s], j & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + half * (C / 4);
#pragma unroll 1
      for (int gb = 0; gb < C / 64; ++gb) {
        float v[16];
        uint4 ur[2], xr[2];
        tmem_ld16(taddr + gb * 16, v);
        const int c0 = half * (C / 4) + gb * 16;
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          const uint32_t o = unit_off(row, (c0 >> 3) + hh);
          ur[hh] = *reinterpret_cast<const uint4*>(bg + o);
          xr[hh] = *reinterpret_cast<const uint4*>(bx + o);
        }
        tmem_ld_wait();
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          const uint32_t o = unit_off(row, (c0 >> 3) + hh);
          float u[8], x[8], d[8];
          unpack8(ur[hh], u);
          unpack8(xr[hh], x);
#pragma unroll
          for (int e = 0; e < 8; ++e) d[e] = p.inverse ? u[e] + x[e] * v[hh * 8 + e] : u[e] - x[e] * v[hh * 8 + e];
          // x^2 on the packed pairs: the product of two bf16 values is exact in fp32, so one HMUL2 gives the same bits as
          // two fp32 multiplies and a pack
          uint4 sq = xr[hh];
          __nv_bfloat162* s2 = reinterpret_cast<__nv_bfloat162*>(&sq);
#pragma unroll
          for (int w2 = 0; w2 < 4; ++w2) s2[w2] = __hmul2(s2[w2], s2[w2]);
          *reinterpret_cast<uint4*>(bg + o) = pack8(d);
          *reinterpret_cast<uint4*>(bx + o) = sq;
        }
      }
      tc_fence_before();
      fence_proxy_async();
      mbar_arrive(&out_full[s]);
      // d(bias) of the producer layer: column sums of the bf16 dx half-tile, while it is being stored.  Rows past P
      // are skipped: the inverse form divides the zero-filled g by the zero-filled r there.
      named_bar(4 + s, GW);
      if (lane < C / 8) {
        const int64_t left = p.P - (int64_t)h * HROWS;
        const int nrows = left < HROWS ? (int)left : HROWS;
        if (gidx * 8 + 8 <= nrows) {
          uint4 rows8[8];
#pragma unroll
          for (int rr2 = 0; rr2 < 8; ++rr2)
            rows8[rr2] = *reinterpret_cast<const uint4*>(bg + unit_off(s * HROWS + gidx * 8 + rr2, lane));
          colsum8_tree(rows8, dsum);
        } else {                        // the ragged end of the tensor: row by row
#pragma unroll
          for (int rr2 = 0; rr2 < 8; ++rr2) {
            if (gidx * 8 + rr2 < nrows) {
              float f[8];
              unpack8(*reinterpret_cast<const uint4*>(bg + unit_off(s * HROWS + gidx * 8 + rr2, lane)), f);
#pragma unroll
              for (int e = 0; e < 8; ++e) dsum[e] += f[e];
            }
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&in_empty[s]);
    }
    // ---- d(gamma) partial of this CTA: TMEM -> dgamma_part[blockIdx] ----
    mbar_wait(dg_done, 0);
    tc_fence_after();
    {   // cross-warp combination of the column sums through the (now idle) t buffer, in a fixed order
      float* sc_t = reinterpret_cast<float*>(bn);
      float* sc_d = sc_t + (G_WORKERS / 32) * C;
      const int widx = s * 8 + gidx;
      if (lane < C / 8) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          sc_t[widx * C + lane * 8 + e] = tsum[e];
          sc_d[widx * C + lane * 8 + e] = dsum[e];
        }
      }
      named_bar(6, G_WORKERS);
      if (wt < C) {
        float a = 0.f, b = 0.f;
#pragma unroll
        for (int k = 0; k < G_WORKERS / 32; ++k) { a += sc_t[k * C + wt]; b += sc_d[k * C + wt]; }
        p.colsum_part[(int64_t)blockIdx.x * C + wt] = a;
        p.dxsum_part[(int64_t)blockIdx.x * C + wt] = b;
      }
    }
    float* dst = p.dgamma_part + (int64_t)blockIdx.x * C * C;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    const bool any = mine > 0;                     // a CTA without work never issued an MMA: its partial is zero
#pragma unroll 1
    for (int c0 = 0; c0 < C / 4; c0 += 16) {
      float v[16];
      tmem_ld16(tmem_base + lane_off + COL_DGA + half * (C / 4) + c0, v);
      tmem_ld_wait();
      float4* d4 = reinterpret_cast<float4*>(dst + (int64_t)row * C + half * (C / 4) + c0);
#pragma unroll
      for (int kk = 0; kk < 4; ++kk)
        d4[kk] = any ? make_float4(v[4 * kk], v[4 * kk + 1], v[4 * kk + 2], v[4 * kk + 3]) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (REM) {
      float v[16];
      tmem_ld16(tmem_base + lane_off + COL_DGB + half * 16, v);
      tmem_ld_wait();
#pragma unroll
      for (int kk = 0; kk < 16; ++kk) dst[(int64_t)(128 + half * 16 + kk) * C + row] = any ? v[kk] : 0.f;
      if (q >= 2) {
        tmem_ld16(tmem_base + lane_off + COL_DGB + 64 + half * 16, v);
        tmem_ld_wait();
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) dst[(int64_t)(128 + half * 16 + kk) * C + 64 + row] = any ? v[kk] : 0.f;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TCOLS);
}

// One launch finishes the backward of a GDN layer from the per-CTA partials, in a fixed summation order:
//   blocks [0, ng)        d(gamma)[i]   = R(gamma_param[i], sgn * sum_s dgamma_part[s][i])     (128 elements per block)
//   blocks [ng, ng+nc)    d(beta)[c]    = R(beta_param[c],  sgn * sum_s colsum_part[s][c])
//   blocks [ng+nc, ..)    d(bias)[c]    = sum_s dxsum_part[s][c]                                (producer layer)
// R(p, g) is the backward of the non-negative reparametrisation (gdn.py:42-62: v = max(p, bound), value v^2 - pedestal,
// LowerBound pass-through rule bound.py:36-46) when the raw parameter is given, the identity otherwise.
struct GdnFinish {
  const float* dgamma_part; const float* colsum_part; const float* dxsum_part;
  const float* gamma_param; const float* beta_param;
  float* dgamma; float* dbeta; float* dbias;
  int nparts, C, ng, nc;
  float sgn, gamma_bound, beta_bound;
};

__device__ __forceinline__ float reparam_bwd(const float* param, int i, float g, float bound) {
  if (!param) return g;
  const float pv = param[i], v = fmaxf(pv, bound);
  const float gp = 2.f * v * g;
  return ((pv >= bound) || (gp < 0.f)) ? gp : 0.f;
}

__global__ void __launch_bounds__(256)
gdn_bwd_finish_kernel(const __grid_constant__ GdnFinish f) {
  __shared__ float4 sm[8][32];
  const int qd = threadIdx.x & 31, g = threadIdx.x >> 5;
  int blk = blockIdx.x;
  const float* part; const float* param; float* out; int n; float scale, bound;
  if (blk < f.ng) { part = f.dgamma_part; param = f.gamma_param; out = f.dgamma; n = f.C * f.C; scale = f.sgn; bound = f.gamma_bound; }
  else if (blk < f.ng + f.nc) { blk -= f.ng; part = f.colsum_part; param = f.beta_param; out = f.dbeta; n = f.C; scale = f.sgn; bound = f.beta_bound; }
  else { blk -= f.ng + f.nc; part = f.dxsum_part; param = nullptr; out = f.dbias; n = f.C; scale = 1.f; bound = 0.f; }
  const int i = (blk * 32 + qd) * 4;
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
  pdl_wait();
  pdl_launch_dependents();
  if (i < n)
    for (int s2 = g; s2 < f.nparts; s2 += 8) {
      const float4 v = __ldcs(reinterpret_cast<const float4*>(part + (int64_t)s2 * n + i));
      a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
  sm[g][qd] = a;
  __syncthreads();
  if (g != 0 || i >= n) return;
#pragma unroll
  for (int k = 1; k < 8; ++k) {
    const float4 v = sm[k][qd];
    a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
  }
  float4 r;
  r.x = reparam_bwd(param, i, a.x * scale, bound);
  r.y = reparam_bwd(param, i + 1, a.y * scale, bound);
  r.z = reparam_bwd(param, i + 2, a.z * scale, bound);
  r.w = reparam_bwd(param, i + 3, a.w * scale, bound);
  *reinterpret_cast<float4*>(out + i) = r;
}

int make_rows_tmap(CUtensorMap* m, const void* base, int64_t P, int C, int box_rows) {
  uint64_t dims[2] = {(uint64_t)C, (uint64_t)P};
  uint64_t strides[1] = {(uint64_t)C * 2};
  uint32_t box[2] = {64, (uint32_t)box_rows};
  uint32_t estr[2] = {1, 1};
  return make_tmap(m, base, 2, dims, strides, box, estr);
}

int gdn_grid(int tiles) { return tiles < hpc_num_sms() ? tiles : hpc_num_sms(); }

}  // namespace

#define GDN_DISPATCH(C_, ...)      \
  switch (C_) {                     \
    case 64: { constexpr int CC = 64; __VA_ARGS__; } break;    \
    case 128: { constexpr int CC = 128; __VA_ARGS__; } break;  \
    default: { constexpr int CC = 192; __VA_ARGS__; } break;   \
  }

static int gdn_check(const char* who, int64_t P, int C, std::initializer_list<const void*> ptrs) {
  HPC_REQUIRE(P > 0, HPC_ERR_BAD_SHAPE, "%s: empty input", who);
  HPC_REQUIRE(C == 64 || C == 128 || C == 192, HPC_ERR_UNSUPPORTED, "%s: C=%d not in {64,128,192}", who, C);
  for (const void* q : ptrs) {
    HPC_REQUIRE(q, HPC_ERR_BAD_SHAPE, "%s: NULL pointer", who);
    HPC_REQUIRE(hpc_aligned16(q), HPC_ERR_BAD_ALIGNMENT, "%s: pointers must be 16B aligned", who);
  }
  return HPC_OK;
}

int hpc_gdn_fwd_umma(const void* x_bf16, const void* gamma_bf16, const float* beta, int64_t P, int C, int inverse,
                     void* y_bf16, void* rnorm_bf16, hpc_stream_t stream) {
  // rnorm_bf16 == NULL: r is not stored (evaluation under no_grad: nothing will run a backward)
  int rc = gdn_check("hpc_gdn_fwd_umma", P, C, {x_bf16, gamma_bf16, beta, y_bf16, rnorm_bf16 ? rnorm_bf16 : y_bf16});
  if (rc) return rc;
  GdnParams kp;
  kp.beta = beta; kp.colsum_part = nullptr; kp.P = P; kp.tiles = (int)((P + 127) / 128); kp.inverse = inverse;
  kp.store_r = rnorm_bf16 ? 1 : 0;
  CUtensorMap tmX, tmY, tmR, tmG;
  if ((rc = make_rows_tmap(&tmX, x_bf16, P, C, 128))) return rc;
  if ((rc = make_rows_tmap(&tmY, y_bf16, P, C, 128))) return rc;
  if ((rc = make_rows_tmap(&tmR, rnorm_bf16 ? rnorm_bf16 : y_bf16, P, C, 128))) return rc;      // unused without r
  if ((rc = make_rows_tmap(&tmG, gamma_bf16, C, C, C))) return rc;
  HpcProfScope prof(HPC_PROF_GDN, (double)P * C * 2 * (rnorm_bf16 ? 3 : 2), stream);
  auto st = (cudaStream_t)stream;
  GDN_DISPATCH(C, {
    static HpcOncePerDevice attr_once;
    int attr_dev;
    if (attr_once.need(&attr_dev)) {
      HPC_CUDA(cudaFuncSetAttribute(gdn_fwd_umma_kernel<CC>, cudaFuncAttributeMaxDynamicSharedMemorySize, GCfg<CC>::SMEM));
      attr_once.mark(attr_dev);
    }
    HPC_CUDA(hpc_launch_pdl(gdn_fwd_umma_kernel<CC>, dim3(gdn_grid(kp.tiles)), dim3(G_THREADS), GCfg<CC>::SMEM, st, tmX, tmY, tmR, tmG, kp));
  });
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

int hpc_gdn_bwd_umma_parts(int64_t P) { return gdn_grid((int)((P + 127) / 128)); }

int hpc_gdn_bwd_umma(const void* g_bf16, const void* x_bf16, const void* rnorm_bf16, const void* gammaT_bf16, int64_t P,
                     int C, int inverse, void* dx_bf16, void* t_bf16, void* xsq_bf16, float* colsum_part,
                     hpc_stream_t stream) {
  int rc = gdn_check("hpc_gdn_bwd_umma", P, C, {g_bf16, x_bf16, rnorm_bf16, gammaT_bf16, dx_bf16, t_bf16, xsq_bf16, colsum_part});
  if (rc) return rc;
  GdnParams kp;
  kp.beta = nullptr; kp.colsum_part = colsum_part; kp.P = P; kp.tiles = (int)((P + 127) / 128); kp.inverse = inverse;
  CUtensorMap tmGr, tmX, tmR, tmDx, tmT, tmSq, tmG;
  if ((rc = make_rows_tmap(&tmGr, g_bf16, P, C, 128))) return rc;
  if ((rc = make_rows_tmap(&tmX, x_bf16, P, C, 128))) return rc;
  if ((rc = make_rows_tmap(&tmR, rnorm_bf16, P, C, 128))) return rc;
  if ((rc = make_rows_tmap(&tmDx, dx_bf16, P, C, 128))) return rc;
  if ((rc = make_rows_tmap(&tmT, t_bf16, P, C, 128))) return rc;
  if ((rc = make_rows_tmap(&tmSq, xsq_bf16, P, C, 128))) return rc;
  if ((rc = make_rows_tmap(&tmG, gammaT_bf16, C, C, C))) return rc;
  HpcProfScope prof(HPC_PROF_GDN, (double)P * C * 2 * 6, stream);
  auto st = (cudaStream_t)stream;
  GDN_DISPATCH(C, {
    static HpcOncePerDevice attr_once;
    int attr_dev;
    if (attr_once.need(&attr_dev)) {
      HPC_CUDA(cudaFuncSetAttribute(gdn_bwd_umma_kernel<CC>, cudaFuncAttributeMaxDynamicSharedMemorySize, GCfg<CC>::SMEM));
      attr_once.mark(attr_dev);
    }
    gdn_bwd_umma_kernel<CC><<<gdn_grid(kp.tiles), G_THREADS, GCfg<CC>::SMEM, st>>>(tmGr, tmX, tmR, tmDx, tmT, tmSq, tmG, kp);
  });
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

size_t hpc_gdn_bwd_fused_umma_workspace_bytes(int64_t P, int C) {
  if (C != 128 && C != 192) return 0;
  return (size_t)gdn_grid((int)((P + 63) / 64)) * (C * C + 2 * C) * sizeof(float);   // >= the single-pipeline grid
}

// rnorm_bf16 != NULL: r is read (4 HBM passes);  rnorm_bf16 == NULL: beta_eff given, r is recomputed in the kernel (3 passes)
static int gdn_bwd_fused_launch(const void* g_bf16, const void* x_bf16, const void* rnorm_bf16, const float* beta_eff,
                                const void* gammaT_bf16, int64_t P, int C, int inverse, void* dx_bf16,
                                const float* gamma_param, float gamma_bound, const float* beta_param, float beta_bound,
                                float* dgamma, float* dbeta, float* dbias, void* workspace, size_t workspace_bytes,
                                hpc_stream_t stream) {
  const bool recompute = rnorm_bf16 == nullptr;
  int rc = gdn_check("hpc_gdn_bwd_fused_umma", P, C,
                     {g_bf16, x_bf16, recompute ? (const void*)beta_eff : rnorm_bf16, gammaT_bf16, dx_bf16, dgamma, dbeta,
                      workspace});
  if (rc) return rc;
  HPC_REQUIRE(C == 128 || C == 192, HPC_ERR_UNSUPPORTED, "hpc_gdn_bwd_fused_umma: C=%d not in {128,192}", C);
  HPC_REQUIRE(workspace_bytes >= hpc_gdn_bwd_fused_umma_workspace_bytes(P, C), HPC_ERR_BAD_SHAPE,
              "hpc_gdn_bwd_fused_umma: workspace too small (%zu)", workspace_bytes);
  HPC_REQUIRE(!dbias || hpc_aligned16(dbias), HPC_ERR_BAD_ALIGNMENT, "hpc_gdn_bwd_fused_umma: dbias must be 16B aligned");
  static int dual = -1;                          // HPC_GDN_BWD_DUAL=0 selects the single-pipeline kernel (A/B measurements)
  if (dual < 0) { const char* e = getenv("HPC_GDN_BWD_DUAL"); dual = e ? (atoi(e) ? 1 : 0) : 1; }
  HPC_REQUIRE(dual || !recompute, HPC_ERR_UNSUPPORTED, "the recompute form of the GDN backward exists for the dual-pipeline "
              "kernel only (HPC_GDN_BWD_DUAL=0 is set)");
  const int grid = dual ? gdn_grid((int)((P + 63) / 64)) : gdn_grid((int)((P + 127) / 128));
  GdnFusedParams kp;
  kp.beta = beta_eff;
  kp.dgamma_part = (float*)workspace;
  kp.colsum_part = kp.dgamma_part + (size_t)grid * C * C;
  kp.dxsum_part = kp.colsum_part + (size_t)grid * C;
  kp.tiles = (int)((P + 127) / 128); kp.inverse = inverse; kp.P = P;
  static int prefetch = -1;
  if (prefetch < 0) { const char* e = getenv("HPC_GDN_PREFETCH"); prefetch = e ? (atoi(e) ? 1 : 0) : 1; }
  kp.prefetch = prefetch;
  static int packed = -1;                        // HPC_GDN_PACKED=0: fp32 arithmetic in the t / u phase (A/B measurements)
  if (packed < 0) { const char* e = getenv("HPC_GDN_PACKED"); packed = e ? (atoi(e) ? 1 : 0) : 1; }
  kp.packed = packed;
  const int box_rows = dual ? 64 : 128;
  CUtensorMap tmGr, tmX, tmR, tmDx, tmG;
  if ((rc = make_rows_tmap(&tmGr, g_bf16, P, C, box_rows))) return rc;
  if ((rc = make_rows_tmap(&tmX, x_bf16, P, C, box_rows))) return rc;
  if ((rc = make_rows_tmap(&tmR, recompute ? x_bf16 : rnorm_bf16, P, C, box_rows))) return rc;      // unused when recomputing
  if ((rc = make_rows_tmap(&tmDx, dx_bf16, P, C, box_rows))) return rc;
  if ((rc = make_rows_tmap(&tmG, gammaT_bf16, C, C, C))) return rc;
  HpcProfScope prof(HPC_PROF_GDN, (double)P * C * 2 * (recompute ? 3 : 4), stream);
  auto st = (cudaStream_t)stream;
#define GDN_BWD_LAUNCH(...)                                                                                     \
  do {                                                                                                          \
    auto kfn = __VA_ARGS__;                                                                                     \
    static HpcOncePerDevice attr_once;                                                                               \
    int attr_dev;                                                                               \
    if (attr_once.need(&attr_dev)) {                                                                                            \
      HPC_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));                   \
      attr_once.mark(attr_dev);                                                                                          \
    }                                                                                                           \
    HPC_CUDA(hpc_launch_pdl(kfn, dim3(grid), dim3(G_THREADS), smem, st, tmGr, tmX, tmR, tmDx, tmG, kp));        \
  } while (0)
  const size_t smem = C == 128 ? GCfg<128>::SMEM : GCfg<192>::SMEM;
  if (dual && recompute) {
    if (C == 128) GDN_BWD_LAUNCH(gdn_bwd_dual_kernel<128, true>); else GDN_BWD_LAUNCH(gdn_bwd_dual_kernel<192, true>);
  } else if (dual) {
    if (C == 128) GDN_BWD_LAUNCH(gdn_bwd_dual_kernel<128, false>); else GDN_BWD_LAUNCH(gdn_bwd_dual_kernel<192, false>);
  } else {
    if (C == 128) GDN_BWD_LAUNCH(gdn_bwd_fused_kernel<128>); else GDN_BWD_LAUNCH(gdn_bwd_fused_kernel<192>);
  }
#undef GDN_BWD_LAUNCH
  HPC_LAUNCH_CHECK();
  GdnFinish f;
  f.dgamma_part = kp.dgamma_part; f.colsum_part = kp.colsum_part; f.dxsum_part = kp.dxsum_part;
  f.gamma_param = gamma_param; f.beta_param = beta_param;
  f.dgamma = dgamma; f.dbeta = dbeta; f.dbias = dbias;
  f.nparts = grid; f.C = C; f.ng = (C * C + 127) / 128; f.nc = (C + 127) / 128;
  f.sgn = inverse ? 0.5f : -0.5f; f.gamma_bound = gamma_bound; f.beta_bound = beta_bound;
  HPC_CUDA(hpc_launch_pdl(gdn_bwd_finish_kernel, dim3(f.ng + f.nc + (dbias ? f.nc : 0)), dim3(256), 0, st, f));
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

int hpc_gdn_bwd_fused_umma(const void* g_bf16, const void* x_bf16, const void* rnorm_bf16, const void* gammaT_bf16,
                           int64_t P, int C, int inverse, void* dx_bf16, const float* gamma_param, float gamma_bound,
                           const float* beta_param, float beta_bound, float* dgamma, float* dbeta, float* dbias,
                           void* workspace, size_t workspace_bytes, hpc_stream_t stream) {
  HPC_REQUIRE(rnorm_bf16, HPC_ERR_BAD_SHAPE, "hpc_gdn_bwd_fused_umma: rnorm is NULL (use hpc_gdn_bwd_fused_rc_umma to "
              "recompute it)");
  return gdn_bwd_fused_launch(g_bf16, x_bf16, rnorm_bf16, nullptr, gammaT_bf16, P, C, inverse, dx_bf16, gamma_param,
                              gamma_bound, beta_param, beta_bound, dgamma, dbeta, dbias, workspace, workspace_bytes, stream);
}

int hpc_gdn_bwd_fused_rc_umma(const void* g_bf16, const void* x_bf16, const float* beta_eff, const void* gammaT_bf16,
                              int64_t P, int C, int inverse, void* dx_bf16, const float* gamma_param, float gamma_bound,
                              const float* beta_param, float beta_bound, float* dgamma, float* dbeta, float* dbias,
                              void* workspace, size_t workspace_bytes, hpc_stream_t stream) {
  HPC_REQUIRE(beta_eff, HPC_ERR_BAD_SHAPE, "hpc_gdn_bwd_fused_rc_umma: beta_eff is NULL");
  return gdn_bwd_fused_launch(g_bf16, x_bf16, nullptr, beta_eff, gammaT_bf16, P, C, inverse, dx_bf16, gamma_param,
                              gamma_bound, beta_param, beta_bound, dgamma, dbeta, dbias, workspace, workspace_bytes, stream);
}

// gdn.py:59-62 for both parameters of a layer + both bf16 operand copies of gamma, one launch:
//   gamma_eff = max(gamma_param, gamma_bound)^2 - pedestal -> gamma_bf16[i][k] (forward B operand) and gammaT_bf16[k][i]
//   beta_eff  = max(beta_param, beta_bound)^2 - pedestal   -> beta_eff[C] fp32
__global__ void gdn_prepare_kernel(const float* __restrict__ gamma_param, const float* __restrict__ beta_param, int C,
                                   float gamma_bound, float beta_bound, float pedestal, float* __restrict__ beta_eff,
                                   __nv_bfloat16* __restrict__ g_bf16, __nv_bfloat16* __restrict__ gT_bf16) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  pdl_wait();
  pdl_launch_dependents();
  if (idx < C * C) {
    const float v = fmaxf(gamma_param[idx], gamma_bound);
    const __nv_bfloat16 h = __float2bfloat16_rn(v * v - pedestal);
    g_bf16[idx] = h;
    gT_bf16[(idx % C) * C + idx / C] = h;
  }
  if (idx < C) {
    const float v = fmaxf(beta_param[idx], beta_bound);
    beta_eff[idx] = v * v - pedestal;
  }
}

int hpc_gdn_prepare(const float* gamma_param, const float* beta_param, int C, float gamma_bound, float beta_bound,
                    float pedestal, float* beta_eff, void* gamma_bf16, void* gammaT_bf16, hpc_stream_t stream) {
  HPC_REQUIRE(gamma_param && beta_param && beta_eff && gamma_bf16 && gammaT_bf16 && C > 0, HPC_ERR_BAD_SHAPE,
              "hpc_gdn_prepare: bad arguments");
  HPC_CUDA(hpc_launch_pdl(gdn_prepare_kernel, dim3((C * C + 255) / 256), dim3(256), 0, (cudaStream_t)stream, gamma_param,
                          beta_param, C, gamma_bound, beta_bound, pedestal, beta_eff, (__nv_bfloat16*)gamma_bf16,
                          (__nv_bfloat16*)gammaT_bf16));
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

