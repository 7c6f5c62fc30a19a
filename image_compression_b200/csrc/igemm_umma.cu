// tcgen05 / TMEM / TMA engine of the implicit-GEMM convolution family (sm_100a).
//
// Forward kernel (conv, transposed-conv phases, dgrads, the GDN gamma contraction):
//   one persistent CTA per SM, 192 threads = warp 0 TMA producer, warp 1 MMA issuer + TMEM owner,
//   warps 2-5 epilogue.  An M tile is a (nb x th x tw) = 128-pixel patch of the iteration grid; for
//   every tap t and 64-channel slice c the producer issues ONE 4-D TMA box over the NHWC input
//   (elementStrides = input stride, out-of-bounds = zero fill == the convolution padding) that lands
//   as a canonical K-major SWIZZLE_128B [128 x 64] bf16 operand, plus a [BLOCK_N x 64] weight box.
//   Accumulators (128 lanes x BLOCK_N fp32 columns) are double-buffered in TMEM so the epilogue of
//   tile i overlaps the MMAs of tile i+1.
//   Builds of the same body (igemm_fwd_body), chosen per launch by fwd_plan():
//     igemm_fwd_umma_kernel          one CTA per tile (grids below 148 tiles; split-K below 74)
//     igemm_fwd_umma_pair_kernel     CTA pairs, tcgen05.mma.cta_group::2, each CTA holds half of the weight tile
//     igemm_fwd_umma_pair2_kernel    ... and TWO M tiles per CTA: the pair kernel was bound by L2 -> SM operand bytes
//     igemm_fwd_umma_pair2s_kernel   ... and one activation box per kernel column (off by default: no faster)
// Weight-gradient kernel (igemm_wgrad_umma_kernel): see its own header further down.
#include <cuda.h>

#include <cstdlib>

#include "igemm_epilogue.cuh"
#include "tmap.cuh"
#include "umma_ptx.cuh"

using namespace umma;

// one contiguous row shared -> global (cp.async.bulk, bulk-group completion); 16-byte aligned, bytes % 16 == 0
__device__ __forceinline__ void bulk_store_row(void* gdst, const void* ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_group_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_group_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_group_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// ---------------------------------------------------------------------------------------
// forward kernel
// ---------------------------------------------------------------------------------------
constexpr int FW_THREADS = 192;
constexpr int FW_BM = 128;   // pixels per tile (UMMA M)
constexpr int FW_BK = 64;    // channels per stage (one 128-byte swizzle row of bf16)

struct FwdKParams {
  int N, GH, GW, is, os, OH, OW, Co, Ci;
  int tw, th, nb;            // tile shape, tw*th*nb == 128
  int tiles_w, tiles_h, tiles_n, tiles_per_phase, total_tiles;
  int interleave;           // see work_item()
  int nphases;
  hpc_igemm_phase phase[HPC_MAX_PHASES];
  int dh[HPC_MAX_TAPS], dw[HPC_MAX_TAPS];
  int kchunks;               // Ci / 64
  int stage_out;             // CTA-pair kernel, plain bias epilogue with the bf16 output only: rows staged in shared
                             // memory and written with one bulk copy each
  int ksplit;                // > 1: work item = (tile, K slice); raw fp32 partials go to ws[slice][N,OH,OW,Co]
  float* ws;
  int64_t out_elems;
  EpiArgs epi;
};

// MT = M tiles per CTA and work item (CTA-pair kernel only).  MT == 2: a CTA holds TWO 128-pixel tiles per stage and both
// of its accumulators belong to the same work item, so every weight box is used for twice the rows.  ncu on the
// 64x64x64-pixel 5x5 layer (r02p): 4.40 GB cross the L2 -> SM crossbar in 408 us = 10.8 TB/s, the chip's L2 output cap,
// with the tensor pipe 63 % busy -- the MT == 1 pair kernel is bound by operand bytes, 43 % of them weights.  MT == 2
// halves the weight bytes per flop (4.40 -> 3.46 GB) at the price of an epilogue that no longer overlaps the next
// tile's MMAs (TMEM holds 2 x 256 columns: both are in use).
// SHARE (with MT == 2): the taps of one kernel column (equal dw) whose dh are consecutive steps of the input stride read
// the same pixels one tile ROW apart, and a row of the tile (tw = 8 or 16 pixels x 128 B) is one or two whole 1 KB
// swizzle groups of the K-major operand: the 2-3 taps of such a group share ONE activation box of th + 1 / th + 2 rows
// and their descriptors start tw x 128 B apart.  A 5 x 5 stride-2 layer then loads 2 boxes (10 + 9 rows) instead of 5
// (8 rows each) per kernel column and channel slice: 52 % fewer activation bytes.  The activation boxes and the weight
// boxes run in separate rings (an activation slot lives for 2-3 weight slots), filled by separate warps.
constexpr int FS_THREADS = 224;        // SHARE: warp 6 loads the weights
constexpr int FS_A_SLOTS = 3, FS_B_SLOTS = 4;
constexpr int FS_A_TILE = (FW_BM + 2 * 16) * 128;     // (th + 2) x tw pixels, tw <= 16: 20 KB per M tile
struct FwdShareSched {                 // K loop of a phase: tap groups (equal dw, dh in consecutive steps of the stride)
  uint8_t gbegin[HPC_MAX_PHASES], ngroups[HPC_MAX_PHASES];
  uint8_t gtap0[HPC_MAX_TAPS], gsize[HPC_MAX_TAPS];   // group g = taps order[gtap0[g]] .. (gsize[g] of them, dh ascending)
  uint8_t order[HPC_MAX_TAPS];
};

template <int BLOCK_N, int NCTA, int MT = 1, bool SHARE = false>
struct FwdCfg {
  static_assert(MT == 1 || NCTA == 2, "two M tiles per CTA: CTA-pair kernel only");
  static_assert(!SHARE || MT == 2, "shared activation boxes: MT == 2 kernel only");
  static constexpr int A_BYTES = FW_BM * FW_BK * 2;        // 16 KB per M tile
  static constexpr int B_BYTES = (BLOCK_N / NCTA) * FW_BK * 2;   // a CTA of a pair holds half of the weight tile
  static constexpr int STAGE_BYTES = MT * A_BYTES + B_BYTES;
  // CTA-pair kernel: one [128 rows x (BLOCK_N bf16 + 16 B pad)] output staging tile (rows leave as single bulk copies)
  static constexpr int STG_PITCH = BLOCK_N * 2 + 16;
  static constexpr int STG_BYTES = NCTA == 2 ? FW_BM * STG_PITCH : 0;
  // MT == 2 needs every byte for its 4th stage (C = 192: 4 x 44 KB + 50 KB staging + 1 KB = 227 KB exactly): no
  // alignment slack (the kernel traps if the dynamic shared memory window is not 1 KB aligned) and a bias area of
  // BLOCK_N floats
  static constexpr int ALIGN_SLACK = MT == 2 ? 0 : 1024;
  static constexpr int BIAS_BYTES = MT == 2 ? BLOCK_N * 4 : 1024;
  static constexpr int FIXED_BYTES = ALIGN_SLACK + 256 /*barriers*/ + BIAS_BYTES;
  static constexpr int BUDGET = NCTA == 2 ? (232448 - FIXED_BYTES - STG_BYTES) : 200 * 1024;
  static constexpr int STAGES = (BUDGET / STAGE_BYTES) > 8 ? 8 : (BUDGET / STAGE_BYTES);
  static constexpr int OPERAND_BYTES = SHARE ? FS_A_SLOTS * MT * FS_A_TILE + FS_B_SLOTS * B_BYTES : STAGES * STAGE_BYTES;
  static constexpr int ACC_STRIDE = BLOCK_N <= 32 ? 32 : (BLOCK_N <= 64 ? 64 : (BLOCK_N <= 128 ? 128 : 256));
  static constexpr int TMEM_COLS = 2 * ACC_STRIDE;          // 64..512, power of two
  static constexpr int SMEM_BYTES = OPERAND_BYTES + FIXED_BYTES + STG_BYTES;
  static_assert(SMEM_BYTES <= 232448, "shared memory budget");
  static_assert((2 * STAGES + 4) * 8 + 4 <= 256 && (2 * (FS_A_SLOTS + FS_B_SLOTS) + 4) * 8 + 4 <= 256, "barrier area");
};

// Work item `w` of the persistent loop -> (phase, tile origin).  NCTA == 1: w is a tile.  NCTA == 2: w is a PAIR of
// consecutive tiles of one phase (so both CTAs run the same tap sequence) and `rank` picks the tile; a phase with an
// odd tile count ends in a phantom tile whose batch origin n0 >= N (TMA zero fill, no rows stored).  MT == 2: w is a
// group of FOUR consecutive tiles, two per CTA.
template <int NCTA, int MT = 1>
__device__ __forceinline__ void tile_coords(const FwdKParams& p, int ph, int idx, int rank, int& n0, int& i0, int& j0, int m = 0) {
  int r = (idx * MT + m) * NCTA + rank;         // MT == 2: accumulator m of the pair = tiles (2 m, 2 m + 1) of the item's four
  const int tj = r % p.tiles_w; r /= p.tiles_w;
  const int ti = r % p.tiles_h; r /= p.tiles_h;
  n0 = r * p.nb; i0 = ti * p.th; j0 = tj * p.tw;
}

// Work item w of a worker (CTA or CTA pair) -> (phase, tile of the phase).  Phase-major by default.  With
// p.interleave (multi-phase = transposed-conv grids of at least one tile per worker and phase) a worker runs the
// phases of ONE spatial tile back to back: the 9/6/6/4-tap phases then mix in time on every SM, so the epilogue of a
// short (4-tap) tile hides under the next long one instead of the whole grid becoming epilogue/store-bound for the
// last phases, and the input pixels the four phases share are read while they are hot.  Returns false for the
// padding items of the last round (same answer in every warp role, so the pipelines stay in step).
__device__ __forceinline__ bool work_item(const FwdKParams& p, int w, int worker, int nworkers, int per_phase,
                                          int ksplit, int& ph, int& idx) {
  if (p.interleave) {
    const int r = w / nworkers;
    ph = r % p.nphases;
    idx = worker + (r / p.nphases) * nworkers;
    return idx < per_phase;
  }
  const int t = w / ksplit;
  ph = t / per_phase;
  idx = t - ph * per_phase;
  return true;
}

// epilogue for 8 consecutive channels of one output pixel (bf16 lp tensors)
__device__ __forceinline__ void load8_bf16(const void* base, int64_t o, float* f) {
  const uint4 u = *reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(base) + o);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x; f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ void store8_bf16(void* base, int64_t o, const float* f) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(base) + o) = u;
}
__device__ __forceinline__ void load8_f32(const void* base, int64_t o, float* f) {
  const float4 a = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(base) + o);
  const float4 b = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(base) + o + 4);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
__device__ __forceinline__ void store8_f32(float* base, int64_t o, const float* f) {
  *reinterpret_cast<float4*>(base + o) = make_float4(f[0], f[1], f[2], f[3]);
  *reinterpret_cast<float4*>(base + o + 4) = make_float4(f[4], f[5], f[6], f[7]);
}

__device__ __forceinline__ void epilogue8(const EpiArgs& e, const float* acc, int64_t o, int c0) {
  float v[8], nrm[8];
  if (e.bias) {
    float b[8];
    load8_f32(e.bias, c0, b);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = acc[i] + b[i];
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = acc[i];
  }
  switch (e.epi) {
    case HPC_EPI_RELU:
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = fmaxf(v[i], 0.f);
      break;
    case HPC_EPI_EXP_CLAMP:
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = fminf(fmaxf(expf(v[i]), 1e-10f), 1e10f);
      break;
    case HPC_EPI_GDN: case HPC_EPI_IGDN: {
      float x[8];
      load8_bf16(e.aux0, o, x);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float r = rsqrtf(v[i]);
        nrm[i] = r;
        v[i] = (e.epi == HPC_EPI_GDN) ? x[i] * r : x[i] * (v[i] * r);
      }
      break;
    }
    case HPC_EPI_GDN_BWD: case HPC_EPI_IGDN_BWD: {
      float g[8], x[8], r[8];
      load8_bf16(e.aux0, o, g);
      load8_bf16(e.aux1, o, x);
      load8_bf16(e.aux2, o, r);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        v[i] = (e.epi == HPC_EPI_GDN_BWD) ? (g[i] * r[i] - x[i] * v[i]) : (g[i] / r[i] + x[i] * v[i]);
      break;
    }
    case HPC_EPI_RELU_MASK: {
      float x[8];
      load8_bf16(e.aux0, o, x);
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = x[i] > 0.f ? v[i] : 0.f;
      break;
    }
    case HPC_EPI_SIGN_ADD: {
      float y[8], a[8];
      load8_f32(e.aux0, o, y);
      load8_f32(e.aux1, o, a);
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = (y[i] > 0.f ? v[i] : (y[i] < 0.f ? -v[i] : 0.f)) + a[i];
      break;
    }
    default: break;
  }
  if (e.out_f32) store8_f32(e.out_f32, o, v);
  if (e.out_lp) store8_bf16(e.out_lp, o, v);
  if (e.out_sq_lp) {
    float s[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] = v[i] * v[i];
    store8_bf16(e.out_sq_lp, o, s);
  }
  if (e.out_abs_lp) {
    float s[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] = fabsf(v[i]);
    store8_bf16(e.out_abs_lp, o, s);
  }
  if (e.out_nrm_lp && (e.epi == HPC_EPI_GDN || e.epi == HPC_EPI_IGDN)) store8_bf16(e.out_nrm_lp, o, nrm);
}

template <int BLOCK_N, int NCTA, int MT, bool SHARE = false>
__device__ __forceinline__ void igemm_fwd_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const FwdKParams& p,
                                               const CUtensorMap* tmA1 = nullptr, const CUtensorMap* tmA2 = nullptr,
                                               const FwdShareSched* sch = nullptr) {
  using Cfg = FwdCfg<BLOCK_N, NCTA, MT, SHARE>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1 KB aligned; pointer arithmetic keeps the
                                                                                  // shared address space (LDS/STS, not generic LD/ST)
  if (Cfg::ALIGN_SLACK == 0 && smem != smem_raw) {         // no slack budgeted: the window itself must be aligned
    if (threadIdx.x == 0) printf("hpcodec: igemm_fwd (MT=2): dynamic shared memory window not 1 KB aligned\n");
    __trap();
  }
  // operand rings.  !SHARE: STAGES x (MT x 16 KB pixels | B_BYTES weights), one full/empty pair per stage.
  // SHARE: FS_A_SLOTS x MT x 20 KB pixels with their own barriers, then FS_B_SLOTS x B_BYTES weights with theirs.
  constexpr int NBAR_A = SHARE ? FS_A_SLOTS : Cfg::STAGES;          // "full_bar/empty_bar" = the activation ring when SHARE
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + (SHARE ? FS_A_SLOTS * MT * FS_A_TILE : Cfg::STAGES * MT * Cfg::A_BYTES);   // multiples of 1 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::OPERAND_BYTES);
  uint64_t* full_bar = bars;                    // [NBAR_A]  TMA -> MMA     (pair: the leader's collects both CTAs)
  uint64_t* empty_bar = bars + NBAR_A;          // [NBAR_A]  MMA -> TMA     (pair: commit multicast to both CTAs)
  uint64_t* tfull_bar = bars + 2 * NBAR_A;      // [2]       MMA -> epilogue (pair: multicast)
  uint64_t* tempty_bar = tfull_bar + 2;         // [2]       epilogue -> MMA (pair: both CTAs arrive at the leader's)
  uint64_t* bfull_bar = tempty_bar + 2;         // [FS_B_SLOTS] SHARE: weight ring
  uint64_t* bempty_bar = bfull_bar + FS_B_SLOTS;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bempty_bar + FS_B_SLOTS);
  float* bias_s = reinterpret_cast<float*>(smem + Cfg::OPERAND_BYTES + 256);   // bias staged once per CTA
  uint8_t* stg = smem + Cfg::OPERAND_BYTES + 256 + Cfg::BIAS_BYTES;             // output staging (pair kernel)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = NCTA == 2 ? (int)cluster_ctarank() : 0;
  const int worker = blockIdx.x / NCTA, nworkers = gridDim.x / NCTA;       // CTA (or CTA pair) index
  const int per_phase = (p.tiles_per_phase + NCTA * MT - 1) / (NCTA * MT);
  const int ksplit = NCTA == 1 ? p.ksplit : 1;              // split-K for grids of a few tiles (single-CTA kernel only)
  const int total_work = p.interleave ? (per_phase + nworkers - 1) / nworkers * nworkers * p.nphases
                                      : per_phase * p.nphases * ksplit;
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int s = 0; s < NBAR_A; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&tfull_bar[s], 1); mbar_init(&tempty_bar[s], 4 * NCTA); }
    if (SHARE) {
      prefetch_tmap(tmA1); prefetch_tmap(tmA2);
      for (int s = 0; s < FS_B_SLOTS; ++s) { mbar_init(&bfull_bar[s], 1); mbar_init(&bempty_bar[s], 1); }
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    if (NCTA == 2) tmem_alloc_pair(tmem_slot, Cfg::TMEM_COLS); else tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
  }
  // everything above overlaps the tail of the previous kernel in the stream (programmatic dependent launch)
  pdl_wait();
  if (p.epi.bias)
    for (int i = threadIdx.x; i < BLOCK_N; i += blockDim.x) bias_s[i] = p.epi.bias[i];
  tc_fence_before();
  if (NCTA == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  pdl_launch_dependents();
  const uint32_t tmem_base = *tmem_slot;

  if (SHARE && (warp == 0 || warp == 6)) {
    // ================================ TMA producers, shared activation boxes ================================
    // warp 0, lanes 0 / 1: the activation box of M tile 0 / 1, one per (tap group, channel slice);
    // warp 6, lane 0: the weight box of every (tap, channel slice), in the order the MMA warp consumes them
    const int nl = warp == 0 ? MT : 1;
    if (lane < nl) {
      int slot = 0; uint32_t phase = 0;
      const int box_row = p.tw * 128;
      for (int w = worker; w < total_work; w += nworkers) {
        int ph, idx, n0, i0, j0;
        if (!work_item(p, w, worker, nworkers, per_phase, ksplit, ph, idx)) continue;
        tile_coords<NCTA, MT>(p, ph, idx, rank, n0, i0, j0, lane);
        for (int g = sch->gbegin[ph]; g < sch->gbegin[ph] + sch->ngroups[ph]; ++g) {
          const int gs = sch->gsize[g], t0 = sch->order[sch->gtap0[g]];
          const CUtensorMap* tm = gs == 1 ? &tmA : (gs == 2 ? tmA1 : tmA2);
          const uint32_t a_bytes = (uint32_t)((p.th + gs - 1) * box_row);
          const int cw = j0 * p.is + p.dw[t0], chh = i0 * p.is + p.dh[t0];
          for (int c = 0; c < p.kchunks; ++c) {
            if (warp == 0) {
              mbar_wait(&empty_bar[slot], phase ^ 1);
              if (rank == 0 && lane == 0) mbar_arrive_expect_tx(&full_bar[slot], 2 * MT * a_bytes);
              __syncwarp((1u << MT) - 1);
              tma_load_4d_pair(smem_a + (slot * MT + lane) * FS_A_TILE, tm, mapa_u32(smem_u32(&full_bar[slot]), 0), c * FW_BK,
                               cw, chh, n0);
              if (++slot == FS_A_SLOTS) { slot = 0; phase ^= 1; }
            } else {
              for (int d = 0; d < gs; ++d) {
                const int t = sch->order[sch->gtap0[g] + d];
                mbar_wait(&bempty_bar[slot], phase ^ 1);
                if (rank == 0) mbar_arrive_expect_tx(&bfull_bar[slot], 2 * Cfg::B_BYTES);
                tma_load_2d_pair(smem_b + slot * Cfg::B_BYTES, &tmB, mapa_u32(smem_u32(&bfull_bar[slot]), 0), c * FW_BK,
                                 t * p.Co + rank * (BLOCK_N / 2));
                if (++slot == FS_B_SLOTS) { slot = 0; phase ^= 1; }
              }
            }
          }
        }
      }
    }
  } else if (SHARE && warp == 1) {
    // ================================ MMA issuer, shared activation boxes (leader CTA) ================================
    if (rank == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(FW_BM * NCTA, BLOCK_N, 0, 0);
      const uint64_t adesc0 = make_smem_desc_sw128(smem_u32(smem_a), 16, 1024);
      const uint64_t bdesc0 = make_smem_desc_sw128(smem_u32(smem_b), 16, 1024);
      const uint32_t row16 = (uint32_t)(p.tw * 128) >> 4;              // one tile row further into the box (16-byte units)
      int sa = 0, sb = 0; uint32_t pa = 0, pb = 0, acc_phase = 0;
      for (int w = worker; w < total_work; w += nworkers) {
        int ph, idx;
        if (!work_item(p, w, worker, nworkers, per_phase, ksplit, ph, idx)) continue;
        mbar_wait(&tempty_bar[0], acc_phase ^ 1);       // both accumulators belong to this work item
        mbar_wait(&tempty_bar[1], acc_phase ^ 1);
        tc_fence_after();
        uint32_t started = 0;
        for (int g = sch->gbegin[ph]; g < sch->gbegin[ph] + sch->ngroups[ph]; ++g) {
          const int gs = sch->gsize[g];
          for (int c = 0; c < p.kchunks; ++c) {
            mbar_wait(&full_bar[sa], pa);
            const uint64_t ad = adesc0 + (uint64_t)(sa * (MT * FS_A_TILE >> 4));
            for (int d = 0; d < gs; ++d) {
              mbar_wait(&bfull_bar[sb], pb);
              tc_fence_after();
              const uint64_t bd = bdesc0 + (uint64_t)(sb * (Cfg::B_BYTES >> 4));
              const uint64_t a0 = ad + (uint64_t)(d * row16);
              if (elect_one()) {
#pragma unroll
                for (int k = 0; k < FW_BK / 16; ++k) {
                  mma_bf16_ss_pair(tmem_base, a0 + 2 * k, bd + 2 * k, idesc, started | (uint32_t)k);
                  mma_bf16_ss_pair(tmem_base + Cfg::ACC_STRIDE, a0 + (FS_A_TILE >> 4) + 2 * k, bd + 2 * k, idesc,
                                   started | (uint32_t)k);
                }
                mma_commit_pair(&bempty_bar[sb]);                      // frees the weight slot (in both CTAs)
                if (d == gs - 1) mma_commit_pair(&empty_bar[sa]);      // ... and after the group's last tap the pixels
              }
              __syncwarp();
              started = 1;
              if (++sb == FS_B_SLOTS) { sb = 0; pb ^= 1; }
            }
            if (++sa == FS_A_SLOTS) { sa = 0; pa ^= 1; }
          }
        }
        if (elect_one()) { mma_commit_pair(&tfull_bar[0]); mma_commit_pair(&tfull_bar[1]); }
        __syncwarp();
        acc_phase ^= 1;
      }
    }
  } else if (warp == 0) {
    // ================================ TMA producer ================================
    // The whole warp runs the loop; lane 0 issues the activation box and lane 1 the weight box of a stage (a TMA
    // instruction costs its issuing thread ~100+ cycles, so the boxes of one stage are issued from different lanes).
    if (lane < 1 + MT) {
      constexpr uint32_t kIssuers = (1u << (1 + MT)) - 1;
      const int mt = lane == 2 ? 1 : 0;          // lane 0: pixels of M tile 0, lane 1: weights, lane 2: pixels of M tile 1
      int stage = 0; uint32_t phase = 0;
      for (int w = worker; w < total_work; w += nworkers) {
        int ph, idx, n0, i0, j0;
        if (!work_item(p, w, worker, nworkers, per_phase, ksplit, ph, idx)) continue;
        tile_coords<NCTA, MT>(p, ph, idx, rank, n0, i0, j0, mt);
        const hpc_igemm_phase P = p.phase[ph];
        const int ksteps = P.ntaps * p.kchunks, sp = w % ksplit;
        const int ks0 = ksteps * sp / ksplit, ks1 = ksteps * (sp + 1) / ksplit;
        for (int ks = ks0; ks < ks1; ++ks) {
          const int t = P.tap_begin + ks / p.kchunks, c = ks % p.kchunks;
          const int cw = j0 * p.is + p.dw[t], chh = i0 * p.is + p.dh[t];
          {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            if (NCTA == 1) {
              if (lane == 0) mbar_arrive_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
              __syncwarp(kIssuers);
              if (lane == 0) tma_load_4d(smem_a + stage * Cfg::A_BYTES, &tmA, &full_bar[stage], c * FW_BK, cw, chh, n0);
              else tma_load_2d(smem_b + stage * Cfg::B_BYTES, &tmB, &full_bar[stage], c * FW_BK, t * p.Co);
            } else {
              // each CTA loads its own pixels and its half of the output channels; all bytes complete on the leader
              if (rank == 0 && lane == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * Cfg::STAGE_BYTES);
              __syncwarp(kIssuers);
              const uint32_t lead = mapa_u32(smem_u32(&full_bar[stage]), 0);
              if (lane != 1) tma_load_4d_pair(smem_a + (stage * MT + mt) * Cfg::A_BYTES, &tmA, lead, c * FW_BK, cw, chh, n0);
              else tma_load_2d_pair(smem_b + stage * Cfg::B_BYTES, &tmB, lead, c * FW_BK, t * p.Co + rank * (BLOCK_N / 2));
            }
            if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer (pair: leader only) ==============
    // The whole warp runs the loop (convergent code keeps the descriptors in uniform registers); one elected lane
    // issues.  Descriptors advance by adding to their address field: 32 bytes (one K = 16 step) = 2 units of 16 B.
    if (rank == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(FW_BM * NCTA, BLOCK_N, 0, 0);
      const uint64_t adesc0 = make_smem_desc_sw128(smem_u32(smem_a), 16, 1024);
      const uint64_t bdesc0 = make_smem_desc_sw128(smem_u32(smem_b), 16, 1024);
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      for (int w = worker; w < total_work; w += nworkers) {
        int ph, idx;
        if (!work_item(p, w, worker, nworkers, per_phase, ksplit, ph, idx)) continue;
        const int kall = p.phase[ph].ntaps * p.kchunks, sp = w % ksplit;
        const int ksteps = kall * (sp + 1) / ksplit - kall * sp / ksplit;
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);     // epilogue(s) have drained this accumulator
        if (MT == 2) mbar_wait(&tempty_bar[1], acc_phase ^ 1);        // (acc == 0 here: both belong to this item)
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * Cfg::ACC_STRIDE;
        for (int ks = 0; ks < ksteps; ++ks) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint64_t ad = adesc0 + (uint64_t)(stage * (MT * Cfg::A_BYTES >> 4));
          const uint64_t bd = bdesc0 + (uint64_t)(stage * (Cfg::B_BYTES >> 4));
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < FW_BK / 16; ++k) {
              if (NCTA == 1) mma_bf16_ss(d_tmem, ad + 2 * k, bd + 2 * k, idesc, (ks | k) != 0);
              else mma_bf16_ss_pair(d_tmem, ad + 2 * k, bd + 2 * k, idesc, (ks | k) != 0);
              if (MT == 2)                     // second M tile of this CTA pair against the same weight half-tiles
                mma_bf16_ss_pair(d_tmem + Cfg::ACC_STRIDE, ad + (Cfg::A_BYTES >> 4) + 2 * k, bd + 2 * k, idesc, (ks | k) != 0);
            }
            // frees the smem stage (in both CTAs) when these MMAs retire
            if (NCTA == 1) mma_commit(&empty_bar[stage]); else mma_commit_pair(&empty_bar[stage]);
          }
          __syncwarp();
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
        }
        if (elect_one()) {
          if (NCTA == 1) mma_commit(&tfull_bar[acc]); else mma_commit_pair(&tfull_bar[acc]);   // accumulator complete
          if (MT == 2) mma_commit_pair(&tfull_bar[1]);
        }
        __syncwarp();
        if (MT == 2) acc_phase ^= 1;
        else if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp < 6) {
    // ================================ epilogue (4 warps) ==========================
    const int q = warp & 3;                 // TMEM lane quarter this warp may access
    const int row = q * 32 + lane;          // tile row == TMEM lane
    EpiArgs epi = p.epi;
    if (epi.bias) epi.bias = bias_s;        // broadcast shared-memory reads instead of 48 global loads per tile
    int acc = 0; uint32_t acc_phase = 0;
    for (int w = worker; w < total_work; w += nworkers) {
      int ph, idx, n0, i0, j0;
      if (!work_item(p, w, worker, nworkers, per_phase, ksplit, ph, idx)) continue;
      const hpc_igemm_phase P = p.phase[ph];
#pragma unroll 1
      for (int m = 0; m < MT; ++m) {            // MT == 2: accumulator m holds M tile m of this work item
      tile_coords<NCTA, MT>(p, ph, idx, rank, n0, i0, j0, m);
      const int jl = row % p.tw, il = (row / p.tw) % p.th, nl = row / (p.tw * p.th);
      const int n = n0 + nl, i = i0 + il, j = j0 + jl;
      const int oh = i * p.os + P.oh_off, ow = j * p.os + P.ow_off;
      const bool valid = (n < p.N) && (i < p.GH) && (j < p.GW) && (oh < p.OH) && (ow < p.OW);
      const int64_t obase = (((int64_t)n * p.OH + oh) * p.OW + ow) * p.Co;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * Cfg::ACC_STRIDE + ((uint32_t)(q * 32) << 16);
      if (NCTA == 2 && p.stage_out) {
        // a thread owns one output ROW (pixel): direct stores are 16-byte pieces of 32 different rows per instruction.
        // Stage the bf16 row in shared memory (pitch BLOCK_N*2 + 16 B: conflict-free) and write it with ONE bulk copy.
        uint8_t* srow = stg + row * Cfg::STG_PITCH;
        bulk_group_wait_read();                  // the previous tile's copy has finished reading this row
        static_assert(BLOCK_N % 32 == 0, "two 16-column loads per round");
#pragma unroll 1
        for (int c0 = 0; c0 < BLOCK_N; c0 += 32) {          // two TMEM loads in flight per wait
          float v[32];
          tmem_ld16(taddr + c0, v);
          tmem_ld16(taddr + c0 + 16, v + 16);
          tmem_ld_wait();
          if (epi.bias) {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] += epi.bias[c0 + i];
          }
          uint4 u[4];
          __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(u);
#pragma unroll
          for (int i = 0; i < 16; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
#pragma unroll
          for (int i = 0; i < 4; ++i) *reinterpret_cast<uint4*>(srow + c0 * 2 + 16 * i) = u[i];
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(mapa_u32(smem_u32(&tempty_bar[acc]), 0));    // accumulator drained
        fence_proxy_async();
        if (valid)
          bulk_store_row(reinterpret_cast<__nv_bfloat16*>(epi.out_lp) + obase, srow, (uint32_t)(BLOCK_N * 2));
        bulk_group_commit();
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        continue;
      }
#pragma unroll 1
      for (int c0 = 0; c0 < BLOCK_N; c0 += 16) {
        float v[16];
        tmem_ld16(taddr + c0, v);
        tmem_ld_wait();
        if (valid) {
          if (ksplit > 1) {           // raw partial of this K slice; bias / epilogue run in splitk_epilogue_kernel
            float* dst = p.ws + (int64_t)(w % ksplit) * p.out_elems;
            store8_f32(dst, obase + c0, v);
            store8_f32(dst, obase + c0 + 8, v + 8);
          } else {
            epilogue8(epi, v, obase + c0, c0);
            epilogue8(epi, v + 8, obase + c0 + 8, c0 + 8);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (NCTA == 1) mbar_arrive(&tempty_bar[acc]);
        else mbar_arrive_cluster(mapa_u32(smem_u32(&tempty_bar[acc]), 0));
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
    if (NCTA == 2 && p.stage_out) bulk_group_wait_all();
  }
  tc_fence_before();
  if (NCTA == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 1) {
    if (NCTA == 2) tmem_dealloc_pair(tmem_base, Cfg::TMEM_COLS); else tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

template <int BLOCK_N>
__global__ void __launch_bounds__(FW_THREADS, 1)
igemm_fwd_umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const __grid_constant__ FwdKParams p) {
  igemm_fwd_body<BLOCK_N, 1, 1>(tmA, tmB, p);
}

// CTA-pair version: cluster of two CTAs = the two SMs of a TPC, one tcgen05.mma.cta_group::2 of M = 256 per k-step.
// Each CTA stages only half of the weight tile, so a pipeline stage is 28 KB instead of 40 KB (C = 192): 7 stages in
// flight and 30% fewer operand bytes per MMA -- the single-CTA kernel is bound by bytes-in-flight x L2 latency.
template <int BLOCK_N>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(FW_THREADS, 1)
igemm_fwd_umma_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                           const __grid_constant__ FwdKParams p) {
  igemm_fwd_body<BLOCK_N, 2, 1>(tmA, tmB, p);
}

// CTA pair, two M tiles per CTA (FwdCfg): M = 512 pixels per work item against one pass over the weights.
template <int BLOCK_N>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(FW_THREADS, 1)
igemm_fwd_umma_pair2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                            const __grid_constant__ FwdKParams p) {
  igemm_fwd_body<BLOCK_N, 2, 2>(tmA, tmB, p);
}

// ... and activation boxes shared between the taps of a kernel column (FwdCfg SHARE): tmA / tmA1 / tmA2 = boxes of
// th, th + 1, th + 2 tile rows.
template <int BLOCK_N>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(FS_THREADS, 1)
igemm_fwd_umma_pair2s_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA1,
                             const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB,
                             const __grid_constant__ FwdKParams p, const __grid_constant__ FwdShareSched sch) {
  igemm_fwd_body<BLOCK_N, 2, 2, true>(tmA, tmB, p, &tmA1, &tmA2, &sch);
}

// Second pass of a split-K launch: sums the K slices of every output pixel in a fixed order and applies the bias /
// epilogue exactly like the in-kernel epilogue does (8 channels per thread).
__global__ void __launch_bounds__(256)
splitk_epilogue_kernel(const float* __restrict__ ws, int ksplit, int64_t out_elems, int Co, const __grid_constant__ EpiArgs epi) {
  const int64_t u = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;      // unit of 8 channels
  const int64_t o = u * 8;
  pdl_wait();
  pdl_launch_dependents();
  if (o >= out_elems) return;
  float acc[8];
  load8_f32(ws, o, acc);
  for (int s = 1; s < ksplit; ++s) {
    float v[8];
    load8_f32(ws + (int64_t)s * out_elems, o, v);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] += v[i];
  }
  epilogue8(epi, acc, o, (int)(o % Co));
}

// ---------------------------------------------------------------------------------------
// host side of the forward kernel
// ---------------------------------------------------------------------------------------

// tap groups of the shared-box kernel: per phase, taps sorted by (dw, dh); a group = up to 3 taps of one dw whose dh
// are consecutive steps of the input stride
static void fwd_share_schedule(const hpc_igemm_fwd_params* p, FwdShareSched* sch) {
  memset(sch, 0, sizeof(*sch));
  int ng = 0, no = 0;
  for (int ph = 0; ph < p->nphases; ++ph) {
    const hpc_igemm_phase& P = p->phase[ph];
    sch->gbegin[ph] = (uint8_t)ng;
    const int o0 = no;
    for (int k = 0; k < P.ntaps; ++k) {
      const int t = P.tap_begin + k;
      int q = no;
      while (q > o0 && (p->dw[sch->order[q - 1]] > p->dw[t] ||
                        (p->dw[sch->order[q - 1]] == p->dw[t] && p->dh[sch->order[q - 1]] > p->dh[t]))) {
        sch->order[q] = sch->order[q - 1];
        --q;
      }
      sch->order[q] = (uint8_t)t;
      ++no;
    }
    for (int k = o0; k < no;) {
      int e = k + 1;
      while (e < no && e - k < 3 && p->dw[sch->order[e]] == p->dw[sch->order[k]] &&
             p->dh[sch->order[e]] == p->dh[sch->order[e - 1]] + p->is)
        ++e;
      sch->gtap0[ng] = (uint8_t)k; sch->gsize[ng] = (uint8_t)(e - k);
      ++ng;
      k = e;
    }
    sch->ngroups[ph] = (uint8_t)(ng - sch->gbegin[ph]);
  }
}

template <int BLOCK_N>
static int launch_fwd_share(const CUtensorMap* tmA, const CUtensorMap& tmB, const FwdKParams& kp, const FwdShareSched& sch,
                            cudaStream_t st) {
  using Cfg = FwdCfg<BLOCK_N, 2, 2, true>;
  static HpcOncePerDevice attr_once;
  int attr_dev;
  if (attr_once.need(&attr_dev)) {
    HPC_CUDA(cudaFuncSetAttribute(igemm_fwd_umma_pair2s_kernel<BLOCK_N>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  Cfg::SMEM_BYTES));
    attr_once.mark(attr_dev);
  }
  const int items = (kp.tiles_per_phase + 3) / 4 * kp.nphases, tpcs = hpc_num_sms() / 2;
  const int grid = 2 * (items < tpcs ? items : tpcs);
  HPC_CUDA(hpc_launch_pdl(igemm_fwd_umma_pair2s_kernel<BLOCK_N>, dim3(grid), dim3(FS_THREADS), Cfg::SMEM_BYTES, st, tmA[0],
                          tmA[1], tmA[2], tmB, kp, sch));
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

template <int BLOCK_N>
static int launch_fwd(const CUtensorMap& tmA, const CUtensorMap& tmB, const FwdKParams& kp, bool pair, int mt, cudaStream_t st) {
  if (pair && mt == 2) {
    using Cfg = FwdCfg<BLOCK_N, 2, 2>;
    static HpcOncePerDevice attr_once;
    int attr_dev;
    if (attr_once.need(&attr_dev)) {
      HPC_CUDA(cudaFuncSetAttribute(igemm_fwd_umma_pair2_kernel<BLOCK_N>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    Cfg::SMEM_BYTES));
      attr_once.mark(attr_dev);
    }
    const int items = (kp.tiles_per_phase + 3) / 4 * kp.nphases, tpcs = hpc_num_sms() / 2;
    const int grid = 2 * (items < tpcs ? items : tpcs);
    HPC_CUDA(hpc_launch_pdl(igemm_fwd_umma_pair2_kernel<BLOCK_N>, dim3(grid), dim3(FW_THREADS), Cfg::SMEM_BYTES, st, tmA, tmB, kp));
  } else if (pair) {
    using Cfg = FwdCfg<BLOCK_N, 2>;
    static HpcOncePerDevice attr_once;
    int attr_dev;
    if (attr_once.need(&attr_dev)) {
      HPC_CUDA(cudaFuncSetAttribute(igemm_fwd_umma_pair_kernel<BLOCK_N>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    Cfg::SMEM_BYTES));
      attr_once.mark(attr_dev);
    }
    const int pairs = (kp.tiles_per_phase + 1) / 2 * kp.nphases, tpcs = hpc_num_sms() / 2;
    const int grid = 2 * (pairs < tpcs ? pairs : tpcs);
    HPC_CUDA(hpc_launch_pdl(igemm_fwd_umma_pair_kernel<BLOCK_N>, dim3(grid), dim3(FW_THREADS), Cfg::SMEM_BYTES, st, tmA, tmB, kp));
  } else {
    using Cfg = FwdCfg<BLOCK_N, 1>;
    static HpcOncePerDevice attr_once;
    int attr_dev;
    if (attr_once.need(&attr_dev)) {
      HPC_CUDA(cudaFuncSetAttribute(igemm_fwd_umma_kernel<BLOCK_N>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    Cfg::SMEM_BYTES));
      attr_once.mark(attr_dev);
    }
    const int work = kp.total_tiles * kp.ksplit;
    const int grid = work < hpc_num_sms() ? work : hpc_num_sms();
    HPC_CUDA(hpc_launch_pdl(igemm_fwd_umma_kernel<BLOCK_N>, dim3(grid), dim3(FW_THREADS), Cfg::SMEM_BYTES, st, tmA, tmB, kp));
  }
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

static int pow2_floor(int v) { int r = 1; while (r * 2 <= v) r *= 2; return r; }

// Split-K factor for grids of a few tiles (the h_a / h_s layers: 8-32 tiles of 27-75 k-steps each leave most SMs
// idle and serialise the K loop): as many slices as fill the SMs, at least 3 k-steps per slice.
static int fwd_ksplit(const hpc_igemm_fwd_params* p, int total_tiles) {
  static int mode = -1;                       // HPC_FWD_KSPLIT=0 disables (A/B measurements)
  if (mode < 0) { const char* e = getenv("HPC_FWD_KSPLIT"); mode = e ? atoi(e) : 1; }
  if (!mode || total_tiles * 2 > hpc_num_sms()) return 1;
  int min_ksteps = 1 << 30;
  for (int i = 0; i < p->nphases; ++i) {
    const int k = p->phase[i].ntaps * (p->Ci / FW_BK);
    if (k < min_ksteps) min_ksteps = k;
  }
  int ks = hpc_num_sms() / total_tiles;
  if (ks > min_ksteps / 3) ks = min_ksteps / 3;
  if (ks > 16) ks = 16;
  return ks < 2 ? 1 : ks;
}

static int fwd_total_tiles(const hpc_igemm_fwd_params* p) {
  int tw = pow2_floor(p->GW < 16 ? p->GW : 16);
  int th = pow2_floor(p->GH < FW_BM / tw ? p->GH : FW_BM / tw);
  int nb = FW_BM / (tw * th);
  return ((p->GW + tw - 1) / tw) * ((p->GH + th - 1) / th) * ((p->N + nb - 1) / nb) * p->nphases;
}

size_t hpc_igemm_fwd_umma_workspace_bytes(const hpc_igemm_fwd_params* p) {
  const int ks = fwd_ksplit(p, fwd_total_tiles(p));
  return ks > 1 ? (size_t)ks * p->N * p->OH * p->OW * p->Co * sizeof(float) : 0;
}

// CTA pairs pay off once every TPC has a pair of tiles; below that the single-CTA kernel keeps more SMs busy.
static bool fwd_use_pair(const FwdKParams& kp) {
  static int mode = -1;                       // HPC_FWD_PAIR=0 / 1 forces the choice (A/B measurements)
  if (mode < 0) { const char* e = getenv("HPC_FWD_PAIR"); mode = e ? (atoi(e) ? 1 : 0) : 2; }
  if (mode != 2) return mode == 1;
  return kp.total_tiles >= hpc_num_sms();
}

static int ntaps_total_all(const hpc_igemm_fwd_params* p) {
  int n = 0;
  for (int i = 0; i < p->nphases; ++i) n += p->phase[i].ntaps;
  return n;
}

// Everything the host decides about a forward launch: tile shape, engine build (single CTA / CTA pair / two M tiles per
// CTA / shared activation boxes), split-K, phase interleave, staged epilogue.  Shared by the launcher and by
// hpc_igemm_fwd_plan_describe (the CPU tests of the host logic).
struct FwdPlan { FwdKParams kp; bool pair, share; int mt, ntaps_total; };

static void fwd_plan(const hpc_igemm_fwd_params* p, FwdPlan* plan) {
  FwdKParams& kp = plan->kp;
  memset(&kp, 0, sizeof(kp));
  kp.N = p->N; kp.GH = p->GH; kp.GW = p->GW; kp.is = p->is; kp.os = p->os; kp.OH = p->OH; kp.OW = p->OW;
  kp.Co = p->Co; kp.Ci = p->Ci;
  kp.tw = pow2_floor(p->GW < 16 ? p->GW : 16);
  kp.th = pow2_floor(p->GH < FW_BM / kp.tw ? p->GH : FW_BM / kp.tw);
  kp.nb = FW_BM / (kp.tw * kp.th);
  kp.tiles_w = (p->GW + kp.tw - 1) / kp.tw;
  kp.tiles_h = (p->GH + kp.th - 1) / kp.th;
  kp.tiles_n = (p->N + kp.nb - 1) / kp.nb;
  kp.tiles_per_phase = kp.tiles_w * kp.tiles_h * kp.tiles_n;
  kp.nphases = p->nphases;
  kp.total_tiles = kp.tiles_per_phase * p->nphases;
  memcpy(kp.phase, p->phase, sizeof(kp.phase));
  memcpy(kp.dh, p->dh, sizeof(kp.dh));
  memcpy(kp.dw, p->dw, sizeof(kp.dw));
  kp.kchunks = p->Ci / FW_BK;
  kp.epi = make_epi_args(p);
  int ntaps_total = 0;
  for (int i = 0; i < p->nphases; ++i) {
    int e = p->phase[i].tap_begin + p->phase[i].ntaps;
    if (e > ntaps_total) ntaps_total = e;
  }

  const bool pair = fwd_use_pair(kp);
  // two M tiles per CTA (FwdCfg): where every CTA pair gets at least one group of four tiles.  HPC_FWD_MT=1 disables.
  int mt = 1;
  {
    static int mode = -1;
    if (mode < 0) { const char* e = getenv("HPC_FWD_MT"); mode = e ? atoi(e) : 2; }
    if (pair && mode == 2 && p->Co == 192 && (kp.tiles_per_phase + 3) / 4 * kp.nphases >= hpc_num_sms() / 2) mt = 2;
  }
  {  // staged bulk-copy epilogue: plain (bias) epilogue whose only output is the bf16 tensor; HPC_FWD_STAGE_OUT=0 disables
    static int want = -1;
    if (want < 0) { const char* e = getenv("HPC_FWD_STAGE_OUT"); want = (e && e[0] == '0') ? 0 : 1; }
    kp.stage_out = want && pair && kp.epi.epi == HPC_EPI_LINEAR && kp.epi.out_lp && !kp.epi.out_f32 && !kp.epi.out_sq_lp &&
                   !kp.epi.out_abs_lp && !kp.epi.out_nrm_lp && hpc_aligned16(kp.epi.out_lp) && (p->Co * 2) % 16 == 0;
  }
  kp.ksplit = 1;
  if (!pair) {
    const int ks = fwd_ksplit(p, kp.total_tiles);
    const size_t need = (size_t)ks * p->N * p->OH * p->OW * p->Co * sizeof(float);
    if (ks > 1 && p->workspace && p->workspace_bytes >= need && hpc_aligned16(p->workspace)) {
      kp.ksplit = ks; kp.ws = (float*)p->workspace; kp.out_elems = (int64_t)p->N * p->OH * p->OW * p->Co;
    }
  }
  {
    static int mode = -1;                        // HPC_FWD_INTERLEAVE=0: phase-major tile order everywhere (A/B measurements)
    if (mode < 0) { const char* e = getenv("HPC_FWD_INTERLEAVE"); mode = e ? atoi(e) : 1; }
    const int per = pair ? (kp.tiles_per_phase + 2 * mt - 1) / (2 * mt) : kp.tiles_per_phase;
    const int workers = pair ? hpc_num_sms() / 2 : hpc_num_sms();
    // whole rounds of (tile x all phases) per worker: only where that padding costs < 4 % (measured: 64x64 -> 128x128
    // transposed conv 463 -> 444 us, but 32x32 -> 64x64 with 3.46 rounds 120 -> 130 us)
    const int rounds = (per + workers - 1) / workers;
    kp.interleave = mode && kp.nphases > 1 && kp.ksplit == 1 && per >= workers &&
                    (int64_t)rounds * workers * 100 <= (int64_t)per * 104;
  }
  // activation boxes shared between the taps of a kernel column (FwdCfg SHARE): the big Co = 192 grids of whole-image
  // tiles.  OFF by default, HPC_FWD_SHARE=1 enables: measured (r02t, same box, back to back) 2.310 ms/step for the
  // forward class with it and 2.271 without -- with two M tiles per CTA the kernel is no longer bound by L2 -> SM bytes
  // (0.84 of the sustained tensor peak), so 38 % fewer of them buy nothing and the second ring costs a little.
  bool share = false;
  {
    static int mode = -1;
    if (mode < 0) { const char* e = getenv("HPC_FWD_SHARE"); mode = (e && e[0] == '1') ? 1 : 0; }
    share = mode && mt == 2 && kp.nb == 1 && kp.tw % 8 == 0 && (kp.th + 2) * kp.tw * 128 <= FS_A_TILE;
  }
  plan->pair = pair; plan->mt = mt; plan->share = share; plan->ntaps_total = ntaps_total;
}

int hpc_igemm_fwd_umma(const hpc_igemm_fwd_params* p, hpc_stream_t stream) {
  HPC_REQUIRE(p->dtype == HPC_BF16, HPC_ERR_UNSUPPORTED, "umma engine: operands must be bf16");
  HPC_REQUIRE(p->Ci % FW_BK == 0, HPC_ERR_UNSUPPORTED, "umma engine: Ci=%d must be a multiple of 64", p->Ci);
  HPC_REQUIRE(p->Co == 64 || p->Co == 128 || p->Co == 192 || p->Co == 256, HPC_ERR_UNSUPPORTED,
              "umma engine: Co=%d not in {64,128,192,256}", p->Co);
  HPC_REQUIRE(p->is >= 1 && p->is <= 8, HPC_ERR_UNSUPPORTED, "umma engine: input stride %d not in [1,8]", p->is);
  HPC_REQUIRE(hpc_aligned16(p->x) && hpc_aligned16(p->w), HPC_ERR_BAD_ALIGNMENT, "umma engine: x/w must be 16B aligned");
  const void* outs[] = {p->out_f32, p->out_lp, p->out_sq_lp, p->out_abs_lp, p->out_nrm_lp, p->aux0, p->aux1, p->aux2,
                        p->bias};
  for (const void* q : outs) HPC_REQUIRE(hpc_aligned16(q), HPC_ERR_BAD_ALIGNMENT, "umma engine: epilogue pointers must be 16B aligned");

  FwdPlan plan;
  fwd_plan(p, &plan);
  FwdKParams& kp = plan.kp;
  const bool pair = plan.pair, share = plan.share;
  const int mt = plan.mt, ntaps_total = plan.ntaps_total;
  CUtensorMap tmAs[3], tmB;
  CUtensorMap& tmA = tmAs[0];
  for (int e = 0; e < (share ? 3 : 1); ++e) {      // boxes of th (+1, +2) tile rows
    uint64_t dims[4] = {(uint64_t)p->Ci, (uint64_t)p->W, (uint64_t)p->H, (uint64_t)p->N};
    uint64_t strides[3] = {(uint64_t)p->Ci * 2, (uint64_t)p->W * p->Ci * 2, (uint64_t)p->H * p->W * p->Ci * 2};
    uint32_t box[4] = {(uint32_t)FW_BK, (uint32_t)(kp.tw * p->is), (uint32_t)((kp.th + e) * p->is), (uint32_t)kp.nb};
    uint32_t estr[4] = {1, (uint32_t)p->is, (uint32_t)p->is, 1};
    int rc = make_tmap(&tmAs[e], p->x, 4, dims, strides, box, estr);
    if (rc) return rc;
  }
  {
    uint64_t dims[2] = {(uint64_t)p->Ci, (uint64_t)ntaps_total * p->Co};
    uint64_t strides[1] = {(uint64_t)p->Ci * 2};
    uint32_t box[2] = {(uint32_t)FW_BK, (uint32_t)(pair ? p->Co / 2 : p->Co)};
    uint32_t estr[2] = {1, 1};
    int rc = make_tmap(&tmB, p->w, 2, dims, strides, box, estr);
    if (rc) return rc;
  }
  auto st = (cudaStream_t)stream;
  // algorithmic flops: every grid pixel sees every tap of its phase once (zero taps of the
  // stride-2 transposed conv are never counted because the phases only hold the live taps)
  HpcProfScope prof(HPC_PROF_IGEMM_FWD,
                    2.0 * p->N * p->GH * p->GW * (double)ntaps_total_all(p) * p->Ci * p->Co, stream);
  int rc;
  if (share) {
    FwdShareSched sch;
    fwd_share_schedule(p, &sch);
    return launch_fwd_share<192>(tmAs, tmB, kp, sch, st);
  }
  switch (p->Co) {
    case 64: rc = launch_fwd<64>(tmA, tmB, kp, pair, 1, st); break;
    case 128: rc = launch_fwd<128>(tmA, tmB, kp, pair, 1, st); break;
    case 192: rc = launch_fwd<192>(tmA, tmB, kp, pair, mt, st); break;
    default: rc = launch_fwd<256>(tmA, tmB, kp, pair, 1, st); break;
  }
  if (rc || kp.ksplit == 1) return rc;
  const int64_t units = kp.out_elems / 8;
  HPC_CUDA(hpc_launch_pdl(splitk_epilogue_kernel, dim3((unsigned)((units + 255) / 256)), dim3(256), 0, st, (const float*)kp.ws,
                          kp.ksplit, kp.out_elems, p->Co, kp.epi));
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}

// Host-side introspection of the forward engine's plan for `p` (no GPU work): out = {pair, M tiles per CTA, shared boxes,
// ksplit, interleave, staged epilogue, tw, th, nb, tiles_per_phase, total_tiles, grid}.  Returns 12 or -HPC_ERR_*.
int hpc_igemm_fwd_plan_describe(const hpc_igemm_fwd_params* p, int* out, int cap) {
  HPC_REQUIRE(p && out && cap >= 12, HPC_ERR_BAD_SHAPE, "hpc_igemm_fwd_plan_describe: needs room for 12 ints");
  if (p->dtype != HPC_BF16 || p->Ci % FW_BK != 0 || !(p->Co == 64 || p->Co == 128 || p->Co == 192 || p->Co == 256)) return -HPC_ERR_UNSUPPORTED;
  FwdPlan plan;
  fwd_plan(p, &plan);
  const FwdKParams& kp = plan.kp;
  int grid;
  if (plan.pair) {
    const int items = (kp.tiles_per_phase + 2 * plan.mt - 1) / (2 * plan.mt) * kp.nphases, tpcs = hpc_num_sms() / 2;
    grid = 2 * (items < tpcs ? items : tpcs);
  } else {
    const int work = kp.total_tiles * kp.ksplit;
    grid = work < hpc_num_sms() ? work : hpc_num_sms();
  }
  const int v[12] = {plan.pair, plan.mt, plan.share, kp.ksplit, kp.interleave, kp.stage_out, kp.tw, kp.th, kp.nb,
                     kp.tiles_per_phase, kp.total_tiles, grid};
  memcpy(out, v, sizeof(v));
  return 12;
}

// ---------------------------------------------------------------------------------------
// weight gradient kernel ("wgrad"): gw[t][r][c] = sum_pixels lo[p+d_t][r] * hi[p'][c]
//
// GEMM view per phase: M = (tap, r) rows, N = c columns, K = pixels.  Both operands come straight from
// the NHWC tensors as MN-major SWIZZLE_128B tiles ([64 pixels] x [64 channels] per TMA box = one UMMA
// "atom"); shifting `lo` by the tap offset is a TMA coordinate, so no im2col is ever materialised.
// A CTA owns one "unit" of 4 consecutive atoms of the phase's (tap, r-chunk) list = two 128-row
// accumulators in TMEM (2 x Cc fp32 columns) and one slice of the pixel range (split-K); it writes its
// partial to the workspace and a second kernel reduces the slices in a fixed order (deterministic).
// ---------------------------------------------------------------------------------------
constexpr int WG_THREADS = 192;
constexpr int WG_P = 64;                       // pixels per stage (K block)
constexpr int WG_ATOM_BYTES = WG_P * 128;      // 8 KB: 64 pixels x 64 bf16
constexpr int WG_MAX_STAGES = 6;               // C = 192: 5 x 42 KB with shared `lo` boxes, 4 x 56 KB without
constexpr int WG_MAX_UNITS = 32;

// A unit = up to 4 atoms = up to four (tap, 64-row r-chunk) blocks of the result = two 128-row accumulators.
// Its `lo` operand arrives in up to 4 TMA boxes.  ncu (r02p, 64x64x64-pixel layer): 4.53 GB cross the L2 -> SM
// crossbar in 400 us = 11.3 TB/s, the chip's L2 output cap, with the tensor pipe 68 % busy: the kernel is bound by
// operand bytes.  Taps of one phase that differ only in dh read the SAME pixels of `lo` one 8-pixel row apart, and
// one row of an 8 x 8 K block is exactly one 1 KB swizzle group of the MN-major operand: atoms of the same
// (dw, r-chunk) therefore share ONE box of 8 + (dh_max - dh_min) rows and their descriptors start 1 KB apart.
// 4 atoms = 18 KB of `lo` instead of 32 KB per stage (42 KB instead of 56 KB with the 24 KB of `hi`).
struct WgUnit {
  uint8_t phase, natoms, nloads, pad;
  uint8_t tap[4], rc[4];          // atom a: result block (tap[a], rows rc[a]*64 ..)
  uint8_t aoff[4];                // atom a starts aoff[a] KB into the stage's `lo` area
  uint8_t lrc[4], lrows[4], loff[4];   // box l: r-chunk, pixel rows (th .. th+2), KB offset in the `lo` area
  int8_t ldw[4], ldh[4];          // box l: its tap shift (dw, smallest dh)
};

struct WgKParams {
  int N, GH, GW, R, HH, HW, Cc, hs;
  int tw, th, nb, tiles_w, tiles_h, tiles_n, kblocks;
  int nunits, nsplits, rchunks, cchunks, ntaps;
  int stages;
  int a_bytes;                                 // `lo` area of a stage (largest unit)
  int epi_smem;                                // stage the accumulators in shared memory and bulk-copy whole rows out
  hpc_igemm_phase phase[HPC_MAX_PHASES];
  int dh[HPC_MAX_TAPS], dw[HPC_MAX_TAPS];
  WgUnit unit[WG_MAX_UNITS];
  float* ws;                                   // [nsplits][ntaps][R][Cc]
};

__global__ void __launch_bounds__(WG_THREADS, 1)
igemm_wgrad_umma_kernel(const __grid_constant__ CUtensorMap tmLo, const __grid_constant__ CUtensorMap tmLo1,
                        const __grid_constant__ CUtensorMap tmLo2, const __grid_constant__ CUtensorMap tmHi,
                        const __grid_constant__ WgKParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1 KB aligned; pointer arithmetic keeps the
                                                                                  // shared address space (LDS/STS, not generic LD/ST)
  const int stage_bytes = p.a_bytes + p.cchunks * WG_ATOM_BYTES;
  const int nstages = p.stages;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + nstages * stage_bytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + WG_MAX_STAGES;
  uint64_t* done_bar = bars + 2 * WG_MAX_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int unit_id = blockIdx.x % p.nunits, split = blockIdx.x / p.nunits;
  const WgUnit U = p.unit[unit_id];
  const hpc_igemm_phase P = p.phase[U.phase];
  const int kb0 = (int)(((int64_t)split * p.kblocks) / p.nsplits);
  const int kb1 = (int)(((int64_t)(split + 1) * p.kblocks) / p.nsplits);
  const int nkb = kb1 - kb0;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmLo); prefetch_tmap(&tmLo1); prefetch_tmap(&tmLo2);
    prefetch_tmap(&tmHi);
    for (int s = 0; s < nstages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(done_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                      // the prologue above overlaps the previous kernel's tail
  pdl_launch_dependents();

  if (warp == 0) {
    // Producer: the whole warp runs the loop and lane l issues box l of the stage (up to 4 `lo` boxes + Cc/64 `hi`
    // atoms), so the TMA instructions of a stage are issued in parallel instead of one after another by one thread.
    int stage = 0; uint32_t phase = 0;
    const bool is_lo = lane < U.nloads, is_hi = lane >= 4 && lane < 4 + p.cchunks;
    uint32_t tx = (uint32_t)p.cchunks * WG_ATOM_BYTES;
    for (int l = 0; l < U.nloads; ++l) tx += (uint32_t)U.lrows[l] * (uint32_t)(p.tw * p.nb) * 128u;
    const int ll = is_lo ? lane : 0;
    const int rc = U.lrc[ll], dwt = U.ldw[ll], dht = U.ldh[ll], extra = U.lrows[ll] - p.th;
    const CUtensorMap* tml = extra == 0 ? &tmLo : (extra == 1 ? &tmLo1 : &tmLo2);
    const int lo_off = U.loff[ll] * 1024;
    for (int kb = kb0; kb < kb1; ++kb) {
      int r = kb;
      const int tj = r % p.tiles_w; r /= p.tiles_w;
      const int ti = r % p.tiles_h; r /= p.tiles_h;
      const int n0 = r * p.nb, i0 = ti * p.th, j0 = tj * p.tw;
      mbar_wait(&empty_bar[stage], phase ^ 1);
      if (lane == 0) mbar_arrive_expect_tx(&full_bar[stage], tx);
      __syncwarp();
      uint8_t* sa = smem + stage * stage_bytes;
      if (is_lo)
        tma_load_4d(sa + lo_off, tml, &full_bar[stage], rc * 64, j0 + dwt, i0 + dht, n0);
      else if (is_hi)
        tma_load_4d(sa + p.a_bytes + (lane - 4) * WG_ATOM_BYTES, &tmHi, &full_bar[stage], (lane - 4) * 64,
                    j0 * p.hs + P.ow_off, i0 * p.hs + P.oh_off, n0);
      if (++stage == nstages) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(128, p.Cc, 1, 1);   // both operands MN-major
      const int nmma = (U.natoms + 1) / 2;
      uint32_t a0off[2], a1off[2];
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        a0off[m] = (uint32_t)U.aoff[2 * m] * 1024u;
        a1off[m] = 2 * m + 1 < U.natoms ? (uint32_t)(U.aoff[2 * m + 1] - U.aoff[2 * m]) * 1024u : 1024u;
      }
      int stage = 0; uint32_t phase = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + stage * stage_bytes);
        const uint32_t b_addr = a_addr + p.a_bytes;
#pragma unroll
        for (int ks = 0; ks < WG_P / 16; ++ks) {
          const uint64_t bd = make_smem_desc_sw128(b_addr + ks * 2048, WG_ATOM_BYTES, 1024);
#pragma unroll
          for (int m = 0; m < 2; ++m) {
            // rows 0..63 of the accumulator = atom 2m, rows 64..127 = atom 2m+1 (leading byte offset = their distance;
            // an odd unit's missing atom reads one row group further: its rows are never stored)
            if (m < nmma) {
              const uint64_t ad = make_smem_desc_sw128(a_addr + a0off[m] + ks * 2048, a1off[m], 1024);
              mma_bf16_ss(tmem_base + m * 256, ad, bd, idesc, (kb | ks) != 0);
            }
          }
        }
        mma_commit(&empty_bar[stage]);
        if (++stage == nstages) { stage = 0; phase ^= 1; }
      }
      mma_commit(done_bar);
    }
  } else if (p.epi_smem) {
    // Epilogue through shared memory: a thread owns one accumulator ROW, so direct global stores would be 16-byte
    // pieces of 32 different rows per instruction.  The operand stages are free once the last MMA has retired:
    // the rows are staged there (pitch Cc*4 + 16 bytes: conflict-free float4 writes) and every row leaves as ONE
    // bulk copy of Cc*4 contiguous bytes into the split-K workspace.
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int pitch = p.Cc * 4 + 16;
    mbar_wait(done_bar, 0);
    tc_fence_after();
    const int nmma = (U.natoms + 1) / 2;
    for (int m = 0; m < nmma; ++m) {
      uint8_t* srow = smem + (size_t)(m * 128 + row) * pitch;
      const uint32_t taddr = tmem_base + m * 256 + ((uint32_t)(q * 32) << 16);
      for (int c0 = 0; c0 < p.Cc; c0 += 16) {
        float v[16];
        if (nkb > 0) {
          tmem_ld16(taddr + c0, v);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = 0.f;
        }
        float4* d = reinterpret_cast<float4*>(srow + c0 * 4);
#pragma unroll
        for (int i = 0; i < 4; ++i) d[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
      }
    }
    fence_proxy_async();                    // generic-proxy writes -> visible to the bulk-copy (async proxy) reads
    // each thread copies out the rows it wrote itself: no cross-thread hand-off is needed
    for (int m = 0; m < nmma; ++m) {
      const int a = 2 * m + row / 64;
      if (a < U.natoms) {
        const int t = U.tap[a], r = U.rc[a] * 64 + (row & 63);
        float* dst = p.ws + (((int64_t)split * p.ntaps + t) * p.R + r) * p.Cc;
        bulk_store_row(dst, smem + (size_t)(m * 128 + row) * pitch, (uint32_t)(p.Cc * 4));
      }
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    mbar_wait(done_bar, 0);
    tc_fence_after();
    const int nmma = (U.natoms + 1) / 2;
    for (int m = 0; m < nmma; ++m) {
      const int a = 2 * m + row / 64;
      const bool valid = a < U.natoms;
      const int t = valid ? U.tap[a] : 0, r = (valid ? U.rc[a] : 0) * 64 + (row & 63);
      float* dst = p.ws + (((int64_t)split * p.ntaps + t) * p.R + r) * p.Cc;
      const uint32_t taddr = tmem_base + m * 256 + ((uint32_t)(q * 32) << 16);
      for (int c0 = 0; c0 < p.Cc; c0 += 16) {
        float v[16];
        if (nkb > 0) {
          tmem_ld16(taddr + c0, v);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = 0.f;
        }
        if (valid) {
          store8_f32(dst, c0, v);
          store8_f32(dst, c0 + 8, v + 8);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// Split-K reduction + output layout in one pass.  A block owns 128 consecutive floats of the packed
// [tap][r][c] result: 32 float4 columns x 8 split lanes; lane g sums splits g, g+8, ... in order and the
// 8 partials are combined in a fixed order (deterministic).  The result goes either to the packed layout
// or straight into the reference weight layout dst[r*sr + c*sc + tap_src[t]] (optionally accumulating).
struct WgOut { int mode, accumulate, R, Cc; int64_t sr, sc; int tap_src[HPC_MAX_TAPS]; };

__global__ void __launch_bounds__(256)
wgrad_reduce_kernel(const float* __restrict__ ws, int nsplits, int64_t n, float scale, float* __restrict__ gw,
                    const __grid_constant__ WgOut o) {
  __shared__ float4 part[8][32];
  const int q = threadIdx.x & 31, g = threadIdx.x >> 5;
  const int64_t i = ((int64_t)blockIdx.x * 32 + q) * 4;
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
  pdl_wait();
  pdl_launch_dependents();
  if (i < n)
    for (int s = g; s < nsplits; s += 8) {
      const float4 v = __ldcs(reinterpret_cast<const float4*>(ws + (int64_t)s * n + i));
      a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
  part[g][q] = a;
  __syncthreads();
  if (g != 0 || i >= n) return;
#pragma unroll
  for (int k = 1; k < 8; ++k) {
    const float4 v = part[k][q];
    a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
  }
  a.x *= scale; a.y *= scale; a.z *= scale; a.w *= scale;
  if (o.mode == 0) {
    float4* d = reinterpret_cast<float4*>(gw + i);
    if (o.accumulate) { const float4 c = *d; a.x += c.x; a.y += c.y; a.z += c.z; a.w += c.w; }
    *d = a;
  } else {
    const int c = (int)(i % o.Cc);
    const int64_t tr = i / o.Cc;
    const int r = (int)(tr % o.R), t = (int)(tr / o.R);
    float* d = gw + r * o.sr + c * o.sc + o.tap_src[t];
    const float v[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) d[k * o.sc] = o.accumulate ? d[k * o.sc] + v[k] : v[k];
  }
}

// Split-K reduction straight into the reference weight layout (out_mode 1) when the taps of one (r, c) pair are
// contiguous there (min(sr, sc) == ntaps: (Co,Ci,kh,kw) / (Ci,Co,kh,kw)).  A block owns an RT x CT tile of (r, c)
// pairs for ALL taps: it sums the K slices in split order (fixed order: deterministic) with 128/32-byte coalesced
// reads, transposes through shared memory (pair-major, odd stride) and writes runs of ntaps consecutive floats --
// the scattered 4-byte stores of wgrad_reduce_kernel cost 8x their bytes in 32-byte sectors.
constexpr int WGR_PAIRS = 128;
__global__ void __launch_bounds__(256)
wgrad_reduce_tiled_kernel(const float* __restrict__ ws, int nsplits, int ntaps, float scale, float* __restrict__ gw,
                          int RT, int CT, const __grid_constant__ WgOut o) {
  __shared__ float sm[WGR_PAIRS * (HPC_MAX_TAPS | 1)];
  __shared__ int inv[HPC_MAX_TAPS];
  const int stride = ntaps | 1;                      // odd: conflict-free reads in the write phase
  const int ctiles = o.Cc / CT;
  const int r0 = (blockIdx.x / ctiles) * RT, c0 = (blockIdx.x % ctiles) * CT;
  if (threadIdx.x < ntaps) inv[o.tap_src[threadIdx.x]] = threadIdx.x;
  pdl_wait();
  pdl_launch_dependents();
  const int q4 = CT / 4;                             // float4 units per tile row
  const int64_t n = (int64_t)ntaps * o.R * o.Cc;
  for (int u = threadIdx.x; u < ntaps * RT * q4; u += 256) {
    const int cq = u % q4, rl = (u / q4) % RT, t = u / (q4 * RT);
    const float* src = ws + ((int64_t)t * o.R + r0 + rl) * o.Cc + c0 + cq * 4;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int s2 = 0; s2 < nsplits; ++s2) {
      const float4 v = __ldcs(reinterpret_cast<const float4*>(src + (int64_t)s2 * n));
      a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
    float* d = sm + (rl * CT + cq * 4) * stride + t;
    d[0] = a.x * scale; d[stride] = a.y * scale; d[2 * stride] = a.z * scale; d[3 * stride] = a.w * scale;
  }
  __syncthreads();
  // pairs ordered so that consecutive pairs are neighbours along the small-stride dimension
  const bool c_inner = o.sc < o.sr;
  for (int e = threadIdx.x; e < RT * CT * ntaps; e += 256) {
    const int k = e % ntaps, pr = e / ntaps;
    const int rl = c_inner ? pr / CT : pr % RT, cl = c_inner ? pr % CT : pr / RT;
    float* d = gw + (int64_t)(r0 + rl) * o.sr + (int64_t)(c0 + cl) * o.sc + k;
    const float v = sm[(rl * CT + cl) * stride + inv[k]];
    *d = o.accumulate ? *d + v : v;
  }
}

static int wgrad_plan(const hpc_igemm_wgrad_params* p, WgKParams* kp) {
  memset(kp, 0, sizeof(*kp));
  kp->N = p->N; kp->GH = p->GH; kp->GW = p->GW; kp->R = p->R; kp->HH = p->HH; kp->HW = p->HW; kp->Cc = p->Cc;
  kp->hs = p->hs;
  kp->tw = pow2_floor(p->GW < 8 ? p->GW : 8);
  kp->th = pow2_floor(p->GH < WG_P / kp->tw ? p->GH : WG_P / kp->tw);
  kp->nb = WG_P / (kp->tw * kp->th);
  kp->tiles_w = (p->GW + kp->tw - 1) / kp->tw;
  kp->tiles_h = (p->GH + kp->th - 1) / kp->th;
  kp->tiles_n = (p->N + kp->nb - 1) / kp->nb;
  kp->kblocks = kp->tiles_w * kp->tiles_h * kp->tiles_n;
  kp->rchunks = p->R / 64; kp->cchunks = p->Cc / 64;
  memcpy(kp->phase, p->phase, sizeof(kp->phase));
  memcpy(kp->dh, p->dh, sizeof(kp->dh));
  memcpy(kp->dw, p->dw, sizeof(kp->dw));
  // `lo` boxes shared between the taps of one (dw, r-chunk) that differ in dh (WgUnit): needs one 8-pixel row of the
  // K block per 1 KB swizzle group.  HPC_WG_SHARE=0 keeps one box per atom (A/B measurements).
  static int share_mode = -1;
  if (share_mode < 0) { const char* e = getenv("HPC_WG_SHARE"); share_mode = (e && e[0] == '0') ? 0 : 1; }
  const bool share = share_mode && kp->nb == 1 && kp->tw == 8 && kp->th == 8;
  const int box_kb = kp->tw * kp->th * kp->nb * 128 / 1024;        // 8: a 64-pixel box
  int nu = 0, nt = 0, amax = 0;
  for (int ph = 0; ph < p->nphases; ++ph) {
    const hpc_igemm_phase& P = p->phase[ph];
    HPC_REQUIRE(P.ntaps >= 1 && P.tap_begin >= 0 && P.tap_begin + P.ntaps <= HPC_MAX_TAPS, HPC_ERR_BAD_SHAPE,
                "umma wgrad: phase %d taps [%d, %d)", ph, P.tap_begin, P.tap_begin + P.ntaps);
    int order[HPC_MAX_TAPS];                      // taps of the phase sorted by (dw, dh)
    for (int k = 0; k < P.ntaps; ++k) {
      const int t = P.tap_begin + k;
      int q = k;
      while (q > 0 && (p->dw[order[q - 1]] > p->dw[t] || (p->dw[order[q - 1]] == p->dw[t] && p->dh[order[q - 1]] > p->dh[t]))) {
        order[q] = order[q - 1];
        --q;
      }
      order[q] = t;
    }
    // atoms in (dw, r-chunk, dh) order, four per unit
    int cur = -1, used = 0;                        // unit being filled, KB of its `lo` area in use
    for (int g0 = 0; g0 < P.ntaps;) {
      int g1 = g0;
      while (g1 < P.ntaps && p->dw[order[g1]] == p->dw[order[g0]]) ++g1;
      for (int rc = 0; rc < kp->rchunks; ++rc)
        for (int k = g0; k < g1; ++k) {
          const int t = order[k];
          if (cur < 0 || kp->unit[cur].natoms == 4) {
            HPC_REQUIRE(nu < WG_MAX_UNITS, HPC_ERR_UNSUPPORTED, "umma wgrad: too many work units");
            cur = nu++;
            kp->unit[cur].phase = (uint8_t)ph;
            used = 0;
          }
          WgUnit& U = kp->unit[cur];
          const int a = U.natoms++;
          U.tap[a] = (uint8_t)t; U.rc[a] = (uint8_t)rc;
          const int l = U.nloads - 1;
          if (share && l >= 0 && U.lrc[l] == rc && U.ldw[l] == p->dw[t] && U.lrows[l] < kp->th + 2 &&
              p->dh[t] == U.ldh[l] + (U.lrows[l] - kp->th) + 1) {
            U.aoff[a] = (uint8_t)(U.loff[l] + (p->dh[t] - U.ldh[l]));      // one pixel row = 1 KB further into the same box
            ++U.lrows[l];
            used += 1;
          } else {
            const int n = U.nloads++;
            U.lrc[n] = (uint8_t)rc; U.ldw[n] = (int8_t)p->dw[t]; U.ldh[n] = (int8_t)p->dh[t];
            U.lrows[n] = (uint8_t)kp->th; U.loff[n] = (uint8_t)used; U.aoff[a] = (uint8_t)used;
            used += box_kb;
          }
          if (used > amax) amax = used;
        }
      g0 = g1;
    }
    nt += P.ntaps;
  }
  kp->a_bytes = amax * 1024;
  kp->nunits = nu; kp->ntaps = nt;
  int ns = hpc_num_sms() / nu;
  if (ns < 1) ns = 1;
  if (ns > kp->kblocks) ns = kp->kblocks;
  kp->nsplits = ns;
  return HPC_OK;
}

static int wgrad_check(const hpc_igemm_wgrad_params* p) {
  HPC_REQUIRE(p->dtype == HPC_BF16, HPC_ERR_UNSUPPORTED, "umma wgrad: operands must be bf16");
  HPC_REQUIRE(p->R % 64 == 0 && p->R <= 256 * 4, HPC_ERR_UNSUPPORTED, "umma wgrad: R=%d must be a multiple of 64", p->R);
  HPC_REQUIRE(p->Cc % 64 == 0 && p->Cc >= 64 && p->Cc <= 256, HPC_ERR_UNSUPPORTED,
              "umma wgrad: Cc=%d must be a multiple of 64 in [64,256]", p->Cc);
  HPC_REQUIRE(p->hs >= 1 && p->hs <= 8, HPC_ERR_UNSUPPORTED, "umma wgrad: hs=%d not in [1,8]", p->hs);
  HPC_REQUIRE(hpc_aligned16(p->lo) && hpc_aligned16(p->hi) && (p->out_mode == 1 || hpc_aligned16(p->gw_f32)),
              HPC_ERR_BAD_ALIGNMENT, "umma wgrad: pointers must be 16B aligned");
  HPC_REQUIRE(p->out_mode == 0 || p->out_mode == 1, HPC_ERR_BAD_SHAPE, "umma wgrad: out_mode %d", p->out_mode);
  return HPC_OK;
}

// Host-side introspection of the work-unit plan (no GPU work; used by the CPU tests of the host logic): writes
// {nunits, nsplits, a_bytes, kblocks} and then, per unit, 35 ints: phase, natoms, nloads, 4 x (tap, r-chunk, offset KB)
// for the atoms and 4 x (r-chunk, dw, dh, rows, offset KB) for the `lo` boxes (-1 = unused slot).
int hpc_igemm_wgrad_plan_describe(const hpc_igemm_wgrad_params* p, int* out, int cap) {
  HPC_REQUIRE(p && out, HPC_ERR_BAD_SHAPE, "hpc_igemm_wgrad_plan_describe: NULL pointer");
  int rc = wgrad_check(p);
  if (rc) return -rc;
  WgKParams kp;
  if ((rc = wgrad_plan(p, &kp))) return -rc;
  const int need = 4 + 35 * kp.nunits;
  HPC_REQUIRE(cap >= need, HPC_ERR_BAD_SHAPE, "hpc_igemm_wgrad_plan_describe: %d ints needed, %d given", need, cap);
  int* o = out;
  *o++ = kp.nunits; *o++ = kp.nsplits; *o++ = kp.a_bytes; *o++ = kp.kblocks;
  for (int u = 0; u < kp.nunits; ++u) {
    const WgUnit& U = kp.unit[u];
    *o++ = U.phase; *o++ = U.natoms; *o++ = U.nloads;
    for (int a = 0; a < 4; ++a) {
      const bool on = a < U.natoms;
      *o++ = on ? U.tap[a] : -1; *o++ = on ? U.rc[a] : -1; *o++ = on ? U.aoff[a] : -1;
    }
    for (int l = 0; l < 4; ++l) {
      const bool on = l < U.nloads;
      *o++ = on ? U.lrc[l] : -1; *o++ = on ? U.ldw[l] : -1; *o++ = on ? U.ldh[l] : -1; *o++ = on ? U.lrows[l] : -1;
      *o++ = on ? U.loff[l] : -1;
    }
  }
  return need;
}

size_t hpc_igemm_wgrad_umma_workspace_bytes(const hpc_igemm_wgrad_params* p) {
  WgKParams kp;
  if (wgrad_check(p) || wgrad_plan(p, &kp)) return 0;
  return (size_t)kp.nsplits * kp.ntaps * p->R * p->Cc * sizeof(float);
}

int hpc_igemm_wgrad_umma(const hpc_igemm_wgrad_params* p, hpc_stream_t stream) {
  int rc = wgrad_check(p);
  if (rc) return rc;
  WgKParams local;
  rc = wgrad_plan(p, &local);
  if (rc) return rc;
  const size_t need = (size_t)local.nsplits * local.ntaps * p->R * p->Cc * sizeof(float);
  HPC_REQUIRE(p->workspace && p->workspace_bytes >= need, HPC_ERR_BAD_SHAPE,
              "umma wgrad: workspace too small (%zu < %zu)", p->workspace_bytes, need);
  HPC_REQUIRE(hpc_aligned16(p->workspace), HPC_ERR_BAD_ALIGNMENT, "umma wgrad: workspace must be 16B aligned");
  local.ws = (float*)p->workspace;
  CUtensorMap tmLo, tmLo1, tmLo2, tmHi;
  {
    uint64_t dims[4] = {(uint64_t)p->R, (uint64_t)p->GW, (uint64_t)p->GH, (uint64_t)p->N};
    uint64_t strides[3] = {(uint64_t)p->R * 2, (uint64_t)p->GW * p->R * 2, (uint64_t)p->GH * p->GW * p->R * 2};
    uint32_t estr[4] = {1, 1, 1, 1};
    CUtensorMap* maps[3] = {&tmLo, &tmLo1, &tmLo2};
    for (int e = 0; e < 3; ++e) {               // boxes of th, th + 1, th + 2 pixel rows (shared by 1, 2, 3 taps)
      uint32_t box[4] = {64, (uint32_t)local.tw, (uint32_t)(local.th + e), (uint32_t)local.nb};
      rc = make_tmap(maps[e], p->lo, 4, dims, strides, box, estr);
      if (rc) return rc;
    }
  }
  {
    uint64_t dims[4] = {(uint64_t)p->Cc, (uint64_t)p->HW, (uint64_t)p->HH, (uint64_t)p->N};
    uint64_t strides[3] = {(uint64_t)p->Cc * 2, (uint64_t)p->HW * p->Cc * 2, (uint64_t)p->HH * p->HW * p->Cc * 2};
    uint32_t box[4] = {64, (uint32_t)(local.tw * p->hs), (uint32_t)(local.th * p->hs), (uint32_t)local.nb};
    uint32_t estr[4] = {1, (uint32_t)p->hs, (uint32_t)p->hs, 1};
    rc = make_tmap(&tmHi, p->hi, 4, dims, strides, box, estr);
    if (rc) return rc;
  }
  const int stage_bytes = local.a_bytes + local.cchunks * WG_ATOM_BYTES;
  local.stages = (232448 - 1024 - 256) / stage_bytes;          // C = 192: 5 x 42 KB (shared boxes) or 4 x 56 KB
  if (local.stages > WG_MAX_STAGES) local.stages = WG_MAX_STAGES;
  HPC_REQUIRE(local.stages >= 2, HPC_ERR_UNSUPPORTED, "umma wgrad: a stage of %d bytes does not fit twice", stage_bytes);
  { const char* e = getenv("HPC_WG_STAGES"); if (e && atoi(e) >= 2 && atoi(e) <= local.stages) local.stages = atoi(e); }
  const int smem_bytes = local.stages * stage_bytes + 1024 + 256;
  {  // staged epilogue when 256 padded rows fit in the operand stages (C <= 192); HPC_WG_EPI_SMEM=0 keeps direct stores
    static int want = -1;
    if (want < 0) { const char* e = getenv("HPC_WG_EPI_SMEM"); want = (e && e[0] == '0') ? 0 : 1; }
    local.epi_smem = want && 256 * (p->Cc * 4 + 16) <= local.stages * stage_bytes;
  }
  static HpcOncePerDevice attr_once;
  int attr_dev;
  if (attr_once.need(&attr_dev)) {
    HPC_CUDA(cudaFuncSetAttribute(igemm_wgrad_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
    attr_once.mark(attr_dev);
  }
  auto st = (cudaStream_t)stream;
  HpcProfScope prof(HPC_PROF_IGEMM_WGRAD, 2.0 * p->N * p->GH * p->GW * (double)local.ntaps * p->R * p->Cc, stream);
  HPC_CUDA(hpc_launch_pdl(igemm_wgrad_umma_kernel, dim3(local.nunits * local.nsplits), dim3(WG_THREADS), smem_bytes, st, tmLo,
                          tmLo1, tmLo2, tmHi, local));
  HPC_LAUNCH_CHECK();
  const int64_t n = (int64_t)local.ntaps * p->R * p->Cc;
  WgOut o;
  memset(&o, 0, sizeof(o));
  o.mode = p->out_mode; o.accumulate = p->out_accumulate; o.R = p->R; o.Cc = p->Cc;
  o.sr = p->out_stride_r; o.sc = p->out_stride_c;
  for (int t = 0; t < local.ntaps && t < HPC_MAX_TAPS; ++t) o.tap_src[t] = p->out_tap_src[t];
  const int64_t smin = o.sr < o.sc ? o.sr : o.sc;
  bool perm = o.mode == 1 && smin == local.ntaps && local.ntaps <= HPC_MAX_TAPS;
  if (perm) {                                  // tap_src must be a permutation of [0, ntaps)
    unsigned long long seen = 0;
    for (int t = 0; t < local.ntaps; ++t)
      if (o.tap_src[t] >= 0 && o.tap_src[t] < local.ntaps) seen |= 1ull << o.tap_src[t];
    perm = seen == (1ull << local.ntaps) - 1;
  }
  if (perm) {
    const int RT = o.sc < o.sr ? 4 : 16, CT = WGR_PAIRS / RT;       // R, Cc are multiples of 64
    HPC_CUDA(hpc_launch_pdl(wgrad_reduce_tiled_kernel, dim3((p->R / RT) * (p->Cc / CT)), dim3(256), 0, st,
                            (const float*)local.ws, local.nsplits, local.ntaps, p->scale, p->gw_f32, RT, CT, o));
  } else {
    HPC_CUDA(hpc_launch_pdl(wgrad_reduce_kernel, dim3((unsigned)((n + 127) / 128)), dim3(256), 0, st, (const float*)local.ws,
                            local.nsplits, n, p->scale, p->gw_f32, o));
  }
  HPC_LAUNCH_CHECK();
  return HPC_OK;
}
