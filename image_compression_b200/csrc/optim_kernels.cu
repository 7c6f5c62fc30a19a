// Fused clip_grad_value_ + AdamW step over all parameter tensors in ONE launch.
// Reference: engine/trainer.py:189-194 (clip_grad_value_, optimizer.step) with the optimizer of
// solver/optim.py:27-39 -- torch.optim.AdamW over 50 single-tensor parameter groups, each with its own lr /
// weight_decay ("bias" rule).  Eager torch runs ~10 foreach/elementwise launches per step over those groups;
// here a block owns one 4096-element chunk of one tensor (multi-tensor apply through a by-value segment table).
// HBM-bound: 28 B per parameter (read p, g, m, v; write p, m, v).
//
// The arithmetic follows torch/optim/adam.py (_multi_tensor_adam, decoupled weight decay) operation by operation
// in fp32, with the scalars the reference derives in double on the host passed in as floats:
//   g   = clamp(g, -clip, clip)                       (clip_grad_value_)
//   p  *= 1 - lr * wd
//   m   = m + (1 - beta1) * (g - m)                   (lerp, weight < 0.5)
//   v   = v * beta2 + (1 - beta2) * g * g             (mul_ then addcmul_)
//   den = sqrt(v) / sqrt(1 - beta2^t) + eps
//   p   = p + (-lr / (1 - beta1^t)) * (m / den)       (addcdiv_)
#include "hpc_common.cuh"

namespace {
constexpr int OPT_CHUNK = 4096;       // elements per block
constexpr int OPT_THREADS = 256;

struct OptTable {
  hpc_adamw_seg seg[HPC_ADAMW_MAX_SEGS];
  int chunk_begin[HPC_ADAMW_MAX_SEGS + 1];      // prefix sums of the chunk counts
  int nseg;
  float beta2, w1, w2, eps, clip;
  int write_clipped;
};

__device__ __forceinline__ void adamw_one(float& p, float& g, float& m, float& v, const hpc_adamw_seg& s, const OptTable& t) {
  if (t.clip > 0.f) g = fminf(fmaxf(g, -t.clip), t.clip);
  if (s.decay != 1.f) p = p * s.decay;
  m = m + t.w1 * (g - m);
  v = v * t.beta2;
  v = v + t.w2 * g * g;
  const float den = sqrtf(v) / s.bc2_sqrt + t.eps;
  p = p + s.neg_step_size * (m / den);
}

__global__ void __launch_bounds__(OPT_THREADS)
adamw_step_kernel(const __grid_constant__ OptTable t) {
  // segment of this block: binary search in the chunk prefix table (<= 6 steps, uniform across the block)
  int lo = 0, hi = t.nseg;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if ((int)blockIdx.x >= t.chunk_begin[mid]) lo = mid; else hi = mid;
  }
  const hpc_adamw_seg s = t.seg[lo];
  const int64_t e0 = (int64_t)((int)blockIdx.x - t.chunk_begin[lo]) * OPT_CHUNK;
  const int64_t e1 = e0 + OPT_CHUNK < s.n ? e0 + OPT_CHUNK : s.n;
  float* __restrict__ P = s.p; float* __restrict__ G = s.g; float* __restrict__ M = s.m; float* __restrict__ V = s.v;
  const bool vec = (((uintptr_t)P | (uintptr_t)G | (uintptr_t)M | (uintptr_t)V) & 15) == 0;
  int64_t i = e0 + (vec ? (int64_t)threadIdx.x * 4 : threadIdx.x);
  if (vec) {
    for (; i + 3 < e1; i += OPT_THREADS * 4) {
      float4 p = *reinterpret_cast<const float4*>(P + i), g = __ldcs(reinterpret_cast<const float4*>(G + i));
      float4 m = *reinterpret_cast<const float4*>(M + i), v = *reinterpret_cast<const float4*>(V + i);
      adamw_one(p.x, g.x, m.x, v.x, s, t); adamw_one(p.y, g.y, m.y, v.y, s, t);
      adamw_one(p.z, g.z, m.z, v.z, s, t); adamw_one(p.w, g.w, m.w, v.w, s, t);
      *reinterpret_cast<float4*>(P + i) = p; *reinterpret_cast<float4*>(M + i) = m; *reinterpret_cast<float4*>(V + i) = v;
      if (t.write_clipped) *reinterpret_cast<float4*>(G + i) = g;
    }
    // tail of the tensor (n % 4 elements): the thread whose vector would straddle the end
    if (i < e1 && i + 3 >= e1)
      for (int64_t j = i; j < e1; ++j) {
        float p = P[j], g = G[j], m = M[j], v = V[j];
        adamw_one(p, g, m, v, s, t);
        P[j] = p; M[j] = m; V[j] = v;
        if (t.write_clipped) G[j] = g;
      }
  } else {
    for (; i < e1; i += OPT_THREADS) {
      float p = P[i], g = G[i], m = M[i], v = V[i];
      adamw_one(p, g, m, v, s, t);
      P[i] = p; M[i] = m; V[i] = v;
      if (t.write_clipped) G[i] = g;
    }
  }
}
}  // namespace

int hpc_adamw_step(const hpc_adamw_seg* segs, int nseg, double beta1, double beta2, double eps, double clip,
                   int write_clipped_grad, hpc_stream_t stream) {
  HPC_REQUIRE(segs && nseg > 0, HPC_ERR_BAD_SHAPE, "hpc_adamw_step: no segments");
  auto st = (cudaStream_t)stream;
  double bytes = 0.0;
  for (int i = 0; i < nseg; ++i) bytes += 28.0 * (double)segs[i].n;
  HpcProfScope prof(HPC_PROF_OPTIM, bytes, stream);
  for (int base = 0; base < nseg; base += HPC_ADAMW_MAX_SEGS) {      // > 64 tensors: several launches
    OptTable t;
    memset(&t, 0, sizeof(t));
    t.nseg = nseg - base < HPC_ADAMW_MAX_SEGS ? nseg - base : HPC_ADAMW_MAX_SEGS;
    int chunks = 0;
    for (int i = 0; i < t.nseg; ++i) {
      const hpc_adamw_seg& s = segs[base + i];
      HPC_REQUIRE(s.p && s.g && s.m && s.v && s.n > 0, HPC_ERR_BAD_SHAPE, "hpc_adamw_step: segment %d is empty or NULL", base + i);
      HPC_REQUIRE((((uintptr_t)s.p | (uintptr_t)s.g | (uintptr_t)s.m | (uintptr_t)s.v) & 3) == 0, HPC_ERR_BAD_ALIGNMENT,
                  "hpc_adamw_step: segment %d is not 4-byte aligned", base + i);
      t.seg[i] = s;
      t.chunk_begin[i] = chunks;
      chunks += (int)((s.n + OPT_CHUNK - 1) / OPT_CHUNK);
    }
    t.chunk_begin[t.nseg] = chunks;
    // torch forms 1 - beta in double and casts the result (1.f - (float)0.999 would be off by 1.3e-5 relative)
    t.beta2 = (float)beta2; t.w1 = (float)(1.0 - beta1); t.w2 = (float)(1.0 - beta2); t.eps = (float)eps;
    t.clip = (float)clip;
    t.write_clipped = write_clipped_grad;
    adamw_step_kernel<<<chunks, OPT_THREADS, 0, st>>>(t);
    HPC_LAUNCH_CHECK();
  }
  return HPC_OK;
}
