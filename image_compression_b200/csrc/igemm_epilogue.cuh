// Epilogue shared by the SIMT and the tcgen05 implicit-GEMM engines.
#pragma once
#include "hpc_common.cuh"

struct EpiArgs {
  int epi;
  const float* bias;
  const void* aux0;
  const void* aux1;
  const void* aux2;
  float* out_f32;
  void* out_lp;
  void* out_sq_lp;
  void* out_abs_lp;
  void* out_nrm_lp;
};

static inline EpiArgs make_epi_args(const hpc_igemm_fwd_params* p) {
  EpiArgs e;
  e.epi = p->epi; e.bias = p->bias; e.aux0 = p->aux0; e.aux1 = p->aux1; e.aux2 = p->aux2;
  e.out_f32 = p->out_f32; e.out_lp = p->out_lp; e.out_sq_lp = p->out_sq_lp; e.out_abs_lp = p->out_abs_lp;
  e.out_nrm_lp = p->out_nrm_lp;
  return e;
}

// value-level epilogue math; `nrm` receives rsqrt(n) for the GDN kinds
template <typename T>
__device__ __forceinline__ float epi_value(const EpiArgs& e, float acc, int64_t o, int co, float* nrm) {
  float v = acc + (e.bias ? e.bias[co] : 0.f);
  switch (e.epi) {
    case HPC_EPI_LINEAR: return v;
    case HPC_EPI_RELU: return fmaxf(v, 0.f);
    case HPC_EPI_EXP_CLAMP: return fminf(fmaxf(expf(v), 1e-10f), 1e10f);
    case HPC_EPI_GDN: {
      float r = rsqrtf(v);
      *nrm = r;
      // reference: x / sqrt(n)  (gdn.py:84-86)
      return ld_as_float<T>((const T*)e.aux0, o) / sqrtf(v);
    }
    case HPC_EPI_IGDN: {
      *nrm = rsqrtf(v);
      return ld_as_float<T>((const T*)e.aux0, o) * sqrtf(v);
    }
    case HPC_EPI_GDN_BWD:
      return ld_as_float<T>((const T*)e.aux0, o) * ld_as_float<T>((const T*)e.aux2, o) -
             ld_as_float<T>((const T*)e.aux1, o) * v;
    case HPC_EPI_IGDN_BWD:
      return ld_as_float<T>((const T*)e.aux0, o) / ld_as_float<T>((const T*)e.aux2, o) +
             ld_as_float<T>((const T*)e.aux1, o) * v;
    case HPC_EPI_RELU_MASK: return ld_as_float<T>((const T*)e.aux0, o) > 0.f ? v : 0.f;
    case HPC_EPI_SIGN_ADD: {
      float y = ((const float*)e.aux0)[o];
      float sg = (y > 0.f) ? 1.f : ((y < 0.f) ? -1.f : 0.f);
      return v * sg + ((const float*)e.aux1)[o];
    }
  }
  return v;
}

template <typename T>
__device__ __forceinline__ void epi_store_scalar(const EpiArgs& e, float acc, int64_t o, int co) {
  float nrm = 0.f;
  float v = epi_value<T>(e, acc, o, co, &nrm);
  if (e.out_f32) e.out_f32[o] = v;
  if (e.out_lp) st_from_float<T>((T*)e.out_lp, o, v);
  if (e.out_sq_lp) st_from_float<T>((T*)e.out_sq_lp, o, v * v);
  if (e.out_abs_lp) st_from_float<T>((T*)e.out_abs_lp, o, fabsf(v));
  if (e.out_nrm_lp && (e.epi == HPC_EPI_GDN || e.epi == HPC_EPI_IGDN)) st_from_float<T>((T*)e.out_nrm_lp, o, nrm);
}
