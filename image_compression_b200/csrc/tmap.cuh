// Host-side TMA tensor-map encoding shared by the tcgen05 kernels (no link-time dependency on libcuda).
#pragma once
#include <cuda.h>

#include "hpc_common.cuh"

// ---------------------------------------------------------------------------------------
// driver entry point for tensor-map encoding (no link-time dependency on libcuda)
// ---------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static inline EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// bf16 tensor map of rank `rank` (dims innermost first), SWIZZLE_128B, zero OOB fill
static inline int make_tmap(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                     const uint32_t* box, const uint32_t* estr) {
  EncodeTiledFn fn = get_encode_fn();
  HPC_REQUIRE(fn, HPC_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), dims, strides_bytes,
                  box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  HPC_REQUIRE(r == CUDA_SUCCESS, HPC_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return HPC_OK;
}


// fp32 tensor map without swizzle (dense box rows in shared memory), zero OOB fill: image patches with their padding
static inline int make_tmap_f32_plain(CUtensorMap* m, const void* base, int rank, const uint64_t* dims,
                                      const uint64_t* strides_bytes, const uint32_t* box) {
  EncodeTiledFn fn = get_encode_fn();
  HPC_REQUIRE(fn, HPC_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  const uint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<void*>(base), dims, strides_bytes, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  HPC_REQUIRE(r == CUDA_SUCCESS, HPC_ERR_CUDA, "cuTensorMapEncodeTiled (fp32) failed with CUresult %d", (int)r);
  return HPC_OK;
}
