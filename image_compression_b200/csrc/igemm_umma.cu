// placeholder until the tcgen05 engine lands
#include "hpc_common.cuh"
int hpc_igemm_fwd_umma(const hpc_igemm_fwd_params* p, hpc_stream_t stream) { HPC_FAIL(HPC_ERR_UNSUPPORTED, "umma engine not built"); }
int hpc_igemm_wgrad_umma(const hpc_igemm_wgrad_params* p, hpc_stream_t stream) { HPC_FAIL(HPC_ERR_UNSUPPORTED, "umma engine not built"); }
size_t hpc_igemm_wgrad_umma_workspace_bytes(const hpc_igemm_wgrad_params* p) { return 0; }
