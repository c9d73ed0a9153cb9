// placeholder, replaced below
#include "nets.cuh"
int mzgo_net_prepare_tc(const mzgo_net_t*, cudaStream_t) { mzgo_set_error("bf16 path not built"); return MZGO_ERR_UNSUPPORTED; }
int mzgo_tc_pack_latent(const mzgo_net_t*, const float*, long, void*, const int32_t*, int, cudaStream_t) { mzgo_set_error("bf16 path not built"); return MZGO_ERR_UNSUPPORTED; }
int mzgo_tc_unpack_latent(const mzgo_net_t*, const void*, const int32_t*, float*, long, int, cudaStream_t) { mzgo_set_error("bf16 path not built"); return MZGO_ERR_UNSUPPORTED; }
int mzgo_tc_heads(const mzgo_net_t*, const void*, const int32_t*, float*, float*, int, int, cudaStream_t) { mzgo_set_error("bf16 path not built"); return MZGO_ERR_UNSUPPORTED; }
int mzgo_tc_recurrent(const mzgo_net_t*, void*, const int32_t*, const int32_t*, const int32_t*, float*, float*, float*, int, int, cudaStream_t) { mzgo_set_error("bf16 path not built"); return MZGO_ERR_UNSUPPORTED; }
