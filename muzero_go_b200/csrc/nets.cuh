// Internal declarations shared by nets_fp32.cu, nets_tc.cu, nets_api.cu and engine.cu.
#pragma once
#include "common.cuh"

struct NetWorkspace {          // byte offsets into mzgo_net_t::workspace
    long wT1, wT2, wT3, wTd;   // float32 conv weights as [Cin][9][Cout]
    long act1, act2;           // float32 [max_batch][64][n*n] representation activations
    long act3;                 // bf16 mode: float32 [max_batch][C][n*n] representation output before packing
    long tc_w;                 // bf16 mode: dynamics conv weights in UMMA canonical layout
    long tc_embbias;           // bf16 mode: float32 [A][9][C] per-action, per-border-class bias (embedding folded in)
    long tc_headw;             // bf16 mode: float32 [3][C] policy / value / reward 1x1 conv weights + 56 head scalars
    long tc_w2, tc_w3;         // bf16 mode: representation conv2 / conv3 weights in UMMA canonical layout
    long tc_act1, tc_act2;     // bf16 mode: bf16 [max_batch][8][PS][8] representation activations (zero padded)
    long total;
};

NetWorkspace mzgo_net_layout(int n, int C, int precision, int max_batch);
int mzgo_net_check(const mzgo_net_t* net);
void mzgo_count_launch();
void mzgo_count_launches(int k);
int64_t mzgo_launch_count_peek();
// bumped by every mzgo_net_prepare of that workspace: kernel-argument constants derived from the weights
// (nets_tc.cu) are stale in anything captured before (the search-loop graph cache of engine.cu keys on it)
unsigned long long mzgo_net_generation(const void* workspace);
void mzgo_net_bump_generation(const void* workspace);

// mzgo_tree_step with the caller's word that the kernel in front of it is the network kernel of the same search (tree.cu)
int mzgo_tree_step_impl(const mzgo_tree_t* tree, const mzgo_search_params_t* p, const float* soft, const float* reward,
                        const float* value, const uint32_t* choice_stream, int stride, int do_backup, int do_select,
                        int after_network, void* stream);

int mzgo_net_prepare_fp32(const mzgo_net_t* net, cudaStream_t st);
int mzgo_rep_fp32(const mzgo_net_t* net, const float* obs, const uint32_t* boards, float* out, const int32_t* out_idx,
                  long out_stride, int B, cudaStream_t st);
int mzgo_heads_fp32(const mzgo_net_t* net, const float* arena, const int32_t* slot, const int32_t* action,
                    long slot_stride, float* reward, float* value, float* policy, int want_softmax, int B,
                    cudaStream_t st);
int mzgo_dyn_fp32(const mzgo_net_t* net, float* arena, const int32_t* in_slot, const int32_t* action,
                  const int32_t* out_slot, int B, cudaStream_t st);
int mzgo_copy_latent_f32(const float* src, const int32_t* src_idx, long src_stride, float* dst, const int32_t* dst_idx,
                         long dst_stride, long count, int B, cudaStream_t st);

// bf16 / tcgen05 path (nets_tc.cu)
int mzgo_net_prepare_tc(const mzgo_net_t* net, cudaStream_t st);
int mzgo_tc_pack_latent(const mzgo_net_t* net, const float* src, long src_stride, void* arena, const int32_t* slot,
                        int B, cudaStream_t st);
int mzgo_tc_unpack_latent(const mzgo_net_t* net, const void* arena, const int32_t* slot, float* dst, long dst_stride,
                          int B, cudaStream_t st);
int mzgo_tc_initial(const mzgo_net_t* net, const float* obs, const uint32_t* boards, void* arena,
                    const int32_t* out_slot, float* value, float* policy, int want_softmax, int B, cudaStream_t st);
int mzgo_tc_recurrent(const mzgo_net_t* net, void* arena, const int32_t* in_slot, const int32_t* action,
                      const int32_t* out_slot, float* reward, float* value, float* policy, int want_softmax, int B,
                      cudaStream_t st);
int mzgo_tc_recurrent_tree(const mzgo_net_t* net, void* arena, const mzgo_tree_t* tree, const mzgo_search_params_t* sp,
                           const uint32_t* choice_stream, int stride, float* reward, float* value, float* policy,
                           int do_select, cudaStream_t st);
