// C-ABI entry points of the network group: dispatch on mzgo_net_t::precision.
//   precision 0: float32 CUDA-core kernels (nets_fp32.cu), slots are float32 [C][n*n]
//   precision 1: bf16 operands / fp32 accumulate on tcgen05 (nets_tc.cu), slots are padded bf16 chunks
#include "nets.cuh"

extern "C" int mzgo_net_prepare(const mzgo_net_t* net, void* stream) {
    if (int e = mzgo_net_check(net)) return e;
    cudaStream_t st = (cudaStream_t)stream;
    mzgo_net_bump_generation(net->workspace);
    if (int e = mzgo_net_prepare_fp32(net, st)) return e;
    if (net->precision == 1) return mzgo_net_prepare_tc(net, st);
    return MZGO_OK;
}

extern "C" int mzgo_net_initial(const mzgo_net_t* net, const float* obs, const uint32_t* boards, void* arena,
                                const int32_t* out_slot, float* value, float* policy, int want_softmax, int B,
                                void* stream) {
    if (int e = mzgo_net_check(net)) return e;
    MZGO_CHECK_ARG((obs != nullptr) != (boards != nullptr), "exactly one of obs / boards must be given");
    MZGO_CHECK_ARG(arena && out_slot && value && policy, "NULL argument");
    MZGO_CHECK_ARG(B >= 0 && B <= net->max_batch, "B=%d outside [0,max_batch=%d]", B, net->max_batch);
    if (B == 0) return MZGO_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const long np = (long)net->n * net->n;
    if (net->precision == 0) {
        const long stride = net->C * np;
        if (int e = mzgo_rep_fp32(net, obs, boards, (float*)arena, out_slot, stride, B, st)) return e;
        return mzgo_heads_fp32(net, (const float*)arena, out_slot, nullptr, stride, nullptr, value, policy,
                               want_softmax, B, st);
    }
    return mzgo_tc_initial(net, obs, boards, arena, out_slot, value, policy, want_softmax, B, st);
}

extern "C" int mzgo_net_recurrent(const mzgo_net_t* net, void* arena, const int32_t* in_slot, const int32_t* action,
                                  const int32_t* out_slot, float* reward, float* value, float* policy,
                                  int want_softmax, int B, void* stream) {
    if (int e = mzgo_net_check(net)) return e;
    MZGO_CHECK_ARG(arena && in_slot && action && out_slot && reward && value && policy, "NULL argument");
    MZGO_CHECK_ARG(B >= 0, "B < 0");
    if (B == 0) return MZGO_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (net->precision == 0) {
        const long stride = (long)net->C * net->n * net->n;
        if (int e = mzgo_dyn_fp32(net, (float*)arena, in_slot, action, out_slot, B, st)) return e;
        return mzgo_heads_fp32(net, (const float*)arena, out_slot, action, stride, reward, value, policy, want_softmax,
                               B, st);
    }
    return mzgo_tc_recurrent(net, arena, in_slot, action, out_slot, reward, value, policy, want_softmax, B, st);
}

extern "C" int mzgo_net_recurrent_tree(const mzgo_net_t* net, void* arena, const mzgo_tree_t* tree,
                                       const mzgo_search_params_t* params, const uint32_t* choice_stream, int stride,
                                       float* reward, float* value, float* policy, int do_select, void* stream) {
    if (int e = mzgo_net_check(net)) return e;
    MZGO_CHECK_ARG(net->precision == 1, "the fused network + tree step needs the bf16 (tensor core) mode");
    MZGO_CHECK_ARG(arena && tree && params && reward && value && policy, "NULL argument");
    MZGO_CHECK_ARG(tree->A == net->A && tree->G >= 0 && tree->G <= net->max_batch, "tree does not match the network (A=%d, G=%d)",
                   tree->A, tree->G);
    MZGO_CHECK_ARG(tree->leaf_in_slot && tree->leaf_action && tree->leaf_out_slot && tree->move_index_dev, "tree has a NULL array");
    if (tree->G == 0) return MZGO_OK;
    return mzgo_tc_recurrent_tree(net, arena, tree, params, choice_stream, stride, reward, value, policy, do_select,
                                  (cudaStream_t)stream);
}

extern "C" int mzgo_net_predict(const mzgo_net_t* net, const void* arena, const int32_t* slot, float* value,
                                float* policy, int want_softmax, int B, void* stream) {
    if (int e = mzgo_net_check(net)) return e;
    MZGO_CHECK_ARG(arena && slot && value && policy, "NULL argument");
    MZGO_CHECK_ARG(B >= 0 && B <= net->max_batch, "B=%d outside [0,max_batch=%d]", B, net->max_batch);
    if (B == 0) return MZGO_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const long cnt = (long)net->C * net->n * net->n;
    if (net->precision == 0)
        return mzgo_heads_fp32(net, (const float*)arena, slot, nullptr, cnt, nullptr, value, policy, want_softmax, B, st);
    NetWorkspace w = mzgo_net_layout(net->n, net->C, net->precision, net->max_batch);
    float* act3 = (float*)((char*)net->workspace + w.act3);
    if (int e = mzgo_tc_unpack_latent(net, arena, slot, act3, cnt, B, st)) return e;
    return mzgo_heads_fp32(net, act3, nullptr, slot, cnt, nullptr, value, policy, want_softmax, B, st);
}

extern "C" int mzgo_net_export_latent(const mzgo_net_t* net, const void* arena, const int32_t* slot, float* out, int B,
                                      void* stream) {
    if (int e = mzgo_net_check(net)) return e;
    MZGO_CHECK_ARG(arena && slot && out && B >= 0, "bad argument");
    if (B == 0) return MZGO_OK;
    const long cnt = (long)net->C * net->n * net->n;
    if (net->precision == 0)
        return mzgo_copy_latent_f32((const float*)arena, slot, cnt, out, nullptr, cnt, cnt, B, (cudaStream_t)stream);
    return mzgo_tc_unpack_latent(net, arena, slot, out, cnt, B, (cudaStream_t)stream);
}

extern "C" int mzgo_net_import_latent(const mzgo_net_t* net, void* arena, const int32_t* slot, const float* in, int B,
                                      void* stream) {
    if (int e = mzgo_net_check(net)) return e;
    MZGO_CHECK_ARG(arena && slot && in && B >= 0, "bad argument");
    if (B == 0) return MZGO_OK;
    const long cnt = (long)net->C * net->n * net->n;
    if (net->precision == 0)
        return mzgo_copy_latent_f32(in, nullptr, cnt, (float*)arena, slot, cnt, cnt, B, (cudaStream_t)stream);
    return mzgo_tc_pack_latent(net, in, cnt, arena, slot, B, (cudaStream_t)stream);
}
