// bf16 / tcgen05 path of the networks: the dynamics conv3x3 (the dominant cost of every simulation,
// /root/reference/main.py:96-106) as an implicit GEMM on the 5th-generation tensor cores, fused with the
// reward head, the prediction heads (main.py:116-125) and the softmax of main.py:298.
//
// Formulation.  A latent lives in the arena as bf16 [C/8][PS][8]: 8-channel chunks, each a run of PS
// zero-padded flat positions with row stride n+1 (pixel (r,c) at q = (r+1)(n+1)+c+1, one shared zero
// column between rows, one zero row above and below).  With that layout
//   * a whole latent is ONE contiguous blob: one cp.async.bulk (TMA bulk copy, SASS UBLKCP) brings it
//     into shared memory exactly in the UMMA "K-major, no swizzle" canonical layout
//     (8 rows x 16 B core matrices, SBO = 128 B, LBO = PS*16 B);
//   * the 9 taps of the convolution are the SAME shared-memory slab read through 9 descriptors whose
//     start address is shifted by (dy*(n+1)+dx)*16 B -- no im2col, the slab is read from L2 once;
//   * M tile = 128 consecutive flat positions, N = 128 output channels, K = 9 taps x 128 channels;
//     accumulators (up to 4 tiles x 128 fp32 columns) live in TMEM.
// The action embedding (main.py:98-100) is linear in the conv, so it is folded at weight-load time into
// a per-action, per-border-class bias table [A][9][C] (border pixels see fewer taps).
//
// Warp roles (192 threads): warp 0 TMA producer, warp 1 TMEM owner + MMA issuer, warps 2-5 epilogue
// (tcgen05.ld -> bias/ReLU -> heads -> bf16 -> coalesced 16 B stores in the arena layout).
#include <cuda_bf16.h>

#include "nets.cuh"

namespace {

constexpr int TC_C = 128;               // latent_dim of the tensor-core path
constexpr int TC_KC = TC_C / 8;         // 8-channel chunks
constexpr int TC_THREADS = 192;
constexpr int TC_STAGES = 4;            // weight k-block ring
constexpr int TC_KBLOCKS = 18;          // 9 taps x 2 halves of 64 input channels
constexpr int TC_BSTAGE_BYTES = 8 * 128 * 16;   // [8 chunks][128 out channels][8 in channels] bf16 = 16 KB
constexpr int TC_MAX_TILES = 4;         // 4 x 128 TMEM columns

__host__ __device__ inline int tc_ps(int n) { return (n + 2) * (n + 1) + 2; }
__host__ __device__ inline int tc_tiles_per_image(int n) { return (n * (n + 1) + 127) / 128; }

// ---------------------------------------------------------------------------------------- PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, bf16 x bf16 -> fp32, M=128 N=128 K=16
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
// K-major, SWIZZLE_NONE shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;   // descriptor version (Blackwell)
    return d;                 // base_offset 0, lbo_mode 0, layout_type 0 = no swizzle
}
// kind::f16 instruction descriptor: D fp32, A/B bf16, both K-major, N=128, M=128
constexpr uint32_t TC_IDESC = (1u << 4) | (1u << 7) | (1u << 10) | ((128u >> 3) << 17) | ((128u >> 4) << 24);

#define TMEM_LD_32X32B_X32(taddr, r)                                                                                  \
    asm volatile(                                                                                                     \
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                                     \
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28," \
        "%29,%30,%31}, [%32];"                                                                                        \
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), \
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),      \
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),     \
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])                   \
        : "r"(taddr))

struct TcArgs {
    const __nv_bfloat16* arena_in;
    __nv_bfloat16* arena_out;
    const int32_t* in_slot;
    const int32_t* action;
    const int32_t* out_slot;
    const __nv_bfloat16* wB;     // [18][8][128][8]
    const float* tbl;            // [A][9][128]: conv bias + embedding contribution per border class
    const float* head_w;         // [3][128]: policy, value, reward 1x1 conv weights
    mzgo_net_t net;              // scalar head parameters are read through the reference pointers
    float* reward;
    float* value;
    float* policy;
    int want_softmax;
    int B;                       // items
    int n, ps, tiles_per_image, images_per_group, num_groups;
};

struct TcSmemLayout {
    uint32_t a_bytes, b_off, tbl_off, headw_off, logit_off, part_off, bar_off, total;
};
__host__ __device__ inline TcSmemLayout tc_smem_layout(int n) {
    TcSmemLayout L;
    const int tiles = tc_tiles_per_image(n), images = TC_MAX_TILES / tiles;
    L.a_bytes = (uint32_t)images * TC_KC * tc_ps(n) * 16;
    L.b_off = (L.a_bytes + 127) / 128 * 128;
    L.tbl_off = L.b_off + TC_STAGES * TC_BSTAGE_BYTES;
    L.headw_off = L.tbl_off + (uint32_t)images * 9 * TC_C * 4;
    L.logit_off = L.headw_off + 3 * TC_C * 4;
    L.part_off = L.logit_off + (uint32_t)images * 384 * 4;
    L.bar_off = L.part_off + TC_MAX_TILES * 4 * 2 * 4;          // [tile][warp][value, reward] partial sums
    L.total = L.bar_off + 16 * 8 + 16;
    return L;
}

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}

// ---------------------------------------------------------------------------------------- the fused kernel
__global__ void __launch_bounds__(TC_THREADS, 1)
k_tc_recurrent(const TcArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    const TcSmemLayout L = tc_smem_layout(a.n);
    uint8_t* sA = smem;
    uint8_t* sB = smem + L.b_off;
    float* s_tbl = reinterpret_cast<float*>(smem + L.tbl_off);       // [images][9][128]
    float* s_hw = reinterpret_cast<float*>(smem + L.headw_off);      // [3][128]
    float* s_logit = reinterpret_cast<float*>(smem + L.logit_off);   // [images][384]
    float* s_part = reinterpret_cast<float*>(smem + L.part_off);     // [tile][warp][2]
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bar_off);
    uint64_t* a_full = bars + 0;
    uint64_t* a_empty = bars + 1;
    uint64_t* acc_full = bars + 2;
    uint64_t* acc_empty = bars + 3;
    uint64_t* b_full = bars + 4;                    // [TC_STAGES]
    uint64_t* b_empty = bars + 4 + TC_STAGES;       // [TC_STAGES]
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 4 + 2 * TC_STAGES);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n = a.n, ps = a.ps, np = n * n;
    const int tpi = a.tiles_per_image, images = a.images_per_group;
    const uint32_t slab_bytes = (uint32_t)TC_KC * ps * 16;
    const int q0 = n + 2;                           // flat position of pixel (0,0)

    if (threadIdx.x == 0) {
        mbar_init(a_full, 1);
        mbar_init(a_empty, 1);
        mbar_init(acc_full, 1);
        mbar_init(acc_empty, 4);
        for (int s = 0; s < TC_STAGES; s++) { mbar_init(b_full + s, 1); mbar_init(b_empty + s, 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < 3 * TC_C; i += TC_THREADS) s_hw[i] = a.head_w[i];
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(s_tmem)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;

    // which images of a group are live (action >= 0): every role evaluates this the same way
    auto live_mask = [&](int grp) -> uint32_t {
        uint32_t m = 0;
        for (int i = 0; i < images; i++) {
            int item = grp * images + i;
            if (item < a.B && a.action[item] >= 0 && a.in_slot[item] >= 0 && a.out_slot[item] >= 0) m |= 1u << i;
        }
        return m;
    };

    if (warp == 0) {
        // ===================================================== TMA producer
        if (lane == 0) {
            uint32_t it = 0, kbc = 0;
            for (int grp = blockIdx.x; grp < a.num_groups; grp += gridDim.x) {
                const uint32_t live = live_mask(grp);
                if (!live) continue;
                mbar_wait(a_empty, (it & 1) ^ 1);
                mbar_expect_tx(a_full, slab_bytes * __popc(live));
                for (int i = 0; i < images; i++) {
                    if (!(live & (1u << i))) continue;
                    const int item = grp * images + i;
                    const uint8_t* src = reinterpret_cast<const uint8_t*>(a.arena_in) + (size_t)a.in_slot[item] * slab_bytes;
                    bulk_g2s(sA + (size_t)i * slab_bytes, src, slab_bytes, a_full);
                }
                for (int kb = 0; kb < TC_KBLOCKS; kb++, kbc++) {
                    const uint32_t s = kbc % TC_STAGES, ph = (kbc / TC_STAGES) & 1;
                    mbar_wait(b_empty + s, ph ^ 1);
                    mbar_expect_tx(b_full + s, TC_BSTAGE_BYTES);
                    bulk_g2s(sB + s * TC_BSTAGE_BYTES, reinterpret_cast<const uint8_t*>(a.wB) + (size_t)kb * TC_BSTAGE_BYTES,
                             TC_BSTAGE_BYTES, b_full + s);
                }
                it++;
            }
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer (one thread)
        if (lane == 0) {
            uint32_t it = 0, kbc = 0;
            const uint32_t sA_addr = smem_u32(sA), sB_addr = smem_u32(sB);
            for (int grp = blockIdx.x; grp < a.num_groups; grp += gridDim.x) {
                const uint32_t live = live_mask(grp);
                if (!live) continue;
                mbar_wait(acc_empty, (it & 1) ^ 1);
                mbar_wait(a_full, it & 1);
                tc_fence_after();
                for (int kb = 0; kb < TC_KBLOCKS; kb++, kbc++) {
                    const uint32_t s = kbc % TC_STAGES, ph = (kbc / TC_STAGES) & 1;
                    mbar_wait(b_full + s, ph);
                    tc_fence_after();
                    const int tap = kb >> 1, half = kb & 1;
                    const int shift = (tap / 3 - 1) * (n + 1) + (tap % 3 - 1);
#pragma unroll
                    for (int ks = 0; ks < 4; ks++) {
                        const uint64_t bdesc = umma_desc(sB_addr + s * TC_BSTAGE_BYTES + ks * 2 * 128 * 16, 128 * 16, 128);
                        const int chunk = half * 8 + ks * 2;
                        for (int t = 0; t < images * tpi; t++) {
                            const int img = t / tpi, tt = t - img * tpi;
                            if (!(live & (1u << img))) continue;
                            const uint32_t aaddr = sA_addr + img * slab_bytes + (uint32_t)chunk * ps * 16 +
                                                   (uint32_t)(q0 + tt * 128 + shift) * 16;
                            umma_bf16(tmem_base + t * 128, umma_desc(aaddr, ps * 16, 128), bdesc, TC_IDESC, (kb | ks) != 0);
                        }
                    }
                    umma_commit(b_empty + s);     // frees the weight stage when these MMAs retire
                }
                umma_commit(a_empty);             // the latent slab may be overwritten
                umma_commit(acc_full);            // accumulators complete
                it++;
            }
        }
    } else {
        // ===================================================== epilogue (4 warps, TMEM lane quarter = warp % 4)
        const int ew = warp - 2, wq = warp & 3;
        const int et = threadIdx.x - 64;              // 0..127
        uint32_t it = 0;
        for (int grp = blockIdx.x; grp < a.num_groups; grp += gridDim.x) {
            const uint32_t live = live_mask(grp);
            if (!live) continue;
            epi_bar_sync();                            // previous group's finalisation has read s_logit / s_part
            for (int i = 0; i < images; i++) {
                if (!(live & (1u << i))) continue;
                const float* src = a.tbl + (size_t)a.action[grp * images + i] * 9 * TC_C;
                for (int k = et; k < 9 * TC_C; k += 128) s_tbl[i * 9 * TC_C + k] = src[k];
            }
            epi_bar_sync();
            mbar_wait(acc_full, it & 1);
            tc_fence_after();
            for (int t = 0; t < images * tpi; t++) {
                const int img = t / tpi, tt = t - img * tpi;
                if (!(live & (1u << img))) continue;
                const int item = grp * images + img;
                const int o = tt * 128 + wq * 32 + lane;            // flat output row of this thread
                const int r = o / (n + 1), c = o - r * (n + 1);
                const bool valid = (r < n) && (c < n);
                const int cls = (r == 0 ? 0 : (r == n - 1 ? 2 : 1)) * 3 + (c == 0 ? 0 : (c == n - 1 ? 2 : 1));
                const int q = q0 + o;
                const bool store_ok = q < ps;
                const float* tb = s_tbl + (img * 9 + (valid ? cls : 4)) * TC_C;
                uint4* dst = reinterpret_cast<uint4*>(a.arena_out) + (size_t)a.out_slot[item] * TC_KC * ps + q;
                float pol = 0.f, val = 0.f, rew = 0.f;
#pragma unroll 1
                for (int cc = 0; cc < 4; cc++) {
                    uint32_t rr[32];
                    TMEM_LD_32X32B_X32(tmem_base + ((uint32_t)(wq * 32) << 16) + t * 128 + cc * 32, rr);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; j++) {
                        const int ch = cc * 32 + j;
                        float x = fmaxf(__uint_as_float(rr[j]) + tb[ch], 0.f);
                        x = valid ? x : 0.f;
                        v[j] = x;
                        pol = fmaf(x, s_hw[ch], pol);
                        val = fmaf(x, s_hw[TC_C + ch], val);
                        rew = fmaf(x, s_hw[2 * TC_C + ch], rew);
                    }
                    if (store_ok) {
#pragma unroll
                        for (int k = 0; k < 4; k++) {
                            uint4 w;
                            w.x = pack_bf16(v[k * 8 + 0], v[k * 8 + 1]);
                            w.y = pack_bf16(v[k * 8 + 2], v[k * 8 + 3]);
                            w.z = pack_bf16(v[k * 8 + 4], v[k * 8 + 5]);
                            w.w = pack_bf16(v[k * 8 + 6], v[k * 8 + 7]);
                            dst[(size_t)(cc * 4 + k) * ps] = w;
                        }
                    }
                }
                if (valid) s_logit[img * 384 + r * n + c] = pol + a.net.pol_b[0];
                val = warp_sum_f(val);
                rew = warp_sum_f(rew);
                if (lane == 0) { s_part[(t * 4 + ew) * 2] = val; s_part[(t * 4 + ew) * 2 + 1] = rew; }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acc_empty);     // TMEM may be overwritten by the next group
            epi_bar_sync();
            // ---- per-image finalisation: value / reward heads and the policy softmax (one warp per image)
            for (int img = ew; img < images; img += 4) {
                if (!(live & (1u << img))) continue;
                const int item = grp * images + img;
                float vs = 0.f, rs = 0.f;
                for (int tt = 0; tt < tpi; tt++)
                    for (int w = 0; w < 4; w++) { vs += s_part[((img * tpi + tt) * 4 + w) * 2]; rs += s_part[((img * tpi + tt) * 4 + w) * 2 + 1]; }
                if (lane == 0) {
                    const float vm = vs / (float)np + a.net.val_b[0];
                    a.value[item] = fmaf(a.net.vfc_w[0], vm, a.net.vfc_b[0]);
                    const float rm = rs / (float)np + a.net.rew_b[0];
                    float rw = a.net.ro_b[0];
                    for (int j = 0; j < 16; j++) rw = fmaf(a.net.ro_w[j], fmaxf(fmaf(a.net.rh_w[j], rm, a.net.rh_b[j]), 0.f), rw);
                    a.reward[item] = rw;
                }
                float* lg = s_logit + img * 384;
                if (lane == 0) lg[np] = a.net.pass_logit[0];
                __syncwarp();
                float* pol_out = a.policy + (size_t)item * (np + 1);
                if (!a.want_softmax) {
                    for (int k = lane; k <= np; k += 32) pol_out[k] = lg[k];
                } else {
                    float m = -INFINITY;
                    for (int k = lane; k <= np; k += 32) m = fmaxf(m, lg[k]);
                    m = warp_max_f(m);
                    float se = 0.f;
                    for (int k = lane; k <= np; k += 32) { float e = expf(lg[k] - m); lg[k] = e; se += e; }
                    se = warp_sum_f(se);
                    for (int k = lane; k <= np; k += 32) pol_out[k] = lg[k] / se;
                }
            }
            it++;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
    }
}

// ---------------------------------------------------------------------------------------- weight preparation
// wB[kb = tap*2 + half][kc][co][e] = bf16(W[co][half*64 + kc*8 + e][tap])
__global__ void k_tc_prep_w(const float* __restrict__ w, __nv_bfloat16* __restrict__ wB) {
    const int total = TC_KBLOCKS * 8 * 128 * 8;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        int e = i & 7, co = (i >> 3) & 127, kc = (i >> 10) & 7, kb = i >> 13;
        int tap = kb >> 1, half = kb & 1;
        int ci = half * 64 + kc * 8 + e;
        wB[i] = __float2bfloat16_rn(w[((size_t)co * TC_C + ci) * 9 + tap]);
    }
}

// tbl[a][cls][co] = bias[co] + sum over the taps that stay on the board for border class cls of
//                   sum_ci W[co][ci][tap] * emb[a][ci]          (conv(latent + emb) = conv(latent) + this)
__global__ void k_tc_prep_tbl(const float* __restrict__ w, const float* __restrict__ bias, const float* __restrict__ emb,
                              float* __restrict__ tbl, int A) {
    const int total = A * 9 * TC_C;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int co = i % TC_C, cls = (i / TC_C) % 9, a = i / (9 * TC_C);
        const int rc = cls / 3, cc = cls % 3;      // 0: first row/col, 1: interior, 2: last row/col
        float tap_sum[9];
#pragma unroll
        for (int t = 0; t < 9; t++) tap_sum[t] = 0.f;
        const float* wr = w + (size_t)co * TC_C * 9;
        const float* er = emb + (size_t)a * TC_C;
        for (int ci = 0; ci < TC_C; ci++) {
            const float e = er[ci];
#pragma unroll
            for (int t = 0; t < 9; t++) tap_sum[t] = fmaf(wr[ci * 9 + t], e, tap_sum[t]);
        }
        float s = bias[co];
#pragma unroll
        for (int t = 0; t < 9; t++) {
            const int dy = t / 3 - 1, dx = t % 3 - 1;
            const bool on = !((rc == 0 && dy < 0) || (rc == 2 && dy > 0) || (cc == 0 && dx < 0) || (cc == 2 && dx > 0));
            if (on) s += tap_sum[t];
        }
        tbl[i] = s;
    }
}

__global__ void k_tc_prep_headw(mzgo_net_t net, float* hw) {
    for (int i = threadIdx.x; i < TC_C; i += blockDim.x) {
        hw[i] = net.pol_w[i];
        hw[TC_C + i] = net.val_w[i];
        hw[2 * TC_C + i] = net.rew_w[i];
    }
}

// float32 [C][n*n] -> arena slot (zero padding written too)
__global__ void k_tc_pack(const float* __restrict__ src, long src_stride, __nv_bfloat16* __restrict__ arena,
                          const int32_t* __restrict__ slot, int n) {
    const int item = blockIdx.y;
    const int s = slot ? slot[item] : item;
    if (s < 0) return;
    const int ps = tc_ps(n), np = n * n, total = TC_KC * ps;
    const float* in = src + (size_t)item * src_stride;
    uint4* out = reinterpret_cast<uint4*>(arena) + (size_t)s * total;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int kc = i / ps, q = i - kc * ps;
        const int o = q - (n + 2);
        float v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (o >= 0) {
            const int r = o / (n + 1), c = o - r * (n + 1);
            if (r < n && c < n) {
#pragma unroll
                for (int e = 0; e < 8; e++) v[e] = in[(size_t)(kc * 8 + e) * np + r * n + c];
            }
        }
        uint4 w;
        w.x = pack_bf16(v[0], v[1]); w.y = pack_bf16(v[2], v[3]); w.z = pack_bf16(v[4], v[5]); w.w = pack_bf16(v[6], v[7]);
        out[i] = w;
    }
}

__global__ void k_tc_unpack(const __nv_bfloat16* __restrict__ arena, const int32_t* __restrict__ slot,
                            float* __restrict__ dst, long dst_stride, int n) {
    const int item = blockIdx.y;
    const int s = slot ? slot[item] : item;
    if (s < 0) return;
    const int ps = tc_ps(n), np = n * n, total = TC_C * np;
    const __nv_bfloat16* in = arena + (size_t)s * TC_KC * ps * 8;
    float* out = dst + (size_t)item * dst_stride;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int ch = i / np, p = i - ch * np;
        const int r = p / n, c = p - r * n;
        const int q = (r + 1) * (n + 1) + c + 1;
        out[i] = __bfloat162float(in[((size_t)(ch >> 3) * ps + q) * 8 + (ch & 7)]);
    }
}

}  // namespace

int mzgo_net_prepare_tc(const mzgo_net_t* net, cudaStream_t st) {
    char* ws = (char*)net->workspace;
    NetWorkspace w = mzgo_net_layout(net->n, net->C, net->precision, net->max_batch);
    k_tc_prep_w<<<144, 256, 0, st>>>(net->dyn_w, (__nv_bfloat16*)(ws + w.tc_w));
    k_tc_prep_tbl<<<296, 128, 0, st>>>(net->dyn_w, net->dyn_b, net->emb, (float*)(ws + w.tc_embbias), net->A);
    k_tc_prep_headw<<<1, 128, 0, st>>>(*net, (float*)(ws + w.tc_headw));
    mzgo_count_launches(3);
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

int mzgo_tc_pack_latent(const mzgo_net_t* net, const float* src, long src_stride, void* arena, const int32_t* slot,
                        int B, cudaStream_t st) {
    dim3 grid(ceil_div(TC_KC * tc_ps(net->n), 256), B);
    k_tc_pack<<<grid, 256, 0, st>>>(src, src_stride, (__nv_bfloat16*)arena, slot, net->n);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

int mzgo_tc_unpack_latent(const mzgo_net_t* net, const void* arena, const int32_t* slot, float* dst, long dst_stride,
                          int B, cudaStream_t st) {
    dim3 grid(ceil_div(TC_C * net->n * net->n, 1024), B);
    k_tc_unpack<<<grid, 256, 0, st>>>((const __nv_bfloat16*)arena, slot, dst, dst_stride, net->n);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

int mzgo_tc_recurrent(const mzgo_net_t* net, void* arena, const int32_t* in_slot, const int32_t* action,
                      const int32_t* out_slot, float* reward, float* value, float* policy, int want_softmax, int B,
                      cudaStream_t st) {
    static int num_sms = 0;
    static bool attr_set = false;
    if (num_sms == 0) {
        int dev = 0;
        MZGO_CUDA(cudaGetDevice(&dev));
        MZGO_CUDA(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
    }
    if (!attr_set) {
        MZGO_CUDA(cudaFuncSetAttribute(k_tc_recurrent, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
        attr_set = true;
    }
    char* ws = (char*)net->workspace;
    NetWorkspace w = mzgo_net_layout(net->n, net->C, net->precision, net->max_batch);
    TcArgs a{};
    a.arena_in = (const __nv_bfloat16*)arena;
    a.arena_out = (__nv_bfloat16*)arena;
    a.in_slot = in_slot; a.action = action; a.out_slot = out_slot;
    a.wB = (const __nv_bfloat16*)(ws + w.tc_w);
    a.tbl = (const float*)(ws + w.tc_embbias);
    a.head_w = (const float*)(ws + w.tc_headw);
    a.net = *net;
    a.reward = reward; a.value = value; a.policy = policy;
    a.want_softmax = want_softmax;
    a.B = B;
    a.n = net->n;
    a.ps = tc_ps(net->n);
    a.tiles_per_image = tc_tiles_per_image(net->n);
    a.images_per_group = TC_MAX_TILES / a.tiles_per_image;
    a.num_groups = ceil_div(B, a.images_per_group);
    const TcSmemLayout L = tc_smem_layout(net->n);
    MZGO_CHECK_ARG(L.total <= 232448, "shared memory layout of %u B does not fit", L.total);
    const int grid = a.num_groups < num_sms ? a.num_groups : num_sms;
    k_tc_recurrent<<<grid, TC_THREADS, L.total, st>>>(a);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}
