// bf16 / tcgen05 path of the networks: every 3x3 convolution with >= 64 input channels -- the dynamics
// conv (the dominant cost of every simulation, /root/reference/main.py:96-106) and conv2 / conv3 of the
// representation network (main.py:80-84) -- as an implicit GEMM on the 5th-generation tensor cores,
// fused with the reward head, the prediction heads (main.py:116-125) and the softmax of main.py:298.
//
// Formulation.  A latent lives in HBM as bf16 [C/8][PS][8]: 8-channel chunks, each a run of PS
// zero-padded flat positions with row stride n+1 (pixel (r,c) at q = (r+1)(n+1)+c+1, one shared zero
// column between rows, one zero row above and below).  With that layout
//   * a whole latent is ONE contiguous blob: cp.async.bulk (TMA bulk copy, SASS UBLKCP) brings it
//     into shared memory exactly in the UMMA "K-major, no swizzle" canonical layout
//     (8 rows x 16 B core matrices, SBO = 128 B, LBO = PS*16 B);
//   * the 9 taps of the convolution are the SAME shared-memory slab read through descriptors whose
//     start address is shifted by (dy*(n+1)+dx)*16 B -- no im2col, the slab is read from L2 once;
//   * M tile = 128 consecutive flat positions, N = Cout output channels, K = 9 taps x Cin channels;
//     accumulators live in four 128-column TMEM slots that rotate from group to group, so the MMAs of
//     the next group start on the free slot while the epilogue still drains the previous ones.
// The action embedding (main.py:98-100) is linear in the conv, so it is folded at weight-load time into
// a per-action, per-border-class bias table [A][9][C] (border pixels see fewer taps).
//
// CTA pairs (cta_group::2).  One 128x128x16 MMA of a single CTA fetches 8 KB of operands per 64 cycles: exactly the
// 128 B/clk of the shared-memory port at exactly the tensor-pipe rate (tools/mma_probe.cu), so every other shared-
// memory byte -- weight ring refill, latent slabs, the epilogue's bias reads -- delayed the tensor pipe (88 cycles per
// MMA measured in round 1).  The kernel therefore runs as clusters of two CTAs on the two SMs of a TPC: one thread of
// the pair's leader issues tcgen05.mma.cta_group::2 (M = 256: 128 rows = one tile of positions from each CTA;
// N = Cout), each CTA holds only ITS half of the output channels of every weight k-block (the tensor cores exchange
// the halves), so operand fetch drops to 6 KB per MMA and the weight ring traffic per CTA halves
// (tools/mma2_probe.cu: operand placement checked against a host matmul, and the cycles per MMA).  Every CTA still
// owns its own images, TMEM accumulators, epilogue and finalisers; the pair walks its groups in lock step.
//
// Warp roles (384 threads): warp 0 TMA producer (latent slab in two 64-channel halves, weight k-blocks
// through an mbarrier ring), warp 1 TMEM owner + MMA issuer in the leader CTA (descriptors precomputed, ~4
// instructions per tcgen05.mma) / relay in the other CTA (tells the leader's barriers that this CTA's operands have
// landed), warps 2-9 epilogue (tcgen05.ld -> bias/ReLU -> heads with the 1x1 weights read as
// kernel-parameter constants -> bf16 -> coalesced 16 B stores in the arena layout).
#include <cuda.h>
#include <cuda_bf16.h>

#include <algorithm>
#include <mutex>
#include <unordered_map>

#include "nets.cuh"
#include "tree_step.cuh"

namespace {

constexpr int TC_C = 128;               // latent_dim of the tensor-core path
constexpr int TC_THREADS = 384;         // producer, MMA issuer / relay, 8 epilogue warps, 2 finalisers (384 x 168 registers)
constexpr int TC_EPI_THREADS = 256;
constexpr int TC_MAX_STAGES = 12;       // weight k-block ring (half a k-block per CTA and stage; as many as fit)
constexpr int TC_SLOTS = 4;             // 4 x 128 TMEM columns
constexpr int TC_SMEM_LIMIT = 232448;   // 227 KB

enum { MODE_PLAIN = 0, MODE_HEADS = 1, MODE_DYN = 2 };
unsigned long long* g_tc_trace = nullptr;

// A latent slab in SHARED memory: [C/8][ps][8] bf16, position q = q0 + r (n + 1) + c of pixel (r, c): rows of n pixels
// and one zero (the right / left padding of neighbouring rows is the same position), q0 >= n + 2 zero positions in
// front (the top padding row and the left padding of row 0) and n + 2 behind the last pixel -- every shifted view of
// the nine taps of the pixels stays inside the slab.
// In GLOBAL memory (the latent arena, the activation buffers) only the positions from the first to the last pixel
// are kept: [C/8][psh][8], position q - q0, tc_plen(n) = n (n + 1) - 1 of the psh in use.  The zero positions in front
// of and behind them live in shared memory alone (written once per launch, never overwritten by the bulk copies):
// 9 % less HBM per slot, which lets one search hold all 2048 19x19 games of a GPU (161 GB of latents) instead of two
// waves of 1024.  q0, ps and psh are multiples of 8 positions (128 B): every chunk of a slab is ONE bulk copy whose
// source and destination are 128-byte aligned, and a warp of the epilogue stores 512 aligned bytes per chunk.
__host__ __device__ inline int tc_q0(int n) { return (n + 2 + 7) / 8 * 8; }
__host__ __device__ inline int tc_plen(int n) { return n * (n + 1) - 1; }
__host__ __device__ inline int tc_ps(int n) { return (tc_q0(n) + tc_plen(n) + n + 2 + 7) / 8 * 8; }
__host__ __device__ inline int tc_psh(int n) { return (tc_plen(n) + 7) / 8 * 8; }
__host__ __device__ inline int tc_tiles_per_image(int n) { return (n * (n + 1) + 127) / 128; }
// A group (the images that share one pass over the weights) fills at most three of the four TMEM
// accumulator slots: with a spare slot the MMAs of the next group start while the epilogue still holds
// the last tiles of the previous one (a full ring stalled the tensor pipe for 3.3 us per group at 9x9).
constexpr int TC_GROUP_TILES = 3;
__host__ __device__ inline int tc_images_per_group(int n) {
    const int k = TC_GROUP_TILES / tc_tiles_per_image(n);
    return k < 1 ? 1 : k;
}

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
// ---- CTA pair plumbing
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {        // every thread of both CTAs
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of the same shared-memory object in CTA `cta` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t saddr, uint32_t cta) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(cta));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// L2 eviction priorities.  A simulation streams 2 x 108 KB of latents per leaf through the L2 exactly once
// (the next reader of a child latent is a later simulation, after > 200 MB of other traffic), while the tree
// kernels re-read the same root rows every simulation: latents go through with evict_first so that the
// trees' working set (about 20 MB per 1024 games) and the weights stay resident in the 126 MB L2.
// One half of an input slab (eight chunks of eight channels, all ps positions) as ONE tensor copy: box
// [8 positions x 8 channels = 128 B][ps / 8][8 chunks] of the map built by tc_make_in_map below; the positions in front
// of and behind the kept run are out of bounds there and arrive as zeros.  c1 = first position / 8 (negative), c2 =
// first chunk.
__device__ __forceinline__ void tensor_g2s_3d(void* dst, const CUtensorMap* map, int c1, int c2, uint64_t* bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2, %3, %4}], [%5], %6;"
        ::"r"(smem_u32(dst)), "l"(map), "r"(0), "r"(c1), "r"(c2), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}
__device__ __forceinline__ uint64_t l2_policy_evict_normal() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void bulk_g2s_hint(void* dst, const void* src, uint32_t bytes, uint64_t* bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
            smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy) : "memory");
}
__device__ __forceinline__ void st_global_v4_hint(void* ptr, uint4 v, uint64_t policy) {
    asm volatile("st.global.L2::cache_hint.v4.b32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(ptr), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w), "l"(policy) : "memory");
}
// Predicated form without a "memory" clobber: nothing in the kernel reads these bytes back, and as a compiler
// barrier (or behind a branch) the store kept the next channel chunk's shared-memory loads from being hoisted.
__device__ __forceinline__ void st_global_v4_hint_if(void* ptr, uint4 v, uint64_t policy, bool ok) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %6, 0;\n\t@p st.global.L2::cache_hint.v4.b32 [%0], {%1, %2, %3, %4}, %5;\n\t}" ::"l"(ptr),
        "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(policy), "r"((uint32_t)ok));
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// completion of all MMAs issued so far -> the barrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                     smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T for the CTA pair, bf16 x bf16 -> fp32, K=16: M = 256 (rows 0-127 from the leader's
// shared memory into the leader's TMEM, rows 128-255 from / into the peer's), N from idesc with each CTA holding N/2
// rows of B.  The descriptors address the same offsets in both CTAs' shared memory.
// Descriptors are passed as (lo, hi) words: only lo (start address, LBO) changes between MMAs.
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                          uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .b64 da, db;\n\t"
        "mov.b64 da, {%1, %2};\n\t"
        "mov.b64 db, {%3, %4};\n\t"
        "setp.ne.b32 p, %6, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %5, p;\n\t"
        "}" ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(acc) : "memory");
}
// K-major, SWIZZLE_NONE shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout):
//   lo = start address >> 4 (14 bits) | leading-dimension byte offset >> 4 << 16
//   hi = stride-dimension byte offset >> 4 | descriptor version 1 at bit 46
__device__ __forceinline__ uint32_t desc_lo(uint32_t saddr, uint32_t lbo_bytes) {
    return ((saddr >> 4) & 0x3FFFu) | (((lbo_bytes >> 4) & 0x3FFFu) << 16);
}
__device__ __forceinline__ uint32_t desc_hi(uint32_t sbo_bytes) { return ((sbo_bytes >> 4) & 0x3FFFu) | (1u << 14); }
// kind::f16 instruction descriptor: D fp32, A/B bf16, both K-major, M=256 (the CTA pair)
__host__ __device__ constexpr uint32_t tc_idesc(int n_cols) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n_cols >> 3) << 17) | ((256u >> 4) << 24);
}

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

struct TcScalars {   // the scalar head parameters (main.py:93-94, :111-114), host copies
    float pol_b, val_b, rew_b, vfc_w, vfc_b, pass_logit, ro_b, pad;
    float rh_w[16], rh_b[16], ro_w[16];
};

struct TcArgs {
    CUtensorMap in_map;          // the input slabs as a tensor [chunk][position / 8][128 B] (tc_make_in_map)
    const __nv_bfloat16* in;     // input slabs: item i at in + in_slot(i) * (CIN/8)*psh*8
    __nv_bfloat16* out;          // output slabs: item i at out + out_slot(i) * (COUT/8)*psh*8
    const int32_t* in_slot;      // optional (identity when NULL)
    const int32_t* out_slot;     // optional (identity when NULL); < 0 skips the item
    const int32_t* action;       // MODE_DYN: table row; < 0 skips the item.  Otherwise optional gate.
    const __nv_bfloat16* wB;     // [9*CIN/64][2 CTAs][8][COUT/2][8]
    const float* tbl;            // MODE_DYN: [A][9][COUT] bias table; otherwise [COUT] bias
    float* reward;
    float* value;
    float* policy;
    int want_softmax;
    int B;                       // items
    int n, q0, ps, psh, plen, tiles_per_image, images_per_group, stages;    // slab geometry: tc_q0 / tc_ps / tc_psh / tc_plen
    unsigned long long* trace;   // optional (tools/tc_trace.py): [grid][64] globaltimer stamps of the roles
    // mzgo_net_recurrent_tree: items are the games' selected leaves (in_slot / action / out_slot = the tree's
    // leaf records), and once the CTA's roles are through, all of its threads run the tree step of the CTA's
    // games (expand + backup, and the next selection when tree_mode == 1): one launch per simulation
    int tree_mode;
    int choice_stride;
    const uint32_t* choice_stream;
    mzgo_tree_t tree;
    mzgo_search_params_t sp;
    TcScalars sc;
    float hw[3 * TC_C];          // policy / value / reward 1x1 conv weights: read as constant-bank operands
};

constexpr int TC_NUM_BARS = 4 + 2 * TC_SLOTS + 2 * TC_MAX_STAGES + 6;
struct TcSmemLayout {
    uint32_t a_bytes, b_off, tbl_off, tbl_stride, pol_off, pol_stride, part_off, bar_off, total;
};
// mode_dyn: per-action bias tables (double buffered, loaded by TMA per group); otherwise one [cout] bias vector
__host__ __device__ inline TcSmemLayout tc_smem_layout(int n, int cin, int cout, int stages, bool mode_dyn) {
    TcSmemLayout L;
    const int images = tc_images_per_group(n);
    L.a_bytes = (uint32_t)images * (cin / 8) * tc_ps(n) * 16;
    L.b_off = (L.a_bytes + 127) / 128 * 128;
    L.tbl_off = L.b_off + (uint32_t)stages * cout * 64;            // a stage = this CTA's half of a weight k-block
    L.tbl_stride = mode_dyn ? (uint32_t)images * 9 * cout * 4 : 0u;                    // per group parity
    L.pol_off = L.tbl_off + (mode_dyn ? 2 * L.tbl_stride : (uint32_t)cout * 4);
    L.pol_stride = (uint32_t)images * 384 * 4;                    // [image][position] logits, per group parity
    L.part_off = L.pol_off + 2 * L.pol_stride;
    L.bar_off = L.part_off + 2 * TC_SLOTS * 4 * 2 * 4;            // [parity][tile][lane quarter][value, reward]
    L.total = L.bar_off + TC_NUM_BARS * 8 + 16;
    return L;
}

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

__device__ __forceinline__ void tc_stamp(const TcArgs& a, int slot) {
    if (a.trace) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)::"memory");     // (ordered against the barriers around it)
        a.trace[(size_t)blockIdx.x * 64 + slot] = t;
    }
}

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}

// ---------------------------------------------------------------------------------------- warp-uniform helpers
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}"
        : "=r"(pred));
    return pred != 0;
}

// CTA b owns the contiguous items [b*B/grid, (b+1)*B/grid) (balanced to within one item) and walks them in
// groups of up to images_per_group.  The two CTAs of a pair walk their groups in LOCK STEP (one MMA covers tile j of
// both): group gi of the pair holds items [lo + gi*images, ...) of each CTA, the pair runs max(groups) groups and a
// group has max(cnt) x tiles_per_image accumulator tiles in both CTAs; a CTA with fewer images treats the rest as
// dead tiles (never loaded, never stored).
struct TcRange {
    int lo, hi, plo, phi;      // this CTA's items, the peer's items
    int ngroups;               // groups of the pair
};
__device__ __forceinline__ TcRange tc_range(const TcArgs& a) {
    TcRange r;
    const long long b = blockIdx.x, pb = blockIdx.x ^ 1u, grid = gridDim.x;
    r.lo = (int)(b * a.B / grid); r.hi = (int)((b + 1) * a.B / grid);
    r.plo = (int)(pb * a.B / grid); r.phi = (int)((pb + 1) * a.B / grid);
    const int m = max(r.hi - r.lo, r.phi - r.plo);
    r.ngroups = (m + a.images_per_group - 1) / a.images_per_group;
    return r;
}
__device__ __forceinline__ int tc_group_cnt(const TcArgs& a, int lo, int hi, int gi) {
    return max(0, min(a.images_per_group, hi - (lo + gi * a.images_per_group)));
}

// The item records of a group: lane i < 8 holds (action, input slot, output slot) of this CTA's image i, lane 8 + i
// those of the PEER's image i (only their liveness is used), -1 beyond the group.  Every role loads them one group
// AHEAD (three independent loads, consumed an iteration later): read at the top of the group they were three
// dependent L2 round trips on the way to the first bulk copy, in front of every accumulator tile's first store, and
// in the launch's tail behind the last epilogue.
struct TcItems {
    int act, isl, osl;
};
__device__ __forceinline__ TcItems tc_load_items(const TcArgs& a, const TcRange& rg, int gi, int lane) {
    TcItems r;
    r.act = r.isl = r.osl = -1;
    const int side = lane >> 3, idx = lane & 7;
    const int item = (side ? rg.plo : rg.lo) + gi * a.images_per_group + idx;
    if (side < 2 && idx < a.images_per_group && gi < rg.ngroups && item < (side ? rg.phi : rg.hi)) {
        MZGO_DCHECK(item >= 0 && item < a.B);
        r.act = a.action ? __ldg(a.action + item) : 0;
        r.isl = a.in_slot ? __ldg(a.in_slot + item) : item;
        r.osl = a.out_slot ? __ldg(a.out_slot + item) : item;
    }
    return r;
}
// Which images of a group are live: bit i = this CTA's image i, bit 8 + i = the peer's image i.  A ballot, so every
// role of BOTH CTAs sees the same pair-uniform "group is live" (mask != 0); `& 0xff` = this CTA's own images.
__device__ __forceinline__ uint32_t tc_live_mask(const TcItems& t) {
    return __ballot_sync(FULL_MASK, (t.act | t.isl | t.osl) >= 0);
}
// accumulator tiles of group gi in each CTA of the pair
__device__ __forceinline__ int tc_group_tiles(const TcArgs& a, const TcRange& rg, int gi) {
    return max(tc_group_cnt(a, rg.lo, rg.hi, gi), tc_group_cnt(a, rg.plo, rg.phi, gi)) * a.tiles_per_image;
}

struct TcCtx {
    uint8_t* s_tbl;        // DYN: [parity][images][9][COUT] float; else [COUT] float
    uint8_t* s_pol;        // [parity][images][384] float
    float* s_part;         // [parity][tile][lane quarter][value, reward]
    uint32_t tbl_stride, pol_stride;
    uint64_t* acc_full;    // [TC_SLOTS]
    uint32_t acc_empty;    // shared::cluster address of the LEADER's acc_empty[0] (the MMA issuer waits there)
    uint64_t* tbl_full;    // [2]
    uint64_t* fin_full;    // [2] epilogue -> finaliser (and producer: bias table buffer free)
    uint64_t* fin_empty;   // [2] finaliser -> epilogue
    uint32_t tmem_base;
};

// ---------------------------------------------------------------------------------------- epilogue role
// 8 warps.  Warp (wq, pi): TMEM lane quarter wq = warp % 4, and of the stream of accumulator tiles it
// takes those whose running index is congruent to pi mod 2 -- the pair (wq,0), (wq,1) pulls two
// consecutive tiles out of TMEM at once (their slots are released as soon as the registers hold them),
// and there is ONE copy of the unrolled channel loop in the instruction cache.
// One thread = one output position x all COUT channels of its tile.
template <int COUT, int MODE>
__device__ __forceinline__ void tc_epilogue_role(const TcArgs& a, const TcCtx& cx, const TcItems& first) {
    constexpr int KC_OUT = COUT / 8;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wq = warp & 3, pi = (warp - 2) >> 2;
    const int n = a.n, psh = a.psh, plen = a.plen;
    const int tpi = a.tiles_per_image;
    const TcRange rg = tc_range(a);
    // geometry of this thread's row in tile tt of an image (tt < 3 for n <= 19): r | c << 5 | cls << 10 | valid << 14
    uint32_t geo[3];
#pragma unroll
    for (int tt = 0; tt < 3; tt++) {
        const int o = tt * 128 + wq * 32 + lane;
        const int r = o / (n + 1), cc = o - r * (n + 1);
        const bool valid = (r < n) && (cc < n);
        const int cls = valid ? (r == 0 ? 0 : (r == n - 1 ? 2 : 1)) * 3 + (cc == 0 ? 0 : (cc == n - 1 ? 2 : 1)) : 4;
        geo[tt] = (uint32_t)(r & 31) | ((uint32_t)(cc & 31) << 5) | ((uint32_t)cls << 10) | ((valid ? 1u : 0u) << 14);
    }
    if (MODE != MODE_DYN) {
        float* sb = reinterpret_cast<float*>(cx.s_tbl);
        for (int k = threadIdx.x - 64; k < COUT; k += TC_EPI_THREADS) sb[k] = a.tbl[k];
        asm volatile("bar.sync 1, 256;" ::: "memory");
    }
    uint32_t tile_ctr = 0, it = 0;
    const uint64_t pol_stream = l2_policy_evict_first();
    // (the next group's records are requested at the END of a group, so that they are not live across the tile
    // loop -- 128 accumulators per thread leave no registers to spare -- and travel during the wait for the
    // next accumulator)
    TcItems nx = first;
    for (int gi = 0; gi < rg.ngroups; gi++, nx = tc_load_items(a, rg, gi, lane)) {
        const int NT = tc_group_tiles(a, rg, gi);                  // tiles of this group in BOTH CTAs of the pair
        const uint32_t live_pair = tc_live_mask(nx), live = live_pair & 0xffu;
        const int my_osl = nx.osl;
        if (!live_pair) continue;
        const uint32_t par = it & 1, use = it >> 1;
        if (MODE == MODE_DYN) mbar_wait(cx.tbl_full + par, use & 1);                 // bias tables of this group have landed
        if (MODE != MODE_PLAIN) mbar_wait(cx.fin_empty + par, (use & 1) ^ 1);        // finaliser is done with this parity's buffers
        const float* s_tbl = reinterpret_cast<const float*>(cx.s_tbl + par * cx.tbl_stride);
        float* s_pol = reinterpret_cast<float*>(cx.s_pol + par * cx.pol_stride);
        float* s_part = cx.s_part + par * (TC_SLOTS * 4 * 2);
        for (int j = 0; j < NT; j++) {
            const uint32_t c = tile_ctr + j;
            if ((c & 1) != (uint32_t)pi) continue;
            const uint32_t slot = c & (TC_SLOTS - 1);
            const int img = j / tpi, tt = j - img * tpi;
            const bool is_live = (live >> img) & 1u;
            mbar_wait(cx.acc_full + slot, (c >> 2) & 1);
            tc_fence_after();
            if (j == 0 && warp == 2 && lane == 0 && it < 4) tc_stamp(a, 10 + 2 * it);
            uint32_t rr[COUT];
            if (is_live) {
                const uint32_t taddr = cx.tmem_base + ((uint32_t)(wq * 32) << 16) + slot * 128;
                TMEM_LD_32X32B_X32(taddr, rr);
                TMEM_LD_32X32B_X32(taddr + 32, (rr + 32));
                if constexpr (COUT == 128) {
                    TMEM_LD_32X32B_X32(taddr + 64, (rr + 64));
                    TMEM_LD_32X32B_X32(taddr + 96, (rr + 96));
                }
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            }
            if (warp == 2 && lane == 0 && it == 2) tc_stamp(a, 18);
            // the accumulator now lives in registers: hand the TMEM slot back to the pair's MMA issuer at once
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(cx.acc_empty + slot * 8);
            if (!is_live) continue;
            const uint32_t g = tt == 0 ? geo[0] : (tt == 1 ? geo[1] : geo[2]);
            const int r = g & 31, cc = (g >> 5) & 31, cls = (g >> 10) & 15;
            const bool valid = (g >> 14) & 1u;
            const int q = tt * 128 + wq * 32 + lane;              // position in the global slab: shared position - q0
            const bool store_ok = q < plen;
            const float* tb = (MODE == MODE_DYN) ? s_tbl + (img * 9 + cls) * COUT : s_tbl;
            const int oslot = __shfl_sync(FULL_MASK, my_osl, img);
            MZGO_DCHECK(oslot >= 0 && img < tc_group_cnt(a, rg.lo, rg.hi, gi) && slot < TC_SLOTS);
            uint4* dst = reinterpret_cast<uint4*>(a.out) + (size_t)oslot * KC_OUT * psh + q;
            float pol = 0.f, val = 0.f, rew = 0.f;
            const uint32_t keep = valid ? 0xffffffffu : 0u;
            float4 nb0 = *reinterpret_cast<const float4*>(tb), nb1 = *reinterpret_cast<const float4*>(tb + 4);
#pragma unroll
            for (int k = 0; k < KC_OUT; k++) {
                // the next chunk's biases are requested one chunk ahead; the compiler barrier at the end of the
                // body keeps the chunks in order (scheduled across chunks the 128 accumulators spill)
                const float4 b0 = nb0, b1 = nb1;
                if (k + 1 < KC_OUT) {
                    nb0 = *reinterpret_cast<const float4*>(tb + (k + 1) * 8);
                    nb1 = *reinterpret_cast<const float4*>(tb + (k + 1) * 8 + 4);
                }
                const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                float x[8];
#pragma unroll
                for (int e = 0; e < 8; e++) {
                    x[e] = fmaxf(__uint_as_float(rr[k * 8 + e]) + bb[e], 0.f);
                    if (MODE != MODE_PLAIN) {
                        pol = fmaf(x[e], a.hw[k * 8 + e], pol);
                        val = fmaf(x[e], a.hw[TC_C + k * 8 + e], val);
                    }
                    if (MODE == MODE_DYN) rew = fmaf(x[e], a.hw[2 * TC_C + k * 8 + e], rew);
                }
                uint4 w;                                             // padding positions stay zero for the next conv
                w.x = pack_bf16(x[0], x[1]) & keep;
                w.y = pack_bf16(x[2], x[3]) & keep;
                w.z = pack_bf16(x[4], x[5]) & keep;
                w.w = pack_bf16(x[6], x[7]) & keep;
                if (MODE == MODE_PLAIN) {
                    if (store_ok) *dst = w;                          // read again by the next layer
                } else {
                    st_global_v4_hint_if(dst, w, pol_stream, store_ok);
                }
                dst += psh;
                asm volatile("" ::: "memory");
            }
            if (warp == 2 && lane == 0 && it == 2) tc_stamp(a, 19);
            if (MODE != MODE_PLAIN) {
                if (valid) s_pol[img * 384 + r * n + cc] = pol + a.sc.pol_b;
                val = warp_sum_f(valid ? val : 0.f);
                if (MODE == MODE_DYN) rew = warp_sum_f(valid ? rew : 0.f);
                if (lane == 0) { s_part[(j * 4 + wq) * 2] = val; s_part[(j * 4 + wq) * 2 + 1] = rew; }
            }
        }
        tile_ctr += NT;
        if (warp == 2 && lane == 0 && it < 4) tc_stamp(a, 11 + 2 * it);
        if (warp == 2 && lane == 0) tc_stamp(a, 56);                 // overwritten by every group: the last one stays
        it++;
        if (MODE != MODE_PLAIN) {
            __syncwarp();
            if (lane == 0) mbar_arrive(cx.fin_full + par);         // release: s_pol / s_part of this warp are visible
        }
    }
}

// ---------------------------------------------------------------------------------------- finaliser role
// Two warps taking turns (warp 10 the even groups, warp 11 the odd ones: the groups' buffers are double
// buffered by the same parity, so a short last group does not queue behind its predecessor): per group the
// value / reward heads (main.py:101-105, :118-120) and the policy softmax of its (up to
// TC_GROUP_TILES) images.  The group's last finaliser pass is the tail of the launch, so the images are worked
// on side by side: lane i runs the head chain of image i, and every softmax pass has the loads of all images
// (four positions per lane and image) in flight before the first dependent instruction.  The per-lane order of
// the additions is the order of a plain loop over k = lane, lane + 32, ...
template <int MODE>
__device__ __forceinline__ void tc_finaliser_role(const TcArgs& a, const TcCtx& cx, const TcItems& first, uint32_t which) {
    constexpr int MI = TC_GROUP_TILES;
    const int lane = threadIdx.x & 31;
    const int n = a.n, np = n * n;
    const int tpi = a.tiles_per_image, images = a.images_per_group;
    const TcRange rg = tc_range(a);
    const float pass_logit = a.sc.pass_logit;
    uint32_t it = 0;
    TcItems nx = first;
    for (int gi = 0; gi < rg.ngroups; gi++) {
        const int base = rg.lo + gi * images;
        const int cnt = tc_group_cnt(a, rg.lo, rg.hi, gi);
        const uint32_t live_pair = tc_live_mask(nx), live = live_pair & 0xffu;
        nx = tc_load_items(a, rg, gi + 1, lane);
        if (!live_pair) continue;
        const uint32_t par = it & 1, use = it >> 1;
        if (par != which) { it++; continue; }           // the other finaliser warp's group
        mbar_wait(cx.fin_full + par, use & 1);
        const bool stamp = lane == 0 && (it == 1 || it == 2);
        const int stamp0 = it == 2 ? 24 : 27;
        if (stamp) tc_stamp(a, stamp0);
        float* s_pol = reinterpret_cast<float*>(cx.s_pol + par * cx.pol_stride);
        const float* s_part = cx.s_part + par * (TC_SLOTS * 4 * 2);
        if (lane < cnt && ((live >> lane) & 1u)) {
            const int img = lane, item = base + img;
            float vs = 0.f, rs = 0.f;
            for (int k = 0; k < tpi * 4; k++) { vs += s_part[(img * tpi * 4 + k) * 2]; rs += s_part[(img * tpi * 4 + k) * 2 + 1]; }
            const float vm = vs / (float)np + a.sc.val_b;
            a.value[item] = fmaf(a.sc.vfc_w, vm, a.sc.vfc_b);
            if (MODE == MODE_DYN) {
                const float rm = rs / (float)np + a.sc.rew_b;
                float rw = a.sc.ro_b;
#pragma unroll
                for (int k = 0; k < 16; k++) rw = fmaf(a.sc.ro_w[k], fmaxf(fmaf(a.sc.rh_w[k], rm, a.sc.rh_b[k]), 0.f), rw);
                a.reward[item] = rw;
            }
        }
        if (stamp) tc_stamp(a, stamp0 + 1);
        bool on[MI];
        float* pol_out[MI];
#pragma unroll
        for (int i = 0; i < MI; i++) {
            on[i] = i < cnt && ((live >> i) & 1u);
            pol_out[i] = a.policy + (size_t)(base + i) * (np + 1);
        }
        float v[MI][4];
        if (!a.want_softmax) {
            for (int k0 = lane; k0 < np; k0 += 128) {
#pragma unroll
                for (int i = 0; i < MI; i++)
#pragma unroll
                    for (int u = 0; u < 4; u++)
                        if (on[i] && k0 + u * 32 < np) v[i][u] = s_pol[i * 384 + k0 + u * 32];
#pragma unroll
                for (int i = 0; i < MI; i++)
#pragma unroll
                    for (int u = 0; u < 4; u++)
                        if (on[i] && k0 + u * 32 < np) pol_out[i][k0 + u * 32] = v[i][u];
            }
#pragma unroll
            for (int i = 0; i < MI; i++)
                if (on[i] && lane == 0) pol_out[i][np] = pass_logit;
        } else {
            // softmax over np board logits + the pass logit (main.py:298); the exponentials are staged in place
            float m[MI], se[MI];
#pragma unroll
            for (int i = 0; i < MI; i++) { m[i] = pass_logit; se[i] = 0.f; }
            for (int k0 = lane; k0 < np; k0 += 128) {
#pragma unroll
                for (int i = 0; i < MI; i++)
#pragma unroll
                    for (int u = 0; u < 4; u++)
                        v[i][u] = (on[i] && k0 + u * 32 < np) ? s_pol[i * 384 + k0 + u * 32] : pass_logit;
#pragma unroll
                for (int i = 0; i < MI; i++)
#pragma unroll
                    for (int u = 0; u < 4; u++) m[i] = fmaxf(m[i], v[i][u]);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
#pragma unroll
                for (int i = 0; i < MI; i++) m[i] = fmaxf(m[i], __shfl_xor_sync(FULL_MASK, m[i], o));
            for (int k0 = lane; k0 < np; k0 += 128) {
#pragma unroll
                for (int i = 0; i < MI; i++)
#pragma unroll
                    for (int u = 0; u < 4; u++)
                        if (on[i] && k0 + u * 32 < np) v[i][u] = s_pol[i * 384 + k0 + u * 32];
#pragma unroll
                for (int i = 0; i < MI; i++)
#pragma unroll
                    for (int u = 0; u < 4; u++)
                        if (on[i] && k0 + u * 32 < np) {
                            const float e = __expf(v[i][u] - m[i]);
                            s_pol[i * 384 + k0 + u * 32] = e;
                            se[i] += e;
                        }
            }
            if (stamp) tc_stamp(a, stamp0 + 2);
            float ep[MI], inv[MI];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
#pragma unroll
                for (int i = 0; i < MI; i++) se[i] += __shfl_xor_sync(FULL_MASK, se[i], o);
#pragma unroll
            for (int i = 0; i < MI; i++) {
                ep[i] = __expf(pass_logit - m[i]);
                inv[i] = 1.0f / (se[i] + ep[i]);
            }
            for (int k0 = lane; k0 < np; k0 += 128) {
#pragma unroll
                for (int i = 0; i < MI; i++)
#pragma unroll
                    for (int u = 0; u < 4; u++)
                        if (on[i] && k0 + u * 32 < np) v[i][u] = s_pol[i * 384 + k0 + u * 32];
#pragma unroll
                for (int i = 0; i < MI; i++)
#pragma unroll
                    for (int u = 0; u < 4; u++)
                        if (on[i] && k0 + u * 32 < np) pol_out[i][k0 + u * 32] = v[i][u] * inv[i];
            }
#pragma unroll
            for (int i = 0; i < MI; i++)
                if (on[i] && lane == 0) pol_out[i][np] = ep[i] * inv[i];
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(cx.fin_empty + par);
        if (lane == 0 && it < 4) tc_stamp(a, 20 + it);
        if (lane == 0) tc_stamp(a, 57);                              // the last group's stays
        it++;
    }
}

// ---------------------------------------------------------------------------------------- MMA issuer role
// The whole warp runs the loop so that every value stays warp-uniform (uniform registers, no per-thread
// serialisation around UTCHMMA); one elected lane issues tcgen05.mma / tcgen05.commit.  NT (tiles per
// group) is a template parameter so that the 4 x NT MMAs of a k-block are straight-line code.
struct TcMmaCtx {
    uint32_t sA, sB;
    uint64_t *a_full, *a_empty, *b_full, *b_empty, *acc_full, *acc_empty;
    uint32_t tmem_base;
};

struct TcMmaState {
    uint32_t it, s, bph, tile_ctr;
    bool leader;
    uint32_t a_lo[TC_SLOTS];     // descriptor of tile j of a group (image j / tpi, tile j % tpi), tap (0,0), chunk 0
};

// All MMAs of one group of NT tiles per CTA (NT = images of the group x tiles per image; MMA j covers tile j of BOTH
// CTAs of the pair).  Runs in the leader CTA only.
template <int CIN, int COUT, int NT>
__device__ __forceinline__ void tc_mma_group(const TcArgs& a, const TcMmaCtx& mx, TcMmaState& st) {
    constexpr int NH = CIN / 64;
    constexpr uint32_t STAGE_BYTES = COUT * 64;             // this CTA's half of a weight k-block: [8 chunks][COUT/2][8]
    constexpr uint32_t IDESC = tc_idesc(COUT);
    const int lane = threadIdx.x & 31;
    const int n = a.n, ps = a.ps;
    const uint32_t a_hi = desc_hi(128), b_hi = desc_hi(128);
    const uint32_t b_lo0 = desc_lo(mx.sB, (COUT / 2) * 16);
    const uint32_t ks_step = 2 * ps;
    const uint32_t tile_ctr = st.tile_ctr, it = st.it;
    uint32_t s = st.s, bph = st.bph;
    const bool leader = st.leader;
    uint32_t d_col[NT];
#pragma unroll
    for (int j = 0; j < NT; j++) d_col[j] = mx.tmem_base + ((tile_ctr + j) & (TC_SLOTS - 1)) * 128;
    // first touch of the TMEM slots of this group: the epilogue must have drained them
#pragma unroll
    for (int j = 0; j < NT; j++) {
        const uint32_t c = tile_ctr + j;
        mbar_wait(mx.acc_empty + (c & (TC_SLOTS - 1)), ((c >> 2) & 1) ^ 1);
    }
    for (int h = 0; h < NH; h++) {
        mbar_wait(mx.a_full + h, it & 1);
        if (h == 0 && it < 4 && lane == 0) tc_stamp(a, 2 + 2 * it);
        if (h == 0 && it < 12 && lane == 0) tc_stamp(a, 32 + 2 * it);      // slots 32..55: every group of a launch
        for (int tap = 0; tap < 9; tap++) {
            mbar_wait(mx.b_full + s, bph);
            tc_fence_after();
            const int dy = (tap * 11) >> 5;                       // tap / 3 for tap < 9
            const int shift = (dy - 1) * (n + 1) + (tap - 3 * dy - 1);
            const uint32_t a_off = (uint32_t)(h * 8 * ps + shift);       // 16-byte units
            const uint32_t b_lo = b_lo0 + s * (STAGE_BYTES >> 4);
            const uint32_t acc0 = (h | tap) != 0;
            if (leader) {
#pragma unroll
                for (int ks = 0; ks < 4; ks++) {
#pragma unroll
                    for (int j = 0; j < NT; j++)
                        umma_bf16(d_col[j], st.a_lo[j] + a_off + ks * ks_step, a_hi, b_lo + ks * COUT, b_hi, IDESC,
                                  ks == 0 ? acc0 : 1u);
                }
                umma_commit(mx.b_empty + s);          // frees the weight stage (in both CTAs) when these MMAs retire
            }
            __syncwarp();
            if (++s == (uint32_t)a.stages) { s = 0; bph ^= 1; }
        }
        if (leader) umma_commit(mx.a_empty + h);      // this half of the input slabs (both CTAs') may be overwritten
    }
    if (leader) {
#pragma unroll
        for (int j = 0; j < NT; j++) umma_commit(mx.acc_full + ((tile_ctr + j) & (TC_SLOTS - 1)));
    }
    if (it < 4 && lane == 0) tc_stamp(a, 3 + 2 * it);
    if (it < 12 && lane == 0) tc_stamp(a, 33 + 2 * it);
    __syncwarp();
    st.s = s;
    st.bph = bph;
    st.tile_ctr = tile_ctr + NT;
    st.it = it + 1;
}

template <int CIN, int COUT>
__device__ __forceinline__ void tc_mma_role(const TcArgs& a, const TcMmaCtx& mx, const TcItems& first) {
    constexpr int KC_IN = CIN / 8;
    const int lane = threadIdx.x & 31;
    const int ps = a.ps, q0 = a.q0;
    const int tpi = a.tiles_per_image;
    const uint32_t slab_bytes = (uint32_t)KC_IN * ps * 16;
    TcMmaState st;
    st.it = st.s = st.bph = st.tile_ctr = 0;
    st.leader = elect_one();
#pragma unroll
    for (int j = 0; j < TC_SLOTS; j++) {
        const int img = j / tpi, tt = j - img * tpi;
        st.a_lo[j] = desc_lo(mx.sA + img * slab_bytes + (uint32_t)(q0 + tt * 128) * 16, ps * 16);
    }
    const TcRange rg = tc_range(a);
    TcItems nx = first;
    for (int gi = 0; gi < rg.ngroups; gi++) {
        const uint32_t live_pair = tc_live_mask(nx);
        nx = tc_load_items(a, rg, gi + 1, lane);
        if (!live_pair) continue;
        switch (tc_group_tiles(a, rg, gi)) {
            case 1: tc_mma_group<CIN, COUT, 1>(a, mx, st); break;
            case 2: tc_mma_group<CIN, COUT, 2>(a, mx, st); break;
            default: tc_mma_group<CIN, COUT, 3>(a, mx, st); break;
        }
    }
}

// ---------------------------------------------------------------------------------------- relay role
// Warp 1 of the pair's SECOND CTA.  The MMA issuer sits in the leader CTA and can only wait on its own barriers; a
// bulk copy signals a barrier of the CTA it writes to.  So this warp waits, in the issuer's order, for this CTA's
// slab halves and weight stages to land and passes each completion on to the leader's barrier of the same name (which
// counts two arrivals per phase: the leader's own producer and this one).  It can never run a phase ahead: a stage is
// only refilled after the leader's commit freed it, i.e. after the leader consumed the phase before.
template <int CIN>
__device__ __forceinline__ void tc_relay_role(const TcArgs& a, uint64_t* a_full, uint64_t* b_full, const TcItems& first) {
    constexpr int NH = CIN / 64;
    const int lane = threadIdx.x & 31;
    const uint32_t ra_full = mapa_u32(smem_u32(a_full), 0), rb_full = mapa_u32(smem_u32(b_full), 0);
    const TcRange rg = tc_range(a);
    uint32_t it = 0, s = 0, bph = 0;
    TcItems nx = first;
    for (int gi = 0; gi < rg.ngroups; gi++) {
        const uint32_t live_pair = tc_live_mask(nx);
        nx = tc_load_items(a, rg, gi + 1, lane);
        if (!live_pair) continue;
        for (int h = 0; h < NH; h++) {
            mbar_wait(a_full + h, it & 1);
            if (lane == 0) mbar_arrive_cluster(ra_full + h * 8);
            for (int tap = 0; tap < 9; tap++) {
                mbar_wait(b_full + s, bph);
                if (lane == 0) mbar_arrive_cluster(rb_full + s * 8);
                if (++s == (uint32_t)a.stages) { s = 0; bph ^= 1; }
            }
        }
        it++;
    }
}

// ---------------------------------------------------------------------------------------- the fused kernel
// Launched as clusters of two CTAs (launch_tc_conv): blockIdx.x ^ 1 is the peer, cluster rank 0 the pair's leader.
template <int CIN, int COUT, int MODE>
__global__ void __launch_bounds__(TC_THREADS, 1)
k_tc_conv(const __grid_constant__ TcArgs a) {
    constexpr int KC_IN = CIN / 8, NH = CIN / 64;
    constexpr uint32_t STAGE_BYTES = COUT * 64;           // this CTA's half of a weight k-block: [8 chunks][COUT/2][8] bf16
    extern __shared__ __align__(128) uint8_t smem[];
    const TcSmemLayout L = tc_smem_layout(a.n, CIN, COUT, a.stages, MODE == MODE_DYN);
    uint8_t* sA = smem;
    uint8_t* sB = smem + L.b_off;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bar_off);
    uint64_t* a_full = bars + 0;                    // [2] input slab halves
    uint64_t* a_empty = bars + 2;                   // [2]
    uint64_t* acc_full = bars + 4;                  // [TC_SLOTS]
    uint64_t* acc_empty = bars + 4 + TC_SLOTS;      // [TC_SLOTS] (the leader's are the ones in use)
    uint64_t* b_full = bars + 4 + 2 * TC_SLOTS;     // [TC_MAX_STAGES]
    uint64_t* b_empty = b_full + TC_MAX_STAGES;     // [TC_MAX_STAGES]
    uint64_t* tbl_full = b_empty + TC_MAX_STAGES;   // [2] per-action bias tables of a group (DYN)
    uint64_t* fin_full = tbl_full + 2;              // [2] epilogue warps -> finaliser / producer
    uint64_t* fin_empty = fin_full + 2;             // [2] finaliser -> epilogue warps
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + TC_NUM_BARS);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ps = a.ps;
    const uint32_t slab_bytes = (uint32_t)KC_IN * ps * 16;
    const uint32_t half_bytes = slab_bytes / NH;                  // 64 channels of a slab in shared memory: one tensor copy
    const uint32_t rank = cluster_ctarank();
    const bool leader_cta = rank == 0;

    if (threadIdx.x == 0) {
        // "operands landed" barriers of the leader count two arrivals per phase: its own producer (with the bytes) and
        // the peer's relay warp; the leader's acc_empty collect the epilogue warps of both CTAs
        const uint32_t full_count = leader_cta ? 2 : 1;
        for (int h = 0; h < 2; h++) { mbar_init(a_full + h, full_count); mbar_init(a_empty + h, 1); }
        for (int s = 0; s < TC_SLOTS; s++) { mbar_init(acc_full + s, 1); mbar_init(acc_empty + s, 8); }
        for (int s = 0; s < TC_MAX_STAGES; s++) { mbar_init(b_full + s, full_count); mbar_init(b_empty + s, 1); }
        for (int p = 0; p < 2; p++) { mbar_init(tbl_full + p, 1); mbar_init(fin_full + p, 8); mbar_init(fin_empty + p, 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {            // the same warp of both CTAs: accumulators at the same TMEM address in both SMs
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(s_tmem)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    if (threadIdx.x == 0) asm volatile("prefetch.tensormap [%0];" ::"l"(&a.in_map) : "memory");     // the map is a kernel parameter: no dependency
    // everything above is private to the pair; what follows reads what the kernel before this one wrote (PDL, common.cuh)
    griddep_wait();
    griddep_launch();
    if (threadIdx.x == 0) tc_stamp(a, 0);
    // the first group's item records travel while the set-up above is published to the pair
    const TcRange rg = tc_range(a);
    const TcItems first = tc_load_items(a, rg, 0, lane);
    tc_fence_before();
    cluster_sync_all();         // both CTAs: barriers initialised, TMEM allocated, before any remote arrive / multicast commit
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;
    if (threadIdx.x == 0) tc_stamp(a, 1);

    if (warp == 0) {
        // ===================================================== TMA producer (lane 0 issues, the warp stays converged)
        // own images' slabs and bias tables, and THIS CTA's half (output channels [rank*COUT/2, ...)) of every k-block
        uint32_t it = 0, s = 0, bph = 0;
        const uint64_t pol_stream = l2_policy_evict_first(), pol_keep = l2_policy_evict_last(), pol_normal = l2_policy_evict_normal();
        const uint8_t* w_half = reinterpret_cast<const uint8_t*>(a.wB) + (size_t)rank * STAGE_BYTES;
        TcItems nx = first;
        for (int gi = 0; gi < rg.ngroups; gi++) {
            const TcItems cur = nx;
            const uint32_t live_pair = tc_live_mask(cur), live = live_pair & 0xffu;
            nx = tc_load_items(a, rg, gi + 1, lane);
            if (!live_pair) continue;
            if (MODE == MODE_DYN) {
                // per-action bias tables of this group -> s_tbl[parity] (free once the epilogue of group it-2 is through)
                const uint32_t par = it & 1, use = it >> 1;
                if (use > 0) mbar_wait(fin_full + par, (use - 1) & 1);
                if (lane == 0) mbar_expect_tx(tbl_full + par, 9u * COUT * 4 * __popc(live));
                __syncwarp();
                if (lane < 8 && ((live >> lane) & 1u))
                    bulk_g2s(smem + L.tbl_off + par * L.tbl_stride + (uint32_t)lane * 9 * COUT * 4,
                             a.tbl + (size_t)cur.act * 9 * COUT, 9u * COUT * 4, tbl_full + par);
            }
            for (int h = 0; h < NH; h++) {
                mbar_wait(a_empty + h, (it & 1) ^ 1);
                if (lane == 0) mbar_expect_tx(a_full + h, half_bytes * __popc(live));
                __syncwarp();
                // one tensor copy per image and half: the kept positions from global memory, zeros around them
                if (lane < 8 && ((live >> lane) & 1u))
                    tensor_g2s_3d(sA + (size_t)lane * slab_bytes + (size_t)h * half_bytes, &a.in_map, -(a.q0 >> 3),
                                  cur.isl * KC_IN + h * 8, a_full + h, MODE == MODE_DYN ? pol_stream : pol_normal);
                for (int tap = 0; tap < 9; tap++) {
                    mbar_wait(b_empty + s, bph ^ 1);
                    if (lane == 0) {
                        mbar_expect_tx(b_full + s, STAGE_BYTES);
                        bulk_g2s_hint(sB + s * STAGE_BYTES, w_half + (size_t)(h * 9 + tap) * 2 * STAGE_BYTES, STAGE_BYTES,
                                      b_full + s, pol_keep);
                    }
                    if (++s == (uint32_t)a.stages) { s = 0; bph ^= 1; }
                }
            }
            it++;
        }
    } else if (warp == 1) {
        if (leader_cta) {
            // ===================================================== MMA issuer of the pair
            TcMmaCtx mx;
            mx.sA = smem_u32(sA); mx.sB = smem_u32(sB);
            mx.a_full = a_full; mx.a_empty = a_empty; mx.b_full = b_full; mx.b_empty = b_empty;
            mx.acc_full = acc_full; mx.acc_empty = acc_empty;
            mx.tmem_base = tmem_base;
            tc_mma_role<CIN, COUT>(a, mx, first);
        } else {
            // ===================================================== relay: this CTA's landed operands -> the leader's barriers
            tc_relay_role<CIN>(a, a_full, b_full, first);
        }
    } else {
        TcCtx cx;
        cx.s_tbl = smem + L.tbl_off;
        cx.s_pol = smem + L.pol_off;
        cx.s_part = reinterpret_cast<float*>(smem + L.part_off);
        cx.tbl_stride = L.tbl_stride;
        cx.pol_stride = L.pol_stride;
        cx.acc_full = acc_full;
        cx.acc_empty = mapa_u32(smem_u32(acc_empty), 0);
        cx.tbl_full = tbl_full;
        cx.fin_full = fin_full;
        cx.fin_empty = fin_empty;
        cx.tmem_base = tmem_base;
        if (warp < 10) tc_epilogue_role<COUT, MODE>(a, cx, first);             // 8 warps: TMEM lane quarter x tile parity
        else if (MODE != MODE_PLAIN) tc_finaliser_role<MODE>(a, cx, first, (uint32_t)warp - 10u);    // heads + softmax per image
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) tc_stamp(a, 30);
    if (MODE == MODE_DYN && a.tree_mode) {
        // ===================================================== tree phase (tree_step.cuh, the code of k_tree_step)
        // Item i of the launch is game i's leaf, and a CTA's items are a contiguous range of games: what the
        // search does with the outputs is CTA-local.  The finalisers' value / reward / policy stores are ordered
        // before these (L2) loads by the barrier above; the scratch aliases the input slabs, which are dead (this
        // CTA's accumulators are complete, so the pair's MMAs have read both CTAs' slabs for the last time).
        const int g_lo = rg.lo, g_hi = rg.hi;
        const int do_select = a.tree_mode == 1;
        const int sub128 = threadIdx.x >> 7;
        const bool v1 = a.sp.variant == MZGO_VARIANT_SELFPLAY;     // a quarter of the step's code belongs to this variant
        if (a.tree.A <= STEP_EPT * 32) {
            StepShared<32>* sh = reinterpret_cast<StepShared<32>*>(sA) + sub128;      // a warp per game
            for (int g = g_lo + warp; g < g_hi; g += TC_THREADS / 32) {
                if (v1) tree_step_body<32, 2, false, true>(a.tree, a.sp, a.policy, a.reward, a.value, a.choice_stream, a.choice_stride, 1,
                                                           do_select, *sh, warp & 3, lane, g, [] {});
                else tree_step_body<32, 2, false, false>(a.tree, a.sp, a.policy, a.reward, a.value, a.choice_stream, a.choice_stride, 1,
                                                         do_select, *sh, warp & 3, lane, g, [] {});
            }
        } else {
            StepShared<128>* sh = reinterpret_cast<StepShared<128>*>(sA) + sub128;    // 128 threads per game
            for (int g = g_lo + sub128; g < g_hi; g += TC_THREADS / 128) {
                if (v1) tree_step_body<128, 2, false, true>(a.tree, a.sp, a.policy, a.reward, a.value, a.choice_stream, a.choice_stride, 1,
                                                            do_select, *sh, 0, threadIdx.x & 127, g, [] {});
                else tree_step_body<128, 2, false, false>(a.tree, a.sp, a.policy, a.reward, a.value, a.choice_stream, a.choice_stride, 1,
                                                          do_select, *sh, 0, threadIdx.x & 127, g, [] {});
            }
        }
        if (threadIdx.x == 0) tc_stamp(a, 31);
    }
    // nobody leaves, and the TMEM goes back, only when the peer is through with this CTA's barriers / shared memory too
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
        if (lane == 0) tc_stamp(a, 58);
    }
}

// ---------------------------------------------------------------------------------------- weight preparation
// wB[kb = half*9 + tap][r][kc][j][e] = bf16(W[r*cout/2 + j][half*64 + kc*8 + e][tap])      (W = [Cout][Cin][3][3])
// r = the CTA of the pair that stages this half of the output channels: each (kb, r) block is one contiguous bulk copy
__global__ void k_tc_prep_w(const float* __restrict__ w, __nv_bfloat16* __restrict__ wB, int cin, int cout) {
    const int total = 9 * cin * cout, ch = cout / 2;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int e = i & 7, j = (i >> 3) % ch, kc = ((i >> 3) / ch) & 7, r = ((i >> 3) / (ch * 8)) & 1, kb = (i >> 3) / (cout * 8);
        const int co = r * ch + j;
        const int half = kb / 9, tap = kb - half * 9;
        const int ci = half * 64 + kc * 8 + e;
        wB[i] = __float2bfloat16_rn(w[((size_t)co * cin + ci) * 9 + tap]);
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

// the parameters the epilogue reads as kernel-argument constants, gathered into one device buffer
__global__ void k_tc_gather_heads(mzgo_net_t net, float* out /* [3*128 + sizeof(TcScalars)/4] */) {
    for (int i = threadIdx.x; i < TC_C; i += blockDim.x) {
        out[i] = net.pol_w[i];
        out[TC_C + i] = net.val_w[i];
        out[2 * TC_C + i] = net.rew_w[i];
    }
    float* sc = out + 3 * TC_C;
    if (threadIdx.x == 0) {
        sc[0] = net.pol_b[0]; sc[1] = net.val_b[0]; sc[2] = net.rew_b[0]; sc[3] = net.vfc_w[0];
        sc[4] = net.vfc_b[0]; sc[5] = net.pass_logit[0]; sc[6] = net.ro_b[0]; sc[7] = 0.f;
    }
    if (threadIdx.x < 16) {
        sc[8 + threadIdx.x] = net.rh_w[threadIdx.x];
        sc[24 + threadIdx.x] = net.rh_b[threadIdx.x];
        sc[40 + threadIdx.x] = net.ro_w[threadIdx.x];
    }
}

// ---------------------------------------------------------------------------------------- representation conv1
// main.py:75/:81: conv3x3 6 -> 64 + ReLU on CUDA cores (0.3 % of the network FLOPs), reading the packed
// board (or float32 GymGo planes) and writing the bf16 slab layout [8][PS][8] that conv2 consumes.
constexpr int C1_THREADS = 256;
__global__ void __launch_bounds__(C1_THREADS)
k_rep_conv1_bf16(const float* __restrict__ obs, const uint32_t* __restrict__ boards, const int32_t* __restrict__ gate,
                 const float* __restrict__ w /* [64][6][3][3] */, const float* __restrict__ bias,
                 __nv_bfloat16* __restrict__ out, int n) {
    __shared__ float s_w[54 * 64];                                  // [ci*9+tap][co]
    __shared__ float s_x[6 * (MZGO_MAX_N + 2) * (MZGO_MAX_N + 2)];  // zero-padded planes
    __shared__ float s_b[64];
    const int item = blockIdx.x;
    if (gate && gate[item] < 0) return;
    const int np = n * n, pw = n + 2, pp = pw * pw, psh = tc_psh(n);
    for (int i = threadIdx.x; i < 54 * 64; i += C1_THREADS) {
        const int co = i & 63, k = i >> 6;
        s_w[i] = w[co * 54 + k];
    }
    if (threadIdx.x < 64) s_b[threadIdx.x] = bias[threadIdx.x];
    const uint32_t* rec = boards ? boards + (size_t)item * MZGO_BOARD_WORDS : nullptr;
    for (int i = threadIdx.x; i < 6 * pp; i += C1_THREADS) {
        const int ci = i / pp, q = i - ci * pp;
        const int r = q / pw - 1, c = q - (q / pw) * pw - 1;
        float v = 0.f;
        if (r >= 0 && r < n && c >= 0 && c < n) {
            if (rec) {
                uint32_t bit;
                if (ci < 2) bit = (rec[ci * n + r] >> c) & 1u;
                else if (ci == 3) bit = (rec[2 * n + r] >> c) & 1u;
                else if (ci == 2) bit = rec[60] & 1u;
                else if (ci == 4) bit = rec[61] & 1u;
                else bit = (rec[61] >> 1) & 1u;
                v = (float)bit;
            } else {
                v = obs[((size_t)item * 6 + ci) * np + r * n + c];
            }
        }
        s_x[i] = v;
    }
    __syncthreads();
    uint4* o = reinterpret_cast<uint4*>(out) + (size_t)item * 8 * psh;
    for (int i = threadIdx.x; i < 8 * np; i += C1_THREADS) {
        const int kc = i / np, p = i - kc * np;
        const int r = p / n, c = p - r * n;
        float acc[8];
#pragma unroll
        for (int e = 0; e < 8; e++) acc[e] = s_b[kc * 8 + e];
#pragma unroll
        for (int ci = 0; ci < 6; ci++) {
#pragma unroll
            for (int tap = 0; tap < 9; tap++) {
                const float x = s_x[ci * pp + (r + tap / 3) * pw + (c + tap % 3)];
                const float4 w0 = *reinterpret_cast<const float4*>(s_w + (ci * 9 + tap) * 64 + kc * 8);
                const float4 w1 = *reinterpret_cast<const float4*>(s_w + (ci * 9 + tap) * 64 + kc * 8 + 4);
                acc[0] = fmaf(w0.x, x, acc[0]); acc[1] = fmaf(w0.y, x, acc[1]);
                acc[2] = fmaf(w0.z, x, acc[2]); acc[3] = fmaf(w0.w, x, acc[3]);
                acc[4] = fmaf(w1.x, x, acc[4]); acc[5] = fmaf(w1.y, x, acc[5]);
                acc[6] = fmaf(w1.z, x, acc[6]); acc[7] = fmaf(w1.w, x, acc[7]);
            }
        }
        uint4 v;
        v.x = pack_bf16(fmaxf(acc[0], 0.f), fmaxf(acc[1], 0.f));
        v.y = pack_bf16(fmaxf(acc[2], 0.f), fmaxf(acc[3], 0.f));
        v.z = pack_bf16(fmaxf(acc[4], 0.f), fmaxf(acc[5], 0.f));
        v.w = pack_bf16(fmaxf(acc[6], 0.f), fmaxf(acc[7], 0.f));
        o[(size_t)kc * psh + r * (n + 1) + c] = v;                  // the zeros between the rows are never rewritten
    }
}

// float32 [C][n*n] -> arena slot (zero padding written too)
__global__ void k_tc_pack(const float* __restrict__ src, long src_stride, __nv_bfloat16* __restrict__ arena,
                          const int32_t* __restrict__ slot, int n) {
    const int item = blockIdx.y;
    const int s = slot ? slot[item] : item;
    if (s < 0) return;
    const int psh = tc_psh(n), np = n * n, total = (TC_C / 8) * psh;
    const float* in = src + (size_t)item * src_stride;
    uint4* out = reinterpret_cast<uint4*>(arena) + (size_t)s * total;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int kc = i / psh, q = i - kc * psh;
        const int o = q;
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
    const int psh = tc_psh(n), np = n * n, total = TC_C * np;
    const __nv_bfloat16* in = arena + (size_t)s * (TC_C / 8) * psh * 8;
    float* out = dst + (size_t)item * dst_stride;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int ch = i / np, p = i - ch * np;
        const int r = p / n, c = p - r * n;
        const int q = r * (n + 1) + c;
        out[i] = __bfloat162float(in[((size_t)(ch >> 3) * psh + q) * 8 + (ch & 7)]);
    }
}

// ---- host copies of the head parameters, keyed by the workspace they were prepared into ----
struct HostHeads {
    float hw[3 * TC_C];
    TcScalars sc;
};
std::mutex g_heads_mu;
std::unordered_map<const void*, HostHeads> g_heads;

int num_sms() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
            n = 148;
    }
    return n;
}

// The input slabs of a launch as a rank-3 tensor map over global memory: [chunk][position / 8][8 positions x 8 channels]
// with tc_psh(n) / 8 rows of 128 B per chunk (stride tc_psh(n) * 16 B between chunks) and a box of eight
// chunks x ALL tc_ps(n) shared-memory positions starting tc_q0(n) positions in front of the first kept one: what lies
// outside [0, tc_psh(n)) is out of bounds and is filled with zeros by the copy engine.  (The positions between
// tc_plen(n) and tc_psh(n) are in bounds: they are zero in memory because nothing ever writes them.)  The chunk
// dimension is unbounded (the library does not know how many slots the caller's arena has).
typedef CUresult (*TcEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int tc_make_in_map(CUtensorMap* map, const void* base, int n) {
    static TcEncodeTiled encode = nullptr;
    static std::mutex mu;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (!encode) {
            void* fn = nullptr;
            cudaDriverEntryPointQueryResult q;
            MZGO_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
            MZGO_CHECK_ARG(fn != nullptr && q == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled is not available in this driver");
            encode = (TcEncodeTiled)fn;
        }
    }
    MZGO_CHECK_ARG(((uintptr_t)base & 15u) == 0, "latent slabs must be 16-byte aligned");
    const cuuint64_t dims[3] = {64, (cuuint64_t)tc_psh(n) / 8, 1ull << 31};
    const cuuint64_t strides[2] = {128, (cuuint64_t)tc_psh(n) * 16};
    const cuuint32_t box[3] = {64, (cuuint32_t)tc_ps(n) / 8, 8};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    MZGO_CHECK_ARG(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with %d (n=%d)", (int)r, n);
    return MZGO_OK;
}

template <int CIN, int COUT, int MODE>
int launch_tc_conv(TcArgs& a, int n, int B, cudaStream_t st) {
    static bool attr_set = false;
    if (!attr_set) {
        MZGO_CUDA(cudaFuncSetAttribute(k_tc_conv<CIN, COUT, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_LIMIT));
        attr_set = true;
    }
    a.B = B;
    a.n = n;
    a.trace = g_tc_trace;
    if (int e = tc_make_in_map(&a.in_map, a.in, n)) return e;
    a.q0 = tc_q0(n);
    a.ps = tc_ps(n);
    a.psh = tc_psh(n);
    a.plen = tc_plen(n);
    a.tiles_per_image = tc_tiles_per_image(n);
    a.images_per_group = tc_images_per_group(n);
    const int kblocks = 9 * CIN / 64;
    const TcSmemLayout L0 = tc_smem_layout(n, CIN, COUT, 0, MODE == MODE_DYN);
    int stages = ((int)TC_SMEM_LIMIT - (int)L0.total) / (COUT * 64);
    if (stages > TC_MAX_STAGES) stages = TC_MAX_STAGES;
    if (stages > kblocks) stages = kblocks;
    MZGO_CHECK_ARG(stages >= 2, "shared memory layout of %u B leaves no room for the weight ring", L0.total);
    a.stages = stages;
    const TcSmemLayout L = tc_smem_layout(n, CIN, COUT, stages, MODE == MODE_DYN);
    // CTA pairs (clusters of two, the two SMs of a TPC): an even grid, at most one CTA per SM
    const int pairs = std::max(1, std::min((B + 1) / 2, num_sms() / 2));
    MZGO_CUDA(mzgo_launch_pdl_cluster(k_tc_conv<CIN, COUT, MODE>, dim3(2 * pairs), dim3(TC_THREADS), L.total, st, 2u, a));
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

}  // namespace
// Debug aid (tools/tc_trace.py): when set, every k_tc_conv launch stamps its role timeline into buf[grid][64].
extern "C" int mzgo_debug_tc_trace(void* buf) {
    g_tc_trace = (unsigned long long*)buf;
    return MZGO_OK;
}
namespace {

int fill_heads(const mzgo_net_t* net, TcArgs& a) {
    std::lock_guard<std::mutex> lk(g_heads_mu);
    auto it = g_heads.find(net->workspace);
    MZGO_CHECK_ARG(it != g_heads.end(), "mzgo_net_prepare was not called for this workspace");
    memcpy(a.hw, it->second.hw, sizeof(a.hw));
    a.sc = it->second.sc;
    return 0;
}

}  // namespace

// Synchronous on `st` (it downloads 1.7 KB of head parameters for the kernel-argument constants).
int mzgo_net_prepare_tc(const mzgo_net_t* net, cudaStream_t st) {
    char* ws = (char*)net->workspace;
    NetWorkspace w = mzgo_net_layout(net->n, net->C, net->precision, net->max_batch);
    k_tc_prep_w<<<144, 256, 0, st>>>(net->dyn_w, (__nv_bfloat16*)(ws + w.tc_w), TC_C, TC_C);
    k_tc_prep_w<<<72, 256, 0, st>>>(net->rep_w2, (__nv_bfloat16*)(ws + w.tc_w2), 64, 64);
    k_tc_prep_w<<<72, 256, 0, st>>>(net->rep_w3, (__nv_bfloat16*)(ws + w.tc_w3), 64, TC_C);
    k_tc_prep_tbl<<<296, 128, 0, st>>>(net->dyn_w, net->dyn_b, net->emb, (float*)(ws + w.tc_embbias), net->A);
    k_tc_gather_heads<<<1, 128, 0, st>>>(*net, (float*)(ws + w.tc_headw));
    mzgo_count_launches(5);
    MZGO_LAUNCH_CHECK();
    // the bf16 activation slabs of the representation network rely on zero padding that is never rewritten
    const size_t slab64 = (size_t)8 * tc_psh(net->n) * 16;
    MZGO_CUDA(cudaMemsetAsync(ws + w.tc_act1, 0, slab64 * net->max_batch, st));
    MZGO_CUDA(cudaMemsetAsync(ws + w.tc_act2, 0, slab64 * net->max_batch, st));
    HostHeads h;
    static_assert(sizeof(TcScalars) == 56 * 4, "TcScalars layout");
    float tmp[3 * TC_C + 56];
    MZGO_CUDA(cudaMemcpyAsync(tmp, ws + w.tc_headw, sizeof(tmp), cudaMemcpyDeviceToHost, st));
    MZGO_CUDA(cudaStreamSynchronize(st));
    memcpy(h.hw, tmp, sizeof(h.hw));
    memcpy(&h.sc, tmp + 3 * TC_C, sizeof(TcScalars));
    std::lock_guard<std::mutex> lk(g_heads_mu);
    g_heads[net->workspace] = h;
    return MZGO_OK;
}

int mzgo_tc_pack_latent(const mzgo_net_t* net, const float* src, long src_stride, void* arena, const int32_t* slot,
                        int B, cudaStream_t st) {
    dim3 grid(ceil_div((TC_C / 8) * tc_psh(net->n), 256), B);
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

// initial_inference in bf16 mode: conv1 on CUDA cores, conv2 / conv3 + prediction heads on tcgen05
int mzgo_tc_initial(const mzgo_net_t* net, const float* obs, const uint32_t* boards, void* arena,
                    const int32_t* out_slot, float* value, float* policy, int want_softmax, int B, cudaStream_t st) {
    char* ws = (char*)net->workspace;
    NetWorkspace w = mzgo_net_layout(net->n, net->C, net->precision, net->max_batch);
    __nv_bfloat16* act1 = (__nv_bfloat16*)(ws + w.tc_act1);
    __nv_bfloat16* act2 = (__nv_bfloat16*)(ws + w.tc_act2);
    k_rep_conv1_bf16<<<B, C1_THREADS, 0, st>>>(obs, boards, out_slot, net->rep_w1, net->rep_b1, act1, net->n);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    TcArgs a{};
    a.in = act1; a.out = act2; a.action = out_slot;          // gate: items without an output slot are skipped
    a.wB = (const __nv_bfloat16*)(ws + w.tc_w2);
    a.tbl = net->rep_b2;
    if (int e = launch_tc_conv<64, 64, MODE_PLAIN>(a, net->n, B, st)) return e;
    TcArgs b{};
    b.in = act2; b.out = (__nv_bfloat16*)arena; b.out_slot = out_slot;
    b.wB = (const __nv_bfloat16*)(ws + w.tc_w3);
    b.tbl = net->rep_b3;
    b.value = value; b.policy = policy; b.want_softmax = want_softmax;
    if (int e = fill_heads(net, b)) return e;
    return launch_tc_conv<64, TC_C, MODE_HEADS>(b, net->n, B, st);
}

int mzgo_tc_recurrent(const mzgo_net_t* net, void* arena, const int32_t* in_slot, const int32_t* action,
                      const int32_t* out_slot, float* reward, float* value, float* policy, int want_softmax, int B,
                      cudaStream_t st) {
    char* ws = (char*)net->workspace;
    NetWorkspace w = mzgo_net_layout(net->n, net->C, net->precision, net->max_batch);
    TcArgs a{};
    a.in = (const __nv_bfloat16*)arena;
    a.out = (__nv_bfloat16*)arena;
    a.in_slot = in_slot; a.action = action; a.out_slot = out_slot;
    a.wB = (const __nv_bfloat16*)(ws + w.tc_w);
    a.tbl = (const float*)(ws + w.tc_embbias);
    a.reward = reward; a.value = value; a.policy = policy;
    a.want_softmax = want_softmax;
    if (int e = fill_heads(net, a)) return e;
    return launch_tc_conv<TC_C, TC_C, MODE_DYN>(a, net->n, B, st);
}

// mzgo_net_recurrent on the tree's selected leaves with the tree step behind it in the same launch (every CTA
// expands + backs up the leaves it evaluated and, if do_select, selects its games' next leaves)
int mzgo_tc_recurrent_tree(const mzgo_net_t* net, void* arena, const mzgo_tree_t* tree, const mzgo_search_params_t* sp,
                           const uint32_t* choice_stream, int stride, float* reward, float* value, float* policy,
                           int do_select, cudaStream_t st) {
    if (int e = ensure_rcp_table()) return e;
    char* ws = (char*)net->workspace;
    NetWorkspace w = mzgo_net_layout(net->n, net->C, net->precision, net->max_batch);
    TcArgs a{};
    a.in = (const __nv_bfloat16*)arena;
    a.out = (__nv_bfloat16*)arena;
    a.in_slot = tree->leaf_in_slot; a.action = tree->leaf_action; a.out_slot = tree->leaf_out_slot;
    a.wB = (const __nv_bfloat16*)(ws + w.tc_w);
    a.tbl = (const float*)(ws + w.tc_embbias);
    a.reward = reward; a.value = value; a.policy = policy;
    a.want_softmax = 1;
    a.tree_mode = do_select ? 1 : 2;
    a.choice_stream = choice_stream;
    a.choice_stride = stride;
    a.tree = *tree;
    a.sp = *sp;
    if (int e = fill_heads(net, a)) return e;
    return launch_tc_conv<TC_C, TC_C, MODE_DYN>(a, net->n, tree->G, st);
}
