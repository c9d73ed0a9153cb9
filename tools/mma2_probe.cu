// Probe for tcgen05.mma.cta_group::2 (a CTA pair, M = 256) with the conv kernel's operand layout (K-major, no swizzle):
//   1. semantics: which rows of A / B each CTA of the pair supplies and where D lands (checked against a host matmul);
//   2. the plumbing the conv kernel needs: tcgen05.alloc.cta_group::2, a remote mbarrier arrive (mapa) from the peer
//      to the leader, tcgen05.commit ... multicast::cluster to both CTAs;
//   3. cycles per MMA, alone and with the other warps of both CTAs hammering shared memory (the conv kernel's
//      epilogue / TMA traffic stands in as plain ld/st.shared), next to the cta_group::1 figure of tools/mma_probe.cu.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/mma2_probe tools/mma2_probe.cu && gpurun_out/mma2_probe
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include <cuda_bf16.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t cta) {
    asm volatile("{\n\t.reg .b32 r;\n\tmapa.shared::cluster.u32 r, %0, %1;\n\tmbarrier.arrive.shared::cluster.b64 _, [r];\n\t}" ::"r"(
                     smem_u32(bar)), "r"(cta) : "memory");
}

constexpr int N = 128, NH = N / 2, KB = 4;   // KB MMAs of K = 16 per "k-block" (K = 64), as in the conv kernel
// per CTA: A [8 chunks][128 rows][8] bf16 = 16 KB (LBO 2048, SBO 128); B half [8 chunks][64 rows][8] = 8 KB (LBO 1024, SBO 128)
struct Out { float d[2][128][N]; long long cycles[4]; };

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(256, 1)
probe2(const __nv_bfloat16* __restrict__ A /*[256][64]*/, const __nv_bfloat16* __restrict__ B /*[128][64]*/, Out* out, int iters,
       int noise) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar_ready, bar_done;
    __shared__ uint32_t s_tmem;
    const uint32_t rank = cluster_rank();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __nv_bfloat16* sA = reinterpret_cast<__nv_bfloat16*>(smem);
    __nv_bfloat16* sB = reinterpret_cast<__nv_bfloat16*>(smem + 16384);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 2;" ::"r"(smem_u32(&bar_ready)));   // own fill + the peer's
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar_done)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // this CTA's operands: rows [128 rank, 128 rank + 128) of A, rows [64 rank, 64 rank + 64) of B
    for (int i = threadIdx.x; i < 128 * 64; i += blockDim.x) {
        const int r = i / 64, k = i % 64;
        sA[((k >> 3) * 128 + r) * 8 + (k & 7)] = A[(size_t)(rank * 128 + r) * 64 + k];
    }
    for (int i = threadIdx.x; i < NH * 64; i += blockDim.x) {
        const int r = i / 64, k = i % 64;
        sB[((k >> 3) * NH + r) * 8 + (k & 7)] = B[(size_t)(rank * NH + r) * 64 + k];
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy fills -> tensor-core (async proxy) reads
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&s_tmem)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync();                                   // barriers initialised and TMEM allocated in both CTAs
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s_tmem;
    // "operands landed": both CTAs tell the LEADER's barrier (the peer through a remote arrive)
    if (threadIdx.x == 0) {
        if (rank == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&bar_ready)) : "memory");
        else mbar_arrive_remote(&bar_ready, 0);
    }
    long long t0 = 0, t1 = 0;
    if (warp == 0 && rank == 0) {
        mbar_wait(&bar_ready, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (lane == 0) {
            // kind::f16: D fp32, A/B bf16, K-major both, N at bits 17.., M >> 4 at bits 24..: M = 256 for the pair
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((256u >> 4) << 24);
            const uint32_t hi = ((128u >> 4) & 0x3FFF) | (1u << 14);
            const uint32_t a_lo0 = ((smem_u32(sA) >> 4) & 0x3FFF) | (((2048u >> 4) & 0x3FFF) << 16);
            const uint32_t b_lo0 = ((smem_u32(sB) >> 4) & 0x3FFF) | ((((uint32_t)NH * 16u >> 4) & 0x3FFF) << 16);
            t0 = clock64();
            for (int it = 0; it < iters; it++) {
#pragma unroll
                for (int ks = 0; ks < KB; ks++) {
                    const uint32_t a_lo = a_lo0 + ks * (2 * 2048 / 16), b_lo = b_lo0 + ks * (2 * NH * 16 / 16);
                    const uint32_t acc = (it | ks) != 0 ? 1u : 0u;
                    const uint32_t d = tmem + (uint32_t)((it & 3) * N) * (iters > 1 ? 1 : 0);
                    asm volatile(
                        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tmov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"
                        "setp.ne.b32 p, %6, 0;\n\t"
                        "tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %5, p;\n\t}" ::"r"(d), "r"(a_lo), "r"(hi), "r"(b_lo),
                        "r"(hi), "r"(idesc), "r"(iters > 1 ? 1u : acc) : "memory");
                }
            }
            // completion of everything above -> bar_done of BOTH CTAs
            asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                             smem_u32(&bar_done)), "h"((uint16_t)3) : "memory");
        }
        __syncwarp();
    }
    if (noise && warp >= 4) {
        // stand-in for the conv kernel's other shared-memory traffic while the MMAs run (warps 4-7 of both CTAs)
        volatile uint32_t* p = reinterpret_cast<volatile uint32_t*>(smem + 32768) + (threadIdx.x - 128);
        uint32_t acc = 0;
        for (int i = 0; i < noise; i++) {
            acc ^= p[(i & 7) * 128] ^ p[(i & 7) * 128 + 1024] ^ p[(i & 7) * 128 + 2048] ^ p[(i & 7) * 128 + 3072];
            if ((i & 3) == 3) p[(i & 7) * 128] = acc;
        }
        if (acc == 0x12345u) out->cycles[3] = 1;
    }
    if (warp < 4) {
        mbar_wait(&bar_done, 0);
        if (warp == 0 && rank == 0 && lane == 0) { t1 = clock64(); out->cycles[noise ? 1 : 0] = t1 - t0; }
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (iters == 1) {
            // each warp reads its 32 TMEM lanes x 128 columns
            for (int c0 = 0; c0 < N; c0 += 8) {
                uint32_t v[8];
                const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + c0;
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                             : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                             : "r"(taddr));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                for (int e = 0; e < 8; e++) out->d[rank][warp * 32 + lane][c0 + e] = __uint_as_float(v[e]);
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync();                                   // nobody leaves (or frees TMEM) while the peer may still use this CTA
    if (warp == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

int main() {
    std::vector<__nv_bfloat16> hA(256 * 64), hB(128 * 64);
    std::vector<float> fA(256 * 64), fB(128 * 64);
    srand(1);
    for (size_t i = 0; i < hA.size(); i++) { fA[i] = (float)(rand() % 7 - 3); hA[i] = __float2bfloat16(fA[i]); }
    for (size_t i = 0; i < hB.size(); i++) { fB[i] = (float)(rand() % 5 - 2); hB[i] = __float2bfloat16(fB[i]); }
    __nv_bfloat16 *dA, *dB;
    Out* dO;
    cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dB, hB.size() * 2); cudaMalloc(&dO, sizeof(Out));
    cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
    cudaMemset(dO, 0, sizeof(Out));
    cudaFuncSetAttribute(probe2, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    probe2<<<2, 256, 96 * 1024>>>(dA, dB, dO, 1, 0);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("probe2 (check): %s\n", cudaGetErrorString(e)); return 1; }
    static Out hO;
    cudaMemcpy(&hO, dO, sizeof(Out), cudaMemcpyDeviceToHost);
    // hypothesis: D[rank][r][c] = sum_k A[128 rank + r][k] * B[c][k], B rows [0,64) from CTA 0 and [64,128) from CTA 1
    long bad = 0;
    for (int rk = 0; rk < 2; rk++)
        for (int r = 0; r < 128; r++)
            for (int c = 0; c < N; c++) {
                float s = 0;
                for (int k = 0; k < 64; k++) s += fA[(rk * 128 + r) * 64 + k] * fB[c * 64 + k];
                if (s != hO.d[rk][r][c]) {
                    if (bad < 5) printf("mismatch rank %d row %d col %d: got %g want %g\n", rk, r, c, hO.d[rk][r][c], s);
                    bad++;
                }
            }
    printf("cta_group::2 M=256 N=128 K=64: %ld mismatches of %d (A rows split by CTA, B rows (N) split by CTA, D rows per CTA)\n", bad,
           2 * 128 * N);
    const int iters = 2000;
    for (int noise : {0, 30000}) {
        for (int rep = 0; rep < 2; rep++) {
            probe2<<<148, 256, 96 * 1024>>>(dA, dB, dO, iters, noise);
            e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("probe2 (timing): %s\n", cudaGetErrorString(e)); return 1; }
        }
        cudaMemcpy(&hO, dO, sizeof(Out), cudaMemcpyDeviceToHost);
        printf("cta_group::2 M=256 N=128 K=16, 74 pairs, %s: %.1f cycles/MMA (per CTA: 128x128x16 per MMA; cta_group::1 needs 64.1)\n",
               noise ? "other warps streaming ld/st.shared" : "nothing else on shared memory",
               (double)hO.cycles[noise ? 1 : 0] / (iters * KB));
    }
    return bad != 0;
}
