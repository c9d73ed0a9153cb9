// Microbenchmark: cycles per tcgen05.mma (kind::f16, bf16 x bf16 -> fp32, cta_group::1, M=128, K=16)
// issued back to back by one thread on garbage shared memory, for several N / layouts / LBO strides.
// Calibrates the floor the conv kernel can reach (shared-memory operand fetch vs tensor pipe).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/mma_probe tools/mma_probe.cu && gpurun_out/mma_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

struct Cfg { int n_cols; int layout; uint32_t a_lbo, a_sbo, b_lbo, b_sbo; int a_step, b_step; int shift_units; int slots; int blocks; };

__global__ void __launch_bounds__(128, 1) probe(Cfg c, int iters, long long* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t s_tmem;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&s_tmem)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s_tmem;
    if (threadIdx.x == 0) {
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(c.n_cols >> 3) << 17) | ((128u >> 4) << 24);
        const uint32_t a_hi = ((c.a_sbo >> 4) & 0x3FFF) | (1u << 14) | ((uint32_t)c.layout << 29);
        const uint32_t b_hi = ((c.b_sbo >> 4) & 0x3FFF) | (1u << 14) | ((uint32_t)c.layout << 29);
        const uint32_t a_lo0 = ((smem_u32(smem) >> 4) & 0x3FFF) | (((c.a_lbo >> 4) & 0x3FFF) << 16);
        const uint32_t b_lo0 = (((smem_u32(smem) + 112 * 1024) >> 4) & 0x3FFF) | (((c.b_lbo >> 4) & 0x3FFF) << 16);
        long long t0 = clock64();
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const uint32_t a_lo = a_lo0 + (uint32_t)((k & 3) * c.a_step + (k >> 2) * c.shift_units);
                const uint32_t b_lo = b_lo0 + (uint32_t)((k & 3) * c.b_step);
                const uint32_t d = tmem + (uint32_t)((k % c.slots) * c.n_cols);
                asm volatile(
                    "{\n\t.reg .b64 da, db;\n\tmov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"
                    "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, 1;\n\t}" ::"r"(d), "r"(a_lo), "r"(a_hi), "r"(b_lo),
                    "r"(b_hi), "r"(idesc) : "memory");
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        uint32_t ok = 0;
        while (!ok) {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok) : "r"(smem_u32(&bar)) : "memory");
        }
        long long t1 = clock64();
        if (blockIdx.x == 0) out[0] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

int main() {
    long long* d_out;
    cudaMalloc(&d_out, 8);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int iters = 500;
    struct Named { const char* name; Cfg c; };
    const int ps = 422;
    Named cfgs[] = {
        // no swizzle, conv-kernel layout: A LBO = ps*16, SBO 128; B LBO = N*16, SBO 128; K step = 2 chunks
        {"noswz N=128 conv layout (1 SM)   ", {128, 0, ps * 16u, 128, 128 * 16u, 128, 2 * ps, 2 * 128, 1, 4, 1}},
        {"noswz N=128 conv layout (148 SMs)", {128, 0, ps * 16u, 128, 128 * 16u, 128, 2 * ps, 2 * 128, 1, 4, 148}},
        {"noswz N=128 aligned shifts       ", {128, 0, ps * 16u, 128, 128 * 16u, 128, 2 * ps, 2 * 128, 8, 4, 148}},
        {"noswz N=128 A LBO=128 SBO=256    ", {128, 0, 128, 256, 128 * 16u, 128, 16, 2 * 128, 0, 4, 148}},
        {"noswz N=128 A,B LBO=128 SBO=256  ", {128, 0, 128, 256, 128, 256, 16, 16, 0, 4, 148}},
        {"noswz N=64  conv layout          ", {64, 0, ps * 16u, 128, 64 * 16u, 128, 2 * ps, 2 * 64, 1, 4, 148}},
        {"noswz N=256 conv layout          ", {256, 0, ps * 16u, 128, 256 * 16u, 128, 2 * ps, 2 * 256, 1, 2, 148}},
        // 128B swizzle, K-major: rows of 128 B, SBO = 1024, K step = 32 B
        {"swz128 N=128                     ", {128, 2, 16, 1024, 16, 1024, 2, 2, 8, 4, 148}},
        {"swz128 N=128 row-shifted A       ", {128, 2, 16, 1024, 16, 1024, 2, 2, 8 * 3, 4, 148}},
        {"swz128 N=256                     ", {256, 2, 16, 1024, 16, 1024, 2, 2, 8, 2, 148}},
        {"swz128 N=64                      ", {64, 2, 16, 1024, 16, 1024, 2, 2, 8, 4, 148}},
    };
    for (auto& nc : cfgs) {
        for (int rep = 0; rep < 2; rep++) {
            probe<<<nc.c.blocks, 128, 200 * 1024>>>(nc.c, iters, d_out);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("%s: %s\n", nc.name, cudaGetErrorString(e)); return 1; }
        }
        long long cyc;
        cudaMemcpy(&cyc, d_out, 8, cudaMemcpyDeviceToHost);
        const double per = (double)cyc / (iters * 8);
        printf("%s  %8.1f cycles/MMA   %6.0f flop/clk/SM  (ideal %d)\n", nc.name, per, 2.0 * 128 * nc.c.n_cols * 16 / per,
               128 * nc.c.n_cols / 256);
    }
    // clock behaviour under sustained tensor load: same stream of MMAs for ~0.05 s / ~1 s on all SMs, wall-clocked
    {
        Cfg c = cfgs[1].c;
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int big : {20000, 400000, 400000}) {
            cudaEventRecord(e0);
            probe<<<148, 128, 200 * 1024>>>(c, big, d_out);
            cudaEventRecord(e1);
            cudaDeviceSynchronize();
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            long long cyc; cudaMemcpy(&cyc, d_out, 8, cudaMemcpyDeviceToHost);
            const double flops = 148.0 * big * 8 * 2.0 * 128 * 128 * 16;
            printf("sustained N=128 x %d iters: %.1f ms, %.0f TFLOP/s, effective SM clock %.0f MHz, %.1f cycles/MMA\n", big, ms,
                   flops / ms / 1e9, cyc / (ms * 1e3), (double)cyc / (big * 8.0));
        }
    }
    return 0;
}
