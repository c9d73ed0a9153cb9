// Shared helpers for the mzgo CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/mzgo.h"

#define FULL_MASK 0xffffffffu

// ---- error plumbing (thread-local message, never throws across the C ABI) ----
void mzgo_set_error(const char* fmt, ...);

#define MZGO_CHECK_ARG(cond, ...)                         \
    do {                                                  \
        if (!(cond)) {                                    \
            mzgo_set_error(__VA_ARGS__);                  \
            return MZGO_ERR_ARG;                          \
        }                                                 \
    } while (0)

#define MZGO_CUDA(call)                                                              \
    do {                                                                             \
        cudaError_t _e = (call);                                                     \
        if (_e != cudaSuccess) {                                                     \
            mzgo_set_error("%s:%d %s: %s", __FILE__, __LINE__, #call,                \
                           cudaGetErrorString(_e));                                  \
            return MZGO_ERR_CUDA;                                                    \
        }                                                                            \
    } while (0)

#define MZGO_LAUNCH_CHECK()  MZGO_CUDA(cudaGetLastError())

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// ---- checked build (python -m muzero_go_b200.build --checked): device-side bounds assertions on every node / edge /
// slot / item index the kernels compute.  compute-sanitizer is closed on this GPU pool (profiles/r02/NOTES.md); the
// whole GPU test suite run against this build (a failed assertion prints file:line and traps the kernel) is the
// substitute.  No code in the normal build.
#ifdef MZGO_CHECKED
#include <assert.h>
#define MZGO_DCHECK(cond) assert(cond)
#else
#define MZGO_DCHECK(cond) ((void)0)
#endif

// ---- programmatic dependent launch (PDL) ----
// The kernels of the search loop (tree step, network) run back to back, each consuming what the one before wrote.
// Launched with the programmatic-serialisation attribute, the next kernel's blocks are scheduled as soon as every
// block of the running kernel has passed griddep_launch() and SM resources free up, do their private set-up
// (barriers, TMEM allocation) and then sit in griddep_wait() until the kernel before them has completed and its
// writes are visible.  Nothing produced by an earlier kernel may be read before griddep_wait().  Both instructions
// are no-ops in a launch without the attribute.  MZGO_NO_PDL=1 switches the attribute off (A/B measurements).
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool mzgo_pdl_enabled();
void mzgo_pdl_disable();

// cluster_x > 1 launches thread-block clusters of that many CTAs along x (grid.x must be a multiple of it).
template <typename... KArgs, typename... Args>
static inline cudaError_t mzgo_launch_pdl_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                                  unsigned cluster_x, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    unsigned na = 0;
    if (cluster_x > 1) {
        at[na].id = cudaLaunchAttributeClusterDimension;
        at[na].val.clusterDim.x = cluster_x;
        at[na].val.clusterDim.y = 1;
        at[na].val.clusterDim.z = 1;
        na++;
    }
    const unsigned base = na;
    if (mzgo_pdl_enabled()) {
        at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[na].val.programmaticStreamSerializationAllowed = 1;
        na++;
    }
    cfg.attrs = at;
    cfg.numAttrs = na;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
    if (e != cudaSuccess && na > base) {         // a driver / stream that refuses the attribute: plain launches from now on
        cudaGetLastError();
        mzgo_pdl_disable();
        cfg.numAttrs = base;
        e = cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
    }
    return e;
}

template <typename... KArgs, typename... Args>
static inline cudaError_t mzgo_launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                          Args&&... args) {
    return mzgo_launch_pdl_cluster(kernel, grid, block, smem, st, 1u, static_cast<Args&&>(args)...);
}

// ---- warp helpers ----
__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o);
    return v;
}
__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o);
    return v;
}
__device__ __forceinline__ float warp_max_f(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(FULL_MASK, v, o));
    return v;
}

// ---- counter-based RNG (Philox4x32-10), keyed by (seed, game), counter = (draw index, stream) ----
__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)M0 * c[0];
        uint64_t p1 = (uint64_t)M1 * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += W0; k1 += W1;
    }
}

__host__ __device__ __forceinline__ uint32_t mzgo_rand_u32(uint64_t seed, uint32_t game, uint32_t stream,
                                                             uint32_t index) {
    uint32_t c[4] = {index, stream, game, 0x6d7a676fu};
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    return c[0];
}
