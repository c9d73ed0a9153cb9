// Arithmetic shared by the tree kernels (tree.cu) and the tree role of the fused network kernel (nets_tc.cu).
// Everything here reproduces the reference's dtypes operation for operation (SURVEY.md 8a "dtype facts").
// Included in an anonymous namespace per translation unit: each unit owns a copy of the reciprocal table.
#pragma once
#include "common.cuh"

namespace {

__device__ __forceinline__ double warp_min_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(FULL_MASK, v, o));
    return v;
}
__device__ __forceinline__ double warp_max_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(FULL_MASK, v, o));
    return v;
}

// numpy's float32 pairwise summation (np.add.reduce on a contiguous array, numpy 2.x):
// blocks of <= 128 use 8 strided accumulators, longer inputs split in halves (multiple of 8).
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }

// numpy's base case: fewer than 8 elements in order, else 8 strided accumulators combined pairwise, then the tail
template <typename T>
__device__ __forceinline__ T np_block_sum(const T* a, int n) {
    if (n < 8) {
        T r = (T)0;
        for (int i = 0; i < n; i++) r = add_rn(r, a[i]);
        return r;
    }
    T r[8];
#pragma unroll
    for (int k = 0; k < 8; k++) r[k] = a[k];
    int i = 8;
    for (; i < n - (n % 8); i += 8) {
#pragma unroll
        for (int k = 0; k < 8; k++) r[k] = add_rn(r[k], a[i + k]);
    }
    T res = add_rn(add_rn(add_rn(r[0], r[1]), add_rn(r[2], r[3])), add_rn(add_rn(r[4], r[5]), add_rn(r[6], r[7])));
    for (; i < n; i++) res = add_rn(res, a[i]);
    return res;
}

// The recursion sum(a, n) = sum(a, n2) + sum(a + n2, n - n2), n2 = n/2 rounded down to a multiple of 8, for
// n > 128, run on an explicit stack: no device function call (ptxas ignores setmaxnreg in kernels with calls).
template <typename T>
__device__ __forceinline__ T np_pairwise_sum(const T* a, int n) {
    constexpr int DEPTH = 12;                   // n <= 128 * 2^11
    const T* ptr[DEPTH];
    int cnt[DEPTH], st[DEPTH];
    T acc[DEPTH];
    int top = 0;
    ptr[0] = a; cnt[0] = n; st[0] = 0;
    T ret = (T)0;
    while (top >= 0) {
        int n2 = cnt[top] / 2;
        n2 -= n2 % 8;
        if (st[top] == 0) {
            if (cnt[top] <= 128 || top == DEPTH - 1) { ret = np_block_sum(ptr[top], cnt[top]); top--; }
            else { st[top] = 1; ptr[top + 1] = ptr[top]; cnt[top + 1] = n2; st[top + 1] = 0; top++; }
        } else if (st[top] == 1) {
            acc[top] = ret;
            st[top] = 2; ptr[top + 1] = ptr[top] + n2; cnt[top + 1] = cnt[top] - n2; st[top + 1] = 0; top++;
        } else {
            ret = add_rn(acc[top], ret);
            top--;
        }
    }
    return ret;
}
__device__ __forceinline__ float np_pairwise_sum_f32(const float* a, int n) { return np_pairwise_sum<float>(a, n); }

// self_play.py:210: policy_child *= valid_mask -- float32 * float64 in float64, rounded back to float32.
// The float64 mask is 0 / 1 on the board and pass_epsilon (or 1.0 when no board move is legal) for the pass.
__device__ __forceinline__ float masked_child_prior(float soft, float mask_f32, bool is_pass, double pass_epsilon) {
    const double m = is_pass ? (mask_f32 == 1.0f ? 1.0 : pass_epsilon) : (double)mask_f32;
    return (float)__dmul_rn((double)soft, m);
}
// self_play.py:211-222.  tot > 0: float32 divide.  tot == 0 (every legal move underflowed): the reference
// switches to a float64 uniform distribution over the mask; it is stored rounded to float32 here (the only
// place this variant departs from the reference's dtypes; unreachable with finite logits).
__device__ __forceinline__ float renorm_child_prior(float pr, float tot, float mask_f32, bool is_pass, double pass_epsilon,
                                                    const float* mask, int A) {
    if (tot > 0.f) return __fdiv_rn(pr, tot);
    double ms = 0.0;
    for (int a = 0; a < A; a++) ms += a == A - 1 ? (mask[a] == 1.0f ? 1.0 : pass_epsilon) : (double)mask[a];
    const double m = is_pass ? (mask_f32 == 1.0f ? 1.0 : pass_epsilon) : (double)mask_f32;
    return (float)(m / ms);
}

// ---- correctly rounded float64 division with a known reciprocal (Markstein 1990; Muller et al., Handbook of
// Floating-Point Arithmetic, 2nd ed., section 4.7): with y = RN(1/b), q0 = RN(a*y) is within 1.5 ulp of a/b,
// one FMA correction makes it faithful, and a second correction q2 = RN(q1 + (a - b*q1)*y) IS RN(a/b) whenever
// nothing under- or overflows (here: |a| in [1e-60, 1e6], b in [1e-17, 1e3]).  Five dependent FP64 operations
// and no branches instead of the ~20-instruction generic division; the PUCT score needs two per edge.
// tests/test_tree_gpu.py::test_fast_division_is_correctly_rounded compares it with __ddiv_rn on 2^27 operands.
__device__ __forceinline__ double div_rn_with_rcp(double a, double b, double y) {
    double q = __dmul_rn(a, y);
    double r = __fma_rn(-b, q, a);
    q = __fma_rn(r, y, q);
    r = __fma_rn(-b, q, a);
    return __fma_rn(r, y, q);
}
// RN(1/d) for d = 0..RCP_TABLE-1 (filled by the host: 1.0 / d in IEEE double is correctly rounded)
constexpr int RCP_TABLE = 1024;
__device__ double g_rcp_table[RCP_TABLE];

// The generic IEEE division, out of line: the tree kernels reach it only outside the proven window of
// div_rn_with_rcp, and inlined at every site it was a sixth of their code (they are instruction-cache sensitive).
__device__ __noinline__ double ddiv_generic(double a, double b) { return __ddiv_rn(a, b); }

__device__ __forceinline__ double div_by_count(double a, int d) {
    if (d < RCP_TABLE) return div_rn_with_rcp(a, (double)d, __ldg(&g_rcp_table[d]));
    return ddiv_generic(a, (double)d);
}

// Fills this translation unit's reciprocal table once per device and host thread.
static int ensure_rcp_table() {
    static thread_local int done_for = -1;
    int dev = 0;
    MZGO_CUDA(cudaGetDevice(&dev));
    if (done_for == dev) return MZGO_OK;
    static double host[RCP_TABLE];
    host[0] = 0.0;
    for (int d = 1; d < RCP_TABLE; d++) host[d] = 1.0 / (double)d;
    MZGO_CUDA(cudaMemcpyToSymbol(g_rcp_table, host, sizeof(host)));
    done_for = dev;
    return MZGO_OK;
}


}  // namespace
