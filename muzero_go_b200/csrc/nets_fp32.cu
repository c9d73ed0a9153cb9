// fp32 network path (CUDA cores): the 1e-5 "reference mode" of the representation / dynamics /
// prediction networks (/root/reference/main.py:72-142) and the representation network of the bf16 mode.
//
// conv3x3: one block = one image x one pixel tile; thread tile 8 output channels x 8 pixels, input
// channels staged through shared memory four at a time (zero-padded image + [ci][tap][co] weights).
#include "common.cuh"
#include "nets.cuh"

namespace {

constexpr int CONV_THREADS = 256;
constexpr int CI_CHUNK = 4;
constexpr int CO_PER_THREAD = 8;
constexpr int PX_PER_THREAD = 8;

struct ConvArgs {
    const float* in;        // item i at in + idx_in(i) * in_stride
    const int32_t* in_idx;  // optional slot indices
    long in_stride;         // floats
    const uint32_t* boards; // optional: input planes expanded from packed boards (Cin == 6)
    const float* emb;       // optional [A][Cin]: added to every in-board pixel (main.py:98-100)
    const int32_t* action;  // optional; action < 0 skips the item
    const float* wT;        // [Cin][9][Cout]
    const float* bias;      // [Cout]
    float* out;             // item i at out + idx_out(i) * out_stride
    const int32_t* out_idx; // optional; < 0 skips the item
    long out_stride;
    int n, Cin, Cout;
};

__global__ void __launch_bounds__(CONV_THREADS)
k_conv3x3_relu(ConvArgs a) {
    extern __shared__ float smem[];
    const int n = a.n, np = n * n, pw = n + 2, pp = pw * pw;
    float* xs = smem;                    // [CI_CHUNK][pp]
    float* ws = smem + CI_CHUNK * pp;    // [CI_CHUNK][9][Cout]
    const int item = blockIdx.x;
    const int act = a.action ? a.action[item] : 0;
    const int oslot = a.out_idx ? a.out_idx[item] : item;
    if (act < 0 || oslot < 0) return;
    const int islot = a.in_idx ? a.in_idx[item] : item;
    const float* in = a.in ? a.in + (long)islot * a.in_stride : nullptr;
    const uint32_t* rec = a.boards ? a.boards + (size_t)item * MZGO_BOARD_WORDS : nullptr;
    const int TC = a.Cout / CO_PER_THREAD;        // threads along output channels
    const int TPIX = CONV_THREADS / TC;           // threads along pixels
    const int tc = threadIdx.x % TC, tp = threadIdx.x / TC;   // tp >= TPIX: spare thread (Cout/8 not a divisor of 256)
    const int pix0 = blockIdx.y * TPIX * PX_PER_THREAD;
    int off[PX_PER_THREAD];
    bool ok[PX_PER_THREAD];
#pragma unroll
    for (int k = 0; k < PX_PER_THREAD; k++) {
        int p = pix0 + tp + k * TPIX;
        ok[k] = p < np && tp < TPIX;
        int r = ok[k] ? p / n : 0, c = ok[k] ? p - r * n : 0;
        off[k] = (r + 1) * pw + (c + 1);
    }
    float acc[CO_PER_THREAD][PX_PER_THREAD];
#pragma unroll
    for (int i = 0; i < CO_PER_THREAD; i++)
#pragma unroll
        for (int k = 0; k < PX_PER_THREAD; k++) acc[i][k] = 0.f;

    for (int c0 = 0; c0 < a.Cin; c0 += CI_CHUNK) {
        __syncthreads();
        for (int i = threadIdx.x; i < CI_CHUNK * pp; i += CONV_THREADS) {
            int ci = i / pp, q = i - ci * pp;
            int r = q / pw - 1, c = q - (q / pw) * pw - 1;
            float v = 0.f;
            int cin = c0 + ci;
            if (cin < a.Cin && r >= 0 && r < n && c >= 0 && c < n) {
                if (rec) {
                    uint32_t bit;
                    if (cin < 2) bit = (rec[cin * n + r] >> c) & 1u;
                    else if (cin == 3) bit = (rec[2 * n + r] >> c) & 1u;
                    else if (cin == 2) bit = rec[60] & 1u;
                    else if (cin == 4) bit = rec[61] & 1u;
                    else bit = (rec[61] >> 1) & 1u;
                    v = (float)bit;
                } else {
                    v = in[(long)cin * np + r * n + c];
                }
                if (a.emb) v += a.emb[(long)act * a.Cin + cin];
            }
            xs[i] = v;
        }
        for (int i = threadIdx.x; i < CI_CHUNK * 9 * a.Cout; i += CONV_THREADS) {
            int ci = i / (9 * a.Cout);
            ws[i] = (c0 + ci < a.Cin) ? a.wT[(long)c0 * 9 * a.Cout + i] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int ci = 0; ci < CI_CHUNK; ci++) {
#pragma unroll
            for (int tap = 0; tap < 9; tap++) {
                const int d = (tap / 3 - 1) * pw + (tap % 3 - 1);
                const float4 w0 = *reinterpret_cast<const float4*>(ws + (ci * 9 + tap) * a.Cout + tc * 8);
                const float4 w1 = *reinterpret_cast<const float4*>(ws + (ci * 9 + tap) * a.Cout + tc * 8 + 4);
                const float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                for (int k = 0; k < PX_PER_THREAD; k++) {
                    const float x = xs[ci * pp + off[k] + d];
#pragma unroll
                    for (int i = 0; i < CO_PER_THREAD; i++) acc[i][k] = fmaf(w[i], x, acc[i][k]);
                }
            }
        }
    }
    float* out = a.out + (long)oslot * a.out_stride;
#pragma unroll
    for (int i = 0; i < CO_PER_THREAD; i++) {
        const int co = tc * 8 + i;
        const float b = a.bias[co];
#pragma unroll
        for (int k = 0; k < PX_PER_THREAD; k++) {
            int p = pix0 + tp + k * TPIX;
            if (ok[k]) out[(long)co * np + p] = fmaxf(acc[i][k] + b, 0.f);
        }
    }
}

// [Cout][Cin][3][3] -> [Cin][9][Cout]
__global__ void k_transpose_conv_w(const float* w, float* wT, int Cin, int Cout) {
    int total = Cin * Cout * 9;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        int co = i % Cout, tap = (i / Cout) % 9, ci = i / (9 * Cout);
        wT[i] = w[((long)co * Cin + ci) * 9 + tap];
    }
}

// Prediction heads (+ the reward head of the dynamics network), main.py:101-105 and :116-125.
// One block per item; x = float32 [C][np].
constexpr int HEAD_THREADS = 128;

__device__ float block_sum(float v, float* red) {
    v = warp_sum_f(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float s = 0.f;
    for (int w = 0; w < HEAD_THREADS / 32; w++) s += red[w];
    return s;
}
__device__ float block_max(float v, float* red) {
    v = warp_max_f(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float s = -INFINITY;
    for (int w = 0; w < HEAD_THREADS / 32; w++) s = fmaxf(s, red[w]);
    return s;
}

__global__ void __launch_bounds__(HEAD_THREADS)
k_heads_fp32(mzgo_net_t net, const float* arena, const int32_t* slot, const int32_t* action, long slot_stride,
             float* reward, float* value, float* policy, int want_softmax) {
    __shared__ float red[HEAD_THREADS / 32];
    __shared__ float logit[MZGO_MAX_N * MZGO_MAX_N + 1];
    const int item = blockIdx.x;
    if (action && action[item] < 0) return;
    const int s = slot ? slot[item] : item;
    if (s < 0) return;
    const int n = net.n, np = n * n, C = net.C, A = np + 1;
    const float* x = arena + (long)s * slot_stride;
    float vsum = 0.f, rsum = 0.f;
    for (int p = threadIdx.x; p < np; p += HEAD_THREADS) {
        float pl = 0.f, vl = 0.f, rl = 0.f;
        for (int c = 0; c < C; c++) {
            float xv = x[(long)c * np + p];
            pl = fmaf(net.pol_w[c], xv, pl);
            vl = fmaf(net.val_w[c], xv, vl);
            rl = fmaf(net.rew_w[c], xv, rl);
        }
        logit[p] = pl + net.pol_b[0];
        vsum += vl + net.val_b[0];
        rsum += rl + net.rew_b[0];
    }
    if (threadIdx.x == 0) logit[np] = net.pass_logit[0];
    vsum = block_sum(vsum, red);
    rsum = block_sum(rsum, red);
    if (threadIdx.x == 0) {
        float vm = vsum / (float)np;
        value[item] = fmaf(net.vfc_w[0], vm, net.vfc_b[0]);
        if (reward) {
            float rm = rsum / (float)np, r = net.ro_b[0];
            for (int j = 0; j < 16; j++) r = fmaf(net.ro_w[j], fmaxf(fmaf(net.rh_w[j], rm, net.rh_b[j]), 0.f), r);
            reward[item] = r;
        }
    }
    __syncthreads();
    float* pol = policy + (long)item * A;
    if (!want_softmax) {
        for (int a = threadIdx.x; a < A; a += HEAD_THREADS) pol[a] = logit[a];
        return;
    }
    float m = -INFINITY;
    for (int a = threadIdx.x; a < A; a += HEAD_THREADS) m = fmaxf(m, logit[a]);
    m = block_max(m, red);
    float se = 0.f;
    for (int a = threadIdx.x; a < A; a += HEAD_THREADS) {
        float e = expf(logit[a] - m);
        logit[a] = e;
        se += e;
    }
    se = block_sum(se, red);
    for (int a = threadIdx.x; a < A; a += HEAD_THREADS) pol[a] = logit[a] / se;
}

__global__ void k_copy_latent_f32(const float* src, const int32_t* src_idx, long src_stride, float* dst,
                                  const int32_t* dst_idx, long dst_stride, long count) {
    const int item = blockIdx.y;
    const int si = src_idx ? src_idx[item] : item, di = dst_idx ? dst_idx[item] : item;
    if (si < 0 || di < 0) return;
    const float* s = src + (long)si * src_stride;
    float* d = dst + (long)di * dst_stride;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < count; i += (long)gridDim.x * blockDim.x) d[i] = s[i];
}

int launch_conv(const ConvArgs& a, int B, cudaStream_t st) {
    const int TC = a.Cout / CO_PER_THREAD;
    const int TPIX = CONV_THREADS / TC;
    const int tile = TPIX * PX_PER_THREAD;
    dim3 grid(B, ceil_div(a.n * a.n, tile));
    size_t smem = (size_t)CI_CHUNK * ((a.n + 2) * (a.n + 2) + 9 * a.Cout) * sizeof(float);
    k_conv3x3_relu<<<grid, CONV_THREADS, smem, st>>>(a);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

}  // namespace

// ---- workspace layout (fp32 part), shared with nets_tc.cu through nets.cuh ----
NetWorkspace mzgo_net_layout(int n, int C, int precision, int max_batch) {
    NetWorkspace w;
    long o = 0;
    auto take = [&](long bytes) { long r = o; o += (bytes + 255) / 256 * 256; return r; };
    w.wT1 = take(6L * 9 * 64 * 4);
    w.wT2 = take(64L * 9 * 64 * 4);
    w.wT3 = take(64L * 9 * C * 4);
    w.wTd = take((long)C * 9 * C * 4);
    // float32 activations of the representation network: fp32 mode only (bf16 mode keeps bf16 slabs)
    w.act1 = precision == 0 ? take((long)max_batch * 64 * n * n * 4) : 0;
    w.act2 = precision == 0 ? take((long)max_batch * 64 * n * n * 4) : 0;
    w.act3 = precision == 1 ? take((long)max_batch * C * n * n * 4) : 0;      // mzgo_net_predict staging
    w.tc_w = precision == 1 ? take((long)C * 9 * C * 2) : 0;
    w.tc_embbias = precision == 1 ? take((long)(n * n + 1) * 9 * C * 4) : 0;
    w.tc_headw = precision == 1 ? take((3L * C + 56) * 4) : 0;
    w.tc_w2 = precision == 1 ? take(64L * 9 * 64 * 2) : 0;
    w.tc_w3 = precision == 1 ? take(64L * 9 * C * 2) : 0;
    const long slab64 = 8L * ((n * (n + 1) - 1 + 7) / 8 * 8) * 16;   // nets_tc.cu: tc_psh positions of 64 bf16 channels
    w.tc_act1 = precision == 1 ? take(slab64 * max_batch) : 0;
    w.tc_act2 = precision == 1 ? take(slab64 * max_batch) : 0;
    w.total = o;
    return w;
}

int mzgo_net_check(const mzgo_net_t* net) {
    MZGO_CHECK_ARG(net != nullptr, "net is NULL");
    MZGO_CHECK_ARG(net->n >= 2 && net->n <= MZGO_MAX_N, "net.n=%d outside [2,%d]", net->n, MZGO_MAX_N);
    MZGO_CHECK_ARG(net->A == net->n * net->n + 1, "net.A=%d != n*n+1", net->A);
    MZGO_CHECK_ARG(net->C >= 8 && net->C % 8 == 0 && net->C <= 256, "net.C=%d must be a multiple of 8 in [8,256]", net->C);
    MZGO_CHECK_ARG(net->precision == 0 || net->precision == 1, "net.precision=%d", net->precision);
    MZGO_CHECK_ARG(net->precision == 0 || net->C == 128, "bf16 mode needs latent_dim 128 (got %d)", net->C);
    MZGO_CHECK_ARG(net->max_batch >= 1, "net.max_batch=%d", net->max_batch);
    MZGO_CHECK_ARG(net->rep_w1 && net->rep_b1 && net->rep_w2 && net->rep_b2 && net->rep_w3 && net->rep_b3 && net->emb &&
                       net->dyn_w && net->dyn_b && net->rew_w && net->rew_b && net->rh_w && net->rh_b && net->ro_w &&
                       net->ro_b && net->val_w && net->val_b && net->vfc_w && net->vfc_b && net->pol_w && net->pol_b &&
                       net->pass_logit,
                   "net has a NULL parameter");
    MZGO_CHECK_ARG(net->workspace != nullptr, "net.workspace is NULL");
    NetWorkspace w = mzgo_net_layout(net->n, net->C, net->precision, net->max_batch);
    MZGO_CHECK_ARG(net->workspace_bytes >= w.total, "net.workspace_bytes=%ld < %ld", (long)net->workspace_bytes, w.total);
    return 0;
}

extern "C" int64_t mzgo_net_workspace_bytes(int n, int C, int precision, int max_batch) {
    return mzgo_net_layout(n, C, precision, max_batch).total;
}

extern "C" int64_t mzgo_net_slot_bytes(int n, int C, int precision) {
    if (precision == 1) return (int64_t)C * (((long)n * (n + 1) - 1 + 7) / 8 * 8) * 2;   // nets_tc.cu: tc_psh positions x C channels, bf16
    return (int64_t)C * n * n * 4;
}

int mzgo_net_prepare_fp32(const mzgo_net_t* net, cudaStream_t st) {
    char* ws = (char*)net->workspace;
    NetWorkspace w = mzgo_net_layout(net->n, net->C, net->precision, net->max_batch);
    k_transpose_conv_w<<<32, 256, 0, st>>>(net->rep_w1, (float*)(ws + w.wT1), 6, 64);
    k_transpose_conv_w<<<64, 256, 0, st>>>(net->rep_w2, (float*)(ws + w.wT2), 64, 64);
    k_transpose_conv_w<<<64, 256, 0, st>>>(net->rep_w3, (float*)(ws + w.wT3), 64, net->C);
    k_transpose_conv_w<<<128, 256, 0, st>>>(net->dyn_w, (float*)(ws + w.wTd), net->C, net->C);
    mzgo_count_launches(4);
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

// representation network into float32 [B][C][np] at `out` (slots via out_idx)
int mzgo_rep_fp32(const mzgo_net_t* net, const float* obs, const uint32_t* boards, float* out, const int32_t* out_idx,
                  long out_stride, int B, cudaStream_t st) {
    char* ws = (char*)net->workspace;
    NetWorkspace w = mzgo_net_layout(net->n, net->C, net->precision, net->max_batch);
    const int np = net->n * net->n;
    ConvArgs a{};
    a.n = net->n;
    a.in = obs; a.in_stride = 6L * np; a.boards = obs ? nullptr : boards;
    a.Cin = 6; a.Cout = 64; a.wT = (float*)(ws + w.wT1); a.bias = net->rep_b1;
    a.out = (float*)(ws + w.act1); a.out_stride = 64L * np; a.out_idx = nullptr;
    if (int e = launch_conv(a, B, st)) return e;
    a.boards = nullptr;
    a.in = (float*)(ws + w.act1); a.in_stride = 64L * np;
    a.Cin = 64; a.wT = (float*)(ws + w.wT2); a.bias = net->rep_b2;
    a.out = (float*)(ws + w.act2);
    if (int e = launch_conv(a, B, st)) return e;
    a.in = (float*)(ws + w.act2);
    a.Cout = net->C; a.wT = (float*)(ws + w.wT3); a.bias = net->rep_b3;
    a.out = out; a.out_stride = out_stride; a.out_idx = out_idx;
    return launch_conv(a, B, st);
}

int mzgo_heads_fp32(const mzgo_net_t* net, const float* arena, const int32_t* slot, const int32_t* action,
                    long slot_stride, float* reward, float* value, float* policy, int want_softmax, int B,
                    cudaStream_t st) {
    k_heads_fp32<<<B, HEAD_THREADS, 0, st>>>(*net, arena, slot, action, slot_stride, reward, value, policy, want_softmax);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

int mzgo_dyn_fp32(const mzgo_net_t* net, float* arena, const int32_t* in_slot, const int32_t* action,
                  const int32_t* out_slot, int B, cudaStream_t st) {
    char* ws = (char*)net->workspace;
    NetWorkspace w = mzgo_net_layout(net->n, net->C, net->precision, net->max_batch);
    const long stride = (long)net->C * net->n * net->n;
    ConvArgs a{};
    a.n = net->n;
    a.in = arena; a.in_idx = in_slot; a.in_stride = stride;
    a.emb = net->emb; a.action = action;
    a.Cin = net->C; a.Cout = net->C; a.wT = (float*)(ws + w.wTd); a.bias = net->dyn_b;
    a.out = arena; a.out_idx = out_slot; a.out_stride = stride;
    return launch_conv(a, B, st);
}

int mzgo_copy_latent_f32(const float* src, const int32_t* src_idx, long src_stride, float* dst, const int32_t* dst_idx,
                         long dst_stride, long count, int B, cudaStream_t st) {
    dim3 grid((unsigned)((count + 1023) / 1024), B);
    k_copy_latent_f32<<<grid, 256, 0, st>>>(src, src_idx, src_stride, dst, dst_idx, dst_stride, count);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}
