// Training-side convolution kernels (the batched trainer, /root/reference/main.py:381-501).
//
// The reference trains with torch autograd on nn.Conv2d (cuDNN on a GPU).  Here the 3x3 convolutions of the
// representation and dynamics networks (main.py:75-77, :90 -- more than 99.9 % of a training step's FLOPs) run
// forward, data-gradient and weight-gradient in these hand-written kernels, float32 on the CUDA cores: the trainer's
// parity contract is float32 (loss and updated weights pinned to the real MuZeroAgent.train at 1e-5,
// tests/golden/train_*.npz), which bf16 / tf32 tensor-core arithmetic cannot meet.  Dense batches, NCHW, pad 1.
//
//   forward      y = [relu](conv(x, W) + b)
//   data grad    dx = conv(dy * (y > 0), W flipped and transposed)       -- the same kernel with another weight view
//   weight grad  dW[co][ci][tap] = sum_{b,p} (dy * (y > 0))[b][co][p] * x[b][ci][p + tap offset],  db[co] = sum_{b,p} ...
//
// ReLU's backward mask is folded into the operand load of both gradient kernels (`gate`), so no masked copy of dy is
// ever written.  Everything is launched on the caller's stream without synchronisation (CUDA-graph capturable: the
// trainer replays its whole step from one graph).
#include <algorithm>
#include <mutex>

#include "common.cuh"
#include "nets.cuh"

namespace {

constexpr int TCONV_THREADS = 256;
constexpr int TCI_CHUNK = 4;

// y[b][co][p] = act(bias[co] + sum_{ci,tap} wT[ci][tap][co] * xin[b][ci][p + d(tap)]),  xin = gate ? x * (gate > 0) : x.
// The pixels of the whole batch form one index space gp = b * n^2 + p; a block owns a tile of TPIX * PX consecutive gp
// (on small boards several images, so that they fill the thread tile), thread tile 8 output channels x PX pixels
// (PX = 8, or 2 when the batch is too small to give every SM a block of 8-pixel tiles); input channels staged through shared memory four at a time (the zero-padded planes of the images the
// tile touches + [ci][tap][co] weights).
template <int PX>
__global__ void __launch_bounds__(TCONV_THREADS)
k_train_conv3x3(const float* __restrict__ x, const float* __restrict__ gate, const float* __restrict__ wT,
                const float* __restrict__ bias, float* __restrict__ y, int relu, int B, int Cin, int Cout, int n, int imgs) {
    extern __shared__ float smem[];
    const int np = n * n, pw = n + 2, pp = pw * pw;
    float* xs = smem;                            // [TCI_CHUNK][imgs][pp]
    float* ws = smem + TCI_CHUNK * imgs * pp;    // [TCI_CHUNK][9][Cout]
    const int TC = Cout / 8;
    const int TPIX = TCONV_THREADS / TC;
    const int tc = threadIdx.x % TC, tp = threadIdx.x / TC;
    const long gp0 = (long)blockIdx.x * TPIX * PX, gp_end = (long)B * np;
    const int b0 = (int)(gp0 / np);              // first image of the tile
    int off[PX];
    bool ok[PX];
#pragma unroll
    for (int k = 0; k < PX; k++) {
        const long gp = gp0 + tp + k * TPIX;
        ok[k] = gp < gp_end && tp < TPIX;
        const int b = ok[k] ? (int)(gp / np) : b0, p = ok[k] ? (int)(gp - (long)b * np) : 0;
        const int r = p / n, c = p - r * n;
        off[k] = (b - b0) * pp + (r + 1) * pw + (c + 1);
    }
    float acc[8][PX];
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int k = 0; k < PX; k++) acc[i][k] = 0.f;
    for (int c0 = 0; c0 < Cin; c0 += TCI_CHUNK) {
        __syncthreads();
        for (int i = threadIdx.x; i < TCI_CHUNK * imgs * pp; i += TCONV_THREADS) {
            const int ci = i / (imgs * pp), rem = i - ci * imgs * pp;
            const int bl = rem / pp, q = rem - bl * pp;
            const int r = q / pw - 1, c = q - (q / pw) * pw - 1;
            float v = 0.f;
            if (c0 + ci < Cin && b0 + bl < B && r >= 0 && r < n && c >= 0 && c < n) {
                const size_t e = ((size_t)(b0 + bl) * Cin + c0 + ci) * np + r * n + c;
                v = x[e];
                if (gate && !(gate[e] > 0.f)) v = 0.f;
            }
            xs[i] = v;
        }
        for (int i = threadIdx.x; i < TCI_CHUNK * 9 * Cout; i += TCONV_THREADS) {
            const int ci = i / (9 * Cout);
            ws[i] = (c0 + ci < Cin) ? wT[(size_t)c0 * 9 * Cout + i] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int ci = 0; ci < TCI_CHUNK; ci++) {
#pragma unroll
            for (int tap = 0; tap < 9; tap++) {
                const int d = (tap / 3 - 1) * pw + (tap % 3 - 1);
                const float4 w0 = *reinterpret_cast<const float4*>(ws + (ci * 9 + tap) * Cout + tc * 8);
                const float4 w1 = *reinterpret_cast<const float4*>(ws + (ci * 9 + tap) * Cout + tc * 8 + 4);
                const float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                for (int k = 0; k < PX; k++) {
                    const float xv = xs[ci * imgs * pp + off[k] + d];
#pragma unroll
                    for (int i = 0; i < 8; i++) acc[i][k] = fmaf(w[i], xv, acc[i][k]);
                }
            }
        }
    }
#pragma unroll
    for (int k = 0; k < PX; k++) {
        if (!ok[k]) continue;
        const long gp = gp0 + tp + k * TPIX;
        const int b = (int)(gp / np), p = (int)(gp - (long)b * np);
        float* yo = y + ((size_t)b * Cout + tc * 8) * np + p;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const float v = acc[i][k] + (bias ? bias[tc * 8 + i] : 0.f);
            yo[(size_t)i * np] = relu ? fmaxf(v, 0.f) : v;
        }
    }
}

// W [Cout][Cin][3][3] -> wT_fwd [Cin][9][Cout] (forward) and wT_bwd [Cout][9][Cin] with the taps flipped (data gradient:
// dx[ci][p] = sum_{co,tap} W[co][ci][tap] * dy[co][p - d(tap)], i.e. a convolution of dy with tap' = 8 - tap)
__global__ void k_train_prep_w(const float* __restrict__ w, float* __restrict__ wf, float* __restrict__ wb, int Cin, int Cout) {
    const int total = Cin * Cout * 9;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int tap = i % 9, ci = (i / 9) % Cin, co = i / (9 * Cin);
        const float v = w[i];
        if (wf) wf[((size_t)ci * 9 + tap) * Cout + co] = v;
        if (wb) wb[((size_t)co * 9 + (8 - tap)) * Cin + ci] = v;
    }
}

// Weight gradient, partial over a slice of the batch: block (co tile of 32, ci tile of 8, batch slice z); thread
// (co_l = tid % 32, ci_l = tid / 32) accumulates the nine taps of dW[co][ci] over the slice's images and pixels from
// shared memory: the gated dy tile transposed to [image][pixel][32 channels] (conflict free), the zero-padded x tile
// [8 channels][image][(n+2)^2] (one address per warp: a broadcast); `ips` images are staged at a time (seven on a
// 6x6 board: the two barriers per stage amortise).  part [Z][Cout][Cin][9], partb [Z][Cout].
constexpr int WG_CO = 32, WG_CI = 8;
__global__ void __launch_bounds__(WG_CO * WG_CI)
k_train_wgrad_partial(const float* __restrict__ x, const float* __restrict__ dy, const float* __restrict__ gate,
                      float* __restrict__ part, float* __restrict__ partb, int B, int Cin, int Cout, int n, int ips) {
    extern __shared__ float smem[];
    const int np = n * n, pw = n + 2, pp = pw * pw;
    float* sg = smem;                             // [ips][np][WG_CO]
    float* sx = sg + ips * np * WG_CO;            // [WG_CI][ips][pp]
    int* soff = reinterpret_cast<int*>(sx + WG_CI * ips * pp);   // [ips * np] padded offset of (image, pixel)
    const int co0 = blockIdx.x * WG_CO, ci0 = blockIdx.y * WG_CI, z = blockIdx.z, Z = gridDim.z;
    const int co_l = threadIdx.x % WG_CO, ci_l = threadIdx.x / WG_CO;
    const int b_lo = (int)((long long)z * B / Z), b_hi = (int)((long long)(z + 1) * B / Z);
    for (int i = threadIdx.x; i < ips * np; i += blockDim.x) {
        const int bl = i / np, p = i - bl * np;
        soff[i] = bl * pp + (p / n + 1) * pw + (p % n + 1);
    }
    float acc[9];
#pragma unroll
    for (int t = 0; t < 9; t++) acc[t] = 0.f;
    float accb = 0.f;
    for (int b = b_lo; b < b_hi; b += ips) {
        const int nb = min(ips, b_hi - b);
        __syncthreads();
        for (int i = threadIdx.x; i < nb * np * WG_CO; i += blockDim.x) {
            const int bl = i / (np * WG_CO), rem = i - bl * np * WG_CO;
            const int c = rem / np, p = rem - c * np;        // coalesced over the pixels of one channel
            float v = 0.f;
            if (co0 + c < Cout) {
                const size_t e = ((size_t)(b + bl) * Cout + co0 + c) * np + p;
                v = dy[e];
                if (gate && !(gate[e] > 0.f)) v = 0.f;
            }
            sg[(bl * np + p) * WG_CO + c] = v;
        }
        for (int i = threadIdx.x; i < WG_CI * nb * pp; i += blockDim.x) {
            const int c = i / (nb * pp), rem = i - c * nb * pp;
            const int bl = rem / pp, q = rem - bl * pp;
            const int r = q / pw - 1, cc = q - (q / pw) * pw - 1;
            float v = 0.f;
            if (ci0 + c < Cin && r >= 0 && r < n && cc >= 0 && cc < n) v = x[((size_t)(b + bl) * Cin + ci0 + c) * np + r * n + cc];
            sx[(c * ips + bl) * pp + q] = v;
        }
        __syncthreads();
        const float* xr = sx + ci_l * ips * pp;
        const int count = nb * np;
        for (int i = 0; i < count; i++) {
            const float g = sg[i * WG_CO + co_l];
            const int o = soff[i];
            accb += g;
#pragma unroll
            for (int t = 0; t < 9; t++) acc[t] = fmaf(g, xr[o + (t / 3 - 1) * pw + (t % 3 - 1)], acc[t]);
        }
    }
    const int co = co0 + co_l, ci = ci0 + ci_l;
    if (co < Cout && ci < Cin) {
        float* o = part + (((size_t)z * Cout + co) * Cin + ci) * 9;
#pragma unroll
        for (int t = 0; t < 9; t++) o[t] = acc[t];
    }
    if (partb && blockIdx.y == 0 && ci_l == 0 && co < Cout) partb[(size_t)z * Cout + co] = accb;
}

// dw[i] = sum_z part[z][i] in a fixed order (deterministic, no atomics); same for the bias gradient
__global__ void k_train_wgrad_reduce(const float* __restrict__ part, const float* __restrict__ partb, float* __restrict__ dw,
                                     float* __restrict__ db, int Z, int total, int Cout) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total + Cout; i += gridDim.x * blockDim.x) {
        float s = 0.f;
        if (i < total) {
            for (int z = 0; z < Z; z++) s += part[(size_t)z * total + i];
            dw[i] = s;
        } else if (db) {
            for (int z = 0; z < Z; z++) s += partb[(size_t)z * Cout + (i - total)];
            db[i - total] = s;
        }
    }
}

int conv_launch(const float* x, const float* gate, const float* wT, const float* bias, float* y, int relu, int B, int Cin,
                int Cout, int n, cudaStream_t st) {
    MZGO_CHECK_ARG(x && wT && y, "conv3x3: NULL tensor");
    MZGO_CHECK_ARG(B >= 0 && Cin >= 1 && Cout >= 8 && Cout % 8 == 0 && Cout <= 256 && n >= 2 && n <= MZGO_MAX_N,
                   "conv3x3: unsupported shape B=%d Cin=%d Cout=%d n=%d (Cout must be a multiple of 8, at most 256)", B, Cin, Cout, n);
    if (B == 0) return MZGO_OK;
    const int TC = Cout / 8, TPIX = TCONV_THREADS / TC, np = n * n;
    const long total = (long)B * np;
    const int px = (total + TPIX * 8 - 1) / (TPIX * 8) >= 296 ? 8 : 2;     // two blocks per SM or smaller tiles
    const int tile = TPIX * px;
    const int imgs = (tile - 1) / np + 2;                  // images a tile of consecutive batch pixels can touch
    const size_t smem = (size_t)TCI_CHUNK * ((size_t)imgs * (n + 2) * (n + 2) + 9 * Cout) * sizeof(float);
    static bool attr_set = false;
    if (!attr_set) {
        MZGO_CUDA(cudaFuncSetAttribute(k_train_conv3x3<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        MZGO_CUDA(cudaFuncSetAttribute(k_train_conv3x3<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        attr_set = true;
    }
    MZGO_CHECK_ARG(smem <= 96 * 1024, "conv3x3: %zu B of shared memory needed", smem);
    const unsigned blocks = (unsigned)((total + tile - 1) / tile);
    if (px == 8) k_train_conv3x3<8><<<blocks, TCONV_THREADS, smem, st>>>(x, gate, wT, bias, y, relu, B, Cin, Cout, n, imgs);
    else k_train_conv3x3<2><<<blocks, TCONV_THREADS, smem, st>>>(x, gate, wT, bias, y, relu, B, Cin, Cout, n, imgs);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

}  // namespace

extern "C" int mzgo_train_conv3x3_prep_w(const float* w, float* wT_fwd, float* wT_bwd, int Cin, int Cout, void* stream) {
    MZGO_CHECK_ARG(w && (wT_fwd || wT_bwd) && Cin >= 1 && Cout >= 1, "prep_w: bad arguments");
    k_train_prep_w<<<ceil_div(Cin * Cout * 9, 256), 256, 0, (cudaStream_t)stream>>>(w, wT_fwd, wT_bwd, Cin, Cout);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

extern "C" int mzgo_train_conv3x3_fwd(const float* x, const float* wT_fwd, const float* bias, float* y, int relu, int B,
                                      int Cin, int Cout, int n, void* stream) {
    return conv_launch(x, nullptr, wT_fwd, bias, y, relu, B, Cin, Cout, n, (cudaStream_t)stream);
}

extern "C" int mzgo_train_conv3x3_bwd_data(const float* dy, const float* y_gate, const float* wT_bwd, float* dx, int B,
                                           int Cin, int Cout, int n, void* stream) {
    // a convolution of the gated dy (Cout channels) with the flipped, transposed weights into Cin channels
    return conv_launch(dy, y_gate, wT_bwd, nullptr, dx, 0, B, Cout, Cin, n, (cudaStream_t)stream);
}

extern "C" int64_t mzgo_train_wgrad_scratch_floats(int B, int Cin, int Cout) {
    const int Z = B < 16 ? (B < 1 ? 1 : B) : 16;
    return (int64_t)Z * ((int64_t)Cout * Cin * 9 + Cout);
}

extern "C" int mzgo_train_conv3x3_bwd_weight(const float* x, const float* dy, const float* y_gate, float* dw, float* db,
                                             float* scratch, int B, int Cin, int Cout, int n, void* stream) {
    MZGO_CHECK_ARG(x && dy && dw && scratch, "bwd_weight: NULL tensor");
    MZGO_CHECK_ARG(B >= 1 && Cin >= 1 && Cout >= 1 && n >= 2 && n <= MZGO_MAX_N, "bwd_weight: unsupported shape");
    cudaStream_t st = (cudaStream_t)stream;
    const int Z = B < 16 ? B : 16;
    const int np = n * n, pp = (n + 2) * (n + 2), total = Cout * Cin * 9;
    const int ips = std::max(1, std::min(8, 300 / np));     // images staged at a time
    const size_t smem = ((size_t)ips * np * WG_CO + (size_t)WG_CI * ips * pp) * sizeof(float) + (size_t)ips * np * sizeof(int);
    static bool attr_set = false;
    if (!attr_set) {
        MZGO_CUDA(cudaFuncSetAttribute(k_train_wgrad_partial, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        attr_set = true;
    }
    float* part = scratch;
    float* partb = scratch + (size_t)Z * total;
    dim3 grid(ceil_div(Cout, WG_CO), ceil_div(Cin, WG_CI), Z);
    k_train_wgrad_partial<<<grid, WG_CO * WG_CI, smem, st>>>(x, dy, y_gate, part, db ? partb : nullptr, B, Cin, Cout, n, ips);
    k_train_wgrad_reduce<<<ceil_div(total + Cout, 256), 256, 0, st>>>(part, partb, dw, db, Z, total, Cout);
    mzgo_count_launches(2);
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

// ---- weight gradient without any allocation (no counterpart in the reference):
// for callers that cannot call cudaMalloc -- code under CUDA-graph capture, tight serving loops, harnesses
// that forbid it.  The partial sums of the two-pass reduction live in a file-scope device array of a fixed maximum
// size; every GPU that loads the library holds these 10 MiB until the process exits.
//
// g_wgrad_scratch, named in HOST code, is a host-side stand-in and not a device address: the kernels get the address
// cudaGetSymbolAddress returns for the CURRENT device, resolved once per device below.
//
// Every call on a device shares the array.  Nothing in it has to start at zero: k_train_wgrad_partial writes every
// partial sum k_train_wgrad_reduce reads (all z < Z, co < Cout, ci < Cin, nine taps; the bias partials whenever db is
// given) before the reduction of the SAME call reads it, in stream order, so the contents an earlier call left behind
// are never read.  Calls that overlap on the GPU (two streams, two host threads) race on the array: serialise them.
constexpr int64_t WGRAD_STATIC_FLOATS = MZGO_TRAIN_WGRAD_STATIC_MAX_FLOATS;
static __device__ __align__(256) float g_wgrad_scratch[WGRAD_STATIC_FLOATS];

namespace {
constexpr int MAX_DEVICES = 64;
float* g_wgrad_scratch_ptr[MAX_DEVICES];   // device address of g_wgrad_scratch on device d (0 = not resolved yet)
std::mutex g_wgrad_scratch_mu;
}  // namespace

extern "C" int64_t mzgo_train_wgrad_static_max_floats(void) { return WGRAD_STATIC_FLOATS; }

extern "C" int mzgo_train_conv3x3_bwd_weight_static(const float* x, const float* dy, const float* y_gate, float* dw, float* db,
                                                    int B, int Cin, int Cout, int n, void* stream) {
    MZGO_CHECK_ARG(B >= 1 && Cin >= 1 && Cout >= 1, "bwd_weight_static: unsupported shape B=%d Cin=%d Cout=%d", B, Cin, Cout);
    const int64_t need = mzgo_train_wgrad_scratch_floats(B, Cin, Cout);
    MZGO_CHECK_ARG(need <= WGRAD_STATIC_FLOATS,
                   "bwd_weight_static: B=%d Cin=%d Cout=%d needs %lld floats of scratch, the built-in array holds %lld "
                   "(use mzgo_train_conv3x3_bwd_weight with a caller-owned buffer)",
                   B, Cin, Cout, (long long)need, (long long)WGRAD_STATIC_FLOATS);
    int dev = -1;
    MZGO_CUDA(cudaGetDevice(&dev));
    MZGO_CHECK_ARG(dev >= 0 && dev < MAX_DEVICES, "bwd_weight_static: device %d out of range", dev);
    float* scratch;
    {
        std::lock_guard<std::mutex> lock(g_wgrad_scratch_mu);
        if (!g_wgrad_scratch_ptr[dev]) {
            float* p = nullptr;
            MZGO_CUDA(cudaGetSymbolAddress((void**)&p, g_wgrad_scratch));     // with device `dev` current
            MZGO_CHECK_ARG(p != nullptr, "bwd_weight_static: no device address for the scratch array on device %d", dev);
            g_wgrad_scratch_ptr[dev] = p;
        }
        scratch = g_wgrad_scratch_ptr[dev];
    }
    return mzgo_train_conv3x3_bwd_weight(x, dy, y_gate, dw, db, scratch, B, Cin, Cout, n, stream);
}
