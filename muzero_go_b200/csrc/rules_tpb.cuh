// Go rules, THREAD per board (included by rules.cu inside its namespace, after hfill / mix64).
//
// The warp-per-board kernels of rules.cu spend ~4,400 warp instructions on a 19x19 ply: with one board per
// warp every shuffle, vote and flood-fill iteration serves one board, and the kernel is issue bound at ~1 % of
// the HBM bandwidth its 324 B per ply would allow.  For LARGE batches of boards up to 9x9 (the playout / step
// entry points switch over at g_tpb_min boards; see tpb_size in rules.cu for why not 19x19) a thread owns a whole
// board instead: the three planes are N rows of uint32 in
// registers (N is a template parameter, every row loop is unrolled), vertical neighbours are the adjacent array
// elements, and one warp instruction advances 32 boards.  The arithmetic is the same bit for bit (the same
// hfill carry trick per row; flood fill sweeps the rows down and up until nothing changes, which reaches the same
// fixed point as the warp version's simultaneous iteration), so the parity tests of the warp kernels apply
// unchanged.  Records travel through shared memory so that global accesses stay coalesced (a thread's record is
// 256 B; the staging rows are padded to 65 words against bank conflicts).
#pragma once

template <int N>
struct TBoard {
    uint32_t bl[N], wh[N], inv[N];
    int turn, pass, done, ply;
};

template <int N>
__device__ __forceinline__ void t_dilate4(const uint32_t (&x)[N], uint32_t (&out)[N]) {
    constexpr uint32_t BM = (1u << N) - 1u;
#pragma unroll
    for (int r = 0; r < N; r++) {
        uint32_t v = (x[r] << 1) | (x[r] >> 1);
        if (r > 0) v |= x[r - 1];
        if (r < N - 1) v |= x[r + 1];
        out[r] = v & BM;
    }
}

// g: seed on entry, the connected component of `mask` containing it on exit
template <int N>
__device__ __forceinline__ void t_flood(uint32_t (&g)[N], const uint32_t (&mask)[N]) {
#pragma unroll
    for (int r = 0; r < N; r++) g[r] = hfill(g[r] & mask[r], mask[r]);
    bool changed = true;
    while (changed) {
        changed = false;
#pragma unroll
        for (int r = 1; r < N; r++) {
            const uint32_t v = g[r - 1] & mask[r] & ~g[r];
            if (v) { g[r] = hfill(g[r] | v, mask[r]); changed = true; }
        }
#pragma unroll
        for (int r = N - 2; r >= 0; r--) {
            const uint32_t v = g[r + 1] & mask[r] & ~g[r];
            if (v) { g[r] = hfill(g[r] | v, mask[r]); changed = true; }
        }
    }
}

// t_flood that stops once the part found so far has `need` (1 or 2) liberties in `empt` (see flood_libs in
// rules.cu).  Returns min(liberties, 2) of what is left in g.
template <int N>
__device__ __forceinline__ int t_flood_libs(uint32_t (&g)[N], const uint32_t (&mask)[N], const uint32_t (&empt)[N], int need) {
#pragma unroll
    for (int r = 0; r < N; r++) g[r] = hfill(g[r] & mask[r], mask[r]);
    while (true) {
        int libs = 0;
#pragma unroll
        for (int r = 0; r < N; r++) {
            uint32_t v = (g[r] << 1) | (g[r] >> 1);
            if (r > 0) v |= g[r - 1];
            if (r < N - 1) v |= g[r + 1];
            libs += __popc(v & empt[r]);
        }
        if (libs > 2) libs = 2;
        if (libs >= need) return libs;
        bool changed = false;
#pragma unroll
        for (int r = 1; r < N; r++) {
            const uint32_t v = g[r - 1] & mask[r] & ~g[r];
            if (v) { g[r] = hfill(g[r] | v, mask[r]); changed = true; }
        }
#pragma unroll
        for (int r = N - 2; r >= 0; r--) {
            const uint32_t v = g[r + 1] & mask[r] & ~g[r];
            if (v) { g[r] = hfill(g[r] | v, mask[r]); changed = true; }
        }
        if (!changed) return libs;
    }
}

template <int N>
__device__ __forceinline__ bool t_any(const uint32_t (&x)[N]) {
    uint32_t o = 0u;
#pragma unroll
    for (int r = 0; r < N; r++) o |= x[r];
    return o != 0u;
}

// lowest set point in row-major order as a one-point board; false if x is empty
template <int N>
__device__ __forceinline__ bool t_first_point(const uint32_t (&x)[N], uint32_t (&pt)[N]) {
    bool found = false;
#pragma unroll
    for (int r = 0; r < N; r++) {
        pt[r] = 0u;
        if (!found && x[r]) { pt[r] = x[r] & (0u - x[r]); found = true; }
    }
    return found;
}

template <int N>
__device__ __forceinline__ int t_count_upto2(const uint32_t (&x)[N]) {
    int c = 0;
#pragma unroll
    for (int r = 0; r < N; r++) c += __popc(x[r]);
    return c > 2 ? 2 : c;
}

// compute_invalid of rules.cu (state_utils.compute_invalid_moves) on a whole board
template <int N>
__device__ __forceinline__ void t_compute_invalid(const uint32_t (&own)[N], const uint32_t (&opp)[N], const uint32_t (&ko)[N],
                                                  uint32_t (&inv)[N]) {
    constexpr uint32_t BM = (1u << N) - 1u;
    uint32_t occ[N], e1[N];
#pragma unroll
    for (int r = 0; r < N; r++) occ[r] = own[r] | opp[r];
#pragma unroll
    for (int r = 0; r < N; r++) {
        const uint32_t l = (occ[r] << 1) | 1u, rr = (occ[r] >> 1) | (1u << (N - 1));
        const uint32_t u = r == 0 ? BM : occ[r - 1], d = r == N - 1 ? BM : occ[r + 1];
        e1[r] = ~occ[r] & BM & l & rr & u & d;
    }
#pragma unroll
    for (int r = 0; r < N; r++) inv[r] = occ[r] | ko[r];
    if (!t_any<N>(e1)) return;
    // one flood for the whole board first (see compute_invalid in rules.cu): stones next to an empty point that is
    // not surrounded, grown through their own colour, are groups with more than one liberty
    uint32_t seeds[N], atari[N], multi[N], empt[N], grp[N], mask[N];
    t_dilate4<N>(e1, seeds);
#pragma unroll
    for (int r = 0; r < N; r++) { empt[r] = ~occ[r] & BM; grp[r] = empt[r] & ~e1[r]; }
    t_dilate4<N>(grp, multi);                                    // multi = points next to a non-surrounded empty point
    uint32_t rem = 0u;
#pragma unroll
    for (int r = 0; r < N; r++) {
        seeds[r] &= occ[r];
        atari[r] = 0u;
        multi[r] = hfill(multi[r] & own[r], own[r]) | hfill(multi[r] & opp[r], opp[r]);
        seeds[r] &= ~multi[r];
        rem |= seeds[r];
    }
    while (rem) {
        bool changed = false;
#pragma unroll
        for (int r = 1; r < N; r++) {
            const uint32_t vo = multi[r - 1] & own[r - 1] & own[r] & ~multi[r];
            const uint32_t vp = multi[r - 1] & opp[r - 1] & opp[r] & ~multi[r];
            if (vo | vp) { multi[r] |= hfill(vo, own[r]) | hfill(vp, opp[r]); changed = true; }
        }
#pragma unroll
        for (int r = N - 2; r >= 0; r--) {
            const uint32_t vo = multi[r + 1] & own[r + 1] & own[r] & ~multi[r];
            const uint32_t vp = multi[r + 1] & opp[r + 1] & opp[r] & ~multi[r];
            if (vo | vp) { multi[r] |= hfill(vo, own[r]) | hfill(vp, opp[r]); changed = true; }
        }
        rem = 0u;
#pragma unroll
        for (int r = 0; r < N; r++) { seeds[r] &= ~multi[r]; rem |= seeds[r]; }
        if (!changed) break;
    }
    while (t_first_point<N>(seeds, grp)) {
        uint32_t hit_own = 0u;
#pragma unroll
        for (int r = 0; r < N; r++) hit_own |= own[r] & grp[r];
#pragma unroll
        for (int r = 0; r < N; r++) mask[r] = hit_own ? own[r] : opp[r];
        const int libs = t_flood_libs<N>(grp, mask, empt, 2);
#pragma unroll
        for (int r = 0; r < N; r++) {
            if (libs == 1) atari[r] |= grp[r];
            else if (libs > 1) multi[r] |= grp[r];
            seeds[r] &= ~grp[r];
        }
    }
    uint32_t a[N], b[N], possible[N], definite[N];
#pragma unroll
    for (int r = 0; r < N; r++) { a[r] = (own[r] & multi[r]) | (opp[r] & atari[r]); b[r] = (own[r] & atari[r]) | (opp[r] & multi[r]); }
    t_dilate4<N>(a, possible);
    t_dilate4<N>(b, definite);
#pragma unroll
    for (int r = 0; r < N; r++) inv[r] = occ[r] | (e1[r] & possible[r] & ~definite[r]) | ko[r];
}

// board_step of rules.cu (gogame.next_state).  Returns 0 ok, -1 illegal, -2 game over (board untouched on error).
template <int N>
__device__ __forceinline__ int t_board_step(TBoard<N>& b, int action) {
    constexpr int NP = N * N;
    if (b.done) return -2;
    if (action < 0 || action > NP) return -1;
    uint32_t own[N], opp[N], ko[N];
#pragma unroll
    for (int r = 0; r < N; r++) { own[r] = b.turn ? b.wh[r] : b.bl[r]; opp[r] = b.turn ? b.bl[r] : b.wh[r]; ko[r] = 0u; }
    int new_pass = 1, new_done = b.done;
    if (action == NP) {
        if (b.pass) new_done = 1;
    } else {
        const int ar = action / N, ac = action - ar * N;
        uint32_t stone[N], adj[N];
        uint32_t bad = 0u;
#pragma unroll
        for (int r = 0; r < N; r++) { stone[r] = (r == ar) ? (1u << ac) : 0u; bad |= b.inv[r] & stone[r]; }
        if (bad) return -1;
        new_pass = 0;
#pragma unroll
        for (int r = 0; r < N; r++) own[r] |= stone[r];
        t_dilate4<N>(stone, adj);
        uint32_t open = 0u;                       // a neighbour that does not hold an opponent stone (before captures)
#pragma unroll
        for (int r = 0; r < N; r++) open |= adj[r] & ~opp[r];
        const bool surrounded = open == 0u;
        constexpr uint32_t BM = (1u << N) - 1u;
        uint32_t cand[N], kill[N], grp[N], empt[N], single_row[N];
#pragma unroll
        for (int r = 0; r < N; r++) { kill[r] = 0u; single_row[r] = 0u; empt[r] = ~(own[r] | opp[r]) & BM; }
        t_dilate4<N>(empt, cand);                 // a neighbour that itself touches an empty point lives
#pragma unroll
        for (int r = 0; r < N; r++) cand[r] = adj[r] & opp[r] & ~cand[r];
        int nkilled = 0;
        bool single = false;
        while (t_first_point<N>(cand, grp)) {
            if (t_flood_libs<N>(grp, opp, empt, 1) == 0) {
                nkilled++;
                single = t_count_upto2<N>(grp) == 1;
#pragma unroll
                for (int r = 0; r < N; r++) { single_row[r] = grp[r]; kill[r] |= grp[r]; }
            }
#pragma unroll
            for (int r = 0; r < N; r++) cand[r] &= ~grp[r];
        }
#pragma unroll
        for (int r = 0; r < N; r++) opp[r] &= ~kill[r];
        if (nkilled == 1 && single && surrounded) {
#pragma unroll
            for (int r = 0; r < N; r++) ko[r] = single_row[r];
        }
    }
    uint32_t inv[N];
    t_compute_invalid<N>(own, opp, ko, inv);
#pragma unroll
    for (int r = 0; r < N; r++) {
        b.bl[r] = b.turn ? opp[r] : own[r];
        b.wh[r] = b.turn ? own[r] : opp[r];
        b.inv[r] = inv[r];
    }
    b.pass = new_pass;
    b.done = new_done;
    b.turn ^= 1;
    b.ply++;
    return 0;
}

template <int N>
__device__ __forceinline__ uint64_t t_board_hash(const TBoard<N>& b) {
    uint64_t h = 0;
#pragma unroll
    for (int r = 0; r < N; r++)
        h += mix64((uint64_t)b.bl[r] | ((uint64_t)b.wh[r] << 20) | ((uint64_t)b.inv[r] << 40) | ((uint64_t)(r + 1) << 59));
    return h + mix64(0xabcdull + (uint64_t)b.turn + 2ull * b.pass + 4ull * b.done);
}

constexpr int TPB_THREADS = 64;              // boards per block
constexpr int TPB_STRIDE = MZGO_BOARD_WORDS + 1;

// block's records <-> shared memory, coalesced
__device__ __forceinline__ void t_stage_in(uint32_t* s, const uint32_t* boards, int first, int count) {
    for (int i = threadIdx.x; i < count * MZGO_BOARD_WORDS; i += TPB_THREADS)
        s[(i / MZGO_BOARD_WORDS) * TPB_STRIDE + (i % MZGO_BOARD_WORDS)] = boards[(size_t)first * MZGO_BOARD_WORDS + i];
    __syncthreads();
}
__device__ __forceinline__ void t_stage_out(const uint32_t* s, uint32_t* boards, int first, int count) {
    __syncthreads();
    for (int i = threadIdx.x; i < count * MZGO_BOARD_WORDS; i += TPB_THREADS)
        boards[(size_t)first * MZGO_BOARD_WORDS + i] = s[(i / MZGO_BOARD_WORDS) * TPB_STRIDE + (i % MZGO_BOARD_WORDS)];
}
template <int N>
__device__ __forceinline__ void t_board_load(TBoard<N>& b, const uint32_t* rec) {
#pragma unroll
    for (int r = 0; r < N; r++) { b.bl[r] = rec[r]; b.wh[r] = rec[N + r]; b.inv[r] = rec[2 * N + r]; }
    b.turn = (int)(rec[60] & 1u);
    b.pass = (int)(rec[61] & 1u);
    b.done = (int)((rec[61] >> 1) & 1u);
    b.ply = (int)rec[62];
}
template <int N>
__device__ __forceinline__ void t_board_store(const TBoard<N>& b, uint32_t* rec) {
#pragma unroll
    for (int r = 0; r < N; r++) { rec[r] = b.bl[r]; rec[N + r] = b.wh[r]; rec[2 * N + r] = b.inv[r]; }
    rec[60] = (uint32_t)b.turn;
    rec[61] = (uint32_t)(b.pass | (b.done << 1));
    rec[62] = (uint32_t)b.ply;
    rec[63] = 0u;
}

template <int N>
__global__ void __launch_bounds__(TPB_THREADS)
k_rules_step_tpb(uint32_t* boards, const int32_t* actions, int32_t* status, int nboards) {
    __shared__ uint32_t s[TPB_THREADS * TPB_STRIDE];
    const int first = blockIdx.x * TPB_THREADS, count = min(TPB_THREADS, nboards - first);
    t_stage_in(s, boards, first, count);
    const int bidx = first + threadIdx.x;
    if (threadIdx.x < count) {
        const int a = actions[bidx];
        int st = 0;
        if (a >= 0) {
            uint32_t* rec = s + threadIdx.x * TPB_STRIDE;
            TBoard<N> b;
            t_board_load<N>(b, rec);
            st = t_board_step<N>(b, a);
            if (st == 0) t_board_store<N>(b, rec);
        }
        if (status) status[bidx] = st;
    }
    t_stage_out(s, boards, first, count);
}

template <int N>
__global__ void __launch_bounds__(TPB_THREADS)
k_rules_playout_tpb(uint32_t* boards, const uint32_t* R, int T, int32_t* actions_out, uint64_t* hash_out, int nboards) {
    constexpr uint32_t BM = (1u << N) - 1u;
    constexpr int NP = N * N;
    __shared__ uint32_t s[TPB_THREADS * TPB_STRIDE];
    const int first = blockIdx.x * TPB_THREADS, count = min(TPB_THREADS, nboards - first);
    t_stage_in(s, boards, first, count);
    const int bidx = first + threadIdx.x;
    if (threadIdx.x < count) {
        uint32_t* rec = s + threadIdx.x * TPB_STRIDE;
        TBoard<N> b;
        t_board_load<N>(b, rec);
        for (int t = 0; t < T; t++) {
            int a = -1;
            if (!b.done) {
                int nvalid = 1;                                 // the pass
#pragma unroll
                for (int r = 0; r < N; r++) nvalid += __popc(~b.inv[r] & BM);
                int k = (int)(R[(size_t)bidx * T + t] % (uint32_t)nvalid);
                bool found = false;
                uint32_t hit_row = 0u;
                int hr = 0;
#pragma unroll
                for (int r = 0; r < N; r++) {
                    const uint32_t valid = ~b.inv[r] & BM;
                    const int c = __popc(valid);
                    if (!found) {
                        if (k < c) { hit_row = valid; hr = r; found = true; }
                        else k -= c;
                    }
                }
                a = found ? hr * N + kth_set_bit(hit_row, k) : NP;
                t_board_step<N>(b, a);
            }
            if (actions_out) actions_out[(size_t)bidx * T + t] = a;
            if (hash_out) hash_out[(size_t)bidx * T + t] = t_board_hash<N>(b);
        }
        t_board_store<N>(b, rec);
    }
    t_stage_out(s, boards, first, count);
}
