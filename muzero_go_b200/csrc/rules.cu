// Go rules engine: warp-per-board bitboards (lane r owns row r of every plane).
//
// Replaces GymGo's numpy/scipy next_state / compute_invalid_moves / areas / winning
// (huangeddie/GymGo @ 3a3d97b, called from /root/reference/main.py:637,681,709; planes read at
// main.py:261).  Group flood fill is an iterated bitboard dilation: horizontal runs are filled
// with a carry-propagation trick inside each lane, vertical neighbours travel by warp shuffle,
// convergence is a warp vote.  State never leaves registers inside a kernel.
#include "common.cuh"

namespace {

struct Board {
    uint32_t bl, wh, inv;  // this lane's row of the black / white / invalid planes
    int turn, pass, done, ply;  // replicated in every lane
};

__device__ __forceinline__ uint32_t row_mask(int n, int lane) { return lane < n ? ((1u << n) - 1u) : 0u; }

__device__ __forceinline__ uint32_t shfl_up0(uint32_t x, int lane) {
    uint32_t v = __shfl_up_sync(FULL_MASK, x, 1);
    return lane > 0 ? v : 0u;
}
__device__ __forceinline__ uint32_t shfl_dn0(uint32_t x, int lane) {
    uint32_t v = __shfl_down_sync(FULL_MASK, x, 1);
    return lane < 31 ? v : 0u;
}

// Fill `g` (subset of `mask`) along the contiguous runs of `mask` in both directions.
// Upward: adding the seed bits to the run makes the carry ripple to the run's end.
__device__ __forceinline__ uint32_t hfill(uint32_t g, uint32_t mask) {
    uint32_t up = (((mask + g) ^ mask) & mask) | g;
    uint32_t rm = __brev(mask), rg = __brev(g);
    uint32_t dn = __brev(((rm + rg) ^ rm) & rm);
    return up | dn;
}

// 4-neighbourhood of x, without x itself, clipped to the board.
__device__ __forceinline__ uint32_t dilate4(uint32_t x, uint32_t bm, int lane) {
    return ((x << 1) | (x >> 1) | shfl_up0(x, lane) | shfl_dn0(x, lane)) & bm;
}

// Connected component of `mask` containing the seed bits (4-connectivity).
__device__ __forceinline__ uint32_t flood(uint32_t seed, uint32_t mask, int lane) {
    uint32_t g = hfill(seed & mask, mask);
    while (true) {
        uint32_t v = (shfl_up0(g, lane) | shfl_dn0(g, lane)) & mask & ~g;
        if (!__any_sync(FULL_MASK, v != 0u)) break;
        g = hfill(g | v, mask);
    }
    return g;
}

// Lowest set point (row-major) of a warp-distributed bitboard: returns lane, sets bit mask.
__device__ __forceinline__ int first_point(uint32_t x, uint32_t* bit) {
    uint32_t lanes = __ballot_sync(FULL_MASK, x != 0u);
    int l = __ffs(lanes) - 1;
    uint32_t row = __shfl_sync(FULL_MASK, x, l);
    *bit = row & (0u - row);
    return l;
}

// min(number of set bits over the whole warp, 2)
__device__ __forceinline__ int count_upto2(uint32_t x) {
    uint32_t lanes = __ballot_sync(FULL_MASK, x != 0u);
    if (lanes == 0u) return 0;
    if (lanes & (lanes - 1u)) return 2;
    uint32_t row = __shfl_sync(FULL_MASK, x, __ffs(lanes) - 1);
    return (row & (row - 1u)) ? 2 : 1;
}

// flood() that stops as soon as the part of the component found so far has `need` liberties (1 or 2) in `empt`:
// a component has at least the liberties of any part of it, and callers only ask "none / exactly one / more".
// Returns min(liberties, 2) of what it returns in *grp -- the whole component unless it stopped early.
__device__ __forceinline__ int flood_libs(uint32_t seed, uint32_t mask, uint32_t empt, uint32_t bm, int need, int lane,
                                          uint32_t* grp) {
    uint32_t g = hfill(seed & mask, mask);
    int libs;
    while (true) {
        const uint32_t lb = dilate4(g, bm, lane) & empt;
        if (need == 1) libs = __any_sync(FULL_MASK, lb != 0u) ? 1 : 0;
        else libs = count_upto2(lb);
        if (libs >= need) break;
        uint32_t v = (shfl_up0(g, lane) | shfl_dn0(g, lane)) & mask & ~g;
        if (!__any_sync(FULL_MASK, v != 0u)) break;
        g = hfill(g | v, mask);
    }
    *grp = g;
    return libs;
}

// state_utils.compute_invalid_moves(state, player, ko): invalid plane for the NEXT mover.
// own = stones of `player` (who just moved), opp = stones of the next mover.
__device__ uint32_t compute_invalid(uint32_t own, uint32_t opp, uint32_t ko_row, int n, int lane) {
    const uint32_t bm = row_mask(n, lane);
    const uint32_t occ = own | opp;
    const uint32_t empt = ~occ & bm;
    // surrounded: all four neighbours occupied, off-board counts as occupied
    uint32_t l = (occ << 1) | 1u;
    uint32_t r = (occ >> 1) | (1u << (n - 1));
    uint32_t u = __shfl_up_sync(FULL_MASK, occ, 1);
    uint32_t d = __shfl_down_sync(FULL_MASK, occ, 1);
    if (lane == 0) u = bm;
    if (lane >= n - 1) d = bm;
    const uint32_t e1 = empt & l & r & u & d;
    if (!__any_sync(FULL_MASK, e1 != 0u)) return occ | ko_row;

    // Classify the stones next to a surrounded empty point p: is p their group's only liberty, or one of several?
    // A group holding a stone next to an empty point that is NOT surrounded has a liberty other than p, so it has
    // several: grow those stones through both colours at once (one flood for the whole board instead of one per
    // group) until every seed is reached or nothing changes.  What is left are groups whose liberties are all
    // surrounded points (eyes); only those are counted one by one.
    const uint32_t seeds = dilate4(e1, bm, lane) & occ;
    const uint32_t open = dilate4(empt & ~e1, bm, lane);
    uint32_t go = hfill(own & open, own), gp = hfill(opp & open, opp);
    uint32_t rem = seeds & ~(go | gp);
    while (__any_sync(FULL_MASK, rem != 0u)) {
        const uint32_t vo = (shfl_up0(go, lane) | shfl_dn0(go, lane)) & own & ~go;
        const uint32_t vp = (shfl_up0(gp, lane) | shfl_dn0(gp, lane)) & opp & ~gp;
        if (!__any_sync(FULL_MASK, (vo | vp) != 0u)) break;
        go = hfill(go | vo, own);
        gp = hfill(gp | vp, opp);
        rem &= ~(go | gp);
    }
    uint32_t atari = 0u, multi = go | gp;
    while (__any_sync(FULL_MASK, rem != 0u)) {
        uint32_t bit;
        int sl = first_point(rem, &bit);
        int is_own = __shfl_sync(FULL_MASK, (own & bit) != 0u, sl);
        // only the stones next to a surrounded empty point matter below, and every one of them is a seed: a
        // group found to have two liberties need not be completed (its other seeds are classified the same way)
        uint32_t grp;
        int libs = flood_libs(lane == sl ? bit : 0u, is_own ? own : opp, empt, bm, 2, lane, &grp);
        if (libs == 1) atari |= grp;
        else if (libs > 1) multi |= grp;
        rem &= ~grp;
    }
    uint32_t possible = dilate4((own & multi) | (opp & atari), bm, lane);
    uint32_t definite = dilate4((own & atari) | (opp & multi), bm, lane);
    return occ | (e1 & possible & ~definite) | ko_row;
}

// gogame.next_state.  Returns 0 ok, -1 illegal, -2 game over (board untouched on error).
// The move is (row r, column c), or the pass when r == n (callers that already hold the coordinates skip the division).
__device__ int board_move(Board& b, int r, int c, int n, int lane) {
    if (b.done) return -2;
    const uint32_t bm = row_mask(n, lane);
    uint32_t own = b.turn ? b.wh : b.bl;
    uint32_t opp = b.turn ? b.bl : b.wh;
    uint32_t ko_row = 0u;
    if (r == n) {
        if (b.pass) b.done = 1;
        b.pass = 1;
    } else {
        MZGO_DCHECK(r >= 0 && r < n && c >= 0 && c < n && n <= MZGO_MAX_N);
        const uint32_t stone = (lane == r) ? (1u << c) : 0u;
        if (__any_sync(FULL_MASK, (b.inv & stone) != 0u)) return -1;
        b.pass = 0;
        own |= stone;
        const uint32_t occ = own | opp;
        const uint32_t adj = dilate4(stone, bm, lane);
        // adj_data: every in-board neighbour holds an opponent stone (before captures)
        const int surrounded = !__any_sync(FULL_MASK, (adj & ~opp) != 0u);
        // update_pieces: opponent groups touching the stone with no liberty (measured once)
        // (a neighbour that itself touches an empty point lives; its group mates, if any are neighbours too, are
        // then found alive by their own flood)
        uint32_t cand = adj & opp & ~dilate4(~occ & bm, bm, lane), kill = 0u;
        int nkilled = 0, single = 0;
        uint32_t single_row = 0u;
        while (__any_sync(FULL_MASK, cand != 0u)) {
            uint32_t bit;
            int sl = first_point(cand, &bit);
            uint32_t grp;                     // complete unless a liberty turned up on the way (then it lives)
            const int libs = flood_libs(lane == sl ? bit : 0u, opp, ~occ & bm, bm, 1, lane, &grp);
            if (libs == 0) {
                nkilled++;
                single = (count_upto2(grp) == 1);
                single_row = grp;
                kill |= grp;
            }
            cand &= ~grp;
        }
        opp &= ~kill;
        if (nkilled == 1 && single && surrounded) ko_row = single_row;
    }
    uint32_t inv = compute_invalid(own, opp, ko_row, n, lane);
    if (b.turn) { b.wh = own; b.bl = opp; } else { b.bl = own; b.wh = opp; }
    b.inv = inv;
    b.turn ^= 1;
    b.ply++;
    return 0;
}

__device__ __forceinline__ int board_step(Board& b, int action, int n, int lane) {
    if (b.done) return -2;
    if (action < 0 || action > n * n) return -1;
    const int r = action / n;           // the pass n*n gives r == n
    return board_move(b, r, action - r * n, n, lane);
}

// position of the k-th (0-based) set bit of v; v has more than k set bits
__device__ __forceinline__ int kth_set_bit(uint32_t v, int k) {
    int pos = 0;
#pragma unroll
    for (int w = 16; w > 0; w >>= 1) {
        const int c = __popc(v & ((1u << w) - 1u));
        if (k >= c) { k -= c; pos += w; v >>= w; }
    }
    return pos;
}

__device__ __forceinline__ void board_load(Board& b, const uint32_t* rec, int n, int lane) {
    b.bl = lane < n ? rec[lane] : 0u;
    b.wh = lane < n ? rec[n + lane] : 0u;
    b.inv = lane < n ? rec[2 * n + lane] : 0u;
    uint4 meta = *reinterpret_cast<const uint4*>(rec + 60);
    b.turn = (int)(meta.x & 1u);
    b.pass = (int)(meta.y & 1u);
    b.done = (int)((meta.y >> 1) & 1u);
    b.ply = (int)meta.z;
}

__device__ __forceinline__ void board_store(const Board& b, uint32_t* rec, int n, int lane) {
    if (lane < n) {
        rec[lane] = b.bl;
        rec[n + lane] = b.wh;
        rec[2 * n + lane] = b.inv;
    }
    if (lane == 0) {
        uint4 meta = make_uint4((uint32_t)b.turn, (uint32_t)(b.pass | (b.done << 1)), (uint32_t)b.ply, 0u);
        *reinterpret_cast<uint4*>(rec + 60) = meta;
    }
}

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
    x ^= x >> 27; x *= 0x94d049bb133111ebull;
    x ^= x >> 31;
    return x;
}

// Order-independent state hash: sum over rows of mix64(black | white<<20 | invalid<<40 | (row+1)<<59)
// plus mix64(meta).  oracle/gymgo_c.c (gymgo_c_hash) computes the same value from GymGo planes.
__device__ uint64_t mzgo_board_hash(const Board& b, int n, int lane) {
    uint64_t h = 0;
    if (lane < n) {
        h = mix64((uint64_t)b.bl | ((uint64_t)b.wh << 20) | ((uint64_t)b.inv << 40) | ((uint64_t)(lane + 1) << 59));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) h += __shfl_xor_sync(FULL_MASK, h, o);
    h += mix64(0xabcdull + (uint64_t)b.turn + 2ull * b.pass + 4ull * b.done);
    return h;
}

#include "rules_tpb.cuh"

// ------------------------------------------------------------------------------------------ kernels
constexpr int WARPS_PER_BLOCK = 4;

__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
k_rules_reset(uint32_t* boards, int nboards) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nboards * MZGO_BOARD_WORDS) boards[i] = 0u;
}

__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
k_rules_step(uint32_t* boards, const int32_t* actions, int32_t* status, int nboards, int n) {
    const int lane = lane_id();
    const int bidx = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
    if (bidx >= nboards) return;
    const int a = actions[bidx];
    int st = 0;
    if (a >= 0) {
        uint32_t* rec = boards + (size_t)bidx * MZGO_BOARD_WORDS;
        Board b;
        board_load(b, rec, n, lane);
        st = board_step(b, a, n, lane);
        if (st == 0) board_store(b, rec, n, lane);
    }
    if (status && lane == 0) status[bidx] = st;
}

__device__ void board_score(const Board& b, int n, int lane, int* black, int* white) {
    const uint32_t bm = row_mask(n, lane);
    const uint32_t empt = ~(b.bl | b.wh) & bm;
    int ba = warp_sum_i(__popc(b.bl)), wa = warp_sum_i(__popc(b.wh));
    uint32_t rem = empt;
    while (__any_sync(FULL_MASK, rem != 0u)) {
        uint32_t bit;
        int sl = first_point(rem, &bit);
        uint32_t region = flood(lane == sl ? bit : 0u, empt, lane);
        uint32_t nb = dilate4(region, bm, lane);
        int tb = __any_sync(FULL_MASK, (nb & b.bl) != 0u);
        int tw = __any_sync(FULL_MASK, (nb & b.wh) != 0u);
        int size = warp_sum_i(__popc(region));
        if (tb && !tw) ba += size;
        else if (tw && !tb) wa += size;
        rem &= ~region;
    }
    *black = ba;
    *white = wa;
}

__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
k_rules_score(const uint32_t* boards, int32_t* areas, int32_t* winner, float komi, int nboards, int n) {
    const int lane = lane_id();
    const int bidx = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
    if (bidx >= nboards) return;
    Board b;
    board_load(b, boards + (size_t)bidx * MZGO_BOARD_WORDS, n, lane);
    int ba, wa;
    board_score(b, n, lane, &ba, &wa);
    if (lane == 0) {
        if (areas) { areas[2 * bidx] = ba; areas[2 * bidx + 1] = wa; }
        if (winner) {
            float d = (float)(ba - wa) - komi;
            winner[bidx] = d > 0.f ? 1 : (d < 0.f ? -1 : 0);
        }
    }
}

template <typename T>
__global__ void k_rules_expand(const uint32_t* boards, T* planes, int nboards, int n) {
    const int np = n * n;
    const int total = nboards * 6 * np;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        int p = i % np;
        int ch = (i / np) % 6;
        int bidx = i / (6 * np);
        const uint32_t* rec = boards + (size_t)bidx * MZGO_BOARD_WORDS;
        int r = p / n, c = p - r * n;
        uint32_t v;
        if (ch < 2) v = (rec[ch * n + r] >> c) & 1u;
        else if (ch == 3) v = (rec[2 * n + r] >> c) & 1u;
        else if (ch == 2) v = rec[60] & 1u;
        else if (ch == 4) v = rec[61] & 1u;
        else v = (rec[61] >> 1) & 1u;
        planes[i] = (T)v;
    }
}

__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
k_rules_pack(const float* planes, uint32_t* boards, int nboards, int n) {
    const int lane = lane_id();
    const int bidx = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
    if (bidx >= nboards) return;
    const int np = n * n;
    const float* pl = planes + (size_t)bidx * 6 * np;
    Board b;
    b.bl = b.wh = b.inv = 0u;
    if (lane < n) {
        for (int c = 0; c < n; c++) {
            b.bl |= (pl[0 * np + lane * n + c] != 0.f) ? (1u << c) : 0u;
            b.wh |= (pl[1 * np + lane * n + c] != 0.f) ? (1u << c) : 0u;
            b.inv |= (pl[3 * np + lane * n + c] != 0.f) ? (1u << c) : 0u;
        }
    }
    b.turn = pl[2 * np] != 0.f;
    b.pass = pl[4 * np] != 0.f;
    b.done = pl[5 * np] != 0.f;
    b.ply = 0;
    uint32_t* rec = boards + (size_t)bidx * MZGO_BOARD_WORDS;
    // clear the unused words so records compare bytewise
    for (int w = lane; w < MZGO_BOARD_WORDS; w += 32) rec[w] = 0u;
    __syncwarp();
    board_store(b, rec, n, lane);
}

__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
k_rules_playout(uint32_t* boards, const uint32_t* R, int T, int32_t* actions_out, uint64_t* hash_out,
                uint8_t* trace_out, int nboards, int n) {
    const int lane = lane_id();
    const int bidx = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
    if (bidx >= nboards) return;
    const int np = n * n;
    uint32_t* rec = boards + (size_t)bidx * MZGO_BOARD_WORDS;
    Board b;
    board_load(b, rec, n, lane);
    const uint32_t bm = row_mask(n, lane);
    for (int t = 0; t < T; t++) {
        int a = -1;
        if (!b.done) {
            uint32_t valid = ~b.inv & bm;
            int cnt = __popc(valid);
            int incl = cnt;  // inclusive prefix sum over lanes
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int v = __shfl_up_sync(FULL_MASK, incl, o);
                if (lane >= o) incl += v;
            }
            int nvalid = __shfl_sync(FULL_MASK, incl, 31) + 1;
            int k = (int)(R[(size_t)bidx * T + t] % (uint32_t)nvalid);
            int excl = incl - cnt;
            uint32_t hit = __ballot_sync(FULL_MASK, k >= excl && k < incl);
            a = np;
            int hl = n, col = 0;
            if (hit) {
                hl = __ffs(hit) - 1;
                col = __shfl_sync(FULL_MASK, kth_set_bit(valid, (k - excl) & 31), hl);
                a = hl * n + col;
            }
            board_move(b, hl, col, n, lane);
        }
        if (actions_out && lane == 0) actions_out[(size_t)bidx * T + t] = a;
        if (hash_out) {
            uint64_t h = mzgo_board_hash(b, n, lane);
            if (lane == 0) hash_out[(size_t)bidx * T + t] = h;
        }
        if (trace_out) {
            uint8_t* tr = trace_out + ((size_t)bidx * T + t) * 6 * np;
            if (lane < n) {
                for (int c = 0; c < n; c++) {
                    tr[0 * np + lane * n + c] = (b.bl >> c) & 1u;
                    tr[1 * np + lane * n + c] = (b.wh >> c) & 1u;
                    tr[2 * np + lane * n + c] = (uint8_t)b.turn;
                    tr[3 * np + lane * n + c] = (b.inv >> c) & 1u;
                    tr[4 * np + lane * n + c] = (uint8_t)b.pass;
                    tr[5 * np + lane * n + c] = (uint8_t)b.done;
                }
            }
        }
    }
    board_store(b, rec, n, lane);
}

}  // namespace

extern void mzgo_count_launch();

// From this many boards on, step and playout run a THREAD per board (rules_tpb.cuh); below, a warp per board
// has more parallelism to offer (1024 self-play games are 1024 warps but only 32 warps of threads).
static int g_tpb_min = 32768;

extern "C" int mzgo_debug_rules_tpb_min(int min_boards) {
    const int old = g_tpb_min;
    if (min_boards >= 0) g_tpb_min = min_boards;
    return old;
}

template <int N>
static void launch_step_tpb(uint32_t* boards, const int32_t* actions, int32_t* status, int nboards, cudaStream_t st) {
    k_rules_step_tpb<N><<<ceil_div(nboards, TPB_THREADS), TPB_THREADS, 0, st>>>(boards, actions, status, nboards);
}
template <int N>
static void launch_playout_tpb(uint32_t* boards, const uint32_t* R, int T, int32_t* actions_out, uint64_t* hash_out,
                               int nboards, cudaStream_t st) {
    k_rules_playout_tpb<N><<<ceil_div(nboards, TPB_THREADS), TPB_THREADS, 0, st>>>(boards, R, T, actions_out, hash_out, nboards);
}
// Measured on 2^20 mid-game boards, one ply each: 9x9 1.55 -> 0.74 ms, but 19x19 4.1 -> 6.2 ms (254 registers per
// thread leave 8 warps per SM, and a warp waits for the board with the most groups to classify): by default the
// thread-per-board kernels serve boards up to 9x9; 13x13 and 19x19 only when the switch is forced to 0 (tests).
static bool tpb_size(int n) { return n == 5 || n == 6 || n == 9 || (g_tpb_min == 0 && (n == 13 || n == 19)); }

static int check_boards(const void* boards, int nboards, int n) {
    MZGO_CHECK_ARG(boards != nullptr, "boards is NULL");
    MZGO_CHECK_ARG(nboards >= 0, "nboards < 0");
    MZGO_CHECK_ARG(n >= 2 && n <= MZGO_MAX_N, "board size %d outside [2,%d]", n, MZGO_MAX_N);
    return 0;
}

extern "C" int mzgo_rules_reset(uint32_t* boards, int nboards, int n, void* stream) {
    if (int e = check_boards(boards, nboards, n)) return e;
    if (nboards == 0) return MZGO_OK;
    int total = nboards * MZGO_BOARD_WORDS;
    k_rules_reset<<<ceil_div(total, 128), 128, 0, (cudaStream_t)stream>>>(boards, nboards);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

extern "C" int mzgo_rules_step(uint32_t* boards, const int32_t* actions, int32_t* status, int nboards, int n,
                               void* stream) {
    if (int e = check_boards(boards, nboards, n)) return e;
    MZGO_CHECK_ARG(actions != nullptr, "actions is NULL");
    if (nboards == 0) return MZGO_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (nboards >= g_tpb_min && tpb_size(n)) {
        switch (n) {
            case 5: launch_step_tpb<5>(boards, actions, status, nboards, st); break;
            case 6: launch_step_tpb<6>(boards, actions, status, nboards, st); break;
            case 9: launch_step_tpb<9>(boards, actions, status, nboards, st); break;
            case 13: launch_step_tpb<13>(boards, actions, status, nboards, st); break;
            default: launch_step_tpb<19>(boards, actions, status, nboards, st); break;
        }
    } else
    k_rules_step<<<ceil_div(nboards, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, 0, st>>>(
        boards, actions, status, nboards, n);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

extern "C" int mzgo_rules_score(const uint32_t* boards, int32_t* areas, int32_t* winner, float komi, int nboards,
                                int n, void* stream) {
    if (int e = check_boards(boards, nboards, n)) return e;
    if (nboards == 0) return MZGO_OK;
    k_rules_score<<<ceil_div(nboards, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, 0, (cudaStream_t)stream>>>(
        boards, areas, winner, komi, nboards, n);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

extern "C" int mzgo_rules_expand_planes(const uint32_t* boards, void* planes, int dtype, int nboards, int n,
                                        void* stream) {
    if (int e = check_boards(boards, nboards, n)) return e;
    MZGO_CHECK_ARG(planes != nullptr, "planes is NULL");
    MZGO_CHECK_ARG(dtype >= 0 && dtype <= 2, "dtype %d not in {0 f32, 1 f64, 2 u8}", dtype);
    if (nboards == 0) return MZGO_OK;
    long total = (long)nboards * 6 * n * n;
    int blocks = (int)((total + 255) / 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == 0) k_rules_expand<float><<<blocks, 256, 0, s>>>(boards, (float*)planes, nboards, n);
    else if (dtype == 1) k_rules_expand<double><<<blocks, 256, 0, s>>>(boards, (double*)planes, nboards, n);
    else k_rules_expand<uint8_t><<<blocks, 256, 0, s>>>(boards, (uint8_t*)planes, nboards, n);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

extern "C" int mzgo_rules_pack_planes(const float* planes, uint32_t* boards, int nboards, int n, void* stream) {
    if (int e = check_boards(boards, nboards, n)) return e;
    MZGO_CHECK_ARG(planes != nullptr, "planes is NULL");
    if (nboards == 0) return MZGO_OK;
    k_rules_pack<<<ceil_div(nboards, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, 0, (cudaStream_t)stream>>>(
        planes, boards, nboards, n);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

extern "C" int mzgo_rules_playout(uint32_t* boards, const uint32_t* R, int T, int32_t* actions_out,
                                  uint64_t* hash_out, uint8_t* trace_out, int nboards, int n, void* stream) {
    if (int e = check_boards(boards, nboards, n)) return e;
    MZGO_CHECK_ARG(R != nullptr, "R is NULL");
    MZGO_CHECK_ARG(T >= 0, "T < 0");
    if (nboards == 0 || T == 0) return MZGO_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (nboards >= g_tpb_min && tpb_size(n) && !trace_out) {
        switch (n) {
            case 5: launch_playout_tpb<5>(boards, R, T, actions_out, hash_out, nboards, st); break;
            case 6: launch_playout_tpb<6>(boards, R, T, actions_out, hash_out, nboards, st); break;
            case 9: launch_playout_tpb<9>(boards, R, T, actions_out, hash_out, nboards, st); break;
            case 13: launch_playout_tpb<13>(boards, R, T, actions_out, hash_out, nboards, st); break;
            default: launch_playout_tpb<19>(boards, R, T, actions_out, hash_out, nboards, st); break;
        }
    } else
    k_rules_playout<<<ceil_div(nboards, WARPS_PER_BLOCK), WARPS_PER_BLOCK * 32, 0, st>>>(
        boards, R, T, actions_out, hash_out, trace_out, nboards, n);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}
