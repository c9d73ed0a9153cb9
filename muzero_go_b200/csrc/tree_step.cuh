// The fused tree step (expand + backup of the evaluated leaf, selection of the next) as device code shared by
// the stand-alone kernel k_tree_step (tree.cu) and the tree phase of the network kernel (nets_tc.cu).
#pragma once
#include "tree_math.cuh"

namespace {

// Trace build (python -m muzero_go_b200.build --define MZGO_TRACE_TREE; tools/tree_trace.py): per-game globaltimer
// stamps of the step's phases.  A stamp takes a loaded value as an operand, so it is taken when that value ARRIVES.
#ifdef MZGO_TRACE_TREE
__device__ unsigned long long* g_tree_trace = nullptr;     // [G][16]
#define TSTAMP(slot, dep)                                                                          \
    do {                                                                                           \
        if (tid == 0 && do_select && g_tree_trace) {                                               \
            unsigned long long _t;                                                                 \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(_t) : "r"((int)(dep)) : "memory");   \
            g_tree_trace[(size_t)g * 16 + (slot)] = _t;                                            \
        }                                                                                          \
    } while (0)
#else
#define TSTAMP(slot, dep) ((void)0)
#endif

constexpr int STEP_THREADS = 128;
constexpr int STEP_EPT = 3;   // edges per thread: A <= 3 * TPG
static_assert(STEP_EPT == 3, "the root-edge patch of tree_step_body selects among three register slots");

__device__ __forceinline__ int npar_root_after(int root_visit0, int act) { return root_visit0 + (act >= 0 ? 1 : 0); }

// TPG == 32: the game's warp.  TPG == 128: the whole block (BAR == 0, stand-alone kernel) or one of the 128-thread
// sub-blocks of the network kernel's tree phase (named barrier BAR + sub-block index).
template <int TPG, int BAR>
__device__ __forceinline__ void step_sync() {
    if (TPG == 32) __syncwarp();
    else if (BAR == 0) __syncthreads();
    else asm volatile("bar.sync %0, 128;" ::"r"(BAR + (int)(threadIdx.x >> 7)) : "memory");
}

template <int TPG>
struct StepShared {
    static constexpr int GPB = STEP_THREADS / TPG, WPG = TPG / 32, NW = STEP_EPT * WPG;
    uint32_t open[GPB][NW];      // ballot words per game (word w covers actions [32w, 32w+32))
    uint32_t live[GPB][WPG];
    unsigned long long kmin[GPB][WPG], kmax[GPB][WPG], kbest[GPB][WPG];   // order-preserving keys of float64 values (dkey)
    int besta[GPB][WPG], bestn[GPB][WPG], bestc[GPB][WPG];
    float pri[GPB][STEP_EPT * TPG];   // self_play.py variant: masked child priors before the renormalisation
    float tot[GPB];
};

// Order-preserving map of the non-NaN float64 values onto uint64: a < b <=> dkey(a) < dkey(b), -0.0 just below +0.0.
// Minimum, maximum and argmax over a warp then take two 32-bit redux.sync operations (high word, then the low word
// among the lanes that hold the winning high word) instead of five rounds of 64-bit shuffles and compares.
__device__ __forceinline__ unsigned long long dkey(double x) {
    const long long b = __double_as_longlong(x);
    return (unsigned long long)(b ^ ((b >> 63) | (long long)0x8000000000000000ull));
}
__device__ __forceinline__ double dkey_inv(unsigned long long k) {
    const long long b = (long long)k;
    return __longlong_as_double(b < 0 ? (b ^ (long long)0x8000000000000000ull) : ~b);
}
__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long k) {
    const uint32_t hi = (uint32_t)(k >> 32), lo = (uint32_t)k;
    const uint32_t mh = __reduce_max_sync(FULL_MASK, hi);
    const uint32_t ml = __reduce_max_sync(FULL_MASK, hi == mh ? lo : 0u);
    return ((unsigned long long)mh << 32) | ml;
}
__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long k) {
    const uint32_t hi = (uint32_t)(k >> 32), lo = (uint32_t)k;
    const uint32_t mh = __reduce_min_sync(FULL_MASK, hi);
    const uint32_t ml = __reduce_min_sync(FULL_MASK, hi == mh ? lo : 0xffffffffu);
    return ((unsigned long long)mh << 32) | ml;
}

// position of the k-th (0-based) set bit of v; v has more than k set bits
__device__ __forceinline__ int nth_set_bit(uint32_t v, int k) {
    int pos = 0;
#pragma unroll
    for (int w = 16; w > 0; w >>= 1) {
        const int c = __popc(v & ((1u << w) - 1u));
        if (k >= c) { k -= c; pos += w; v >>= w; }
    }
    return pos;
}

// The terms of the PUCT score (main.py:328-358) of ONE edge that do not depend on the level's min-max normalisation:
// q = value_sum / visits (0 for an unvisited edge) and u = c P sqrt(N) / (1 + n).  sq = sqrt(N) of the level.
__device__ __forceinline__ void edge_terms(int nv, double cp, double vs, double sq, double* q, double* u) {
    *q = 0.0;
    if (nv > 0) *q = (fabs(vs) < 1e6) ? div_by_count(vs, nv) : ddiv_generic(vs, (double)nv);
    const double num = __dmul_rn(cp, sq);
    *u = (num < 1e6) ? div_by_count(num, 1 + nv) : ddiv_generic(num, (double)(1 + nv));
}
// sqrt(N) of a level whose parent has npar visits: main.py:354 / self_play.py:297-300
__device__ __forceinline__ double level_sqrt(int npar, bool v1) { return __dsqrt_rn((double)(v1 ? max(1, npar) : npar + 1)); }

template <int TPG>
__device__ __forceinline__ void level_terms(const int (&nv)[STEP_EPT], const double (&cp)[STEP_EPT], const double (&vs)[STEP_EPT],
                                            uint32_t pos, double sq, double (&q)[STEP_EPT], double (&u)[STEP_EPT]) {
#pragma unroll
    for (int e = 0; e < STEP_EPT; e++) {
        q[e] = 0.0; u[e] = 0.0;
        if ((pos >> e) & 1u) edge_terms(nv[e], cp[e], vs[e], sq, &q[e], &u[e]);
    }
}

// One PUCT level (main.py:318-358) on register-resident edge data: visits nv, value sums vs, c * prior cp, child links
// ch, the live set pos (prior > 0), sq = sqrt(N).  The edge terms q / u (level_terms) are computed here unless the
// caller has them (have_terms, uniform for the warp: the root level, in front of the wait) -- and only by warps that
// hold no unexpanded edge, i.e. not on a level that ends in a draw.  r = this step's random number (one unexpanded
// edge is drawn per step at most).  Returns 0: an unexpanded edge was drawn (*out_a), 1: descend through *out_a into child *out_c (which
// has *out_n visits), 2: nothing selectable.
// TPG == 128: two block barriers per PUCT level (ballots + the warps' q range; the warps' best edges), one per level
// that ends in a draw; no barrier between levels -- every scratch array is written before one barrier of a level and
// read before the next one, so a warp that runs ahead into the next level cannot overwrite what another still reads.
template <int TPG, int BAR>
__device__ __forceinline__ int step_level(StepShared<TPG>& sh, int sub, int tid, const int (&nv)[STEP_EPT],
                                          const double (&cp)[STEP_EPT], const double (&vs)[STEP_EPT], double sq,
                                          bool have_terms, double (&q)[STEP_EPT], double (&u)[STEP_EPT],
                                          const int (&ch)[STEP_EPT], uint32_t pos, uint32_t r, int* out_a, int* out_n,
                                          int* out_c) {
    constexpr int WPG = TPG / 32, NW = STEP_EPT * WPG;
    const int wig = tid >> 5, lane = tid & 31;
    uint32_t ow[NW];                             // TPG == 32: the ballots themselves; else filled from shared memory
    uint32_t mine = 0u;
#pragma unroll
    for (int e = 0; e < STEP_EPT; e++) {
        const bool opn = ((pos >> e) & 1u) && nv[e] == 0;
        const uint32_t b = __ballot_sync(FULL_MASK, opn);
        if (WPG == 1) ow[e] = b;
        else if (lane == 0) sh.open[sub][e * WPG + wig] = b;
        mine |= b;
    }
    uint32_t all_live = __ballot_sync(FULL_MASK, pos != 0u);
    // the range of q over the live edges (min-max normalisation set = prior > 0, main.py:328-339); a warp that holds
    // an unexpanded edge knows that the level ends in a draw and skips it
    unsigned long long kmin = dkey(INFINITY), kmax = dkey(-INFINITY);
    if (WPG == 1 && mine == 0u) {
        // A node with ONE live edge, already visited (late positions of small boards: only the pass is legal, and the
        // tree is a chain as deep as the number of simulations): PUCT over one candidate picks it whenever its score
        // is a number above -inf, i.e. whenever its statistics are finite -- no division, no normalisation, no argmax.
        uint32_t lb[STEP_EPT];
        int live_n = 0;
#pragma unroll
        for (int e = 0; e < STEP_EPT; e++) { lb[e] = __ballot_sync(FULL_MASK, (pos >> e) & 1u); live_n += __popc(lb[e]); }
        if (live_n == 1) {
            const int e_o = lb[0] ? 0 : (lb[1] ? 1 : 2);
            const int lane_o = __ffs(lb[0] | lb[1] | lb[2]) - 1;
            const int nv_o = __shfl_sync(FULL_MASK, e_o == 0 ? nv[0] : e_o == 1 ? nv[1] : nv[2], lane_o);
            const int ch_o = __shfl_sync(FULL_MASK, e_o == 0 ? ch[0] : e_o == 1 ? ch[1] : ch[2], lane_o);
            const double vs_o = __shfl_sync(FULL_MASK, e_o == 0 ? vs[0] : e_o == 1 ? vs[1] : vs[2], lane_o);
            const double cp_o = __shfl_sync(FULL_MASK, e_o == 0 ? cp[0] : e_o == 1 ? cp[1] : cp[2], lane_o);
            if (fabs(vs_o) < INFINITY && fabs(cp_o) < INFINITY && fabs(sq) < INFINITY) {
                *out_a = e_o * TPG + lane_o;
                *out_n = nv_o;
                *out_c = ch_o;
                return 1;
            }
        }
    }
    if (mine == 0u) {
        if (!have_terms) level_terms<TPG>(nv, cp, vs, pos, sq, q, u);
        double qmin = INFINITY, qmax = -INFINITY;
#pragma unroll
        for (int e = 0; e < STEP_EPT; e++) {
            if ((pos >> e) & 1u) { qmin = fmin(qmin, q[e]); qmax = fmax(qmax, q[e]); }
        }
        kmin = warp_min_u64(dkey(qmin));
        kmax = warp_max_u64(dkey(qmax));
    }
    if (WPG > 1) {
        if (lane == 0) { sh.live[sub][wig] = all_live; sh.kmin[sub][wig] = kmin; sh.kmax[sub][wig] = kmax; }
        step_sync<TPG, BAR>();
#pragma unroll
        for (int w = 0; w < NW; w++) ow[w] = sh.open[sub][w];
#pragma unroll
        for (int w = 0; w < WPG; w++) {
            all_live |= sh.live[sub][w];
            kmin = min(kmin, sh.kmin[sub][w]);
            kmax = max(kmax, sh.kmax[sub][w]);
        }
    }
    int n_open = 0;
#pragma unroll
    for (int w = 0; w < NW; w++) n_open += __popc(ow[w]);
    if (n_open > 0) {
        // random.choice(unexpanded), main.py:322-325
        int k = (int)(r % (uint32_t)n_open);
        int w_sel = -1;
        uint32_t word = 0u;
#pragma unroll
        for (int w = 0; w < NW; w++) {
            const int c = __popc(ow[w]);
            if (w_sel < 0) {
                if (k < c) { w_sel = w; word = ow[w]; }
                else k -= c;
            }
        }
        // word w = e * WPG + wig covers the edges e * TPG + wig * 32 + [0, 32): the same numbering as the action index
        *out_a = w_sel * 32 + nth_set_bit(word, k);
        return 0;
    }
    if (!all_live) return 2;                     // nothing selectable: simulation dropped / terminal leaf
    const double qmin = dkey_inv(kmin), qmax = dkey_inv(kmax);
    const bool norm = qmax > qmin;
    const double range = __dsub_rn(qmax, qmin);
    // 1/range once per level; a range outside the proven window of div_rn_with_rcp takes the generic division
    const bool fast = norm && range > 1e-17 && range < 1e3 && fabs(qmin) < 1e6 && fabs(qmax) < 1e6;
    const double rrange = fast ? __drcp_rn(range) : 0.0;
    double best = -INFINITY;
    int best_a = -1, best_n = 0, best_c = -1;
#pragma unroll
    for (int e = 0; e < STEP_EPT; e++) {
        if ((pos >> e) & 1u) {
            const double dq = __dsub_rn(q[e], qmin);
            const double qn = norm ? (fast ? div_rn_with_rcp(dq, range, rrange) : ddiv_generic(dq, range)) : q[e];
            const double score = __dadd_rn(qn, u[e]);
            if (score > best) { best = score; best_a = e * TPG + tid; best_n = nv[e]; best_c = ch[e]; }   // e ascending = action ascending
        }
    }
    // argmax, ties to the lowest action: key 0 = no candidate (NaN and -inf scores never win, as in the reference's
    // `score > best_score` with best_score = -inf); -0.0 + 0.0 = +0.0 makes equal scores equal keys
    unsigned long long kb = best_a >= 0 ? dkey(__dadd_rn(best, 0.0)) : 0ull;
    const unsigned long long kw = warp_max_u64(kb);
    const uint32_t aw = __reduce_min_sync(FULL_MASK, (kb == kw && best_a >= 0) ? (uint32_t)best_a : 0xffffffffu);
    best_a = (int)aw;                            // -1 when the warp has no candidate
    best_n = __shfl_sync(FULL_MASK, best_n, (int)(aw & 31u));   // the winning edge e * TPG + wig * 32 + lane sits in lane aw % 32
    best_c = __shfl_sync(FULL_MASK, best_c, (int)(aw & 31u));
    kb = kw;
    if (WPG > 1) {
        if (lane == 0) { sh.kbest[sub][wig] = kb; sh.besta[sub][wig] = best_a; sh.bestn[sub][wig] = best_n; sh.bestc[sub][wig] = best_c; }
        step_sync<TPG, BAR>();
        kb = sh.kbest[sub][0]; best_a = sh.besta[sub][0]; best_n = sh.bestn[sub][0]; best_c = sh.bestc[sub][0];
#pragma unroll
        for (int w = 1; w < WPG; w++) {
            const unsigned long long ok = sh.kbest[sub][w];
            const int oa = sh.besta[sub][w], on = sh.bestn[sub][w], oc = sh.bestc[sub][w];
            const bool take = (oa >= 0) && (best_a < 0 || ok > kb || (ok == kb && oa < best_a));
            if (take) { kb = ok; best_a = oa; best_n = on; best_c = oc; }
        }
    }
    if (best_a < 0) return 2;                    // only NaN / -inf scores: the reference falls out of the loop too
    *out_a = best_a;
    *out_n = best_n;
    *out_c = best_c;
    return 1;
}

// The step of game g by TPG threads (tid = 0..TPG-1; sub = index of the game's scratch in sh).
// wait_outputs() returns once soft / reward / value of game g are readable (a no-op in the stand-alone kernel;
// in the network kernel another warp of the CTA produces them, hence the ld.global.cg reads).  Everything
// that does not depend on the network outputs is loaded BEFORE the wait.
template <int TPG, int BAR, bool EARLY, bool V1, class Wait>
__device__ __forceinline__ void tree_step_body(const mzgo_tree_t& t, const mzgo_search_params_t& p, const float* soft,
                                               const float* reward, const float* value, const uint32_t* choice_stream,
                                               int stride, int do_backup, int do_select, StepShared<TPG>& sh, int sub,
                                               int tid, int g, Wait wait_outputs) {
    constexpr bool early_loads = EARLY;
    constexpr bool v1 = V1;                      // p.variant == MZGO_VARIANT_SELFPLAY, resolved by the caller (the variant's code is a quarter of the step)
    const int A = t.A, M = t.M;
    int* path = reinterpret_cast<int*>(t.prior + (size_t)g * M * A);   // scratch row of the root (its priors are float64, elsewhere)
    const size_t rbase = (size_t)g * M * A;      // edges of the root

    // ---------------------------------------------------------------- round trip 1: every load of TREE data
    // Nothing here was written by the kernel before this one (the network kernel only writes latents and its
    // outputs; the trees were last written by the previous tree step, which completed before the network kernel
    // could start), so in the stand-alone kernel all of it is issued BEFORE the grid-dependency wait: the blocks
    // become resident while the network kernel's last CTAs drain and spend that wait fetching.  Only when the caller
    // vouches for that order (early_loads: the steps mzgo_search_run enqueues behind a network kernel): the first step
    // of a search follows the root preparation and a step called back to back with another one follows a tree step,
    // both of which DO write tree data -- those wait first.
    if (!early_loads) wait_outputs();
    const int act = do_backup ? t.leaf_action[g] : -1;
    const int par = t.leaf_parent[g];
    const int nw = t.leaf_node[g];
    const int depth_b = t.leaf_depth[g];
    const int is_active = t.active[g];
    const uint32_t move_index = (uint32_t)*t.move_index_dev;   // written by the root preparation of this move
    const int pk0 = path[0];                     // first edge of the stored path (an edge of the root)
    const int pk_raw = tid < A ? path[tid] : -1;
    const int root_visit0 = t.root_visit[g];
    const uint32_t ctr0 = t.choice_ctr[g];
    const int nn0 = t.n_nodes[g];
    const float* mask = t.root_mask + (size_t)g * A;
    float mk[STEP_EPT], sf[STEP_EPT];
    int nv_r[STEP_EPT], ch_r[STEP_EPT];
    double pr_r[STEP_EPT], vs_r[STEP_EPT];
#pragma unroll
    for (int e = 0; e < STEP_EPT; e++) {
        const int a = e * TPG + tid;
        mk[e] = 0.f; sf[e] = 0.f; nv_r[e] = 0; ch_r[e] = -1; pr_r[e] = 0.0; vs_r[e] = 0.0;
        if (a < A) {
            mk[e] = mask[a];
            if (do_select) {
                nv_r[e] = t.edge_visit[rbase + a];
                pr_r[e] = t.root_prior[(size_t)g * A + a];
                vs_r[e] = t.edge_vsum[rbase + a];
                ch_r[e] = (int)t.child[rbase + a];
            }
        }
    }
    MZGO_DCHECK(g >= 0 && g < t.G && A <= STEP_EPT * TPG && nn0 >= 0 && nn0 <= M);
    MZGO_DCHECK(act < A && (act < 0 || (par >= 0 && par < nn0 && nw == nn0 && nw < M && depth_b >= 1)));
    MZGO_DCHECK(!(act >= 0 && depth_b <= A && tid < depth_b) || ((pk_raw >> 9) < nn0 && (pk_raw & 511) < A && pk_raw >= 0));
    TSTAMP(2, act + nv_r[0] + ch_r[STEP_EPT - 1] + pk_raw);

    // the statistics of the path edges this step backs up: a second, dependent batch of tree data
    const int pk_mine = (act >= 0 && depth_b <= A && tid < depth_b) ? pk_raw : -1;
    int ev_mine = 0;
    double es_mine = 0.0;
    if (pk_mine >= 0) {
        const size_t e = ((size_t)g * M + (pk_mine >> 9)) * A + (pk_mine & 511);
        ev_mine = t.edge_visit[e];
        es_mine = t.edge_vsum[e];
    }
    const double root_vsum0 = t.root_vsum[g];
    const int sims_done0 = t.sims_done[g];
    const long long depth_sum0 = t.depth_sum[g];

    // ---------------------------------------------------------------- work that needs tree data only
    // (behind the engine's network kernel it runs while that kernel drains, in front of the wait)
    // the step's one random number: random.choice(unexpanded) ends the descent, main.py:322-325
    const size_t stream_off = (size_t)g * stride;
    uint32_t rnd = 0u;
    if (do_select) rnd = choice_stream ? choice_stream[stream_off + ctr0]
                                       : mzgo_rand_u32(p.seed, (uint32_t)(g + p.game_base), 0x20000u + move_index, ctr0);
    // behind the wait shadow (early_loads) also the root level's edge terms, from the statistics as they are BEFORE this
    // step's backup: the backup changes one root edge (the first edge of the path), whose terms are redone below from
    // the patched statistics
    int npar = root_visit0 + (act >= 0 ? 1 : 0);
    double cp[STEP_EPT], q_r[STEP_EPT], u_r[STEP_EPT];
    uint32_t pos = 0u;
    double sq = 0.0;
    if (do_select) {
        sq = level_sqrt(npar, v1);
#pragma unroll
        for (int e = 0; e < STEP_EPT; e++) {
            cp[e] = __dmul_rn(p.c_puct, pr_r[e]);    // root: float64 priors (main.py:276-282)
            if (e * TPG + tid < A && pr_r[e] > 0.0) pos |= 1u << e;
        }
        if (early_loads) level_terms<TPG>(nv_r, cp, vs_r, pos, sq, q_r, u_r);   // else: by the level itself, if it needs them
    }

    // The stores of expand + backup that need no network output: the new node's empty statistics, the visit counts
    // along the path, the child / parent links and the per-game counters.  EARLY: issued here, in front of the wait
    // (nobody reads or writes them while the network kernel runs); else as part of expand + backup below.
    auto stores_without_outputs = [&]() {
        if (act < 0) return;
        const size_t nbase = ((size_t)g * M + nw) * A;
#pragma unroll
        for (int e = 0; e < STEP_EPT; e++) {
            const int a = e * TPG + tid;
            if (a < A) {
                t.edge_visit[nbase + a] = 0;
                t.edge_vsum[nbase + a] = 0.0;
                t.child[nbase + a] = (int16_t)-1;
            }
        }
        if (pk_mine >= 0) t.edge_visit[((size_t)g * M + (pk_mine >> 9)) * A + (pk_mine & 511)] = ev_mine + 1;
        if (tid == 0) {
            t.child[((size_t)g * M + par) * A + act] = (int16_t)nw;
            t.parent[(size_t)g * M + nw] = par * 512 + act;
            t.n_nodes[g] = nw + 1;
            t.root_visit[g] = root_visit0 + 1;
            t.sims_done[g] = sims_done0 + 1;
            t.depth_sum[g] = depth_sum0 + depth_b;
        }
    };
    if (EARLY) stores_without_outputs();

    // ---------------------------------------------------------------- the network outputs of this game's leaf
    if (early_loads) wait_outputs();
    TSTAMP(1, 0);
    const float rew_g = do_backup ? __ldcg(reward + g) : 0.f, val_g = do_backup ? __ldcg(value + g) : 0.f;
    if (do_backup) {
#pragma unroll
        for (int e = 0; e < STEP_EPT; e++) {
            const int a = e * TPG + tid;
            if (a < A) sf[e] = __ldcg(soft + (size_t)g * A + a);
        }
    }

    // ---------------------------------------------------------------- expand + backup
    // backup_value = reward.item() + discount * value.item()   (main.py:311), float64
    const double bv = __dadd_rn((double)rew_g, __dmul_rn(p.discount, (double)val_g));
    if (act >= 0 && v1) {
        // policy_child *= valid_mask; /= sum (float32 pairwise)   self_play.py:210-214
#pragma unroll
        for (int e = 0; e < STEP_EPT; e++) {
            const int a = e * TPG + tid;
            if (a < A) sh.pri[sub][a] = masked_child_prior(sf[e], mk[e], a == A - 1, p.pass_epsilon);
        }
    }
    if (v1) {                                    // uniform for the block: p is a kernel parameter
        step_sync<TPG, BAR>();
        if (tid == 0 && act >= 0) sh.tot[sub] = np_pairwise_sum<float>(sh.pri[sub], A);
        step_sync<TPG, BAR>();
    }
    if (act >= 0) {
        const int depth = depth_b;
        const size_t nbase = ((size_t)g * M + nw) * A;
#pragma unroll
        for (int e = 0; e < STEP_EPT; e++) {
            const int a = e * TPG + tid;
            if (a < A) {
                // prior_child = policy_child[a] if valid_mask[a] > 0 else 0   (the ROOT's mask, main.py:303)
                float pr = mk[e] > 0.f ? sf[e] : 0.f;
                if (v1) pr = renorm_child_prior(sh.pri[sub][a], sh.tot[sub], mk[e], a == A - 1, p.pass_epsilon, mask, A);
                t.prior[nbase + a] = pr;
            }
        }
        if (!EARLY) stores_without_outputs();
        // self_play.py:340-343: node i levels above the new leaf gets value * (-1)^i; edge d of the path holds the
        // statistics of the node at depth d + 1, i.e. i = depth - 1 - d, and the root has i = depth
        if (depth <= A) {
            if (pk_mine >= 0) {
                const size_t e = ((size_t)g * M + (pk_mine >> 9)) * A + (pk_mine & 511);
                t.edge_vsum[e] = __dadd_rn(es_mine, (v1 && ((depth - 1 - tid) & 1)) ? -bv : bv);
            }
            for (int d = tid + TPG; d < depth; d += TPG) {
                const int pk = path[d];
                MZGO_DCHECK(pk >= 0 && (pk >> 9) < M && (pk & 511) < A);
                const size_t e = ((size_t)g * M + (pk >> 9)) * A + (pk & 511);
                t.edge_visit[e] += 1;
                t.edge_vsum[e] = __dadd_rn(t.edge_vsum[e], (v1 && ((depth - 1 - d) & 1)) ? -bv : bv);
            }
        }
        if (tid == 0) {
            if (depth > A) {                     // path longer than the scratch row: walk the parent links
                int node = nw;
                double sv = bv;
                while (node != 0) {
                    MZGO_DCHECK(node > 0 && node < M);
                    const int pk = t.parent[(size_t)g * M + node];
                    const int pn = pk >> 9, pa = pk & 511;
                    MZGO_DCHECK(pk >= 0 && pn < M && pa < A);
                    const size_t e = ((size_t)g * M + pn) * A + pa;
                    t.edge_visit[e] += 1;
                    t.edge_vsum[e] = __dadd_rn(t.edge_vsum[e], sv);
                    if (v1) sv = -sv;
                    node = pn;
                }
            }
            t.root_vsum[g] = __dadd_rn(root_vsum0, (v1 && (depth & 1)) ? -bv : bv);
        }
        // the root level fetched above predates this backup: patch the one root edge on the path (the same
        // float64 addition the backup thread performs on the same operands) and the new child link
        const int a_first = pk0 & 511;           // stored paths always start with an edge of the root
        const double sv0 = (v1 && ((depth - 1) & 1)) ? -bv : bv;
        if (do_select) {
            const int e_first = a_first / TPG;
            if (a_first - e_first * TPG == tid) {    // one thread of the game: its edge e_first is the path's root edge
                const int nv1 = (e_first == 0 ? nv_r[0] : e_first == 1 ? nv_r[1] : nv_r[2]) + 1;
                const double vs1 = __dadd_rn(e_first == 0 ? vs_r[0] : e_first == 1 ? vs_r[1] : vs_r[2], sv0);
                const double cp1 = e_first == 0 ? cp[0] : e_first == 1 ? cp[1] : cp[2];
                double q1 = 0.0, u1 = 0.0;
                if (early_loads && ((pos >> e_first) & 1u)) edge_terms(nv1, cp1, vs1, sq, &q1, &u1);
#pragma unroll
                for (int e = 0; e < STEP_EPT; e++)
                    if (e == e_first) { nv_r[e] = nv1; vs_r[e] = vs1; q_r[e] = q1; u_r[e] = u1; }
            }
#pragma unroll
            for (int e = 0; e < STEP_EPT; e++)
                if (par == 0 && e * TPG + tid == act) ch_r[e] = nw;
        }
    }
    TSTAMP(3, ev_mine + __float_as_int(sf[0]));
    if (!do_select) return;
    step_sync<TPG, BAR>();                            // the statistics written above are read below by other threads
    TSTAMP(4, 0);

    // ---------------------------------------------------------------- select
    if (!is_active) {
        if (tid == 0) { t.leaf_parent[g] = -1; t.leaf_action[g] = -1; }
        return;
    }
    int node = 0, depth = 1;
    int sel_parent = -1, sel_action = -1;
    // The level loop works on ONE set of register arrays (nv_r / ch_r / q_r / u_r): they hold the prefetched and
    // patched root level on entry and are reloaded for every level below, so there is a single copy of the
    // level code (the kernel is instruction-cache sensitive).
    while (true) {
        int oa = -1, on = 0, oc = -1;
        const int rc = step_level<TPG, BAR>(sh, sub, tid, nv_r, cp, vs_r, sq, EARLY, q_r, u_r, ch_r, pos, rnd,
                                            &oa, &on, &oc);
        TSTAMP(min(5 + 2 * (depth - 1), 13), rc + oa);
        if (rc == 0) { sel_parent = node; sel_action = oa; break; }
        if (rc == 2 || oc <= 0) break;           // terminal leaf / dropped simulation (node keeps the leaf)
        MZGO_DCHECK(oa >= 0 && oa < A && oc < M && oc < (act >= 0 ? nw + 1 : nn0));
        if (tid == 0 && depth <= A) path[depth - 1] = node * 512 + oa;
        node = oc; npar = on; depth++;
        const size_t base = ((size_t)g * M + node) * A;
        float prf[STEP_EPT];
        // all loads of the level first, so that they are in flight together (and the level's sqrt under them)
#pragma unroll
        for (int e = 0; e < STEP_EPT; e++) {
            const int a = e * TPG + tid;
            nv_r[e] = 0; prf[e] = 0.f; vs_r[e] = 0.0; ch_r[e] = -1;
            if (a < A) {
                nv_r[e] = t.edge_visit[base + a];
                prf[e] = t.prior[base + a];
                vs_r[e] = t.edge_vsum[base + a];
                ch_r[e] = (int)t.child[base + a];
            }
        }
        sq = level_sqrt(npar, v1);
        TSTAMP(min(6 + 2 * (depth - 2), 14), nv_r[0] + ch_r[STEP_EPT - 1] + __float_as_int(prf[0]));
        pos = 0u;
#pragma unroll
        for (int e = 0; e < STEP_EPT; e++) {
            cp[e] = (double)__fmul_rn((float)p.c_puct, prf[e]);   // c*P rounds in float32 (NEP 50)
            if (prf[e] > 0.f) pos |= 1u << e;
        }
        if (EARLY) level_terms<TPG>(nv_r, cp, vs_r, pos, sq, q_r, u_r);
    }
    if (tid == 0) {
        const int nn = act >= 0 ? nw + 1 : nn0;
        if (sel_action < 0 && v1) {
            // terminal leaf: backpropagate(path, 0) -- a visit for every node on the path (self_play.py:188-190)
            int nd = node;
            while (nd != 0) {
                MZGO_DCHECK(nd > 0 && nd < M);
                const int pk = t.parent[(size_t)g * M + nd];
                MZGO_DCHECK(pk >= 0 && (pk >> 9) < M && (pk & 511) < A);
                t.edge_visit[((size_t)g * M + (pk >> 9)) * A + (pk & 511)] += 1;
                nd = pk >> 9;
            }
            t.root_visit[g] = npar_root_after(root_visit0, act) + 1;
        }
        if (sel_action >= 0) t.choice_ctr[g] = ctr0 + 1u;
        if (sel_action >= 0 && nn >= M) sel_action = -1;   // arena full: drop (cannot happen when M > simulations)
        MZGO_DCHECK(sel_action < A && (sel_action < 0 || (sel_parent >= 0 && sel_parent < nn && nn < M)));
        if (sel_action >= 0 && depth <= A) path[depth - 1] = sel_parent * 512 + sel_action;
        t.leaf_parent[g] = sel_action >= 0 ? sel_parent : -1;
        t.leaf_action[g] = sel_action;
        t.leaf_node[g] = nn;
        t.leaf_depth[g] = depth;
        t.leaf_in_slot[g] = sel_action >= 0 ? g * M + sel_parent : -1;
        t.leaf_out_slot[g] = sel_action >= 0 ? g * M + nn : -1;
        if (sel_action < 0) t.dropped[g] += 1;
    }
    TSTAMP(15, 0);
}


}  // namespace
