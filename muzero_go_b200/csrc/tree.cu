// MCTS trees as structure-of-arrays, one warp per game.
//
// Replaces MCTS.run / select_leaf / backpropagate of /root/reference/main.py:247-368 (and play.py:104-216,
// the same algorithm with c_puct=10.0).  The arithmetic keeps the reference's dtypes so that visit counts
// are bit-exact when both sides consume identical float32 network outputs (SURVEY.md 8a "dtype facts"):
//   root priors float64, child priors float32, PUCT score float64 with every operation individually
//   rounded (no FMA contraction), ties broken towards the lowest action index.
#include "tree_step.cuh"

namespace {

constexpr int TREE_WARPS = 4;
constexpr int MAX_EPL = 12;  // edges per lane: A <= 384

// ---- device Dirichlet sampler (production path; distribution parity only) ----
__device__ __forceinline__ float u01(uint32_t x) { return ((float)(x >> 8) + 0.5f) * (1.0f / 16777216.0f); }

__device__ float sample_gamma(float alpha, uint64_t seed, uint32_t game, uint32_t stream, uint32_t base) {
    // Marsaglia-Tsang for alpha+1, boosted by U^(1/alpha)
    const float d = alpha + 1.0f - 1.0f / 3.0f;
    const float c = rsqrtf(9.0f * d);
    uint32_t ctr = base;
    float out = d;
    for (int it = 0; it < 64; it++) {
        uint32_t r[4] = {ctr++, stream, game, 0x67616d6du};
        philox4x32_10(r, (uint32_t)seed, (uint32_t)(seed >> 32));
        float u1 = u01(r[0]), u2 = u01(r[1]), u3 = u01(r[2]);
        float x = sqrtf(-2.0f * __logf(u1)) * __cosf(6.28318530718f * u2);
        float v = 1.0f + c * x;
        if (v <= 0.f) continue;
        v = v * v * v;
        if (__logf(u3) < 0.5f * x * x + d - d * v + d * __logf(v)) { out = d * v; break; }
    }
    uint32_t r[4] = {ctr, stream, game, 0x626f6f73u};
    philox4x32_10(r, (uint32_t)seed, (uint32_t)(seed >> 32));
    return out * __powf(u01(r[0]), 1.0f / alpha);
}

// ------------------------------------------------------------------------------------------ root prep
// main.py:259-288.  One warp per game.  prior row 0 of the tree ([g][0][A], otherwise unused) is scratch.
__global__ void __launch_bounds__(TREE_WARPS * 32)
k_tree_reset_roots(mzgo_tree_t t, mzgo_search_params_t p, const float* __restrict__ root_soft,
                   const uint32_t* __restrict__ boards, int n, const double* __restrict__ noise,
                   uint32_t move_index) {
    const int lane = lane_id();
    const int g = blockIdx.x * TREE_WARPS + (threadIdx.x >> 5);
    if (g >= t.G) return;
    if (lane == 0) {
        t.leaf_parent[g] = -1;
        t.leaf_action[g] = -1;
        if (g == 0) *t.move_index_dev = (int32_t)move_index;
    }
    if (!t.active[g]) return;
    const int A = t.A, np = n * n;
    const uint32_t* rec = boards + (size_t)g * MZGO_BOARD_WORDS;
    // valid mask, main.py:261-267
    int nvalid = 0;
    for (int a = lane; a < np; a += 32) {
        int r = a / n, c = a - r * n;
        nvalid += !((rec[2 * n + r] >> c) & 1u);
    }
    nvalid = warp_sum_i(nvalid);
    const double pass_prior = nvalid > 0 ? p.pass_epsilon : 1.0;
    float* scratch = t.prior + (size_t)g * t.M * A;
    float* mask = t.root_mask + (size_t)g * A;
    for (int a = lane; a < A; a += 32) {
        double m;
        if (a < np) {
            int r = a / n, c = a - r * n;
            m = ((rec[2 * n + r] >> c) & 1u) ? 0.0 : 1.0;
        } else {
            m = pass_prior;
        }
        mask[a] = (float)m;
        // policy *= valid_mask : float32 * float64 computed in float64, stored as float32 (main.py:270)
        scratch[a] = (float)__dmul_rn((double)root_soft[(size_t)g * A + a], m);
    }
    __syncwarp();
    float total = 0.f;
    if (lane == 0) total = np_pairwise_sum_f32(scratch, A);   // policy.sum(), main.py:271
    total = __shfl_sync(FULL_MASK, total, 0);
    const float one_minus_eps = (float)(1.0 - p.dirichlet_epsilon);
    // device Dirichlet: gamma draws normalised by their sum
    float gsum = 0.f;
    float gam[MAX_EPL];
    if (!noise) {
#pragma unroll
        for (int j = 0; j < MAX_EPL; j++) {
            int a = j * 32 + lane;
            gam[j] = 0.f;
            if (a < A) gam[j] = sample_gamma(p.dirichlet_alpha, p.seed, (uint32_t)(g + p.game_base), 0x10000u + move_index, (uint32_t)a * 80u);
            gsum += gam[j];
        }
        gsum = warp_sum_f(gsum);
    }
    double* rp = t.root_prior + (size_t)g * A;
    // self_play.py:163: a masked policy that sums to zero becomes uniform over the mask, in float64
    double msum = 0.0;
    if (p.variant == MZGO_VARIANT_SELFPLAY) {
        // sum of the float64 mask, numpy pairwise order (edge_vsum row 0 is free until the root is cleared below)
        double* dscr = t.edge_vsum + (size_t)g * t.M * A;
        for (int a = lane; a < A; a += 32) dscr[a] = a < np ? (double)mask[a] : pass_prior;
        __syncwarp();
        if (lane == 0) msum = np_pairwise_sum<double>(dscr, A);
        msum = __shfl_sync(FULL_MASK, msum, 0);
        __syncwarp();
    }
#pragma unroll
    for (int j = 0; j < MAX_EPL; j++) {
        int a = j * 32 + lane;
        if (a >= A) continue;
        float pol = scratch[a];
        const double m = a < np ? (double)mask[a] : pass_prior;
        double nz = noise ? noise[(size_t)g * A + a] : (double)(gam[j] / gsum);
        double scaled;
        if (total > 0.f) {
            pol = __fdiv_rn(pol, total);                                   // main.py:272-273
            // (1 - eps) * policy stays float32; + eps * dirichlet promotes to float64 (main.py:63-64)
            scaled = (double)__fmul_rn(one_minus_eps, pol);
        } else if (p.variant == MZGO_VARIANT_SELFPLAY) {
            scaled = __dmul_rn(1.0 - p.dirichlet_epsilon, __ddiv_rn(m, msum));   // float64 uniform over the mask
        } else {
            scaled = (double)__fmul_rn(one_minus_eps, pol);
        }
        double mixed = __dadd_rn(scaled, __dmul_rn(p.dirichlet_epsilon, nz));
        if (p.variant == MZGO_VARIANT_SELFPLAY) mixed = __dmul_rn(mixed, m);   // policy *= valid_mask, self_play.py:172
        rp[a] = mask[a] > 0.f ? mixed : 0.0;                               // main.py:281
    }
    if (p.variant == MZGO_VARIANT_SELFPLAY) {
        // renormalise after the noise, float64 pairwise sum (self_play.py:173-178)
        __syncwarp();
        double tot2 = 0.0;
        if (lane == 0) tot2 = np_pairwise_sum<double>(rp, A);
        tot2 = __shfl_sync(FULL_MASK, tot2, 0);
        for (int a = lane; a < A; a += 32) {
            const double m = a < np ? (double)mask[a] : pass_prior;
            rp[a] = tot2 > 0.0 ? __ddiv_rn(rp[a], tot2) : __ddiv_rn(m, msum);
        }
        __syncwarp();
    }
    // fresh root: no children, no visits
    size_t base = (size_t)g * t.M * A;
    for (int a = lane; a < A; a += 32) {
        t.edge_visit[base + a] = 0;
        t.edge_vsum[base + a] = 0.0;
        t.child[base + a] = (int16_t)-1;
    }
    if (lane == 0) {
        t.parent[(size_t)g * t.M] = -1;
        t.root_visit[g] = 0;
        t.root_vsum[g] = 0.0;
        t.n_nodes[g] = 1;
        t.choice_ctr[g] = 0u;
        t.depth_sum[g] = 0;
        t.sims_done[g] = 0;
        t.dropped[g] = 0;
    }
}

// ------------------------------------------------------------------------------------------ select
// main.py:318-363.  Descends from the root until a node with an unexpanded edge (prior > 0, no child).
__global__ void __launch_bounds__(TREE_WARPS * 32)
k_tree_select(mzgo_tree_t t, mzgo_search_params_t p, const uint32_t* __restrict__ choice_stream, int stride) {
    const int lane = lane_id();
    const int g = blockIdx.x * TREE_WARPS + (threadIdx.x >> 5);
    if (g >= t.G) return;
    if (!t.active[g]) {
        if (lane == 0) { t.leaf_parent[g] = -1; t.leaf_action[g] = -1; }
        return;
    }
    const int A = t.A, M = t.M;
    int node = 0, depth = 1;
    int npar = t.root_visit[g];
    int sel_parent = -1, sel_action = -1;
    while (true) {
        const size_t base = ((size_t)g * M + node) * A;
        double cp[MAX_EPL];   // c_puct * prior, float64 (root) or float32 widened (below the root)
        int nv[MAX_EPL];
        uint32_t live = 0u;   // bit j: my edge j has prior > 0
        uint32_t open_b[MAX_EPL];
        int n_open = 0;
#pragma unroll
        for (int j = 0; j < MAX_EPL; j++) {
            int a = j * 32 + lane;
            bool pos = false, opn = false;
            cp[j] = 0.0; nv[j] = 0;
            if (a < A) {
                nv[j] = t.edge_visit[base + a];
                if (node == 0) {
                    double pr = t.root_prior[(size_t)g * A + a];
                    pos = pr > 0.0;
                    cp[j] = __dmul_rn(p.c_puct, pr);
                } else {
                    float pr = t.prior[base + a];
                    pos = pr > 0.f;
                    cp[j] = (double)__fmul_rn((float)p.c_puct, pr);   // c*P rounds in float32 (NEP 50)
                }
                opn = pos && nv[j] == 0;
            }
            if (pos) live |= 1u << j;
            open_b[j] = __ballot_sync(FULL_MASK, opn);
            n_open += __popc(open_b[j]);
        }
        if (n_open > 0) {
            // random.choice(unexpanded), main.py:322-325
            uint32_t ctr = t.choice_ctr[g];
            uint32_t r = choice_stream ? choice_stream[(size_t)g * stride + ctr]
                                       : mzgo_rand_u32(p.seed, (uint32_t)(g + p.game_base), 0x20000u + (uint32_t)t.move_index, ctr);
            int k = (int)(r % (uint32_t)n_open);
            int a_sel = -1;
#pragma unroll
            for (int j = 0; j < MAX_EPL; j++) {
                int c = __popc(open_b[j]);
                if (a_sel < 0) {
                    if (k < c) a_sel = j * 32 + (int)__fns(open_b[j], 0, k + 1);
                    else k -= c;
                }
            }
            sel_parent = node;
            sel_action = a_sel;
            if (lane == 0) t.choice_ctr[g] = ctr + 1u;
            break;
        }
        if (!__any_sync(FULL_MASK, live != 0u)) break;   // nothing selectable: simulation dropped
        // Q of every live edge, min-max normalisation set = prior > 0 (main.py:328-339)
        double q[MAX_EPL];
        double qmin = INFINITY, qmax = -INFINITY;
#pragma unroll
        for (int j = 0; j < MAX_EPL; j++) {
            q[j] = 0.0;
            if (live & (1u << j)) {
                int a = j * 32 + lane;
                if (nv[j] > 0) q[j] = __ddiv_rn(t.edge_vsum[base + a], (double)nv[j]);
                qmin = fmin(qmin, q[j]);
                qmax = fmax(qmax, q[j]);
            }
        }
        qmin = warp_min_d(qmin);
        qmax = warp_max_d(qmax);
        // main.py:354 sqrt(N+1); self_play.py:297-300 sqrt(max(1, N))
        const double sq = __dsqrt_rn((double)(p.variant == MZGO_VARIANT_SELFPLAY ? max(1, npar) : npar + 1));
        const bool norm = qmax > qmin;
        const double range = __dsub_rn(qmax, qmin);
        double best = -INFINITY;
        int best_a = -1, best_n = 0;
#pragma unroll
        for (int j = 0; j < MAX_EPL; j++) {
            if (live & (1u << j)) {
                double qn = norm ? __ddiv_rn(__dsub_rn(q[j], qmin), range) : q[j];
                double u = __ddiv_rn(__dmul_rn(cp[j], sq), (double)(1 + nv[j]));
                double score = __dadd_rn(qn, u);
                if (score > best) { best = score; best_a = j * 32 + lane; best_n = nv[j]; }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            double ob = __shfl_xor_sync(FULL_MASK, best, o);
            int oa = __shfl_xor_sync(FULL_MASK, best_a, o);
            int on = __shfl_xor_sync(FULL_MASK, best_n, o);
            bool take = (oa >= 0) && (best_a < 0 || ob > best || (ob == best && oa < best_a));
            if (take) { best = ob; best_a = oa; best_n = on; }
        }
        if (best_a < 0) break;   // only NaN / -inf scores: the reference falls out of the loop too
        MZGO_DCHECK(best_a < A);
        node = (int)t.child[base + best_a];
        MZGO_DCHECK(node < t.n_nodes[g]);
        npar = best_n;
        depth++;
    }
    if (lane == 0) {
        int nn = t.n_nodes[g];
        if (sel_action < 0 && p.variant == MZGO_VARIANT_SELFPLAY) {
            // terminal leaf: backpropagate(path, 0) -- a visit for every node on the path (self_play.py:188-190)
            int nd = node;
            while (nd != 0) {
                MZGO_DCHECK(nd > 0 && nd < M);
                const int pk = t.parent[(size_t)g * M + nd];
                MZGO_DCHECK(pk >= 0 && (pk >> 9) < M && (pk & 511) < A);
                t.edge_visit[((size_t)g * M + (pk >> 9)) * A + (pk & 511)] += 1;
                nd = pk >> 9;
            }
            t.root_visit[g] += 1;
        }
        if (sel_action >= 0 && nn >= M) sel_action = -1;   // arena full: drop (cannot happen when M > simulations)
        t.leaf_parent[g] = sel_action >= 0 ? sel_parent : -1;
        t.leaf_action[g] = sel_action;
        t.leaf_node[g] = nn;
        t.leaf_depth[g] = depth;
        t.leaf_in_slot[g] = sel_action >= 0 ? g * M + sel_parent : -1;
        t.leaf_out_slot[g] = sel_action >= 0 ? g * M + nn : -1;
        if (sel_action < 0) t.dropped[g] += 1;
    }
}

// ------------------------------------------------------------------------------------------ expand + backup
// main.py:300-313 (expansion) and :365-368 (backpropagate).
__global__ void __launch_bounds__(TREE_WARPS * 32)
k_tree_expand_backup(mzgo_tree_t t, mzgo_search_params_t p, const float* __restrict__ soft,
                     const float* __restrict__ reward, const float* __restrict__ value) {
    const int lane = lane_id();
    const int g = blockIdx.x * TREE_WARPS + (threadIdx.x >> 5);
    if (g >= t.G) return;
    const int act = t.leaf_action[g];
    if (act < 0) return;
    const int A = t.A, M = t.M;
    const int par = t.leaf_parent[g];
    const int nw = t.leaf_node[g];
    MZGO_DCHECK(act < A && par >= 0 && par < nw && nw < M && nw == t.n_nodes[g]);
    const size_t nbase = ((size_t)g * M + nw) * A;
    const float* mask = t.root_mask + (size_t)g * A;
    const bool v1 = p.variant == MZGO_VARIANT_SELFPLAY;
    for (int a = lane; a < A; a += 32) {
        // prior_child = policy_child[a] if valid_mask[a] > 0 else 0   (the ROOT's mask, main.py:303)
        float pr = mask[a] > 0.f ? soft[(size_t)g * A + a] : 0.f;
        if (v1) pr = masked_child_prior(soft[(size_t)g * A + a], mask[a], a == A - 1, p.pass_epsilon);
        t.prior[nbase + a] = pr;
        t.edge_visit[nbase + a] = 0;
        t.edge_vsum[nbase + a] = 0.0;
        t.child[nbase + a] = (int16_t)-1;
    }
    if (v1) {
        // policy_child /= policy_child.sum(): float32 pairwise sum, float32 divide (self_play.py:211-214)
        __syncwarp();
        float tot = 0.f;
        if (lane == 0) tot = np_pairwise_sum<float>(t.prior + nbase, A);
        tot = __shfl_sync(FULL_MASK, tot, 0);
        __syncwarp();
        for (int a = lane; a < A; a += 32)
            t.prior[nbase + a] = renorm_child_prior(t.prior[nbase + a], tot, mask[a], a == A - 1, p.pass_epsilon, mask, A);
    }
    if (lane == 0) {
        t.child[((size_t)g * M + par) * A + act] = (int16_t)nw;
        t.parent[(size_t)g * M + nw] = par * 512 + act;
        t.n_nodes[g] = nw + 1;
        // backup_value = reward.item() + discount * value.item()   (main.py:311), float64
        const double bv = __dadd_rn((double)reward[g], __dmul_rn(p.discount, (double)value[g]));
        double sv = bv;                                      // self_play.py:340-343: +value at the new leaf, then alternating
        int node = nw;
        while (node != 0) {
            MZGO_DCHECK(node > 0 && node < M);
            int pk = t.parent[(size_t)g * M + node];
            int pn = pk >> 9, pa = pk & 511;
            MZGO_DCHECK(pk >= 0 && pn < M && pa < A);
            size_t e = ((size_t)g * M + pn) * A + pa;
            t.edge_visit[e] += 1;
            t.edge_vsum[e] = __dadd_rn(t.edge_vsum[e], sv);
            if (v1) sv = -sv;
            node = pn;
        }
        t.root_visit[g] += 1;
        t.root_vsum[g] = __dadd_rn(t.root_vsum[g], sv);
        t.sims_done[g] += 1;
        t.depth_sum[g] += t.leaf_depth[g];
    }
}

// ------------------------------------------------------------------------------------------ fused step
// One launch per simulation: expand + backup of the leaf evaluated by the previous network call
// (main.py:300-313, :365-368), then the selection of the next leaf (main.py:318-363).  TPG threads
// work on one game (128 when A > 96, else one warp), so a PUCT level is one coalesced pass over the
// edge arrays with up to three edges per thread, and the backup updates all edges of the selection
// path in parallel: the path is written by the selection into the scratch row ([g][0][A] of `prior`).
// Arithmetic is the same, operation for operation, as in k_tree_select / k_tree_expand_backup.
//
// Memory round trips (the kernel is latency bound: a few KB per game, dependent levels):
//   1. all TREE data that does not depend on another load is fetched at the top in one batch: the leaf record, the
//      mask, the stored path, the per-game counters AND the root level of the next selection (visits, priors, value
//      sums, child links); the root statistics this very step changes (one edge of the backup path, the child link
//      when the leaf hangs off the root) are patched in registers.  Behind the engine's own network kernel this batch
//      and the next are issued BEFORE the grid-dependency wait (the blocks are resident while that kernel drains);
//   2. the statistics of the path edges to back up (needed by the backup's stores only); in front of the wait too:
//      the step's random number, the edge terms of the root level (all edges but the one the backup changes) and
//      every store that needs no network output (the new node's empty statistics, visit counts, links, counters);
//   3. behind the wait: the network outputs of the leaf (value, reward, softmax);
//   4. one batch per level below the root (visits, priors, value sums and child links together).
// (the step itself lives in tree_step.cuh: the network kernel of nets_tc.cu runs the same code as a warp-group role)

// EARLY: behind the engine's own network kernel (see round trip 1 above; also the edge terms of every level are
// computed as soon as its statistics arrive, which measures faster at 19x19, while the lazy form -- only on levels
// that do not end in a draw -- is the faster one in the tree phase of the 9x9 network kernel)
template <int TPG, bool EARLY, bool V1>
__global__ void __launch_bounds__(STEP_THREADS, 7)
k_tree_step(mzgo_tree_t t, mzgo_search_params_t p, const float* soft, const float* reward, const float* value,
            const uint32_t* __restrict__ choice_stream, int stride, int do_backup, int do_select) {
    constexpr int GPB = STEP_THREADS / TPG;     // games per block
    __shared__ StepShared<TPG> sh;
    const int sub = threadIdx.x / TPG, tid = threadIdx.x % TPG;
    const int g = blockIdx.x * GPB + sub;
#ifdef MZGO_TRACE_TREE
    if (g < t.G) TSTAMP(0, 0);
#endif
    griddep_launch();
    if (g >= t.G) { griddep_wait(); return; }    // TPG == 128: the whole block; TPG == 32: whole warps
    // the grid-dependency wait (PDL, common.cuh) sits INSIDE the step, in front of the first read of what the kernel
    // before this one wrote (the network outputs): the tree data is fetched while that kernel drains
    tree_step_body<TPG, 0, EARLY, V1>(t, p, soft, reward, value, choice_stream, stride, do_backup, do_select, sh, sub, tid, g,
                                  [] { griddep_wait(); });
}

// main.py:666-669
__global__ void k_tree_visit_counts(mzgo_tree_t t, int32_t* out) {
    const long total = (long)t.G * t.A;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        int g = (int)(i / t.A), a = (int)(i - (long)g * t.A);
        int v = 0;
        if (t.active[g] && t.root_mask[i] > 0.f) v = t.edge_visit[(size_t)g * t.M * t.A + a];
        out[i] = v;
    }
}

// main.py:670-677 (greedy) / self_play.py:389-400 (temperature sampling)
__global__ void __launch_bounds__(TREE_WARPS * 32)
k_tree_choose_action(mzgo_tree_t t, mzgo_search_params_t p, float temperature, const float* temperature_per_game,
                     uint32_t move_index, int32_t* action, float* policy, float* root_value) {
    const int lane = lane_id();
    const int g = blockIdx.x * TREE_WARPS + (threadIdx.x >> 5);
    if (g >= t.G) return;
    const int A = t.A;
    if (!t.active[g]) {
        if (lane == 0) { action[g] = -1; if (root_value) root_value[g] = 0.f; }
        if (policy) for (int a = lane; a < A; a += 32) policy[(size_t)g * A + a] = 0.f;
        return;
    }
    if (temperature_per_game) temperature = temperature_per_game[g];
    const size_t base = (size_t)g * t.M * A;
    const float* mask = t.root_mask + (size_t)g * A;
    int vc[MAX_EPL];
    int sum = 0, best = -1, best_a = -1, nvalid = 0;
#pragma unroll
    for (int j = 0; j < MAX_EPL; j++) {
        int a = j * 32 + lane;
        vc[j] = 0;
        if (a < A) {
            bool ok = mask[a] > 0.f;
            nvalid += ok;
            if (ok) vc[j] = t.edge_visit[base + a];
            sum += vc[j];
            if (vc[j] > best) { best = vc[j]; best_a = a; }
        }
    }
    sum = warp_sum_i(sum);
    nvalid = warp_sum_i(nvalid);
    int chosen = -1;
    if (sum > 0) {
        if (temperature > 0.f) {
            // sample proportionally to visits^(1/T)  (self_play.py:391-400)
            float w[MAX_EPL], wsum = 0.f;
#pragma unroll
            for (int j = 0; j < MAX_EPL; j++) {
                w[j] = vc[j] > 0 ? __powf((float)vc[j], 1.0f / temperature) : 0.f;
                wsum += w[j];
            }
            float tot = warp_sum_f(wsum);
            float incl = wsum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                float v = __shfl_up_sync(FULL_MASK, incl, o);
                if (lane >= o) incl += v;
            }
            float target = u01(mzgo_rand_u32(p.seed, (uint32_t)(g + p.game_base), 0x30000u + move_index, 0u)) * tot;
            uint32_t hit = __ballot_sync(FULL_MASK, wsum > 0.f && target < incl);
            int hl = hit ? __ffs(hit) - 1 : 31 - __clz(__ballot_sync(FULL_MASK, wsum > 0.f));
            int mine = -1;
            float acc = incl - wsum;
#pragma unroll
            for (int j = 0; j < MAX_EPL; j++) {
                if (w[j] > 0.f) {
                    acc += w[j];
                    if (mine < 0 || target >= acc - w[j]) mine = j * 32 + lane;
                }
            }
            chosen = __shfl_sync(FULL_MASK, mine, hl);
        } else {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                int ob = __shfl_xor_sync(FULL_MASK, best, o);
                int oa = __shfl_xor_sync(FULL_MASK, best_a, o);
                if (oa >= 0 && (best_a < 0 || ob > best || (ob == best && oa < best_a))) { best = ob; best_a = oa; }
            }
            chosen = best_a;   // np.argmax: first maximum
        }
        if (policy) {
#pragma unroll
            for (int j = 0; j < MAX_EPL; j++) {
                int a = j * 32 + lane;
                if (a < A) policy[(size_t)g * A + a] = (float)((double)vc[j] / (double)sum);
            }
        }
    } else {
        // uniform over np.nonzero(valid_mask) (main.py:674-677)
        int k = (int)(mzgo_rand_u32(p.seed, (uint32_t)(g + p.game_base), 0x40000u + move_index, 0u) % (uint32_t)max(nvalid, 1));
#pragma unroll
        for (int j = 0; j < MAX_EPL; j++) {
            int a = j * 32 + lane;
            uint32_t b = __ballot_sync(FULL_MASK, a < A && mask[a < A ? a : 0] > 0.f);
            int c = __popc(b);
            if (chosen < 0) {
                if (k < c) chosen = j * 32 + (int)__fns(b, 0, k + 1);
                else k -= c;
            }
        }
        if (policy) {
#pragma unroll
            for (int j = 0; j < MAX_EPL; j++) {
                int a = j * 32 + lane;
                if (a < A) policy[(size_t)g * A + a] = a == chosen ? 1.f : 0.f;
            }
        }
    }
    if (lane == 0) {
        action[g] = chosen;
        if (root_value) {
            int rv = t.root_visit[g];
            root_value[g] = rv > 0 ? (float)(t.root_vsum[g] / (double)rv) : 0.f;
        }
    }
}

}  // namespace

extern void mzgo_count_launch();

// compares div_rn_with_rcp / div_by_count with the generic IEEE division on pseudo-random operands
__global__ void k_fastdiv_check(unsigned long long seed, long long per_thread, unsigned long long* mismatches) {
    const unsigned long long tid = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
    unsigned long long bad = 0;
    for (long long i = 0; i < per_thread; i++) {
        uint32_t c[4] = {(uint32_t)i, (uint32_t)(i >> 32), (uint32_t)tid, (uint32_t)(tid >> 32)};
        philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
        // numerator: random mantissa, exponent in [-200, 19] (about 1e-60 .. 1e6), random sign
        const unsigned long long mant = (((unsigned long long)c[0] << 32) | c[1]) & 0x000FFFFFFFFFFFFFull;
        const int ea = (int)(c[2] % 220u) - 200;
        double a = __longlong_as_double((long long)(((unsigned long long)(1023 + ea) << 52) | mant));
        if (c[3] & 1u) a = -a;
        // (1) small integer divisors
        const int d = 1 + (int)((c[3] >> 1) % (uint32_t)(RCP_TABLE - 1));
        if (__double_as_longlong(div_by_count(a, d)) != __double_as_longlong(__ddiv_rn(a, (double)d))) bad++;
        // (2) arbitrary divisors in [1e-17, 1e3) with a correctly rounded reciprocal
        uint32_t c2[4] = {(uint32_t)i, 0x9e3779b9u, (uint32_t)tid, 0x7f4a7c15u};
        philox4x32_10(c2, (uint32_t)seed, (uint32_t)(seed >> 32));
        const unsigned long long mb = (((unsigned long long)c2[0] << 32) | c2[1]) & 0x000FFFFFFFFFFFFFull;
        const int eb = (int)(c2[2] % 66u) - 56;                    // 2^-56 .. 2^9
        const double b = __longlong_as_double((long long)(((unsigned long long)(1023 + eb) << 52) | mb));
        if (__double_as_longlong(div_rn_with_rcp(a, b, __drcp_rn(b))) != __double_as_longlong(__ddiv_rn(a, b))) bad++;
    }
    if (bad) atomicAdd(mismatches, bad);
}

// Test hook: number of operand pairs (of 2 * blocks * 256 * per_thread) on which the fast division differs
// from __ddiv_rn.  Synchronous.
extern "C" long long mzgo_debug_fastdiv_check(unsigned long long seed, int blocks, long long per_thread) {
    if (ensure_rcp_table()) return -1;
    unsigned long long* dev_bad = nullptr;
    if (cudaMalloc(&dev_bad, 8) != cudaSuccess) return -1;
    cudaMemset(dev_bad, 0, 8);
    k_fastdiv_check<<<blocks, 256>>>(seed, per_thread, dev_bad);
    unsigned long long bad = 0;
    const cudaError_t e = cudaMemcpy(&bad, dev_bad, 8, cudaMemcpyDeviceToHost);
    cudaFree(dev_bad);
    return e == cudaSuccess ? (long long)bad : -1;
}

static int check_tree(const mzgo_tree_t* t) {
    MZGO_CHECK_ARG(t != nullptr, "tree is NULL");
    MZGO_CHECK_ARG(t->G >= 0, "tree.G < 0");
    MZGO_CHECK_ARG(t->A >= 2 && t->A <= 32 * 12, "tree.A=%d outside [2,384]", t->A);
    MZGO_CHECK_ARG(t->M >= 2 && t->M <= 32767, "tree.M=%d outside [2,32767]", t->M);
    MZGO_CHECK_ARG(t->root_prior && t->root_mask && t->prior && t->edge_visit && t->edge_vsum && t->child &&
                       t->parent && t->root_visit && t->root_vsum && t->n_nodes && t->leaf_parent &&
                       t->leaf_action && t->leaf_node && t->leaf_depth && t->leaf_in_slot && t->leaf_out_slot && t->choice_ctr && t->depth_sum &&
                       t->sims_done && t->dropped && t->active && t->move_index_dev,
                   "tree has a NULL array");
    return 0;
}

extern "C" int mzgo_tree_reset_roots(const mzgo_tree_t* tree, const mzgo_search_params_t* p, const float* root_soft,
                                     const uint32_t* boards, int n, const double* noise, uint32_t move_index,
                                     void* stream) {
    if (int e = check_tree(tree)) return e;
    MZGO_CHECK_ARG(p && root_soft && boards, "NULL argument");
    MZGO_CHECK_ARG(n * n + 1 == tree->A, "board size %d does not match tree.A=%d", n, tree->A);
    if (tree->G == 0) return MZGO_OK;
    k_tree_reset_roots<<<ceil_div(tree->G, TREE_WARPS), TREE_WARPS * 32, 0, (cudaStream_t)stream>>>(
        *tree, *p, root_soft, boards, n, noise, move_index);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

extern "C" int mzgo_tree_select(const mzgo_tree_t* tree, const mzgo_search_params_t* p, const uint32_t* choice_stream,
                                int stride, void* stream) {
    if (int e = check_tree(tree)) return e;
    MZGO_CHECK_ARG(p != nullptr, "params is NULL");
    if (tree->G == 0) return MZGO_OK;
    k_tree_select<<<ceil_div(tree->G, TREE_WARPS), TREE_WARPS * 32, 0, (cudaStream_t)stream>>>(*tree, *p, choice_stream,
                                                                                               stride);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

extern "C" int mzgo_tree_expand_backup(const mzgo_tree_t* tree, const mzgo_search_params_t* p, const float* soft,
                                       const float* reward, const float* value, void* stream) {
    if (int e = check_tree(tree)) return e;
    MZGO_CHECK_ARG(p && soft && reward && value, "NULL argument");
    if (tree->G == 0) return MZGO_OK;
    k_tree_expand_backup<<<ceil_div(tree->G, TREE_WARPS), TREE_WARPS * 32, 0, (cudaStream_t)stream>>>(*tree, *p, soft,
                                                                                                      reward, value);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

// Fused simulation step (engine path): expand + backup of the previous leaf, then the next selection.
// after_network != 0: the caller enqueued the network kernel of the same search directly in front of this step (and the
// tree step before that), so the step may fetch its tree data before the grid-dependency wait (tree_step.cuh).
int mzgo_tree_step_impl(const mzgo_tree_t* tree, const mzgo_search_params_t* p, const float* soft, const float* reward,
                        const float* value, const uint32_t* choice_stream, int stride, int do_backup, int do_select,
                        int after_network, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    const int early = (after_network && do_backup) ? 1 : 0;
    if (int e = check_tree(tree)) return e;
    if (int e = ensure_rcp_table()) return e;
    MZGO_CHECK_ARG(p && (!do_backup || (soft && reward && value)), "NULL argument");
    if (tree->G == 0) return MZGO_OK;
    const dim3 block(STEP_THREADS);
    const bool small = tree->A <= STEP_EPT * 32, v1 = p->variant == MZGO_VARIANT_SELFPLAY;
    const dim3 grid(small ? ceil_div(tree->G, STEP_THREADS / 32) : tree->G);
#define MZGO_STEP_LAUNCH(TPG, E, V)                                                                                          \
    MZGO_CUDA(mzgo_launch_pdl(k_tree_step<TPG, E, V>, grid, block, 0, st, *tree, *p, soft, reward, value, choice_stream, stride, \
                              do_backup, do_select))
    if (small) {
        if (early) { if (v1) MZGO_STEP_LAUNCH(32, true, true); else MZGO_STEP_LAUNCH(32, true, false); }
        else { if (v1) MZGO_STEP_LAUNCH(32, false, true); else MZGO_STEP_LAUNCH(32, false, false); }
    } else {
        if (early) { if (v1) MZGO_STEP_LAUNCH(128, true, true); else MZGO_STEP_LAUNCH(128, true, false); }
        else { if (v1) MZGO_STEP_LAUNCH(128, false, true); else MZGO_STEP_LAUNCH(128, false, false); }
    }
#undef MZGO_STEP_LAUNCH
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

extern "C" int mzgo_tree_step(const mzgo_tree_t* tree, const mzgo_search_params_t* p, const float* soft,
                              const float* reward, const float* value, const uint32_t* choice_stream, int stride,
                              int do_backup, int do_select, void* stream) {
    return mzgo_tree_step_impl(tree, p, soft, reward, value, choice_stream, stride, do_backup, do_select, 0, stream);
}

// Trace build only: arm / disarm the per-game stamps of the tree step (buf = device [G][16] uint64, or NULL).
extern "C" int mzgo_debug_tree_trace(void* buf) {
#ifdef MZGO_TRACE_TREE
    unsigned long long* p = (unsigned long long*)buf;
    MZGO_CUDA(cudaMemcpyToSymbol(g_tree_trace, &p, sizeof(p)));
    return MZGO_OK;
#else
    (void)buf;
    mzgo_set_error("mzgo_debug_tree_trace needs the trace build (python -m muzero_go_b200.build --define MZGO_TRACE_TREE)");
    return MZGO_ERR_UNSUPPORTED;
#endif
}

extern "C" int mzgo_tree_visit_counts(const mzgo_tree_t* tree, int32_t* out, void* stream) {
    if (int e = check_tree(tree)) return e;
    MZGO_CHECK_ARG(out != nullptr, "out is NULL");
    if (tree->G == 0) return MZGO_OK;
    long total = (long)tree->G * tree->A;
    int blocks = (int)((total + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_tree_visit_counts<<<blocks, 256, 0, (cudaStream_t)stream>>>(*tree, out);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

extern "C" int mzgo_tree_choose_action(const mzgo_tree_t* tree, const mzgo_search_params_t* p, float temperature,
                                       const float* temperature_per_game, uint32_t move_index, int32_t* action, float* policy, float* root_value,
                                       void* stream) {
    if (int e = check_tree(tree)) return e;
    MZGO_CHECK_ARG(p && action, "NULL argument");
    if (tree->G == 0) return MZGO_OK;
    k_tree_choose_action<<<ceil_div(tree->G, TREE_WARPS), TREE_WARPS * 32, 0, (cudaStream_t)stream>>>(
        *tree, *p, temperature, temperature_per_game, move_index, action, policy, root_value);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}
