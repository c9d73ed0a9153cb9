// Batched self-play driver: G games in lock step, launches only (no host synchronisation).
//
// Replaces the per-move body of /root/reference/main.py:656-702 (MCTS(...).run(obs), visit-count action
// choice, env.step, trajectory append, two-pass / move-cap termination) and the game epilogue
// main.py:703-711.  The host loop below only enqueues kernels; per simulation it enqueues
// select -> recurrent inference -> expand/backup for the whole batch of games.
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <vector>

#include "nets.cuh"

namespace {

// ---- CUDA graph of the search loop.  One move enqueues 2 * simulations + 1 dependent launches; at 9x9 the
// gaps between them cost 8 % of the step (tools/graph_probe.py), so mzgo_selfplay_run replays a captured
// graph of mzgo_search_run instead.  Everything the loop reads per move comes from device memory (the move
// index through tree.move_index_dev), so one graph serves every move of an engine.  Key: the descriptor bytes
// (with the by-value move index cleared) and the weight generation (head parameters are kernel-argument
// constants, nets_tc.cu).  The first move of a new key runs un-captured (it may set function attributes).
struct SearchGraph {
    mzgo_selfplay_t key;
    unsigned long long generation;
    int device;
    int uses;
    int launches;
    cudaGraphExec_t exec;
};
std::mutex g_graph_mu;
std::vector<SearchGraph> g_graphs;
long long g_graph_replays = 0;
constexpr size_t MAX_GRAPHS = 8;

constexpr int SP_THREADS = 128;

// Start of a move: retire / restart finished games, set the per-game sampling temperature.
__global__ void k_sp_begin_move(mzgo_selfplay_t sp) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= sp.tree.G) return;
    uint32_t* rec = sp.boards + (size_t)g * MZGO_BOARD_WORDS;
    bool over = ((rec[61] >> 1) & 1u) || sp.pass_streak[g] >= 2 || sp.move_count[g] >= sp.max_game_moves;
    if (over && sp.restart_finished) {
        for (int w = 0; w < MZGO_BOARD_WORDS; w++) rec[w] = 0u;   // GoEnv.reset, main.py:637
        sp.move_count[g] = 0;
        sp.pass_streak[g] = 0;
        over = false;
    }
    sp.tree.active[g] = over ? 0 : 1;
    sp.out_slot[g] = over ? -1 : g * sp.tree.M;   // the root latent lives in slot (g, node 0)
    sp.game_temperature[g] = (sp.move_count[g] < sp.temperature_moves) ? sp.temperature : 0.f;
}

// After env.step: bookkeeping of main.py:690-711 and the trajectory row of this move.
__global__ void k_sp_end_move(mzgo_selfplay_t sp, int t) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    const int G = sp.tree.G;
    if (g >= G) return;
    const size_t row = (size_t)t * G + g;
    if (!sp.tree.active[g]) {
        sp.traj_action[row] = -1;
        sp.traj_reward[row] = 0.f;
        sp.traj_done[row] = 0;
        sp.traj_winner[row] = 0;
        return;
    }
    const int a = sp.action[g];
    const uint32_t* rec = sp.boards + (size_t)g * MZGO_BOARD_WORDS;
    const int pass_action = sp.n * sp.n;
    const bool env_done = (rec[61] >> 1) & 1u;
    int mc = sp.move_count[g] + 1;
    int ps = (a == pass_action) ? sp.pass_streak[g] + 1 : 0;
    sp.move_count[g] = mc;
    sp.pass_streak[g] = ps;
    const bool done = env_done || ps >= 2;
    const bool capped = !done && mc >= sp.max_game_moves;
    // GymGo reward_method='real': winner once the game has ended, else 0; env.winner() likewise (main.py:709)
    const int winner = env_done ? sp.score_winner[g] : 0;
    float reward = env_done ? (float)winner : 0.f;
    if (capped) reward += -0.5f;                                    // main.py:703-706
    sp.traj_action[row] = a;
    sp.traj_reward[row] = reward;
    sp.traj_done[row] = (done || capped) ? 1 : 0;
    sp.traj_winner[row] = (int8_t)winner;
    unsigned long long* c = (unsigned long long*)sp.counters;
    atomicAdd(c + 0, 1ull);
    atomicAdd(c + 1, (unsigned long long)sp.tree.sims_done[g]);
    atomicAdd(c + 3, (unsigned long long)sp.tree.dropped[g]);
    atomicAdd(c + 4, (unsigned long long)sp.tree.depth_sum[g]);
    if (done || capped) {
        atomicAdd(c + 2, 1ull);
        if (winner > 0) atomicAdd(c + 5, 1ull);
        if (winner < 0) atomicAdd(c + 6, 1ull);
        if (capped) atomicAdd(c + 7, 1ull);
    }
}

// Trajectory rows in the wire layout of the replay exchange (muzero_go_b200/dist.py, ROW_FIELDS): per (move, game)
// a fixed-stride record {board 256 B, action i32, policy A x f32, root value f32, reward f32, done u8, winner i8}.
// One warp per (row, game); out = [nrows][G_total][stride] bytes, this engine's games at [g_offset, g_offset + G).
__global__ void k_sp_pack_rows(mzgo_selfplay_t sp, int t0, int nrows, uint8_t* out, int g_offset, int G_total) {
    const int G = sp.tree.G, A = sp.tree.A;
    const int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (w >= nrows * G) return;
    const int r = w / G, g = w - r * G;
    const size_t row = (size_t)(t0 + r) * G + g;
    const size_t stride = 256 + 4 + 4 * (size_t)A + 4 + 4 + 1 + 1;
    uint8_t* dst = out + ((size_t)r * G_total + g_offset + g) * stride;
    // the stride is even but not a multiple of 4 in general: 16-bit stores keep every field at its byte offset
    uint16_t* d16 = reinterpret_cast<uint16_t*>(dst);
    const uint16_t* board = reinterpret_cast<const uint16_t*>(sp.traj_boards + row * MZGO_BOARD_WORDS);
    for (int i = lane; i < 128; i += 32) d16[i] = board[i];
    const uint16_t* pol = reinterpret_cast<const uint16_t*>(sp.traj_policy + row * A);
    for (int i = lane; i < 2 * A; i += 32) d16[130 + i] = pol[i];
    if (lane == 0) {
        const uint32_t act = (uint32_t)sp.traj_action[row];
        d16[128] = (uint16_t)act; d16[129] = (uint16_t)(act >> 16);
        const uint32_t rv = __float_as_uint(sp.traj_root_value[row]), rw = __float_as_uint(sp.traj_reward[row]);
        uint16_t* tail = d16 + 130 + 2 * A;
        tail[0] = (uint16_t)rv; tail[1] = (uint16_t)(rv >> 16);
        tail[2] = (uint16_t)rw; tail[3] = (uint16_t)(rw >> 16);
        tail[4] = (uint16_t)sp.traj_done[row] | (uint16_t)((uint16_t)(uint8_t)sp.traj_winner[row] << 8);
    }
}

int check_sp(const mzgo_selfplay_t* sp) {
    MZGO_CHECK_ARG(sp != nullptr, "selfplay is NULL");
    MZGO_CHECK_ARG(sp->n * sp->n + 1 == sp->tree.A && sp->net.n == sp->n, "board size mismatch");
    MZGO_CHECK_ARG(sp->num_simulations >= 0 && sp->num_simulations < sp->tree.M, "tree.M=%d must exceed simulations=%d",
                   sp->tree.M, sp->num_simulations);
    MZGO_CHECK_ARG(sp->boards && sp->arena && sp->root_policy && sp->root_value_net && sp->leaf_policy &&
                       sp->leaf_reward && sp->leaf_value && sp->out_slot && sp->action && sp->move_count &&
                       sp->pass_streak && sp->game_temperature && sp->score_winner && sp->counters,
                   "selfplay has a NULL array");
    MZGO_CHECK_ARG(sp->net.max_batch >= sp->tree.G, "net.max_batch=%d < G=%d", sp->net.max_batch, sp->tree.G);
    return 0;
}

}  // namespace

extern "C" int mzgo_search_run(const mzgo_selfplay_t* sp, const uint32_t* choice_stream, int stride, void* stream) {
    if (int e = check_sp(sp)) return e;
    const mzgo_tree_t* t = &sp->tree;
    cudaStream_t st = (cudaStream_t)stream;
    static const bool unfused = getenv("MZGO_NO_FUSED_TREE") != nullptr;      // A/B switch for measurements
    if (sp->net.precision == 1 && !unfused && sp->num_simulations > 0 && t->A <= 96) {
        // bf16 mode, boards up to 9x9 (a warp per game in the tree step): ONE launch per simulation.  The first
        // selection, then S x [recurrent inference with the expansion + backup of its leaf and the next selection
        // behind it in the same kernel].  On larger boards a game takes 128 threads and a CTA's seven games
        // three rounds, which costs more than the launch it saves (19x19: 8.7 vs 9.3 M sims/s): two launches.
        if (int e = mzgo_tree_step(t, &sp->params, sp->leaf_policy, sp->leaf_reward, sp->leaf_value, choice_stream, stride,
                                   0, 1, st))
            return e;
        for (int s = 0; s < sp->num_simulations; s++)
            if (int e = mzgo_net_recurrent_tree(&sp->net, sp->arena, t, &sp->params, choice_stream, stride, sp->leaf_reward,
                                                sp->leaf_value, sp->leaf_policy, s + 1 < sp->num_simulations, stream))
                return e;
        return MZGO_OK;
    }
    // per simulation: [expand/backup of the previous leaf + next selection] -> recurrent inference
    for (int s = 0; s < sp->num_simulations; s++) {
        // (from the second step on the kernel in front of the step is this loop's own network kernel)
        if (int e = mzgo_tree_step_impl(t, &sp->params, sp->leaf_policy, sp->leaf_reward, sp->leaf_value, choice_stream, stride,
                                        s > 0, 1, s > 0, st))
            return e;
        if (int e = mzgo_net_recurrent(&sp->net, sp->arena, t->leaf_in_slot, t->leaf_action, t->leaf_out_slot,
                                       sp->leaf_reward, sp->leaf_value, sp->leaf_policy, 1, t->G, stream))
            return e;
    }
    if (sp->num_simulations > 0)
        if (int e = mzgo_tree_step_impl(t, &sp->params, sp->leaf_policy, sp->leaf_reward, sp->leaf_value, choice_stream, stride,
                                        1, 0, 1, st))
            return e;
    return MZGO_OK;
}

// The search of one move through the graph cache (falls back to plain launches when capture is not possible).
static int search_run_cached(const mzgo_selfplay_t* sp, cudaStream_t st) {
    static const bool disabled = getenv("MZGO_NO_GRAPH") != nullptr;      // A/B switch for measurements
    if (disabled || sp->num_simulations == 0) return mzgo_search_run(sp, nullptr, 0, st);
    mzgo_selfplay_t key;
    memcpy(&key, sp, sizeof(key));
    key.tree.move_index = 0;
    int device = 0;
    MZGO_CUDA(cudaGetDevice(&device));
    const unsigned long long gen = mzgo_net_generation(sp->net.workspace);
    std::lock_guard<std::mutex> lk(g_graph_mu);
    SearchGraph* e = nullptr;
    for (auto& c : g_graphs)
        if (c.device == device && memcmp(&c.key, &key, sizeof(key)) == 0) e = &c;
    if (e && e->generation != gen) {
        // the same engine after a weight update (a training step, a broadcast): the old graph carries the old head
        // constants.  Drop it and capture again on THIS call -- the kernels' attributes are set already, so a train /
        // self-play loop pays one capture per weight version and never runs a move un-captured.
        if (e->exec) cudaGraphExecDestroy(e->exec);
        e->exec = nullptr;
        e->generation = gen;
        if (e->launches >= 0) e->launches = 0;
    }
    if (!e) {
        if (g_graphs.size() >= MAX_GRAPHS) {
            if (g_graphs.front().exec) cudaGraphExecDestroy(g_graphs.front().exec);
            g_graphs.erase(g_graphs.begin());
        }
        SearchGraph c;
        memcpy(&c.key, &key, sizeof(key));
        c.generation = gen; c.device = device; c.uses = 0; c.launches = 0; c.exec = nullptr;
        g_graphs.push_back(c);
        e = &g_graphs.back();
    }
    e->uses++;
    if (e->uses == 1) return mzgo_search_run(sp, nullptr, 0, st);
    if (!e->exec && e->launches >= 0) {
        // captured on a private stream (the caller's may be the legacy default stream, which cannot capture);
        // the instantiated graph is then launched into the caller's stream like any other work
        static thread_local cudaStream_t cap = nullptr;
        cudaGraph_t graph = nullptr;
        const int64_t before = mzgo_launch_count_peek();
        if ((!cap && cudaStreamCreateWithFlags(&cap, cudaStreamNonBlocking) != cudaSuccess) ||
            cudaStreamBeginCapture(cap, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
            cudaGetLastError();
            e->launches = -1;                                   // do not try again
            return mzgo_search_run(sp, nullptr, 0, st);
        }
        const int rc = mzgo_search_run(sp, nullptr, 0, cap);
        const cudaError_t ce = cudaStreamEndCapture(cap, &graph);
        const int captured = (int)(mzgo_launch_count_peek() - before);
        mzgo_count_launches(-captured);                         // nothing ran yet
        if (rc != MZGO_OK || ce != cudaSuccess || !graph ||
            cudaGraphInstantiate(&e->exec, graph, 0) != cudaSuccess) {
            cudaGetLastError();
            if (graph) cudaGraphDestroy(graph);
            e->exec = nullptr;
            e->launches = -1;
            return mzgo_search_run(sp, nullptr, 0, st);
        }
        cudaGraphDestroy(graph);
        e->launches = captured;
    }
    if (!e->exec) return mzgo_search_run(sp, nullptr, 0, st);
    MZGO_CUDA(cudaGraphLaunch(e->exec, st));
    g_graph_replays++;
    mzgo_count_launches(e->launches);
    return MZGO_OK;
}

// Searches served by a replayed graph so far (bench.py reports it next to gpu_launches).
extern "C" long long mzgo_graph_replays(void) { return g_graph_replays; }

extern "C" int mzgo_selfplay_pack_rows(const mzgo_selfplay_t* sp, int t0, int nrows, void* out, int g_offset, int G_total,
                                       void* stream) {
    if (int e = check_sp(sp)) return e;
    MZGO_CHECK_ARG(out != nullptr && nrows >= 0 && t0 >= 0 && t0 + nrows <= sp->traj_capacity, "rows [%d,%d) outside capacity %d",
                   t0, t0 + nrows, sp->traj_capacity);
    MZGO_CHECK_ARG(g_offset >= 0 && g_offset + sp->tree.G <= G_total, "games [%d,%d) outside the %d games of the buffer", g_offset,
                   g_offset + sp->tree.G, G_total);
    MZGO_CHECK_ARG(((uintptr_t)out & 1) == 0, "row buffer must be 2-byte aligned");
    const int warps = nrows * sp->tree.G;
    if (warps == 0) return MZGO_OK;
    k_sp_pack_rows<<<ceil_div(warps, SP_THREADS / 32), SP_THREADS, 0, (cudaStream_t)stream>>>(*sp, t0, nrows, (uint8_t*)out, g_offset,
                                                                                              G_total);
    mzgo_count_launch();
    MZGO_LAUNCH_CHECK();
    return MZGO_OK;
}

extern "C" int mzgo_selfplay_run(const mzgo_selfplay_t* sp_in, int num_moves, int traj_offset,
                                 uint32_t first_move_index, void* stream) {
    if (int e = check_sp(sp_in)) return e;
    MZGO_CHECK_ARG(num_moves >= 0 && traj_offset >= 0 && traj_offset + num_moves <= sp_in->traj_capacity,
                   "trajectory rows [%d,%d) outside capacity %d", traj_offset, traj_offset + num_moves,
                   sp_in->traj_capacity);
    MZGO_CHECK_ARG(sp_in->traj_boards && sp_in->traj_action && sp_in->traj_policy && sp_in->traj_root_value &&
                       sp_in->traj_reward && sp_in->traj_done && sp_in->traj_winner,
                   "selfplay has a NULL trajectory array");
    mzgo_selfplay_t sp = *sp_in;
    cudaStream_t st = (cudaStream_t)stream;
    const int G = sp.tree.G, A = sp.tree.A;
    if (G == 0) return MZGO_OK;
    const int blocks = ceil_div(G, SP_THREADS);
    for (int i = 0; i < num_moves; i++) {
        const int t = traj_offset + i;
        const uint32_t move_index = first_move_index + (uint32_t)i;
        sp.tree.move_index = (int32_t)move_index;
        k_sp_begin_move<<<blocks, SP_THREADS, 0, st>>>(sp);
        mzgo_count_launch();
        MZGO_LAUNCH_CHECK();
        // observation of this move (trajectory['observations'], main.py:640-645): the packed position
        MZGO_CUDA(cudaMemcpyAsync(sp.traj_boards + (size_t)t * G * MZGO_BOARD_WORDS, sp.boards,
                                  (size_t)G * MZGO_BOARD_WORDS * 4, cudaMemcpyDeviceToDevice, st));
        // MCTS.run: initial inference + root preparation (main.py:256-288)
        if (int e = mzgo_net_initial(&sp.net, nullptr, sp.boards, sp.arena, sp.out_slot, sp.root_value_net,
                                     sp.root_policy, 1, G, stream))
            return e;
        if (int e = mzgo_tree_reset_roots(&sp.tree, &sp.params, sp.root_policy, sp.boards, sp.n, nullptr, move_index,
                                          stream))
            return e;
        // the simulations (main.py:291-314)
        if (int e = search_run_cached(&sp, st)) return e;
        // action + policy target (main.py:659-677), then env.step (main.py:681)
        if (int e = mzgo_tree_choose_action(&sp.tree, &sp.params, 0.f, sp.game_temperature, move_index, sp.action,
                                            sp.traj_policy + (size_t)t * G * A, sp.traj_root_value + (size_t)t * G,
                                            stream))
            return e;
        if (int e = mzgo_rules_step(sp.boards, sp.action, nullptr, G, sp.n, stream)) return e;
        if (int e = mzgo_rules_score(sp.boards, nullptr, sp.score_winner, sp.komi, G, sp.n, stream)) return e;
        k_sp_end_move<<<blocks, SP_THREADS, 0, st>>>(sp, t);
        mzgo_count_launch();
        MZGO_LAUNCH_CHECK();
    }
    return MZGO_OK;
}
