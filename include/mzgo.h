/* mzgo.h -- C ABI of libmzgo.so, the B200 (sm_100a) self-play hot path for MuZero-Go.
 *
 * The reference (Sir-Teo/MuZero-Go) is pure Python and has no FFI; every entry point below
 * names the reference code (file:line under /root/reference) whose arithmetic it replaces.
 * Python binds these with ctypes (muzero_go_b200/_lib.py).
 *
 * Conventions
 *   - every function returns 0 (MZGO_OK) or a negative MZGO_ERR_* code; mzgo_last_error()
 *     returns a thread-local message.  Nothing throws or exits across this boundary.
 *   - pointers marked "device" are CUDA device pointers owned by the CALLER (PyTorch tensors);
 *     the library allocates no device memory.  Descriptor structs (mzgo_tree_t, mzgo_net_t,
 *     mzgo_selfplay_t) live on the HOST and only carry device pointers and sizes.
 *   - `stream` is a cudaStream_t passed as void*.  All calls are asynchronous on that stream
 *     unless documented otherwise.
 *   - there is no CPU fallback: without a CUDA device every compute call fails with
 *     MZGO_ERR_CUDA.
 */
#ifndef MZGO_H
#define MZGO_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MZGO_OK 0
#define MZGO_ERR_ARG (-1)
#define MZGO_ERR_CUDA (-2)
#define MZGO_ERR_UNSUPPORTED (-3)

#define MZGO_ABI_VERSION 1

const char* mzgo_last_error(void);
int mzgo_abi_version(void);
/* 1 when a CUDA device with compute capability 10.x is usable, else 0 (never fails). */
int mzgo_device_ok(void);
/* sizeof of the descriptor structs as compiled: 0 mzgo_tree_t, 1 mzgo_search_params_t, 2 mzgo_net_t,
 * 3 mzgo_selfplay_t (lets a binding check its struct mirrors); -1 for an unknown index. */
int64_t mzgo_sizeof(int which);

/* ------------------------------------------------------------------------------------------
 * 1. Go rules engine (replaces GymGo: call sites main.py:628,637,681,709; planes main.py:261)
 *
 * Packed board record: MZGO_BOARD_WORDS uint32 per board (256 B, one coalesced warp load):
 *   words [0,n)      black stones, one row per word, bit c = column c
 *   words [n,2n)     white stones
 *   words [2n,3n)    invalid-move plane for the player to move (GymGo INVD_CHNL)
 *   word  60         turn (0 black, 1 white)          (GymGo TURN_CHNL)
 *   word  61         bit0 previous move was a pass    (PASS_CHNL), bit1 game over (DONE_CHNL)
 *   word  62         ply counter (not part of GymGo's state)
 *   word  63         reserved (0)
 * Board sizes 2 <= n <= 19.  Actions: a = row*n + col, pass = n*n (play.py:261-278).
 * ------------------------------------------------------------------------------------------ */
#define MZGO_BOARD_WORDS 64
#define MZGO_MAX_N 19

/* GoEnv.reset (main.py:637): all boards empty, black to move. */
int mzgo_rules_reset(uint32_t* boards /*device*/, int nboards, int n, void* stream);

/* gogame.next_state via GoEnv.step (main.py:681).  actions[b] < 0 leaves board b untouched.
 * status[b] (optional, may be NULL): 0 ok / skipped, -1 illegal move (board untouched),
 * -2 game already over (board untouched). */
int mzgo_rules_step(uint32_t* boards /*device*/, const int32_t* actions /*device*/,
                    int32_t* status /*device, optional*/, int nboards, int n, void* stream);

/* gogame.areas + gogame.winning (GoEnv.winner, main.py:709).  areas[b] = {black, white};
 * winner[b] = sign(black - white - komi) from black's point of view, regardless of the done flag. */
int mzgo_rules_score(const uint32_t* boards /*device*/, int32_t* areas /*device [nboards][2]*/,
                     int32_t* winner /*device [nboards]*/, float komi, int nboards, int n, void* stream);

/* Packed record -> GymGo planes [nboards][6][n][n] (what env.reset/env.step return, main.py:637/:681).
 * dtype: 0 = float32, 1 = float64, 2 = uint8. */
int mzgo_rules_expand_planes(const uint32_t* boards /*device*/, void* planes /*device*/, int dtype,
                             int nboards, int n, void* stream);

/* GymGo planes (float32 [nboards][6][n][n], the `observation` of MCTS.run, main.py:255) -> packed. */
int mzgo_rules_pack_planes(const float* planes /*device*/, uint32_t* boards /*device*/, int nboards,
                           int n, void* stream);

/* BASELINE.json configs[2]: random playout of T plies per board, all in registers.  Ply t of board b
 * plays the k-th (ascending action index) entry of {valid points} U {pass}, k = R[b*T+t] % n_valid;
 * finished boards stop.  Optional outputs (NULL to skip): actions_out [nboards][T] int32 (-1 = not
 * played), hash_out [nboards][T] uint64 (state hash after every ply, see mzgo_board_hash in
 * rules.cu), trace_out [nboards][T][6][n][n] uint8 planes after every ply. */
int mzgo_rules_playout(uint32_t* boards /*device*/, const uint32_t* R /*device*/, int T,
                       int32_t* actions_out, uint64_t* hash_out, uint8_t* trace_out, int nboards, int n,
                       void* stream);

/* ------------------------------------------------------------------------------------------
 * 2. MCTS trees, structure-of-arrays, one tree per concurrent game
 *    (replaces MCTSNode / MCTS.run / select_leaf / backpropagate, main.py:145-155, :247-368).
 *
 * Node 0 is the root.  A node's per-edge statistics live with the PARENT
 * (edge_visit[a] == child.visit_count, edge_vsum[a] == child.value_sum, always equal in the
 * reference because every visit of a child passes through its edge), so PUCT reads are contiguous.
 * ------------------------------------------------------------------------------------------ */
typedef struct mzgo_tree {
    int32_t G;            /* concurrent games (trees) */
    int32_t A;            /* action size n*n+1, <= 384 */
    int32_t M;            /* node capacity per tree (>= simulations + 1), <= 32767 */
    int32_t move_index;   /* keys the device RNG together with params.seed; set by the caller per move */
    double* root_prior;   /* device [G][A]    float64 root priors (main.py:276-282) */
    float* root_mask;     /* device [G][A]    valid_mask of main.py:261-267 (0/1, pass = pass_epsilon or 1) */
    float* prior;         /* device [G][M][A] float32 child priors (main.py:303); row 0 unused */
    int32_t* edge_visit;  /* device [G][M][A] */
    double* edge_vsum;    /* device [G][M][A] */
    int16_t* child;       /* device [G][M][A] node index of the child, -1 = not expanded */
    int32_t* parent;      /* device [G][M]    parent_node * 512 + action, -1 for the root */
    int32_t* root_visit;  /* device [G] */
    double* root_vsum;    /* device [G] */
    int32_t* n_nodes;     /* device [G] */
    int32_t* leaf_parent; /* device [G] node chosen by the last select, -1 = simulation dropped / idle */
    int32_t* leaf_action; /* device [G] */
    int32_t* leaf_node;   /* device [G] index the new child will get */
    int32_t* leaf_depth;  /* device [G] depth of the new child (root children = 1) */
    int32_t* leaf_in_slot;  /* device [G] latent slot of leaf_parent = g*M + leaf_parent (-1 when dropped / idle) */
    int32_t* leaf_out_slot; /* device [G] latent slot of the new child = g*M + leaf_node */
    uint32_t* choice_ctr; /* device [G] random.choice calls made so far (main.py:324) */
    int64_t* depth_sum;   /* device [G] sum of leaf depths over completed simulations */
    int32_t* sims_done;   /* device [G] completed simulations */
    int32_t* dropped;     /* device [G] simulations dropped because nothing was selectable (main.py:359-363) */
    uint8_t* active;      /* device [G] 0 = tree idle (game finished) */
    int32_t* move_index_dev; /* device [1]: mzgo_tree_reset_roots stores its move_index here and the tree-step
                                kernel keys its RNG with it, so a captured CUDA graph of the search loop can be
                                replayed for every move (mzgo_selfplay_run) */
} mzgo_tree_t;

/* Search variants.  MAIN: main.py:247-368 and play.py:104-216.  SELFPLAY: self_play.py:142-343
 * (root priors re-masked and renormalised after the noise, child priors masked and renormalised,
 * sqrt(max(1, N_parent)), sign-alternating backup, terminal leaves back 0 up along the path). */
#define MZGO_VARIANT_MAIN 0
#define MZGO_VARIANT_SELFPLAY 1

typedef struct mzgo_search_params {
    double c_puct;             /* main.py:248 (2), play.py:116 (10.0), self_play.py:143 (2.5) */
    double discount;           /* main.py:37, used at :311 */
    double dirichlet_epsilon;  /* main.py:44 */
    double pass_epsilon;       /* main.py:52 */
    float dirichlet_alpha;     /* main.py:45 (device sampler only) */
    int32_t c_puct_is_int;     /* 1: c*P exact in P's dtype (main.py); 0: float c, rounds in float32 below the root */
    uint64_t seed;             /* device RNG key (Philox) when no stream / noise is injected */
    int32_t variant;           /* MZGO_VARIANT_* */
    int32_t game_base;         /* global id of game 0 of this tree set: the device RNG is keyed by (seed, game_base + g,
                                  move, draw), so a job's games play the same whichever way they are sharded over
                                  ranks or waves */
} mzgo_search_params_t;

/* Root preparation, main.py:259-288.  root_soft: float32 softmax of the initial policy logits.
 * The valid mask comes from the packed boards' invalid plane (main.py:261-267).
 * noise: float64 [G][A] replacing np.random.dirichlet (main.py:63), or NULL to sample on device
 * (Philox + Marsaglia-Tsang gamma; distribution parity only).  move_index keys the device sampler.
 * Clears all per-tree counters and marks trees with active[g]!=0 ready. */
int mzgo_tree_reset_roots(const mzgo_tree_t* tree, const mzgo_search_params_t* p,
                          const float* root_soft /*device [G][A]*/, const uint32_t* boards /*device*/,
                          int n, const double* noise /*device, optional*/, uint32_t move_index,
                          void* stream);

/* select_leaf, main.py:318-363.  choice_stream: uint32 [G][stride] replacing random.choice as
 * unexpanded[stream[g][i] % len] for the i-th call of game g, or NULL for the device RNG.
 * Writes leaf_parent/leaf_action/leaf_node (leaf_action = -1 when the simulation is dropped). */
int mzgo_tree_select(const mzgo_tree_t* tree, const mzgo_search_params_t* p,
                     const uint32_t* choice_stream /*device, optional*/, int stride, void* stream);

/* Expansion + backpropagate, main.py:300-313 and :365-368.  soft: float32 [G][A] softmax of the
 * child's policy logits; reward/value: float32 [G].  Child priors = soft where root_mask > 0. */
int mzgo_tree_expand_backup(const mzgo_tree_t* tree, const mzgo_search_params_t* p,
                            const float* soft /*device*/, const float* reward /*device*/,
                            const float* value /*device*/, void* stream);

/* Fused simulation step (what mzgo_search_run launches once per simulation): expansion + backup of the
 * leaf picked by the previous step (do_backup != 0; arguments as mzgo_tree_expand_backup), then the
 * selection of the next leaf (do_select != 0; arguments as mzgo_tree_select).  Same arithmetic as the
 * two separate entry points. */
int mzgo_tree_step(const mzgo_tree_t* tree, const mzgo_search_params_t* p, const float* soft /*device*/,
                   const float* reward /*device*/, const float* value /*device*/,
                   const uint32_t* choice_stream /*device, optional*/, int stride, int do_backup, int do_select,
                   void* stream);

/* Visit-count extraction, main.py:666-669: out[g][a] = child visit count where root_mask > 0. */
int mzgo_tree_visit_counts(const mzgo_tree_t* tree, int32_t* out /*device [G][A]*/, void* stream);

/* Action choice + policy target, main.py:670-677 (temperature <= 0: argmax, first maximum wins) and
 * self_play.py:389-400 (temperature > 0: sample from visits^(1/T)).  Falls back to a uniform random
 * valid action when every count is zero.  policy: float32 [G][A] = counts / sum; root_value [G] =
 * root value_sum / visit_count.  Idle trees get action -1. */
int mzgo_tree_choose_action(const mzgo_tree_t* tree, const mzgo_search_params_t* p, float temperature,
                            const float* temperature_per_game /*device [G], optional, overrides temperature*/,
                            uint32_t move_index, int32_t* action /*device [G]*/,
                            float* policy /*device [G][A], optional*/,
                            float* root_value /*device [G], optional*/, void* stream);

/* ------------------------------------------------------------------------------------------
 * 3. Networks (replaces RepresentationNetwork / DynamicsNetwork / PredictionNetwork /
 *    MuZeroNet.initial_inference / recurrent_inference, main.py:72-142)
 *
 * Weights are the reference's own float32 parameters (state_dict layout), read in
 * place from the caller's tensors.  mzgo_net_prepare derives the kernel-side layouts into a
 * caller-provided workspace (mzgo_net_workspace_bytes).
 *
 * Latent arena: `slots` of one latent each, addressed by int32 slot indices.
 *   precision 0 (fp32 mode): slot = float32 [C][n*n]          (NCHW, the reference layout)
 *   precision 1 (bf16 mode): slot = bf16 [C/8][PSH][8]: the n(n+1)-1 flat positions r(n+1)+c from the first to the
 *                            last pixel (row stride n+1: one zero between two rows), PSH = that count rounded up to
 *                            a multiple of 8; the padding around them exists in shared memory only (DESIGN.md
 *                            "latent arena").  The arena must start out ZEROED (the zeros between the rows and
 *                            behind the last pixel are never rewritten) and be 16-byte aligned; use
 *                            mzgo_net_slot_bytes, mzgo_net_import_latent / _export_latent rather than the layout.
 * ------------------------------------------------------------------------------------------ */
typedef struct mzgo_net {
    int32_t n;          /* board size */
    int32_t C;          /* latent_dim (main.py:29); bf16 mode needs C == 128 */
    int32_t A;          /* n*n+1 == rows of the action embedding */
    int32_t precision;  /* 0 fp32 (CUDA cores, 1e-5 parity), 1 bf16 operands / fp32 accumulate (tcgen05) */
    int32_t max_batch;  /* largest B passed to mzgo_net_initial (sizes the workspace) */
    int32_t reserved;
    /* reference parameters, float32, device, contiguous, state_dict shapes */
    const float* rep_w1; const float* rep_b1;   /* [64][6][3][3], [64]   main.py:75 */
    const float* rep_w2; const float* rep_b2;   /* [64][64][3][3], [64]  main.py:76 */
    const float* rep_w3; const float* rep_b3;   /* [C][64][3][3], [C]    main.py:77 */
    const float* emb;                           /* [A][C]                main.py:89 */
    const float* dyn_w; const float* dyn_b;     /* [C][C][3][3], [C]     main.py:90 */
    const float* rew_w; const float* rew_b;     /* [1][C][1][1], [1]     main.py:91 */
    const float* rh_w; const float* rh_b;       /* [16][1], [16]         main.py:93 */
    const float* ro_w; const float* ro_b;       /* [1][16], [1]          main.py:94 */
    const float* val_w; const float* val_b;     /* [1][C][1][1], [1]     main.py:111 */
    const float* vfc_w; const float* vfc_b;     /* [1][1], [1]           main.py:112 */
    const float* pol_w; const float* pol_b;     /* [1][C][1][1], [1]     main.py:113 */
    const float* pass_logit;                    /* [1]                   main.py:114 */
    void* workspace;                            /* device, mzgo_net_workspace_bytes() bytes, 256-B aligned */
    int64_t workspace_bytes;
} mzgo_net_t;

int64_t mzgo_net_workspace_bytes(int n, int C, int precision, int max_batch);
/* bytes of one latent slot for this board size / width / precision */
int64_t mzgo_net_slot_bytes(int n, int C, int precision);
/* Re-derive kernel-side weight layouts after the parameters changed (load_state_dict, broadcast). */
int mzgo_net_prepare(const mzgo_net_t* net, void* stream);

/* initial_inference, main.py:134-137.  Input is either GymGo planes (obs, float32 [B][6][n][n]) or
 * packed boards (obs == NULL).  Writes the latent of item b into slot out_slot[b] (skipped when
 * out_slot[b] < 0), value[b], and policy[b][A] = logits, or their softmax when want_softmax != 0. */
int mzgo_net_initial(const mzgo_net_t* net, const float* obs /*device, optional*/,
                     const uint32_t* boards /*device, optional*/, void* arena /*device*/,
                     const int32_t* out_slot /*device [B]*/, float* value /*device [B]*/,
                     float* policy /*device [B][A]*/, int want_softmax, int B, void* stream);

/* recurrent_inference, main.py:139-142.  Item b reads slot in_slot[b] and action[b], writes slot
 * out_slot[b]; items with action[b] < 0 are skipped.  reward/value [B], policy [B][A] as above. */
int mzgo_net_recurrent(const mzgo_net_t* net, void* arena /*device*/, const int32_t* in_slot,
                       const int32_t* action, const int32_t* out_slot, float* reward, float* value,
                       float* policy, int want_softmax, int B, void* stream);

/* bf16 mode only: mzgo_net_recurrent on every game's selected leaf (in_slot / action / out_slot are the tree's
 * leaf_in_slot / leaf_action / leaf_out_slot) with the search fused behind it in the same kernel: as soon as a
 * game's outputs are final, a warp group of the same CTA runs mzgo_tree_step for it (expansion + backup and, when
 * do_select != 0, the selection of its next leaf; same arithmetic, main.py:294-363).  One launch per
 * simulation; this is what mzgo_search_run enqueues in bf16 mode.  reward / value / policy (softmax) still
 * receive the network outputs. */
int mzgo_net_recurrent_tree(const mzgo_net_t* net, void* arena /*device*/, const mzgo_tree_t* tree,
                            const mzgo_search_params_t* params, const uint32_t* choice_stream /*device, optional*/,
                            int stride, float* reward /*device [G]*/, float* value /*device [G]*/,
                            float* policy /*device [G][A]*/, int do_select, void* stream);

/* PredictionNetwork.forward alone (main.py:116-125) on the latents in slot[b]: value [B], policy [B][A]. */
int mzgo_net_predict(const mzgo_net_t* net, const void* arena /*device*/, const int32_t* slot /*device [B]*/,
                     float* value, float* policy, int want_softmax, int B, void* stream);

/* Slot <-> reference tensor layout (float32 [B][C][n][n]); identity copy in fp32 mode. */
int mzgo_net_export_latent(const mzgo_net_t* net, const void* arena, const int32_t* slot,
                           float* out /*device [B][C][n][n]*/, int B, void* stream);
int mzgo_net_import_latent(const mzgo_net_t* net, void* arena, const int32_t* slot,
                           const float* in /*device [B][C][n][n]*/, int B, void* stream);

/* ------------------------------------------------------------------------------------------
 * 4. Batched self-play (replaces the per-move loop main.py:656-702 / self_play.py:453-526 for
 *    G concurrent games in lock step; slot of (game g, node i) = g*M + i)
 * ------------------------------------------------------------------------------------------ */
typedef struct mzgo_selfplay {
    mzgo_tree_t tree;
    mzgo_search_params_t params;
    mzgo_net_t net;
    int32_t n;                 /* board size */
    int32_t num_simulations;   /* main.py:32 */
    int32_t max_game_moves;    /* main.py:53: int(1.5*n*n); self_play.py:461: n*n */
    int32_t temperature_moves; /* self_play.py:29 (0 = always greedy, main.py:671) */
    float temperature;         /* self_play.py:30 */
    float komi;                /* main.py:628 komi=0 */
    int32_t restart_finished;  /* 1: a finished game is replaced by a fresh one at once */
    int32_t traj_capacity;     /* moves recordable per call of mzgo_selfplay_run (T) */
    uint32_t* boards;          /* device [G][64] current positions */
    void* arena;               /* device [G*M] latent slots */
    float* root_policy;        /* device [G][A] scratch: initial softmax */
    float* root_value_net;     /* device [G]    scratch: initial value */
    float* leaf_policy;        /* device [G][A] scratch: recurrent softmax */
    float* leaf_reward;        /* device [G] */
    float* leaf_value;         /* device [G] */
    int32_t* in_slot;          /* device [G] scratch */
    int32_t* out_slot;         /* device [G] scratch */
    int32_t* action;           /* device [G] scratch: chosen move */
    int32_t* move_count;       /* device [G] moves played in the current game */
    int32_t* pass_streak;      /* device [G] consecutive passes (main.py:696-702) */
    float* game_temperature;   /* device [G] scratch: sampling temperature of the current move */
    int32_t* score_winner;     /* device [G] scratch: gogame.winning of the current position */
    /* trajectory records, main.py:640-645 / self_play.py:415-450, indexed [t][g] */
    uint32_t* traj_boards;     /* device [T][G][64] position BEFORE the move */
    int32_t* traj_action;      /* device [T][G]  (-1 = no move recorded) */
    float* traj_policy;        /* device [T][G][A] visit-count policy target (main.py:672) */
    float* traj_root_value;    /* device [T][G] */
    float* traj_reward;        /* device [T][G] env reward of the move (GymGo 'real': winner at the end) */
    uint8_t* traj_done;        /* device [T][G] 1 when this move ended the game */
    int8_t* traj_winner;       /* device [T][G] env.winner() after a game-ending move (main.py:709) */
    int64_t* counters;         /* device [8]: 0 moves, 1 simulations, 2 games finished, 3 dropped sims,
                                  4 sum of leaf depths, 5 black wins, 6 white wins, 7 move-cap endings */
} mzgo_selfplay_t;

/* Plays `num_moves` lock-step moves (one MCTS.run + env.step per active game each), recording into
 * trajectory rows [traj_offset, traj_offset+num_moves).  first_move_index keys the RNG.
 * Launches only; the caller synchronises the stream. */
int mzgo_selfplay_run(const mzgo_selfplay_t* sp, int num_moves, int traj_offset,
                      uint32_t first_move_index, void* stream);

/* Trajectory rows [t0, t0 + nrows) of this engine in the wire layout of the trajectory gather (north_star (4): "a
 * trajectory gather to the replay buffer"; SURVEY 8e: fixed-stride records): per (row, game) the bytes
 *   board record 256 | action int32 | policy A x float32 | root value float32 | reward float32 | done uint8 | winner int8
 * (stride 270 + 4A), written to out[r][g_offset + g], out = [nrows][G_total][stride] bytes (device, 2-byte aligned).
 * The waves of a WaveGroup pack into one buffer with their game offsets; the buffer is what is all-gathered. */
int mzgo_selfplay_pack_rows(const mzgo_selfplay_t* sp, int t0, int nrows, void* out /*device*/, int g_offset, int G_total,
                            void* stream);

/* One search only (the body of MCTS.run after the root is prepared): `num_simulations` rounds of
 * select -> recurrent inference -> expand/backup on sp->tree.  choice_stream as in mzgo_tree_select. */
int mzgo_search_run(const mzgo_selfplay_t* sp, const uint32_t* choice_stream, int stride, void* stream);

/* ------------------------------------------------------------------------------------------
 * 5. Training-side 3x3 convolutions (MuZeroAgent.train, main.py:381-501, whose forward and
 *    backward run on nn.Conv2d / autograd in the reference -- main.py:75-77, :90).  Dense float32 NCHW batches,
 *    padding 1; forward, data gradient and weight gradient; ReLU's backward mask is folded into the gradient
 *    kernels' operand loads (y_gate = the layer's own output; NULL = no ReLU).  float32 on the CUDA cores: the
 *    trainer's parity contract is 1e-5 against the real MuZeroAgent.train.  All launches asynchronous on `stream`,
 *    CUDA-graph capturable.
 * ------------------------------------------------------------------------------------------ */
/* w [Cout][Cin][3][3] -> wT_fwd [Cin][9][Cout] (forward operand) and / or wT_bwd [Cout][9][Cin], taps flipped (data
 * gradient operand); either may be NULL. */
int mzgo_train_conv3x3_prep_w(const float* w /*device*/, float* wT_fwd /*device*/, float* wT_bwd /*device*/, int Cin,
                              int Cout, void* stream);
/* y [B][Cout][n][n] = conv3x3(x [B][Cin][n][n]) + bias (optional), then ReLU when relu != 0.  Cout % 8 == 0. */
int mzgo_train_conv3x3_fwd(const float* x, const float* wT_fwd, const float* bias, float* y, int relu, int B, int Cin,
                           int Cout, int n, void* stream);
/* dx [B][Cin][n][n] = data gradient of the layer for dy [B][Cout][n][n] (masked by y_gate > 0 when given).  Cin % 8 == 0. */
int mzgo_train_conv3x3_bwd_data(const float* dy, const float* y_gate, const float* wT_bwd, float* dx, int B, int Cin,
                                int Cout, int n, void* stream);
/* dw [Cout][Cin][3][3], db [Cout] (optional) = weight / bias gradient; scratch: mzgo_train_wgrad_scratch_floats floats.
 * Deterministic (two passes, no atomics). */
int64_t mzgo_train_wgrad_scratch_floats(int B, int Cin, int Cout);
int mzgo_train_conv3x3_bwd_weight(const float* x, const float* dy, const float* y_gate, float* dw, float* db,
                                  float* scratch, int B, int Cin, int Cout, int n, void* stream);
/* The same weight gradient for callers that cannot allocate (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments and results as mzgo_train_conv3x3_bwd_weight (bit for bit: same kernels, same order of summation),
 * minus `scratch`: the partial sums live in a device array built into the library, MZGO_TRAIN_WGRAD_STATIC_MAX_FLOATS
 * floats (10 MiB, held by every GPU that loads the library until the process exits; enough for the reference's
 * largest layer, 129 -> 128 channels, at any batch size).  Shapes with mzgo_train_wgrad_scratch_floats(B, Cin, Cout)
 * beyond that are refused with MZGO_ERR_ARG and a message naming both numbers.  The call itself allocates nothing
 * (no cudaMalloc / cudaFree / allocating library routine) and launches on `stream` without synchronising.  All calls
 * on one device share the array: nothing in it needs to start at zero (each call writes every partial sum before
 * it reads it), but calls that OVERLAP on the GPU -- two streams, two host threads -- race on it; serialise them. */
#define MZGO_TRAIN_WGRAD_STATIC_MAX_FLOATS (10 * 1024 * 1024 / 4)
int64_t mzgo_train_wgrad_static_max_floats(void);
int mzgo_train_conv3x3_bwd_weight_static(const float* x, const float* dy, const float* y_gate, float* dw, float* db, int B,
                                         int Cin, int Cout, int n, void* stream);

/* Test / profiling hooks.  mzgo_debug_fastdiv_check: operand pairs on which the tree kernels' 5-operation
 * correctly rounded division differs from the generic IEEE division (must be 0; synchronous).
 * mzgo_debug_tc_trace: when buf != NULL every tensor-core conv launch stamps its role timeline into
 * buf[grid][64] (uint64 globaltimer values, tools/tc_trace.py). */
long long mzgo_debug_fastdiv_check(unsigned long long seed, int blocks, long long per_thread);
int mzgo_debug_tc_trace(void* buf /*device*/);

/* Trace build only (python -m muzero_go_b200.build --define MZGO_TRACE_TREE): when buf != NULL every tree step stamps
 * its phases per game into buf[G][16] (uint64 globaltimer values, tools/tree_trace.py); MZGO_ERR_UNSUPPORTED otherwise. */
int mzgo_debug_tree_trace(void* buf /*device*/);

/* mzgo_rules_step / mzgo_rules_playout run a warp per board below this many boards and a thread per board from
 * it on (board sizes 5, 6, 9, where it is the faster one on large batches; 13 and 19 only when the value is 0;
 * same results bit for bit).  Returns the previous value; a negative argument only reads it.  Default 32768. */
int mzgo_debug_rules_tpb_min(int min_boards);

/* Number of searches mzgo_selfplay_run served by replaying its captured CUDA graph of the search loop
 * (the graph is captured on the second move of an engine; MZGO_NO_GRAPH=1 in the environment disables it). */
long long mzgo_graph_replays(void);

/* Number of kernels launched by this library in the calling thread since the last reset
 * (bench.py's gpu_launches). */
int64_t mzgo_launch_count(int reset);

#ifdef __cplusplus
}
#endif
#endif /* MZGO_H */
