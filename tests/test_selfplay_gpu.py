"""GPU: the drop-in MCTS.run / MuZeroAgent surface and the batched self-play engine."""
import numpy as np
import pytest
import torch

from oracle import gymgo, mcts_ref
from oracle.nets_ref import MuZeroNetRef

pytestmark = pytest.mark.gpu


def make_net(n, c, seed, precision="fp32"):
    import muzero_go_b200 as mz
    torch.manual_seed(seed)
    ref = MuZeroNetRef(c, n * n + 1).eval()
    with torch.no_grad():     # spread the outputs so the search goes deeper than the root's children
        ref.prediction.policy_conv.weight.mul_(8.0)
        ref.prediction.value_fc.weight.fill_(3.0)
    net = mz.MuZeroNet(c, n * n + 1, precision=precision)
    net.load_state_dict(ref.state_dict())
    return ref, net.cuda().eval()


@pytest.mark.parametrize("variant", ["main", "self_play"])
def test_mcts_run_drop_in_matches_oracle_search_with_same_network_outputs(variant):
    """MCTS(net, A, sims).run(obs) on the GPU vs the oracle search driven by the SAME float32 network
    outputs: the oracle's callbacks evaluate our CUDA fp32 network (so both sides consume identical
    numbers, as north_star specifies) and apply the same softmax kernel.  This path runs the fused
    tree-step kernel (mzgo_search_run)."""
    import muzero_go_b200 as mz
    from oracle.make_golden import midgame_obs
    n, c, sims = 6, 32, 150
    A = n * n + 1
    _, net = make_net(n, c, 21)
    obs = midgame_obs(n, 7, 3)
    rng = np.random.default_rng(0)
    noise = rng.dirichlet([0.3] * A)
    stream = rng.integers(0, 2 ** 32, size=sims + 8, dtype=np.uint64).astype(np.uint32)
    mcts = mz.MCTS(net, A, sims, variant=variant)
    root = mcts.run(obs, noise=noise, choice_stream=stream)
    if variant == "self_play":                      # self_play.py:233-237 returns a triple
        root, vc_ret, rv_ret = root

    def initial(o):
        lat, v, p = net._run_initial(torch.from_numpy(np.asarray(o, dtype=np.float32))[None].cuda(), want_softmax=True)
        return lat, v.cpu().numpy().reshape(-1)[0], p.cpu().numpy()[0]

    def recurrent(lat, a):
        nxt, r, v, p = net._run_recurrent(lat, torch.tensor([a]).cuda(), want_softmax=True)
        return nxt, r.cpu().numpy().reshape(-1)[0], v.cpu().numpy().reshape(-1)[0], p.cpu().numpy()[0]

    if variant == "self_play":
        sc = mz.SearchConfig.for_variant("self_play")
        tree, mask = mcts_ref.run_mcts_selfplay(obs, initial, recurrent, sims, noise, stream, c_puct=sc.c_puct,
                                                dirichlet_epsilon=sc.dirichlet_epsilon, pass_epsilon=sc.pass_epsilon)
        assert np.array_equal(vc_ret, [tree.visit[int(c_)] if c_ >= 0 else 0 for c_ in tree.child[0]])
        assert rv_ret == pytest.approx(tree.q(0), rel=1e-12)
    else:
        tree, mask = mcts_ref.run_mcts(obs, initial, recurrent, sims, noise, stream)
    got = np.array([ch["node"].visit_count if ch["node"] is not None else 0 for ch in root.children.values()])
    assert np.array_equal(got, tree.root_visit_counts(mask))
    assert root.visit_count == tree.visit[0] == sims
    assert root.value() == pytest.approx(tree.q(0), rel=1e-12)
    assert max(tree.depths) >= 2
    priors = np.array([float(ch["prior"]) for ch in root.children.values()])
    assert np.array_equal(priors, np.asarray(tree.prior[0], dtype=np.float64))
    # a child view exposes the reference's fields too
    a_best = int(np.argmax(got))
    child = root.children[a_best]["node"]
    assert child.visit_count == got[a_best] and tuple(child.latent.shape) == (c, n, n)
    assert sum(ch["node"].visit_count for ch in child.children.values() if ch["node"] is not None) == child.visit_count - 1


def test_agent_plays_a_legal_game_against_the_oracle_env():
    import muzero_go_b200 as mz
    n = 5
    agent = mz.MuZeroAgent(n, 16, n * n + 1, 20)
    env = gymgo.GoEnv(n)
    obs, done, moves = env.reset(), False, 0
    while not done and moves < 30:
        if moves % 2 == 0:
            a = agent.select_action(obs)
        else:
            a, pol, rv = agent.select_action(obs, 1.0)
            assert pol.shape == (n * n + 1,) and abs(pol.sum() - 1) < 1e-9
        assert env.valid_moves()[a] == 1
        obs, _, done, _ = env.step(a)
        moves += 1
    assert moves > 4


@pytest.mark.parametrize("n,c,G,S,precision", [(5, 16, 64, 12, "fp32"), (9, 128, 32, 20, "fp32"), (9, 128, 32, 20, "bf16"),
                                               (13, 128, 6, 16, "bf16")])
def test_engine_invariants(n, c, G, S, precision):
    import muzero_go_b200 as mz
    _, net = make_net(n, c, 5, precision)
    T = 2 * int(n * n * 1.5) + 6
    eng = mz.SelfPlayEngine(net, n, G, S, traj_capacity=T, restart_finished=True)
    eng.run(T)
    torch.cuda.synchronize()
    ctr = eng.read_counters()
    assert ctr["moves"] == G * T
    assert ctr["simulations"] + ctr["dropped_simulations"] == G * T * S
    assert ctr["games_finished"] >= G          # every slot finished at least one game and restarted
    act = eng.traj_action.cpu().numpy()
    pol = eng.traj_policy.cpu().numpy()
    done = eng.traj_done.cpu().numpy()
    rows = eng.traj_boards.cpu().numpy().view(np.uint32)
    assert act.min() >= 0 and act.max() <= n * n
    np.testing.assert_allclose(pol.sum(axis=2), 1.0, atol=1e-5)
    # replay every recorded game through the oracle rules: positions before each move must match
    for g in range(min(G, 8)):
        st = gymgo.init_state(n)
        count, passes = 0, 0
        for t in range(T):
            r = rows[t, g]
            black = np.array([[(r[i] >> j) & 1 for j in range(n)] for i in range(n)])
            white = np.array([[(r[n + i] >> j) & 1 for j in range(n)] for i in range(n)])
            inv = np.array([[(r[2 * n + i] >> j) & 1 for j in range(n)] for i in range(n)])
            assert np.array_equal(black, st[0]) and np.array_equal(white, st[1]) and np.array_equal(inv, st[3]), (g, t)
            a = int(act[t, g])
            assert gymgo.valid_moves(st)[a] == 1
            assert pol[t, g, a] > 0
            st = gymgo.next_state(st, a)
            count += 1
            passes = passes + 1 if a == n * n else 0
            over = bool(gymgo.game_ended(st)) or passes >= 2 or count >= int(n * n * 1.5)
            assert bool(done[t, g]) == over
            if over:
                if gymgo.game_ended(st):
                    assert int(eng.traj_winner[t, g]) == int(gymgo.winning(st))
                st, count, passes = gymgo.init_state(n), 0, 0


@pytest.mark.parametrize("n,c,G,S,precision,variant", [(5, 16, 16, 10, "fp32", "main"), (5, 128, 16, 40, "bf16", "main"),
                                                       (9, 128, 150, 120, "bf16", "main"), (9, 128, 40, 100, "bf16", "self_play"),
                                                       (13, 128, 9, 200, "bf16", "main"), (19, 128, 150, 24, "bf16", "main"),
                                                       (19, 128, 5, 430, "bf16", "main")])
def test_engine_equals_stepwise_api(n, c, G, S, precision, variant):
    """mzgo_selfplay_run is the composition of the exported pieces: same seed => same moves as driving
    net/tree/rules call by call from Python.  The engine runs the fused tree-step kernel and replays the search
    loop from a CUDA graph; the stepwise side uses the stand-alone select / recurrent / expand_backup entry
    points (tensor-core networks at 9x9, 13x13 and 19x19, both search variants)."""
    import muzero_go_b200 as mz
    from muzero_go_b200 import _lib
    _, net = make_net(n, c, 9, precision)
    cfg = mz.SearchConfig.for_variant(variant, seed=1234)
    eng = mz.SelfPlayEngine(net, n, G, S, search=cfg, traj_capacity=6, restart_finished=True)
    eng.run(6)
    torch.cuda.synchronize()
    if S > n * n + 1 + 50:   # more simulations than root edges: the compared searches reach below the root's children
        ctr = eng.read_counters()
        assert ctr["leaf_depth_sum"] > ctr["simulations"] + 6 * G * 20, ctr
    # step by step
    A, M = n * n + 1, S + 1
    dev = torch.device("cuda")
    boards = mz.GoBoards(G, n).reset()
    tree = mz.Tree(G, A, M)
    arena = net.new_arena(G * M)
    d = net.descriptor(G)
    L, s = _lib.load(), _lib.stream_ptr()
    roots = (torch.arange(G, dtype=torch.int32, device=dev) * M).contiguous()
    val = torch.empty(G, device=dev)
    pol = torch.empty((G, A), device=dev)
    rew = torch.empty(G, device=dev)
    for t in range(6):
        _lib.check(L.mzgo_net_initial(d, None, _lib.ptr(boards.records), _lib.ptr(arena), _lib.ptr(roots),
                                      _lib.ptr(val), _lib.ptr(pol), 1, G, s))
        tree.reset_roots(cfg, pol, boards, n, noise=None, move_index=t)
        for _ in range(S):
            tree.select(cfg, None)
            _lib.check(L.mzgo_net_recurrent(d, _lib.ptr(arena), _lib.ptr(tree.leaf_in_slot), _lib.ptr(tree.leaf_action),
                                            _lib.ptr(tree.leaf_out_slot), _lib.ptr(rew), _lib.ptr(val), _lib.ptr(pol),
                                            1, G, s))
            tree.expand_backup(cfg, pol, rew, val)
        action, policy, _ = tree.choose_action(cfg, 0.0, move_index=t)
        assert torch.equal(action, eng.traj_action[t]), t
        assert torch.equal(policy, eng.traj_policy[t])
        boards.step(action)


def test_games_are_invariant_to_sharding_and_waves():
    """The device RNG is keyed by (seed, global game id, move, draw) (SearchConfig.game_base), so a job's games
    play the same moves whether they run as one engine, as waves that share search buffers, or as shards on
    different ranks.  Also checks graph replay against plain launches (first move of every engine)."""
    import dataclasses
    import muzero_go_b200 as mz
    n, c, G, S, T = 5, 16, 12, 12, 5
    _, net = make_net(n, c, 3)
    cfg = mz.SearchConfig(seed=77)
    whole = mz.SelfPlayEngine(net, n, G, S, search=cfg, traj_capacity=T)
    whole.run(T)
    group = mz.WaveGroup(net, n, G, S, waves=3, search=cfg, traj_capacity=T)
    group.run(T)
    torch.cuda.synchronize()
    rows = group.trajectory_rows(0, T)
    assert tuple(rows["action"].shape) == (T, G) and group.read_counters()["moves"] == G * T
    assert group.engines[1].arena.data_ptr() == group.engines[0].arena.data_ptr()
    assert torch.equal(rows["action"], whole.traj_action)
    assert torch.equal(rows["policy"], whole.traj_policy)
    assert torch.equal(rows["boards"], whole.traj_boards)
    # the same waves spread over two CUDA streams (two sets of search buffers, two network workspaces), whose kernels
    # overlap: still the same games, move for move, and the same trajectory rows
    both = mz.WaveGroup(net, n, G, S, waves=4, streams=2, search=cfg, traj_capacity=T)
    both.run(T)
    torch.cuda.synchronize()
    rows2 = both.trajectory_rows(0, T)
    assert both.engines[2].arena.data_ptr() == both.engines[0].arena.data_ptr() != both.engines[1].arena.data_ptr()
    assert both.read_counters()["moves"] == G * T
    for k in ("action", "policy", "boards", "root_value", "reward", "done", "winner"):
        assert torch.equal(rows2[k], rows[k]), k
    # the engines' own row packing (mzgo_selfplay_pack_rows, one launch per wave) writes the wire layout of the
    # trajectory gather byte for byte
    from muzero_go_b200 import dist as mzdist
    wire = torch.zeros((T, G, mzdist.row_stride(n * n + 1)), dtype=torch.uint8, device="cuda")
    both.pack_rows(0, T, wire)
    assert torch.equal(wire, mzdist.pack_rows(rows2))
    ring = mzdist.RowRing(G, n * n + 1, 2, "cuda")
    ring.write_from(1, both, 3)
    assert torch.equal(ring.buf[1], wire[3]) and torch.equal(ring.rows(1, 2)["policy"][0], rows2["policy"][3])
    for r in range(2):                               # two "ranks" of 6 games each
        shard = mz.SelfPlayEngine(net, n, G // 2, S, search=dataclasses.replace(cfg, game_base=r * (G // 2)), traj_capacity=T)
        shard.run(T)
        torch.cuda.synchronize()
        lo, hi = r * (G // 2), (r + 1) * (G // 2)
        assert torch.equal(shard.traj_action, whole.traj_action[:, lo:hi])
        assert torch.equal(shard.traj_policy, whole.traj_policy[:, lo:hi])


@pytest.mark.parametrize("n,G,S,variant", [(5, 33, 30, "main"), (9, 300, 60, "main"), (9, 64, 60, "self_play"),
                                           (13, 20, 40, "main"), (19, 151, 12, "main"), (5, 2000, 6, "main"),
                                           (13, 700, 5, "self_play"), (19, 6, 440, "main"), (19, 4, 420, "self_play")])
def test_recurrent_tree_equals_separate_calls(n, G, S, variant):
    """mzgo_net_recurrent_tree (network + tree phase in one launch: a warp per game up to 9x9, 128 threads per
    game above; fewer games than SMs, more games per CTA than it has warps / 128-thread groups) leaves the same trees as
    select / mzgo_net_recurrent / expand_backup: visit counts, value sums, priors and the chosen moves, bit-exact."""
    import muzero_go_b200 as mz
    from muzero_go_b200 import _lib
    _, net = make_net(n, 128, 5, "bf16")
    cfg = mz.SearchConfig.for_variant(variant, seed=99)
    A, M = n * n + 1, S + 1
    dev = torch.device("cuda")
    L, s = _lib.load(), _lib.stream_ptr()
    d = net.descriptor(G)
    roots = (torch.arange(G, dtype=torch.int32, device=dev) * M).contiguous()
    boards = mz.GoBoards(G, n).reset()
    boards.step(torch.randint(0, n * n, (G,), dtype=torch.int32, device=dev) % (n * n))     # different positions per game

    def search(fused):
        tree = mz.Tree(G, A, M)
        arena = net.new_arena(G * M)
        val, rew, pol = torch.empty(G, device=dev), torch.empty(G, device=dev), torch.empty((G, A), device=dev)
        _lib.check(L.mzgo_net_initial(d, None, _lib.ptr(boards.records), _lib.ptr(arena), _lib.ptr(roots), _lib.ptr(val),
                                      _lib.ptr(pol), 1, G, s))
        tree.reset_roots(cfg, pol, boards, n, noise=None, move_index=3)
        p = cfg.to_struct()
        if fused:
            _lib.check(L.mzgo_tree_step(tree.desc, p, _lib.ptr(pol), _lib.ptr(rew), _lib.ptr(val), None, 0, 0, 1, s))
            for i in range(S):
                _lib.check(L.mzgo_net_recurrent_tree(d, _lib.ptr(arena), tree.desc, p, None, 0, _lib.ptr(rew), _lib.ptr(val),
                                                     _lib.ptr(pol), 1 if i + 1 < S else 0, s))
        else:
            for _ in range(S):
                tree.select(cfg, None)
                _lib.check(L.mzgo_net_recurrent(d, _lib.ptr(arena), _lib.ptr(tree.leaf_in_slot), _lib.ptr(tree.leaf_action),
                                                _lib.ptr(tree.leaf_out_slot), _lib.ptr(rew), _lib.ptr(val), _lib.ptr(pol), 1, G, s))
                tree.expand_backup(cfg, pol, rew, val)
        action, policy, root_value = tree.choose_action(cfg, 0.0, move_index=3)
        torch.cuda.synchronize()
        return tree, action, policy, root_value

    ta, aa, pa, va = search(True)
    tb, ab, pb, vb = search(False)
    if S > A + 50:
        assert int(ta.depth_sum.sum()) > int(ta.sims_done.sum()) + G * 20      # leaves below the root's children
    assert torch.equal(aa, ab) and torch.equal(pa, pb) and torch.equal(va, vb)
    for name in ("edge_visit", "edge_vsum", "child", "root_visit", "root_vsum", "n_nodes", "sims_done", "depth_sum"):
        assert torch.equal(getattr(ta, name), getattr(tb, name)), name
    # (node 0's float32 prior row is the fused step's path scratch: the root's priors are the float64 root_prior)
    assert torch.equal(ta.prior[:, 1:], tb.prior[:, 1:]) and torch.equal(ta.root_prior, tb.root_prior)


@pytest.mark.parametrize("n,c,S,G,precision,variant,min_depth", [
    (19, 128, 450, 3, "bf16", "main", 2),     # k_tree_step<128> + k_tc_conv, two launches per simulation (the headline path)
    (19, 128, 800, 2, "bf16", "main", 3),     # ... at the headline simulation count, late positions
    (19, 128, 430, 2, "bf16", "self_play", 3),
    (9, 128, 400, 3, "bf16", "main", 3),      # mzgo_net_recurrent_tree: the tree phase of the network kernel
    (9, 128, 300, 2, "bf16", "self_play", 2),
    (13, 128, 300, 2, "bf16", "main", 2),
    (19, 32, 430, 2, "fp32", "main", 2),      # fp32 networks + k_tree_step<128>
])
def test_engine_search_matches_oracle_search_below_the_root(n, c, S, G, precision, variant, min_depth):
    """mzgo_search_run -- exactly what mzgo_selfplay_run enqueues for a move: the fused tree step / the network
    kernel with its tree phase -- against the ORACLE search (oracle/mcts_ref.py, pinned to the real reference MCTS)
    at simulation counts that take the leaves below the root's children (19x19 needs > 362 simulations for that).
    Both sides consume identical float32 network outputs: the oracle's callbacks evaluate our network on the
    single item (bf16 latents round-trip exactly through the float32 export), with injected Dirichlet noise and
    random.choice streams.  G different positions in one batch.  Bit-exact: root priors, visit counts, value sums
    of the root's children, leaf depths in order, dropped simulations, node counts."""
    import muzero_go_b200 as mz
    from muzero_go_b200 import _lib
    from oracle.make_golden import midgame_obs
    A = n * n + 1
    dev = torch.device("cuda")
    _, net = make_net(n, c, 31, precision)
    cfg = mz.SearchConfig.for_variant(variant, seed=4)
    rng = np.random.default_rng(n * 1000 + S)
    # min_depth 3 on 19x19: late positions (few legal points), so that 800 simulations reach the grandchildren
    lo, hi = (n * n // 3, n * n) if not (n == 19 and min_depth >= 3) else (330, 420)
    obs = [midgame_obs(n, int(rng.integers(lo, hi)), int(rng.integers(1 << 30))) for _ in range(G)]
    noise = np.stack([rng.dirichlet([0.3] * A) for _ in range(G)])
    streams = rng.integers(0, 2 ** 32, size=(G, S + 8), dtype=np.uint64).astype(np.uint32)
    eng = mz.SelfPlayEngine(net, n, G, S, search=cfg, restart_finished=False, traj_capacity=1)
    obs_t = torch.from_numpy(np.stack(obs).astype(np.float32)).to(dev)
    eng.boards.load_planes(obs_t)
    d = eng.descriptor()
    L, s = _lib.load(), _lib.stream_ptr()
    roots = (torch.arange(G, dtype=torch.int32, device=dev) * eng.M).contiguous()
    eng.tree.active.fill_(1)
    _lib.check(L.mzgo_net_initial(d.net, _lib.ptr(obs_t), None, _lib.ptr(eng.arena), _lib.ptr(roots),
                                  _lib.ptr(eng.root_value_net), _lib.ptr(eng.root_policy), 1, G, s))
    eng.tree.reset_roots(cfg, eng.root_policy, eng.boards, n, noise=noise, move_index=0)
    stream_t = mz.Tree.stream_tensor(streams, dev)
    _lib.check(L.mzgo_search_run(d, _lib.ptr(stream_t), stream_t.shape[1], s))
    torch.cuda.synchronize()
    t = eng.tree
    vc = t.visit_counts().cpu().numpy()

    def initial(o):
        lat, v, p = net._run_initial(torch.from_numpy(np.asarray(o, dtype=np.float32))[None].cuda(), want_softmax=True)
        return lat, v.cpu().numpy().reshape(-1)[0], p.cpu().numpy()[0]

    def recurrent(lat, a):
        nxt, r, v, p = net._run_recurrent(lat, torch.tensor([a]).cuda(), want_softmax=True)
        return nxt, r.cpu().numpy().reshape(-1)[0], v.cpu().numpy().reshape(-1)[0], p.cpu().numpy()[0]

    run_ref = mcts_ref.run_mcts_selfplay if variant == "self_play" else mcts_ref.run_mcts
    deepest = 0
    for k in range(G):
        ref, mask = run_ref(obs[k], initial, recurrent, S, noise[k], streams[k], c_puct=cfg.c_puct, discount=cfg.discount,
                            dirichlet_epsilon=cfg.dirichlet_epsilon, pass_epsilon=cfg.pass_epsilon)
        assert np.array_equal(t.root_prior[k].cpu().numpy(), np.asarray(ref.prior[0], dtype=np.float64)), k
        assert np.array_equal(vc[k], ref.root_visit_counts(mask)), k
        assert np.array_equal(t.edge_vsum[k, 0].cpu().numpy(), ref.root_child_value_sums()), k
        assert int(t.root_visit[k]) == ref.visit[0] and int(t.dropped[k]) == ref.dropped, k
        assert int(t.n_nodes[k]) == len(ref.visit), k
        assert int(t.depth_sum[k]) == sum(ref.depths) and int(t.sims_done[k]) == len(ref.depths), k
        deepest = max(deepest, max(ref.depths))
    assert deepest >= min_depth, deepest
