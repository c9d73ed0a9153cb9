"""GPU: the SoA MCTS kernels (through the C ABI) against (a) the golden fixtures produced by the real
reference MCTS (oracle/make_golden.py) and (b) the oracle restatement on fresh random tables, with
identical float32 network outputs injected on both sides.  Everything is bit-exact."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import mcts_ref

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
MAIN_LIKE = sorted(glob.glob(os.path.join(GOLDEN, "mcts_*.npz")))   # main.py, play.py and self_play.py fixtures


DRIVERS = ["standalone", "step"]


def run_gpu_search(n, obs_list, init_soft, soft, rewards, values, noise, streams, sims, cfg, driver="standalone"):
    """G independent searches in lock step; soft/rewards/values are per-game call-indexed tables
    ([G][sims][A], [G][sims], [G][sims]), consumed in the order each game's simulations complete.

    driver "standalone": mzgo_tree_select + mzgo_tree_expand_backup (k_tree_select / k_tree_expand_backup).
    driver "step": mzgo_tree_step(do_backup, do_select) -- k_tree_step<32> (A <= 96) / k_tree_step<128>, the kernel
    mzgo_search_run launches once per simulation (and the code of the network kernel's tree phase), with the table
    row of the leaf it selected written into the output buffers before the step that backs it up."""
    import muzero_go_b200 as mz
    G, A = len(obs_list), n * n + 1
    dev = torch.device("cuda")
    boards = mz.GoBoards(G, n).load_planes(np.stack(obs_list).astype(np.float32))
    tree = mz.Tree(G, A, sims + 1)
    tree.reset_roots(cfg, torch.from_numpy(np.stack(init_soft)).to(dev), boards, n, noise=np.stack(noise))
    stream_t = mz.Tree.stream_tensor(np.stack(streams), dev)
    soft_t = torch.from_numpy(np.stack(soft)).to(dev)            # [G][sims][A]
    rew_t = torch.from_numpy(np.stack(rewards)).to(dev)
    val_t = torch.from_numpy(np.stack(values)).to(dev)
    calls = torch.zeros(G, dtype=torch.int64, device=dev)
    gi = torch.arange(G, device=dev)
    depths = [[] for _ in range(G)]
    zs, zr = torch.zeros((G, A), device=dev), torch.zeros(G, device=dev)
    if driver == "step":
        tree.step(cfg, zs, zr, zr, stream_t, do_backup=False, do_select=True)
    for s in range(sims):
        if driver == "standalone":
            tree.select(cfg, stream_t)
        live = tree.leaf_action >= 0
        idx = calls.clamp(max=sims - 1)
        d = tree.leaf_depth.cpu().numpy()
        for g in np.nonzero(live.cpu().numpy())[0]:
            depths[g].append(int(d[g]))
        out = (soft_t[gi, idx].contiguous(), rew_t[gi, idx].contiguous(), val_t[gi, idx].contiguous())
        if driver == "standalone":
            tree.expand_backup(cfg, *out)
        else:
            tree.step(cfg, *out, stream_t, do_backup=True, do_select=s + 1 < sims)
        calls += live.long()
    return tree, depths


def cfg_from_fixture(g):
    import muzero_go_b200 as mz
    c = float(g["c_puct"])
    return mz.SearchConfig(c_puct=int(c) if bool(g["c_puct_is_int"]) else c, discount=float(g["discount"]),
                           dirichlet_epsilon=float(g["dirichlet_epsilon"]), pass_epsilon=float(g["pass_epsilon"]),
                           variant="self_play" if str(g["variant"]) == "self_play" else "main")


@pytest.mark.parametrize("driver", DRIVERS)
@pytest.mark.parametrize("path", MAIN_LIKE, ids=os.path.basename)
def test_visit_counts_bit_exact_vs_reference_fixture(path, driver):
    g = np.load(path)
    n, sims = int(g["n"]), int(g["sims"])
    cfg = cfg_from_fixture(g)
    # two copies of the same game in one batch: lock-step batching must not couple games
    tree, depths = run_gpu_search(n, [g["obs"]] * 2, [g["init_soft"]] * 2, [g["soft"]] * 2, [g["rewards"]] * 2,
                                  [g["values"]] * 2, [g["noise"]] * 2, [g["stream"]] * 2, sims, cfg, driver)
    if sims >= 400 and n >= 9:
        assert max(depths[0]) >= 2          # the fixture reaches below the root's children (19x19/800: depth 2, 9x9/800: 3)
    for k in range(2):
        assert np.array_equal(tree.root_prior[k].cpu().numpy(), g["root_prior"]), "root priors (float64) differ"
        assert np.array_equal(tree.edge_visit[k, 0].cpu().numpy().astype(np.int64), g["visit"])
        assert np.array_equal(tree.visit_counts()[k].cpu().numpy().astype(np.int64), g["visit"] * (g["root_prior"] > 0))
        assert int(tree.root_visit[k]) == int(g["root_visit"])
        assert float(tree.root_vsum[k]) == float(g["root_value_sum"])
        assert np.array_equal(tree.edge_vsum[k, 0].cpu().numpy(), g["child_value_sum"])
        assert np.array_equal(np.array(depths[k]), g["sel_depth"][g["sel_action"] >= 0])
        assert int(tree.dropped[k]) == int((g["sel_action"] < 0).sum())


@pytest.mark.parametrize("n,sims,G,c_puct,variant", [(5, 120, 6, 2, "main"), (9, 260, 4, 2, "main"), (9, 150, 3, 10.0, "main"),
                                                     (19, 500, 2, 2, "main"), (5, 150, 6, 2.5, "self_play"),
                                                     (9, 260, 4, 2.5, "self_play"), (19, 450, 2, 2.5, "self_play")])
@pytest.mark.parametrize("driver", DRIVERS)
def test_batched_search_matches_oracle_on_random_tables(n, sims, G, c_puct, variant, driver):
    """Different positions / tables per game; also exercises games whose simulations get dropped
    (main.py) or end in a terminal leaf (self_play.py)."""
    import muzero_go_b200 as mz
    from oracle.make_golden import midgame_obs
    A = n * n + 1
    rng = np.random.default_rng(1000 + n + sims)
    cfg = mz.SearchConfig(c_puct=c_puct, discount=0.99, dirichlet_epsilon=0.25, pass_epsilon=0.05, variant=variant)
    run_ref = mcts_ref.run_mcts_selfplay if variant == "self_play" else mcts_ref.run_mcts
    obs, init_soft, soft, rew, val, noise, streams, trees = [], [], [], [], [], [], [], []
    for k in range(G):
        o = midgame_obs(n, plies=int(rng.integers(0, n * n)), seed=int(rng.integers(1 << 30)))
        if k == G - 1:           # an almost-finished position: few legal points, some simulations dropped
            o = o.copy()
            o[3] = 1
            o[3].flat[int(rng.integers(n * n))] = 0
        il = rng.normal(0, 1.5, A).astype(np.float32)
        lg = rng.normal(0, 2.0, (sims, A)).astype(np.float32)
        obs.append(o)
        init_soft.append(torch.softmax(torch.from_numpy(il).view(1, -1), dim=1).numpy()[0])
        soft.append(torch.softmax(torch.from_numpy(lg), dim=1).numpy())
        rew.append(rng.normal(0, 0.1, sims).astype(np.float32))
        val.append(rng.normal(0, 0.5, sims).astype(np.float32))
        noise.append(rng.dirichlet([0.3] * A))
        streams.append(rng.integers(0, 2 ** 32, size=sims + 8, dtype=np.uint64).astype(np.uint32))
        calls = [0]

        def recurrent(latent, action, k=k, calls=calls):
            i = calls[0]
            calls[0] += 1
            return None, rew[k][i], val[k][i], soft[k][i]
        t, mask = run_ref(o, lambda ob, k=k: (None, np.float32(0), init_soft[k]), recurrent, sims, noise[k],
                          streams[k], c_puct=c_puct, discount=0.99, dirichlet_epsilon=0.25, pass_epsilon=0.05)
        trees.append((t, mask))
    tree, depths = run_gpu_search(n, obs, init_soft, soft, rew, val, noise, streams, sims, cfg, driver)
    if n == 19:
        assert max(max(d) for d in depths if d) >= 2     # deeper levels of k_tree_step<128> at A = 362
    vc = tree.visit_counts().cpu().numpy()
    for k, (t, mask) in enumerate(trees):
        assert np.array_equal(tree.root_prior[k].cpu().numpy(), np.asarray(t.prior[0], dtype=np.float64))
        assert np.array_equal(vc[k], t.root_visit_counts(mask)), f"game {k}"
        assert np.array_equal(tree.edge_vsum[k, 0].cpu().numpy(), t.root_child_value_sums())
        assert int(tree.root_visit[k]) == t.visit[0] and int(tree.dropped[k]) == t.dropped
        assert depths[k] == t.depths
        assert int(tree.n_nodes[k]) == len(t.visit)
    # action choice, main.py:670-677
    action, policy, root_value = tree.choose_action(cfg)
    for k, (t, mask) in enumerate(trees):
        a_ref, pol_ref = mcts_ref.choose_action(t.root_visit_counts(mask), mask)
        if t.root_visit_counts(mask).sum() > 0:
            assert int(action[k]) == a_ref
            np.testing.assert_allclose(policy[k].cpu().numpy(), pol_ref, rtol=1e-6, atol=1e-7)
        else:
            assert mask[int(action[k])] > 0
        assert float(root_value[k]) == pytest.approx(t.q(0), rel=1e-6, abs=1e-7)


def test_idle_trees_are_left_alone():
    import muzero_go_b200 as mz
    n, G, sims = 5, 4, 20
    A = n * n + 1
    dev = torch.device("cuda")
    cfg = mz.SearchConfig()
    boards = mz.GoBoards(G, n).reset()
    tree = mz.Tree(G, A, sims + 1)
    tree.active[1] = 0
    soft = torch.full((G, A), 1.0 / A, device=dev)
    tree.reset_roots(cfg, soft, boards, n, noise=None, move_index=3)     # device Dirichlet sampler
    rp = tree.root_prior.cpu().numpy()
    assert np.all(rp[0] > 0) and abs(rp[0].sum() - 1.0) < 1e-3 and np.all(rp[1] == 0)
    z = torch.zeros(G, device=dev)
    for _ in range(sims):
        tree.select(cfg, None)                                           # device RNG for random.choice
        tree.expand_backup(cfg, soft, z, z)
    assert tree.root_visit.cpu().tolist() == [sims, 0, sims, sims]
    action, policy, _ = tree.choose_action(cfg)
    assert int(action[1]) == -1 and float(policy[1].sum()) == 0.0
    assert all(0 <= int(action[k]) < A for k in (0, 2, 3))
    # Dirichlet noise: the mixed-in part sums to epsilon and differs between games
    assert np.all(rp[[0, 2, 3]] >= 0)
    assert not np.allclose(rp[0], rp[2])


def test_fast_division_is_correctly_rounded():
    """The tree-step kernel divides with a known reciprocal (five FP64 operations, Markstein's correction);
    bit-exact visit counts need it to equal the IEEE division everywhere: 2^27 operand pairs, two divisor kinds."""
    from muzero_go_b200 import _lib
    lib = _lib.load()
    _lib.require_device()
    for seed in (1, 2):
        assert lib.mzgo_debug_fastdiv_check(seed, 1024, 256) == 0


@pytest.mark.parametrize("n,live_points", [(5, 0), (9, 0), (9, 1), (13, 0), (19, 0)])
@pytest.mark.parametrize("driver", DRIVERS)
def test_chains_of_single_live_edges_match_the_oracle(n, live_points, driver):
    """Late positions: only the pass (live_points = 0) or the pass and one point are legal, so every node of the tree
    has one or two live edges and the tree is a chain about as deep as the number of simulations.  The warp-per-game
    step walks a node whose single live edge is visited without the PUCT arithmetic (tree_step.cuh): depths, visit
    counts and value sums must still be those of the reference's loop, bit for bit."""
    import muzero_go_b200 as mz
    from oracle.make_golden import midgame_obs
    A, sims, G = n * n + 1, 64, 3
    rng = np.random.default_rng(77 + n + live_points)
    cfg = mz.SearchConfig(c_puct=2.0, discount=0.99, dirichlet_epsilon=0.25, pass_epsilon=0.05, variant="main")
    obs, init_soft, soft, rew, val, noise, streams, trees = [], [], [], [], [], [], [], []
    for k in range(G):
        o = midgame_obs(n, plies=n * n // 2, seed=int(rng.integers(1 << 30))).copy()
        o[3] = 1                                              # no board point is legal ...
        for p in rng.choice(n * n, size=live_points, replace=False):
            o[3].flat[int(p)] = 0                             # ... but these
        obs.append(o)
        init_soft.append(torch.softmax(torch.from_numpy(rng.normal(0, 1.5, A).astype(np.float32)).view(1, -1), dim=1).numpy()[0])
        soft.append(torch.softmax(torch.from_numpy(rng.normal(0, 2.0, (sims, A)).astype(np.float32)), dim=1).numpy())
        rew.append(rng.normal(0, 0.1, sims).astype(np.float32))
        val.append(rng.normal(0, 0.5, sims).astype(np.float32))
        noise.append(rng.dirichlet([0.3] * A))
        streams.append(rng.integers(0, 2 ** 32, size=sims + 8, dtype=np.uint64).astype(np.uint32))
        calls = [0]

        def recurrent(latent, action, k=k, calls=calls):
            i = calls[0]
            calls[0] += 1
            return None, rew[k][i], val[k][i], soft[k][i]
        trees.append(mcts_ref.run_mcts(o, lambda ob, k=k: (None, np.float32(0), init_soft[k]), recurrent, sims, noise[k],
                                       streams[k], c_puct=2.0, discount=0.99, dirichlet_epsilon=0.25, pass_epsilon=0.05))
    tree, depths = run_gpu_search(n, obs, init_soft, soft, rew, val, noise, streams, sims, cfg, driver)
    if live_points == 0:
        assert all(max(d) >= sims - 2 for d in depths), [max(d) for d in depths]     # one node per level
    vc = tree.visit_counts().cpu().numpy()
    for k, (t, mask) in enumerate(trees):
        assert int(mask.sum() > 0) and int((mask > 0).sum()) == live_points + 1
        assert depths[k] == t.depths, f"game {k}"
        assert np.array_equal(vc[k], t.root_visit_counts(mask)), f"game {k}"
        assert np.array_equal(tree.edge_vsum[k, 0].cpu().numpy(), t.root_child_value_sums())
        assert int(tree.root_visit[k]) == t.visit[0] and int(tree.dropped[k]) == t.dropped
        assert int(tree.n_nodes[k]) == len(t.visit)
