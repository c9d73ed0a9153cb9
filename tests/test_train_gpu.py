"""SURVEY 8(f) ranks 1-2 on the GPU: self-play rows -> reference-schema episodes -> replay -> batched training
step -> self-play with the updated weights."""
import numpy as np
import pytest
import torch

from oracle import gymgo

pytestmark = pytest.mark.gpu


def test_selfplay_to_training_loop():
    import muzero_go_b200 as mz
    n, c, G, S, T = 5, 16, 24, 12, 40
    A = n * n + 1
    torch.manual_seed(4)
    net = mz.MuZeroNet(c, A).cuda()
    eng = mz.SelfPlayEngine(net, n, G, S, search=mz.SearchConfig(seed=5), traj_capacity=T, restart_finished=True)
    eng.run(T)
    torch.cuda.synchronize()
    episodes, carry = mz.episodes_from_rows(eng.trajectory_rows(0, T), n)
    assert len(episodes) >= G // 2 and len(carry) == G
    for ep in episodes[:8]:
        L = len(ep["actions"])
        assert len(ep["observations"]) == L + 2 and len(ep["rewards"]) == L + 1 and len(ep["policies"]) == L
        st = gymgo.init_state(n)
        for t in range(L):                       # the record is a legal game of the oracle rules, state by state
            assert np.array_equal(ep["observations"][t], st), t
            assert abs(ep["policies"][t].sum() - 1.0) < 1e-5 and ep["policies"][t][ep["actions"][t]] > 0
            st = gymgo.next_state(st, ep["actions"][t])
        assert np.array_equal(ep["observations"][L], st) and np.array_equal(ep["observations"][L + 1], st)
        if gymgo.game_ended(st):
            assert ep["rewards"][-1] == float(gymgo.winning(st))
    rb = mz.MultiVersionReplayBuffer(64, 1, True)
    for ep in episodes:
        rb.add(ep)
    tr = mz.BatchedTrainer(net, A)
    before = [p.detach().clone() for p in net.parameters()]
    np.random.seed(0)
    losses = [tr.train(rb, 8) for _ in range(3)]
    assert all(l is not None and np.isfinite(l) for l in losses)
    assert any(not torch.equal(a, b) for a, b in zip(before, net.parameters()))
    # the engine picks the new weights up (same seed, different network => the games differ somewhere)
    eng2 = mz.SelfPlayEngine(net, n, G, S, search=mz.SearchConfig(seed=5), traj_capacity=T, restart_finished=True)
    eng2.run(T)
    torch.cuda.synchronize()
    assert eng2.read_counters()["moves"] == G * T
    assert not torch.equal(eng2.traj_policy, eng.traj_policy)


def test_graph_replayed_step_equals_eager_step():
    """The CUDA-graph-replayed training step (capture once per batch shape, warm-up undone) moves the weights like
    the eager step does: three steps on fresh batches, same losses and weights to 1e-6; the learning-rate
    schedule acts on the captured optimizer."""
    import copy
    import muzero_go_b200 as mz
    from muzero_go_b200.trainer import Batch, BatchedTrainer, TrainConfig
    n, c, B, K = 6, 32, 16, 10
    A = n * n + 1
    dev = torch.device("cuda")
    torch.manual_seed(2)
    net_a = mz.MuZeroNet(c, A).to(dev)
    net_b = copy.deepcopy(net_a)
    cfg = TrainConfig(lr_step_size=2, lr_gamma=0.5)
    eager, graphed = BatchedTrainer(net_a, A, cfg, cuda_graph=False), BatchedTrainer(net_b, A, cfg, cuda_graph=True)
    g = torch.Generator(device="cpu").manual_seed(7)

    def batch():
        pol = torch.rand(B, K + 1, A, generator=g)
        return Batch((torch.rand(B, 6, n, n, generator=g) < 0.3).float().to(dev), torch.randint(0, A, (B, K), generator=g).to(dev),
                     (torch.rand(B, K, generator=g) - 0.5).to(dev), (torch.rand(B, K + 1, generator=g) * 2 - 1).to(dev),
                     (pol / pol.sum(-1, keepdim=True)).to(dev))
    w0 = [p.detach().clone() for p in net_b.parameters()]
    for step in range(4):
        b = batch()
        la, _ = eager.train_step(b)
        if step == 0:
            graphed._graphs[tuple(tuple(getattr(b, f).shape) for f in ("observations", "actions", "target_reward", "target_value", "target_policy"))] = graphed._capture(b)
            assert all(torch.equal(p, q) for p, q in zip(net_b.parameters(), w0))      # capturing changed nothing
        lb, _ = graphed.train_step(b)
        assert torch.allclose(la, lb, rtol=1e-5, atol=1e-6), step
        for p, q in zip(net_a.parameters(), net_b.parameters()):
            assert torch.allclose(p, q, rtol=0, atol=2e-6), step
        eager._schedule()
        graphed._schedule()
        assert abs(eager.lr - graphed.lr) < 1e-12
    assert abs(graphed.lr - 0.25e-4) < 1e-12 and len(graphed._graphs) == 1


def test_device_replay_batch_equals_host_batch():
    """DeviceReplay (games as padded device tensors, gathers, rules-kernel planes) builds the batch that
    trainer.build_batch builds from the reference-schema dicts of the same games and start indices: actions,
    reward targets and policies exactly, value targets to 1e-6 (their bootstrap forward runs in another batch)."""
    import muzero_go_b200 as mz
    from muzero_go_b200.device_replay import DeviceReplay, train_on_device
    from muzero_go_b200.trainer import TrainConfig, build_batch
    n, c, G, S, T = 5, 16, 20, 10, 38
    A = n * n + 1
    torch.manual_seed(6)
    net = mz.MuZeroNet(c, A).cuda()
    eng = mz.SelfPlayEngine(net, n, G, S, search=mz.SearchConfig(seed=11), traj_capacity=T, restart_finished=True)
    eng.run(T)
    torch.cuda.synchronize()
    rows = eng.trajectory_rows(0, T)
    # fed in two pieces: games that straddle the cut are completed by the second call
    dr = DeviceReplay(64, n, device="cuda")
    dr.add_rows({k: v[:17] for k, v in rows.items()})
    dr.add_rows({k: v[17:] for k, v in rows.items()})
    episodes, _ = mz.episodes_from_rows(rows, n)
    assert len(dr) == len(episodes) >= G
    # match the stored games to the host episodes by content (the ring stores them in completion order)
    key = lambda acts: tuple(int(a) for a in acts)
    by_actions = {key(ep["actions"]): ep for ep in episodes}
    slots = torch.arange(len(dr), device="cuda")
    lengths = dr.length[slots].tolist()
    eps = [by_actions[key(dr.actions[e, :L].tolist())] for e, L in zip(slots.tolist(), lengths)]
    cfg = TrainConfig(unroll_steps=10)
    starts = [0, lengths[1] - 1] + [(7 * i) % L for i, L in enumerate(lengths[2:])]     # first move, last move, others
    starts_t = torch.tensor(starts, device="cuda")
    got = dr.build_batch(net, slots, starts_t, cfg)
    want = build_batch(net, eps, starts, cfg, A, torch.device("cuda"))
    assert torch.equal(got.observations, want.observations) and torch.equal(got.actions, want.actions)
    assert torch.equal(got.target_reward, want.target_reward) and torch.equal(got.target_policy, want.target_policy)
    assert torch.allclose(got.target_value, want.target_value, rtol=0, atol=1e-6)
    # sampling + a training step entirely on the device
    tr = mz.BatchedTrainer(net, A, cfg)
    gen = torch.Generator(device="cuda").manual_seed(3)
    drawn = dr.sample(8, gen)
    assert drawn is not None and len(set(drawn[0].tolist())) == 8 and bool((drawn[1] < dr.length[drawn[0]]).all())
    assert dr.sample(10 ** 6) is None
    losses = [train_on_device(tr, dr, 8, gen) for _ in range(3)]
    assert all(np.isfinite(l) for l in losses)
    assert sorted(set(dr.priority[: len(dr)].tolist())) != [1.0]          # the losses were written back as priorities


def _random_batch(g, B, K, n, A, dev):
    from muzero_go_b200.trainer import Batch
    pol = torch.rand(B, K + 1, A, generator=g)
    return Batch((torch.rand(B, 6, n, n, generator=g) < 0.3).float().to(dev), torch.randint(0, A, (B, K), generator=g).to(dev),
                 (torch.rand(B, K, generator=g) - 0.5).to(dev), (torch.rand(B, K + 1, generator=g) * 2 - 1).to(dev),
                 (pol / pol.sum(-1, keepdim=True)).to(dev))


def test_forward_follows_graph_replayed_training_steps():
    """A replayed CUDA graph rewrites the parameters without touching tensor._version: after
    several replay-only steps the libmzgo forward (prepared weight layouts, head constants, cached search graph)
    must be the forward of the CURRENT parameters -- checked against the torch forward the trainer differentiates,
    after every step, and through a SelfPlayEngine built before the steps."""
    import muzero_go_b200 as mz
    from muzero_go_b200 import trainer as T
    n, c, B, K = 6, 32, 8, 10
    A = n * n + 1
    dev = torch.device("cuda")
    torch.manual_seed(12)
    net = mz.MuZeroNet(c, A).to(dev)
    cfg = T.TrainConfig(learning_rate=3e-3)                     # large steps: stale weights show at once
    tr = T.BatchedTrainer(net, A, cfg, cuda_graph=True)
    eng = mz.SelfPlayEngine(net, n, 16, 8, search=mz.SearchConfig(seed=2), traj_capacity=4)
    g = torch.Generator(device="cpu").manual_seed(5)
    obs = (torch.rand(5, 6, n, n, generator=g) < 0.3).float().to(dev)
    act = torch.randint(0, A, (5,), generator=g).to(dev)
    prev = None
    for step in range(5):
        tr.train_step(_random_batch(g, B, K, n, A, dev))
        assert len(tr._graphs) == 1
        with torch.no_grad():
            lat_t = T.representation(net, obs)
            val_t, log_t = T.prediction(net, lat_t)
            nxt_t, rew_t = T.dynamics(net, lat_t, act)
            val2_t, log2_t = T.prediction(net, nxt_t)
        lat, val, log = net.initial_inference(obs)
        nxt, rew, val2, log2 = net.recurrent_inference(lat_t, act)
        for got, want in ((lat, lat_t), (val, val_t), (log, log_t), (nxt, nxt_t), (rew, rew_t), (val2, val2_t), (log2, log2_t)):
            assert torch.allclose(got, want, rtol=1e-4, atol=1e-5), step
        if prev is not None:
            assert not torch.allclose(lat, prev, rtol=1e-3, atol=1e-4)      # the steps do move the forward
        prev = lat.clone()
        # the engine (descriptor and search graph cached from earlier moves) searches with the new weights: its root
        # value is the network's value of the position
        eng.run(1)
        torch.cuda.synchronize()
    # root_value_net holds the initial-inference value the last move's search started from (position: traj row 0)
    gb = mz.GoBoards(eng.G, n, dev)
    gb.records.copy_(eng.traj_boards[0])
    with torch.no_grad():
        want_v, _ = T.prediction(net, T.representation(net, gb.planes(torch.float32)))
    assert torch.allclose(eng.root_value_net, want_v.flatten(), rtol=1e-4, atol=1e-5)


def test_second_batch_shape_keeps_earlier_graphs_valid():
    """capturing a graph for a second batch shape must not move Adam's moment tensors from under
    the first graph.  Alternating two shapes equals the eager trainer on the same batches."""
    import copy
    import muzero_go_b200 as mz
    from muzero_go_b200.trainer import BatchedTrainer, TrainConfig
    n, c, K = 5, 16, 10
    A = n * n + 1
    dev = torch.device("cuda")
    torch.manual_seed(3)
    net_a = mz.MuZeroNet(c, A).to(dev)
    net_b = copy.deepcopy(net_a)
    cfg = TrainConfig()
    eager, graphed = BatchedTrainer(net_a, A, cfg, cuda_graph=False), BatchedTrainer(net_b, A, cfg, cuda_graph=True)
    g = torch.Generator(device="cpu").manual_seed(9)
    for step, B in enumerate([8, 8, 4, 8, 4, 4, 8]):
        b = _random_batch(g, B, K, n, A, dev)
        la, _ = eager.train_step(b)
        lb, _ = graphed.train_step(b)
        assert torch.allclose(la, lb, rtol=1e-5, atol=1e-6), step
        for p, q in zip(net_a.parameters(), net_b.parameters()):
            assert torch.allclose(p, q, rtol=0, atol=3e-6), step
    assert len(graphed._graphs) == 2


@pytest.mark.parametrize("B,cin,cout,n", [(3, 6, 64, 6), (5, 64, 64, 9), (2, 64, 128, 19), (4, 128, 128, 6), (3, 16, 16, 5),
                                          (20, 128, 128, 9), (1, 96, 96, 13)])
def test_native_conv_kernels_match_float64_autograd(B, cin, cout, n):
    """csrc/train.cu (forward, data gradient, weight / bias gradient with ReLU's mask folded in) through the C ABI
    against torch autograd in float64: every output within 1e-5 of the tensor's scale (float32 accumulation order)."""
    from muzero_go_b200.train_ops import conv3x3_relu
    g = torch.Generator().manual_seed(B * 1000 + cin + n)
    x = torch.randn(B, cin, n, n, generator=g)
    w = torch.randn(cout, cin, 3, 3, generator=g) / (3.0 * cin ** 0.5)
    b = torch.randn(cout, generator=g) * 0.1
    dy = torch.randn(B, cout, n, n, generator=g)
    need_dx = cin % 8 == 0            # the observation planes (6 channels) never need a gradient (main.py:397)
    xr, wr, br = (t.double().requires_grad_(True) for t in (x, w, b))
    yr = torch.relu(torch.nn.functional.conv2d(xr, wr, br, padding=1))
    yr.backward(dy.double())
    xg = x.cuda().requires_grad_(need_dx)
    wg, bg = (t.cuda().requires_grad_(True) for t in (w, b))
    y = conv3x3_relu(xg, wg, bg)
    y.backward(dy.cuda())
    checks = [(y, yr, "y"), (wg.grad, wr.grad, "dw"), (bg.grad, br.grad, "db")] + ([(xg.grad, xr.grad, "dx")] if need_dx else [])
    for got, want, name in checks:
        err = float((got.detach().cpu().double() - want.detach()).abs().max() / (want.detach().abs().max() + 1e-12))
        assert err < 1e-5, (name, err)
    assert float((y == 0).float().mean()) > 0.2          # the mask was exercised


@pytest.mark.parametrize("name", ["N5_C16_mid", "N5_C16_tail", "N6_C32_head"])
def test_reference_training_step_through_the_native_kernels(name):
    """tests/golden/train_*.npz -- the REAL MuZeroAgent.train on a one-trajectory buffer -- with the network on the
    GPU, where every 3x3 convolution runs forward and backward in csrc/train.cu: same loss (1e-5 relative) and same
    weights after the clipped Adam step (2e-6 absolute), eager and as a replayed CUDA graph."""
    import os
    from muzero_go_b200.networks import MuZeroNet
    from muzero_go_b200.trainer import BatchedTrainer, TrainConfig, build_batch
    from tests.test_trainer import GOLDEN, load_case
    z, n, c, traj, w0, w1 = load_case(os.path.join(GOLDEN, f"train_{name}.npz"))
    A = n * n + 1
    for graph in (False, True):
        net = MuZeroNet(c, A)
        net.load_state_dict(w0)
        net = net.cuda()
        tr = BatchedTrainer(net, A, TrainConfig(unroll_steps=int(z["unroll_steps"]), discount=float(z["discount"])),
                            cuda_graph=graph)
        batch = build_batch(net, [traj], [int(z["start_index"])], tr.cfg, A, torch.device("cuda"))
        per_traj, _ = tr.train_step(batch)
        assert abs(per_traj.item() - float(z["loss"])) <= 1e-5 * abs(float(z["loss"])), graph
        sd = net.state_dict()
        for k in w1:
            assert torch.allclose(sd[k].cpu(), w1[k], rtol=0, atol=2e-6), (k, graph)


def test_reference_order_batch_on_the_gpu():
    """tests/golden/trainbatch_N5_C16_B3.npz (the REAL MuZeroAgent.train, batch of 3: three clipped Adam steps on
    accumulating gradients) reproduced on the GPU through the native convolution kernels."""
    import os
    from muzero_go_b200.networks import MuZeroNet
    from muzero_go_b200.trainer import BatchedTrainer, TrainConfig
    from tests.test_trainer import GOLDEN, load_batch_case
    z, trajs, w0, w1 = load_batch_case(os.path.join(GOLDEN, "trainbatch_N5_C16_B3.npz"))
    n, c = int(z["n"]), int(z["c"])
    A = n * n + 1
    order, starts = [int(i) for i in z["order"]], [int(s) for s in z["starts"]]
    net = MuZeroNet(c, A)
    net.load_state_dict(w0)
    net = net.cuda()
    tr = BatchedTrainer(net, A, TrainConfig(unroll_steps=int(z["unroll_steps"]), discount=float(z["discount"])))
    per_traj = tr.train_reference_order([trajs[i] for i in order], starts)
    assert abs(float(per_traj.sum()) / len(order) - float(z["loss"])) <= 1e-5 * abs(float(z["loss"]))
    sd = net.state_dict()
    for k in w1:
        assert torch.allclose(sd[k].cpu(), w1[k], rtol=0, atol=4e-6), k
