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
