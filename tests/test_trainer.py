"""batched trainer and replay buffers against the REAL reference.

tests/golden/train_*.npz hold the outcome of the unmodified `MuZeroAgent.train` (main.py:381-501) on a
one-trajectory buffer (loss, weights before and after its clipped Adam step); replay_sampling.npz the draw order
of the real buffers (main.py:158-244) under a seeded numpy RNG.  Tolerances: losses 1e-5 relative (fp32, the
batched forward may pick another convolution algorithm), weight updates 2e-6 absolute (one Adam step of 1e-4).
"""
import glob
import os
import pickle
import random

import numpy as np
import pytest
import torch

from oracle import nets_ref, train_ref

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
TRAIN = sorted(glob.glob(os.path.join(GOLDEN, "train_*.npz")))


def load_case(path):
    z = np.load(path)
    n, c = int(z["n"]), int(z["c"])
    T = len(z["actions"])
    traj = {"observations": [z["observations"][i] for i in range(T + 2)], "actions": [int(a) for a in z["actions"]],
            "rewards": [float(r) for r in z["rewards"]], "policies": [z["policies"][i] for i in range(T)]}
    w0 = {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("w0.")}
    w1 = {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("w1.")}
    return z, n, c, traj, w0, w1


@pytest.mark.parametrize("path", TRAIN, ids=[os.path.basename(p)[6:-4] for p in TRAIN])
def test_oracle_loss_matches_reference(path):
    z, n, c, traj, w0, _ = load_case(path)
    net = nets_ref.MuZeroNetRef(c, n * n + 1)
    net.load_state_dict(w0)
    loss, parts = train_ref.trajectory_loss(net, traj, int(z["start_index"]), n * n + 1, int(z["unroll_steps"]),
                                            float(z["discount"]))
    assert abs(loss.item() - float(z["loss"])) <= 1e-5 * abs(float(z["loss"]))
    assert abs(sum(parts.values()) - float(z["loss"])) <= 1e-5 * abs(float(z["loss"]))


@pytest.mark.parametrize("path", TRAIN, ids=[os.path.basename(p)[6:-4] for p in TRAIN])
def test_batch_of_one_is_the_reference_step(path):
    """batch_size 1: same loss, and the same weights after zero_grad / backward / clip 1.0 / Adam(1e-4)."""
    from muzero_go_b200.networks import MuZeroNet
    from muzero_go_b200.trainer import BatchedTrainer, TrainConfig, build_batch
    z, n, c, traj, w0, w1 = load_case(path)
    A = n * n + 1
    net = MuZeroNet(c, A)
    net.load_state_dict(w0)
    tr = BatchedTrainer(net, A, TrainConfig(unroll_steps=int(z["unroll_steps"]), discount=float(z["discount"])))
    batch = build_batch(net, [traj], [int(z["start_index"])], tr.cfg, A, torch.device("cpu"))
    per_traj, parts = tr.train_step(batch)
    assert abs(per_traj.item() - float(z["loss"])) <= 1e-5 * abs(float(z["loss"]))
    sd = net.state_dict()
    moved = 0.0
    for k in w1:
        assert torch.allclose(sd[k], w1[k], rtol=0, atol=2e-6), k
        moved = max(moved, float((w1[k] - w0[k]).abs().max()))
    assert moved > 5e-5            # the step did move the weights (Adam's first step is ~lr per element)


def test_batch_losses_are_the_per_trajectory_losses():
    """A real batch: every entry of the batched loss vector is the oracle's step_loss of that trajectory (start
    indices in the middle, at the last move and with the unroll running past the end of the game)."""
    from muzero_go_b200.networks import MuZeroNet
    from muzero_go_b200.trainer import BatchedTrainer, build_batch
    cases = [load_case(p) for p in TRAIN if "N5_C16" in p]
    _, n, c, _, w0, _ = cases[0]
    A = n * n + 1
    trajs = [cs[3] for cs in cases] * 2
    starts = [4, 9, len(trajs[0]["actions"]) - 1, 0]
    ref = nets_ref.MuZeroNetRef(c, A)
    ref.load_state_dict(w0)
    want = [train_ref.trajectory_loss(ref, t, s, A)[0].item() for t, s in zip(trajs, starts)]
    net = MuZeroNet(c, A)
    net.load_state_dict(w0)
    tr = BatchedTrainer(net, A)
    per_traj, parts = tr.losses(build_batch(net, trajs, starts, tr.cfg, A, torch.device("cpu")))
    got = per_traj.detach().numpy()
    assert np.allclose(got, want, rtol=1e-5, atol=0), (got, want)
    assert np.allclose(sum(v.detach().numpy() for v in parts.values()), got, rtol=1e-6)


def test_replay_sampling_matches_reference():
    from muzero_go_b200.replay import MultiVersionReplayBuffer
    z = np.load(os.path.join(GOLDEN, "replay_sampling.npz"))
    np.random.seed(int(z["seed"]))
    rb = MultiVersionReplayBuffer(5, 2, True)
    log = []

    def draw(k):
        s, m = rb.sample(k)
        log.append([x["id"] for x in s] + [b * 100 + l for b, l in m])
        return m
    for i in range(7):
        rb.add({"id": i})
    rb.update_priorities(draw(3), [4.0, 0.5, 2.0])
    rb.add_version()
    for i in range(7, 10):
        rb.add({"id": i})
    for _ in range(3):
        rb.update_priorities(draw(4), [0.25] * 4)
    rb.add_version()
    rb.add({"id": 10})
    draw(2)
    for got, want in zip(log, z["log"]):
        assert got == [int(v) for v in want if v >= 0]
    assert rb.sample(50) == ([], [])                       # main.py:215-216
    plain = MultiVersionReplayBuffer(3, 1, False)
    for i in range(5):
        plain.add(i)
    random.seed(0)
    s, m = plain.sample(2)
    assert m is None and set(s) <= {2, 3, 4} and len(plain) == 3


def test_train_loop_priorities_and_schedule(tmp_path):
    """MuZeroAgent.train's bookkeeping (main.py:383-386, :495-499): None until a batch fits, the batch-average
    loss written back as the sampled trajectories' priority, one scheduler step per call."""
    from muzero_go_b200.networks import MuZeroNet
    from muzero_go_b200.replay import MultiVersionReplayBuffer, calculate_returns, load_episodes, save_episodes
    from muzero_go_b200.trainer import BatchedTrainer, TrainConfig
    _, n, c, traj, w0, _ = load_case(TRAIN[0])
    A = n * n + 1
    net = MuZeroNet(c, A)
    net.load_state_dict(w0)
    tr = BatchedTrainer(net, A, TrainConfig(lr_step_size=2, lr_gamma=0.5))
    rb = MultiVersionReplayBuffer(8, 1, True)
    rb.add(traj)
    assert tr.train(rb, 2) is None
    rb.add(dict(traj))
    np.random.seed(1)
    l1 = tr.train(rb, 2, rng=random.Random(3))
    l2 = tr.train(rb, 2, rng=random.Random(3))
    assert rb.buffers[0].priorities == [l2, l2] and np.isfinite(l1) and l2 < l1      # same samples: the step helped
    assert abs(tr.optimizer.param_groups[0]["lr"] - 0.5e-4) < 1e-12
    # the trajectory file of self_play.py:561-583
    ep = dict(traj, values=[0.0] * len(traj["actions"]), final_reward=traj["rewards"][-1])
    path = os.path.join(tmp_path, "batch.pkl")
    assert save_episodes(path, [ep], discount=0.5) == 1
    rec = load_episodes(path)[0]
    assert sorted(rec) == ["actions", "final_reward", "observations", "policies", "returns", "rewards", "values"]
    T = len(traj["actions"])
    assert len(rec["rewards"]) == T and len(rec["returns"]) == T
    assert rec["returns"][-1] == rec["rewards"][-1] + 0.5 * rec["final_reward"]
    assert calculate_returns([1.0, 2.0], 4.0, 0.5) == [1.0 + 0.5 * (2.0 + 0.5 * 4.0), 2.0 + 0.5 * 4.0]
    with open(path, "rb") as f:
        assert isinstance(pickle.load(f), list)


def load_batch_case(path):
    z = np.load(path)
    B = int(z["batch"])
    trajs = []
    for i in range(B):
        T = len(z[f"t{i}.actions"])
        trajs.append({"observations": [z[f"t{i}.observations"][k] for k in range(T + 2)],
                      "actions": [int(a) for a in z[f"t{i}.actions"]], "rewards": [float(r) for r in z[f"t{i}.rewards"]],
                      "policies": [z[f"t{i}.policies"][k] for k in range(T)]})
    w0 = {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("w0.")}
    w1 = {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("w1.")}
    return z, trajs, w0, w1


def test_reference_order_reproduces_a_reference_batch():
    """tests/golden/trainbatch_*.npz: the REAL MuZeroAgent.train with batch_size 3 -- zero_grad once, then per trajectory
    backward on accumulating gradients, clip, Adam step, bootstraps with the weights as updated so far
    (main.py:391-481).  `train_reference_order` reproduces the average loss and the weights after the three steps;
    the default one-step-per-batch `train_step` on the same batch does NOT land on those weights (the documented
    difference)."""
    from muzero_go_b200.networks import MuZeroNet
    from muzero_go_b200.trainer import BatchedTrainer, TrainConfig, build_batch
    z, trajs, w0, w1 = load_batch_case(os.path.join(GOLDEN, "trainbatch_N5_C16_B3.npz"))
    n, c = int(z["n"]), int(z["c"])
    A = n * n + 1
    order, starts = [int(i) for i in z["order"]], [int(s) for s in z["starts"]]
    cfg = TrainConfig(unroll_steps=int(z["unroll_steps"]), discount=float(z["discount"]))
    net = MuZeroNet(c, A)
    net.load_state_dict(w0)
    tr = BatchedTrainer(net, A, cfg)
    per_traj = tr.train_reference_order([trajs[i] for i in order], starts)
    assert abs(float(per_traj.sum()) / len(order) - float(z["loss"])) <= 1e-5 * abs(float(z["loss"]))
    sd = net.state_dict()
    for k in w1:
        assert torch.allclose(sd[k], w1[k], rtol=0, atol=4e-6), k
    net2 = MuZeroNet(c, A)
    net2.load_state_dict(w0)
    tr2 = BatchedTrainer(net2, A, cfg)
    tr2.train_step(build_batch(net2, [trajs[i] for i in order], starts, cfg, A, torch.device("cpu")))
    far = max(float((net2.state_dict()[k] - w1[k]).abs().max()) for k in w1)
    assert far > 5e-5            # one step of 1e-4 against three
