"""CPU: the oracle restatements against the golden vectors produced by the real
reference (oracle/make_golden.py).  This is what pins oracle/nets_ref.py and
oracle/mcts_ref.py to /root/reference/main.py and play.py."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import mcts_ref
from oracle.nets_ref import MuZeroNetRef

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def build_ref_net(g):
    torch.manual_seed(int(g["seed"]))
    net = MuZeroNetRef(int(g["c"]), int(g["n"]) ** 2 + 1).eval()
    with torch.no_grad():
        net.prediction.pass_logit.fill_(float(g["pass_logit"]))
        ws = float(g["weight_scale"]) if "weight_scale" in g else 1.0
        es = float(g["emb_scale"]) if "emb_scale" in g else 1.0
        if ws != 1.0 or es != 1.0:               # the *_scaled fixtures (oracle/make_golden.py:make_nets)
            for p in net.parameters():
                p.mul_(ws)
            net.dynamics.action_embedding.weight.mul_(es)
    return net


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "nets_*.npz"))), ids=os.path.basename)
def test_nets_oracle_matches_reference(path):
    g = np.load(path)
    net = build_ref_net(g)
    sd = net.state_dict()
    # same construction order => same random-init weights as the reference's MuZeroNet
    assert float(sum(v.double().sum() for v in sd.values())) == pytest.approx(float(g["weight_sum"]), abs=1e-9)
    assert float(sum(v.double().abs().sum() for v in sd.values())) == pytest.approx(float(g["weight_abs_sum"]), abs=1e-9)
    with torch.no_grad():
        latent, value, logits = net.initial_inference(torch.from_numpy(g["obs"]))
        nxt, reward, value2, logits2 = net.recurrent_inference(latent, torch.from_numpy(g["actions"]))
    for got, key in ((latent, "latent"), (value, "value"), (logits, "logits"), (nxt, "next_latent"),
                     (reward, "reward"), (value2, "value2"), (logits2, "logits2")):
        np.testing.assert_allclose(got.numpy(), g[key], rtol=1e-5, atol=1e-6, err_msg=key)


def run_oracle_on_fixture(g):
    soft, rewards, values = g["soft"], g["rewards"], g["values"]
    calls = [0]

    def initial(obs):
        return None, np.float32(0), g["init_soft"]

    def recurrent(latent, action):
        i = calls[0]
        calls[0] += 1
        return None, rewards[i], values[i], soft[i]

    c = float(g["c_puct"])
    c_puct = int(c) if bool(g["c_puct_is_int"]) else c
    run = mcts_ref.run_mcts_selfplay if str(g["variant"]) == "self_play" else mcts_ref.run_mcts
    tree, mask = run(g["obs"], initial, recurrent, int(g["sims"]), g["noise"], g["stream"],
                     c_puct=c_puct, discount=float(g["discount"]), dirichlet_epsilon=float(g["dirichlet_epsilon"]),
                     pass_epsilon=float(g["pass_epsilon"]))
    return tree, mask


MAIN_LIKE = sorted(glob.glob(os.path.join(GOLDEN, "mcts_*.npz")))


@pytest.mark.parametrize("path", MAIN_LIKE, ids=os.path.basename)
def test_mcts_oracle_bit_exact_vs_reference(path):
    g = np.load(path)
    tree, mask = run_oracle_on_fixture(g)
    ch = tree.child[0]
    visits = np.array([tree.visit[int(c)] if c >= 0 else 0 for c in ch])
    assert np.array_equal(np.asarray(tree.prior[0], dtype=np.float64), g["root_prior"])   # bit-exact float64
    assert np.array_equal(visits, g["visit"])
    assert tree.visit[0] == int(g["root_visit"])
    assert float(tree.value_sum[0]) == float(g["root_value_sum"])
    assert np.array_equal(tree.root_child_value_sums(), g["child_value_sum"])
    assert np.array_equal(np.array(tree.depths), g["sel_depth"][g["sel_action"] >= 0])
    assert tree.dropped == int((g["sel_action"] < 0).sum())


def test_choose_action_matches_main():
    vc = np.array([0, 3, 7, 7, 1])
    mask = np.array([1, 1, 1, 1, 0.05])
    a, pol = mcts_ref.choose_action(vc, mask)
    assert a == 2 and pol.sum() == pytest.approx(1.0)
    a, pol = mcts_ref.choose_action(np.zeros(5, dtype=int), np.array([0, 1, 0, 1, 0.05]), fallback_u32=4)
    assert a == 3 and pol[3] == 1.0
