"""Generate tests/golden/*.npz from the REAL reference (TEST INFRASTRUCTURE).

Run in the build container only (needs /root/reference):

    python -m oracle.make_golden

Every fixture is produced by the unmodified reference classes imported through
oracle/ref_loader.py; the only interventions are the documented RNG patches
(np.random.dirichlet -> supplied noise, random.choice -> stream-indexed pick)
and table/recording network objects standing in for `muzero_net`.

Fixtures
  nets_N{n}_C{c}.npz   reference MuZeroNet (main.py:127-142) outputs on fixed inputs
  mcts_*.npz           reference MCTS visit counts with injected network outputs
"""
import os
import random
import sys

import numpy as np
import torch

from . import gymgo, ref_loader

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def midgame_obs(n, plies, seed):
    rng = np.random.default_rng(seed)
    st = gymgo.init_state(n)
    for _ in range(plies):
        if gymgo.game_ended(st):
            break
        v = np.nonzero(gymgo.valid_moves(st))[0]
        a = int(v[rng.integers(len(v) - 1)]) if len(v) > 1 else int(v[0])
        st = gymgo.next_state(st, a)
    return st


# --------------------------------------------------------------------------- nets
def make_nets(n, c, batch, seed, weight_scale=1.0, emb_scale=1.0, tag=""):
    """weight_scale / emb_scale: every parameter times weight_scale, then the action embedding times emb_scale
    (after the pass_logit fill).  The default init leaves latents and head outputs close to zero; the scaled
    fixtures pin the kernels where activations are O(1)."""
    main = ref_loader.configure(ref_loader.load("main"), n)
    A = n * n + 1
    torch.manual_seed(seed)
    net = main.MuZeroNet(c, A).eval()
    # move biases / pass_logit / fc weights off their degenerate init so every term is exercised
    with torch.no_grad():
        net.prediction.pass_logit.fill_(0.37)
        if weight_scale != 1.0 or emb_scale != 1.0:
            for p in net.parameters():
                p.mul_(weight_scale)
            net.dynamics.action_embedding.weight.mul_(emb_scale)
    obs = np.stack([midgame_obs(n, 5 + 7 * i, seed + i) for i in range(batch)]).astype(np.float32)
    rng = np.random.default_rng(seed)
    actions = rng.integers(0, A, size=batch).astype(np.int64)
    actions[0] = A - 1            # pass
    actions[-1] = 0               # corner
    with torch.no_grad():
        latent, value, logits = net.initial_inference(torch.from_numpy(obs))
        nxt, reward, value2, logits2 = net.recurrent_inference(latent, torch.from_numpy(actions))
        nxt2, reward3, value3, logits3 = net.recurrent_inference(nxt, torch.from_numpy(actions[::-1].copy()))
    sd = net.state_dict()
    checksum = float(sum(v.double().sum() for v in sd.values()))
    abs_checksum = float(sum(v.double().abs().sum() for v in sd.values()))
    np.savez_compressed(
        os.path.join(OUT, f"nets_N{n}_C{c}{tag}.npz"),
        n=n, c=c, seed=seed, pass_logit=0.37, weight_scale=weight_scale, emb_scale=emb_scale, obs=obs, actions=actions,
        weight_sum=checksum, weight_abs_sum=abs_checksum,
        latent=latent.numpy(), value=value.numpy(), logits=logits.numpy(),
        next_latent=nxt.numpy(), reward=reward.numpy(), value2=value2.numpy(), logits2=logits2.numpy(),
        reward3=reward3.numpy(), value3=value3.numpy(), logits3=logits3.numpy(),
        next_latent2_sum=float(nxt2.double().sum()))
    print("nets", n, c, "weights", checksum)


# --------------------------------------------------------------------------- mcts
class TableNet:
    """Stands in for `muzero_net`: returns pre-drawn outputs by call index."""

    def __init__(self, init_logits, logits, rewards, values):
        self.init_logits = torch.from_numpy(init_logits).view(1, -1)
        self.logits = torch.from_numpy(logits)
        self.rewards = torch.from_numpy(rewards)
        self.values = torch.from_numpy(values)
        self.calls = 0

    def eval(self):
        return self

    def initial_inference(self, obs):
        return torch.zeros(1, 1), torch.zeros(1, 1), self.init_logits

    def recurrent_inference(self, latent, action):
        i = self.calls
        self.calls += 1
        return torch.zeros(1, 1), self.rewards[i].view(1, 1), self.values[i].view(1, 1), self.logits[i:i + 1]


class RecordingNet:
    """Wraps a real reference net and records what the search consumed."""

    def __init__(self, net):
        self.net = net
        self.logits, self.rewards, self.values = [], [], []
        self.init_logits = None

    def eval(self):
        self.net.eval()
        return self

    def initial_inference(self, obs):
        latent, value, logits = self.net.initial_inference(obs)
        self.init_logits = logits.detach().numpy()[0].copy()
        return latent, value, logits

    def recurrent_inference(self, latent, action):
        out = self.net.recurrent_inference(latent, action)
        self.rewards.append(np.float32(out[1].item()))
        self.values.append(np.float32(out[2].item()))
        self.logits.append(out[3].detach().numpy()[0].copy())
        return out


def run_reference_mcts(modname, n, sims, obs, net, noise, stream, c_puct=None):
    mod = ref_loader.configure(ref_loader.load(modname), n)
    A = n * n + 1
    trace_a, trace_d = [], []
    calls = [0]

    def fake_choice(seq):
        k = int(stream[calls[0]]) % len(seq)
        calls[0] += 1
        return seq[k]

    old_dir, old_choice = np.random.dirichlet, random.choice
    np.random.dirichlet = lambda alpha: np.array(noise, dtype=np.float64)
    random.choice = fake_choice
    try:
        mcts = mod.MCTS(net, A, sims) if c_puct is None else mod.MCTS(net, A, sims, c_puct=c_puct)
        orig_select = mcts.select_leaf

        def traced(*args):
            path, action = orig_select(*args)
            trace_a.append(-1 if action is None else action)
            trace_d.append(len(path))
            return path, action
        mcts.select_leaf = traced
        with torch.no_grad():
            out = mcts.run(obs)
    finally:
        np.random.dirichlet, random.choice = old_dir, old_choice
    root = out[0] if isinstance(out, tuple) else out
    visit = np.array([ch['node'].visit_count if ch['node'] is not None else 0 for ch in root.children.values()],
                     dtype=np.int64)
    vsum = np.array([float(ch['node'].value_sum) if ch['node'] is not None else 0.0
                     for ch in root.children.values()], dtype=np.float64)
    priors = np.array([float(ch['prior']) for ch in root.children.values()], dtype=np.float64)
    cfg = mod.config
    return dict(visit=visit, child_value_sum=vsum, root_prior=priors, root_visit=root.visit_count,
                root_value_sum=float(root.value_sum), sel_action=np.array(trace_a, dtype=np.int32),
                sel_depth=np.array(trace_d, dtype=np.int32), c_puct=float(mcts.c_puct),
                c_puct_is_int=isinstance(mcts.c_puct, int), discount=cfg.discount,
                dirichlet_epsilon=cfg.dirichlet_epsilon, pass_epsilon=cfg.pass_epsilon)


def make_mcts(tag, modname, n, sims, seed, mode, plies, c=16, c_puct=None, logit_scale=2.0, value_scale=0.5):
    A = n * n + 1
    rng = np.random.default_rng(seed)
    obs = midgame_obs(n, plies, seed)
    noise = rng.dirichlet([0.3] * A)
    stream = rng.integers(0, 2 ** 32, size=sims + 8, dtype=np.uint64).astype(np.uint32)
    if mode == "table":
        init_logits = rng.normal(0, 1.5, A).astype(np.float32)
        logits = rng.normal(0, logit_scale, (sims, A)).astype(np.float32)
        rewards = rng.normal(0, 0.1, sims).astype(np.float32)
        values = rng.normal(0, value_scale, sims).astype(np.float32)
        net = TableNet(init_logits, logits, rewards, values)
        res = run_reference_mcts(modname, n, sims, obs, net, noise, stream, c_puct)
    else:  # a real random-init reference net, outputs recorded
        mod = ref_loader.configure(ref_loader.load(modname), n)
        torch.manual_seed(seed)
        real = mod.MuZeroNet(c, A).eval()
        with torch.no_grad():   # spread the outputs a little so the tree goes deep
            real.prediction.policy_conv.weight.mul_(8.0)
            real.prediction.value_fc.weight.fill_(3.0)
        net = RecordingNet(real)
        res = run_reference_mcts(modname, n, sims, obs, net, noise, stream, c_puct)
        init_logits = net.init_logits.astype(np.float32)
        logits = np.stack(net.logits).astype(np.float32)
        rewards = np.array(net.rewards, dtype=np.float32)
        values = np.array(net.values, dtype=np.float32)
    # the injection boundary: post-softmax float32 priors, computed exactly like main.py:259/:298
    init_soft = torch.softmax(torch.from_numpy(init_logits).view(1, -1), dim=1).numpy()[0]
    soft = np.stack([torch.softmax(torch.from_numpy(l).view(1, -1), dim=1).numpy()[0] for l in logits])
    np.savez_compressed(os.path.join(OUT, f"mcts_{tag}.npz"), variant=modname, n=n, sims=sims, obs=obs,
                        noise=noise, stream=stream, init_soft=init_soft, soft=soft, rewards=rewards,
                        values=values, **res)
    print("mcts", tag, "root_visit", res["root_visit"], "max child", res["visit"].max(),
          "depth hist", np.bincount(res["sel_depth"]))


# --------------------------------------------------------------------------- trainer
def synthetic_trajectory(n, moves, seed):
    """A main.py-schema trajectory (main.py:636-711) from a random legal game of the oracle rules: observations
    (moves + 2: the terminal state is appended twice, :690 and :710), actions, visit-count-like policies,
    per-move rewards with the winner appended (:709)."""
    rng = np.random.default_rng(seed)
    A = n * n + 1
    st = gymgo.init_state(n)
    traj = {"observations": [st.copy()], "actions": [], "rewards": [], "policies": []}
    for _ in range(moves):
        valid = gymgo.valid_moves(st)
        counts = rng.integers(0, 40, size=A) * (valid > 0)
        if counts.sum() == 0:
            counts = (valid > 0).astype(np.int64)
        a = int(np.argmax(counts))
        traj["policies"].append(counts / counts.sum())
        traj["actions"].append(a)
        st = gymgo.next_state(st, a)
        traj["rewards"].append(float(rng.integers(-1, 2)) * 0.25)
        traj["observations"].append(st.copy())
        if gymgo.game_ended(st):
            break
    traj["rewards"].append(float(gymgo.winning(st)))
    traj["observations"].append(st.copy())
    return traj


def make_train(tag, n, c, moves, start_index, seed):
    """The REAL MuZeroAgent.train (main.py:381-501) on a buffer holding one trajectory, batch_size 1, with
    random.randint patched to `start_index`: the returned avg_loss is that trajectory's step_loss, and the
    network afterwards has taken one clipped Adam step on it."""
    main = ref_loader.configure(ref_loader.load("main"), n)
    A = n * n + 1
    main.config.latent_dim = c
    main.device = torch.device("cpu")
    torch.manual_seed(seed)
    agent = main.MuZeroAgent(n, c, A, 8)
    with torch.no_grad():
        agent.net.prediction.pass_logit.fill_(0.21)
        agent.net.prediction.value_fc.weight.fill_(0.8)
    before = {k: v.detach().clone().numpy() for k, v in agent.net.state_dict().items()}
    traj = synthetic_trajectory(n, moves, seed)
    rb = main.MultiVersionReplayBuffer(8, 1, True)
    rb.add(traj)
    real_randint, real_wandb = main.random.randint, main.wandb
    main.random.randint = lambda lo, hi: start_index
    main.wandb = type("NoWandb", (), {"log": staticmethod(lambda *a, **k: None)})     # logging only (main.py:488-500)
    agent.net.train()
    try:
        loss = agent.train(rb, 1)
    finally:
        main.random.randint, main.wandb = real_randint, real_wandb
    after = {k: v.detach().clone().numpy() for k, v in agent.net.state_dict().items()}
    out = {"n": n, "c": c, "start_index": start_index, "loss": np.float64(loss),
           "priority_after": np.float64(rb.buffers[0].priorities[0]),
           "lr_after": np.float64(agent.optimizer.param_groups[0]["lr"]),
           "observations": np.stack(traj["observations"]), "actions": np.asarray(traj["actions"], dtype=np.int64),
           "rewards": np.asarray(traj["rewards"], dtype=np.float64), "policies": np.stack(traj["policies"]),
           "unroll_steps": main.config.unroll_steps, "discount": main.config.discount}
    for k, v in before.items():
        out["w0." + k] = v
    for k, v in after.items():
        out["w1." + k] = v
    np.savez_compressed(os.path.join(OUT, f"train_{tag}.npz"), **out)
    print(f"train_{tag}: T={len(traj['actions'])} start={start_index} loss={loss:.6f}")


def make_train_batch(tag, n, c, moves, starts, seed):
    """The REAL MuZeroAgent.train (main.py:381-501) with batch_size = len(starts) > 1: zero_grad ONCE (:391), then for
    every sampled trajectory backward on gradients that keep accumulating, clip_grad_norm_ in place and an Adam step
    (:478-480), the n-step bootstrap of each trajectory computed with the weights as updated so far.  The sampled
    order is recorded (the buffer draws with np.random.choice), random.randint is patched to `starts` in call order."""
    main = ref_loader.configure(ref_loader.load("main"), n)
    A = n * n + 1
    main.config.latent_dim = c
    main.device = torch.device("cpu")
    torch.manual_seed(seed)
    agent = main.MuZeroAgent(n, c, A, 8)
    with torch.no_grad():
        agent.net.prediction.pass_logit.fill_(0.21)
        agent.net.prediction.value_fc.weight.fill_(0.8)
    before = {k: v.detach().clone().numpy() for k, v in agent.net.state_dict().items()}
    B = len(starts)
    trajs = [synthetic_trajectory(n, moves + 3 * i, seed + i) for i in range(B)]
    for i, t in enumerate(trajs):
        t["id"] = i
    rb = main.MultiVersionReplayBuffer(8, 1, True)
    for t in trajs:
        rb.add(t)
    order = []
    real_sample = rb.sample

    def recording_sample(k):
        batch, idx = real_sample(k)
        order.extend(int(t["id"]) for t in batch)
        return batch, idx
    rb.sample = recording_sample
    calls = [0]

    def fake_randint(lo, hi):
        v = starts[calls[0]]
        calls[0] += 1
        assert lo <= v <= hi
        return v
    real_randint, real_wandb = main.random.randint, main.wandb
    main.random.randint = fake_randint
    main.wandb = type("NoWandb", (), {"log": staticmethod(lambda *a, **k: None)})
    np.random.seed(seed)
    agent.net.train()
    try:
        loss = agent.train(rb, B)
    finally:
        main.random.randint, main.wandb = real_randint, real_wandb
    after = {k: v.detach().clone().numpy() for k, v in agent.net.state_dict().items()}
    out = {"n": n, "c": c, "batch": B, "starts": np.asarray(starts, dtype=np.int64), "order": np.asarray(order, dtype=np.int64),
           "loss": np.float64(loss), "unroll_steps": main.config.unroll_steps, "discount": main.config.discount}
    for i, t in enumerate(trajs):
        out[f"t{i}.observations"] = np.stack(t["observations"])
        out[f"t{i}.actions"] = np.asarray(t["actions"], dtype=np.int64)
        out[f"t{i}.rewards"] = np.asarray(t["rewards"], dtype=np.float64)
        out[f"t{i}.policies"] = np.stack(t["policies"])
    for k, v in before.items():
        out["w0." + k] = v
    for k, v in after.items():
        out["w1." + k] = v
    np.savez_compressed(os.path.join(OUT, f"trainbatch_{tag}.npz"), **out)
    print(f"trainbatch_{tag}: order={order} starts={list(starts)} loss={loss:.6f}")


def make_replay(seed):
    """Sampling order of the REAL replay buffers (main.py:158-244) under a fixed numpy seed: FIFO eviction,
    priority updates, multi-version mapping."""
    main = ref_loader.load("main")
    np.random.seed(seed)
    rb = main.MultiVersionReplayBuffer(5, 2, True)
    log = []
    for i in range(7):                       # capacity 5: items 0 and 1 are evicted
        rb.add({"id": i})
    s, m = rb.sample(3)
    log.append([x["id"] for x in s] + [b * 100 + l for b, l in m])
    rb.update_priorities(m, [4.0, 0.5, 2.0])
    rb.add_version()
    for i in range(7, 10):
        rb.add({"id": i})
    for _ in range(3):
        s, m = rb.sample(4)
        log.append([x["id"] for x in s] + [b * 100 + l for b, l in m])
        rb.update_priorities(m, [0.25] * 4)
    rb.add_version()                         # deque(maxlen=2): the first version's buffer falls out
    rb.add({"id": 10})
    s, m = rb.sample(2)
    log.append([x["id"] for x in s] + [b * 100 + l for b, l in m])
    flat = np.full((len(log), 8), -1, dtype=np.int64)
    for i, row in enumerate(log):
        flat[i, :len(row)] = row
    np.savez_compressed(os.path.join(OUT, "replay_sampling.npz"), seed=seed, log=flat)
    print("replay_sampling:", flat.tolist())


def make_arena(seed):
    """The REAL SelfPlayEvaluator.evaluate (main.py:595-608) with play_game replaced by recorded per-game results:
    win rate (with the reference's inversion of the odd games) and both Elo ratings over successive calls."""
    main = ref_loader.load("main")
    rng = np.random.default_rng(seed)
    real_wandb = main.wandb
    main.wandb = type("NoWandb", (), {"log": staticmethod(lambda *a, **k: None)})
    rows = []
    try:
        ev = main.SelfPlayEvaluator(None, None, None, num_games=10)
        for p in (0.5, 0.9, 0.2, 0.8, 0.95):
            results = (rng.random(10) < p).astype(np.int64)
            calls = []

            def fake_play_game(starting_player=0, _r=results, _c=calls):
                _c.append(starting_player)
                return int(_r[len(_c) - 1])
            ev.play_game = fake_play_game
            wr, elo = ev.evaluate()
            assert calls == [i % 2 for i in range(10)]
            rows.append(list(results) + [wr, ev.current_elo, ev.best_elo])
    finally:
        main.wandb = real_wandb
    np.savez_compressed(os.path.join(OUT, "arena_evaluate.npz"), rows=np.asarray(rows, dtype=np.float64),
                        threshold=main.config.evaluation_win_threshold, elo_k=main.config.elo_k,
                        initial_elo=main.config.initial_elo)
    print("arena_evaluate:", [r[10:] for r in rows])


def main():
    if not ref_loader.available():
        sys.exit("reference not available here")
    os.makedirs(OUT, exist_ok=True)
    make_nets(9, 128, 4, seed=0)
    make_nets(19, 128, 2, seed=1)
    make_nets(6, 32, 3, seed=2)
    make_nets(9, 128, 3, seed=3, weight_scale=1.7, emb_scale=0.3, tag="_scaled")
    make_nets(19, 128, 2, seed=4, weight_scale=1.7, emb_scale=0.3, tag="_scaled")
    make_mcts("main_N9_S50_table", "main", 9, 50, seed=10, mode="table", plies=12)
    make_mcts("main_N9_S800_table", "main", 9, 800, seed=11, mode="table", plies=30)
    make_mcts("main_N9_S400_net", "main", 9, 400, seed=12, mode="net", plies=8)
    make_mcts("main_N6_S256_table", "main", 6, 256, seed=13, mode="table", plies=6)
    make_mcts("main_N19_S800_table", "main", 19, 800, seed=14, mode="table", plies=150)
    make_mcts("main_N5_S300_endgame", "main", 5, 300, seed=15, mode="table", plies=40)
    make_mcts("play_N9_S300_table", "play", 9, 300, seed=16, mode="table", plies=20)
    make_mcts("selfplay_N9_S300_table", "self_play", 9, 300, seed=17, mode="table", plies=20)
    # 19x19 below the root's children (k_tree_step<128>'s deeper-level loop at A = 362): a late position with few
    # legal points and peaked policies
    make_mcts("main_N19_S600_deep", "main", 19, 600, seed=18, mode="table", plies=300, logit_scale=4.0, value_scale=1.0)
    make_mcts("selfplay_N19_S450_deep", "self_play", 19, 450, seed=19, mode="table", plies=280, logit_scale=4.0,
              value_scale=1.0)
    make_train("N5_C16_mid", 5, 16, 30, start_index=4, seed=20)
    make_train("N5_C16_tail", 5, 16, 14, start_index=9, seed=21)     # unroll runs past the end of the game
    make_train("N6_C32_head", 6, 32, 40, start_index=0, seed=22)
    make_train_batch("N5_C16_B3", 5, 16, 16, starts=[2, 7, 0], seed=23)
    make_replay(seed=5)
    make_arena(seed=6)


if __name__ == "__main__":
    main()
