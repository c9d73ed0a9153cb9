"""CPU restatement of the reference trainer's per-trajectory loss (TEST INFRASTRUCTURE).

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this package.  Plain torch fp32 (the
losses are floating point; the tolerance is written in the tests).  Pinned against the REAL reference by
tests/golden/train_*.npz, which oracle/make_golden.py produces by running the unmodified `MuZeroAgent.train`
(main.py:381-501) on a one-trajectory buffer with `random.randint` patched to a fixed start index.

Follows:
  reward_value_transform        main.py:67-69
  compute_target_value          main.py:503-522
  trajectory_loss               main.py:396-480 (one iteration of the `for trajectory in batch` loop)
"""
import numpy as np
import torch
import torch.nn.functional as F



def reward_value_transform(x, epsilon=0.001):
    """h(x) = sign(x) (sqrt(|x| + 1) - 1) + eps x   (main.py:67-69)"""
    return torch.sign(x) * (torch.sqrt(torch.abs(x) + 1) - 1) + epsilon * x


def compute_target_value(net, trajectory, index, unroll_steps, discount):
    """main.py:503-522: discounted sum of up to `unroll_steps` recorded rewards from `index`, plus the discounted
    value of a FRESH initial inference of observation index + unroll_steps if that is still inside the game."""
    target, factor = 0.0, 1.0
    T = len(trajectory["rewards"])
    for k in range(unroll_steps):
        j = index + k
        if j < T:
            target += factor * trajectory["rewards"][j]
            factor *= discount
        else:
            break
    if index + unroll_steps < T:
        obs = torch.from_numpy(np.asarray(trajectory["observations"][index + unroll_steps], dtype=np.float32))[None]
        with torch.no_grad():
            _, value, _ = net.initial_inference(obs)
        target += factor * value.item()
    return target


def trajectory_loss(net, trajectory, start_index, action_size, unroll_steps=10, discount=0.99, value_w=1.0,
                    policy_w=1.0, reward_w=1.0, use_value_transform=True):
    """step_loss of main.py:396-476 for one trajectory and a given start index.  `net` is an
    oracle.nets_ref.MuZeroNetRef (the torch fp32 restatement of main.py:72-142).  Returns (loss, dict of the three weighted sums)."""
    uniform = np.ones(action_size) / action_size

    def vloss(value, target):
        t = torch.tensor(target, dtype=torch.float32)
        if use_value_transform:
            return F.mse_loss(reward_value_transform(value.squeeze()), reward_value_transform(t))
        return F.mse_loss(value.squeeze(), t)

    def ploss(logits, target):
        t = torch.tensor(np.asarray(target), dtype=torch.float32)
        return F.kl_div(F.log_softmax(logits, dim=1), t, reduction="batchmean")

    obs = torch.from_numpy(np.asarray(trajectory["observations"][start_index], dtype=np.float32))[None]
    latent, value, logits = net.initial_inference(obs)
    tv = compute_target_value(net, trajectory, start_index, unroll_steps, discount)
    tp = trajectory["policies"][start_index] if start_index < len(trajectory["policies"]) else uniform
    v, p = vloss(value, tv), ploss(logits, tp)
    loss = value_w * v + policy_w * p
    parts = {"value": value_w * v.item(), "policy": policy_w * p.item(), "reward": 0.0}
    for k in range(1, unroll_steps + 1):
        i = start_index + k - 1
        action = trajectory["actions"][i] if i < len(trajectory["actions"]) else action_size - 1
        latent, reward, value, logits = net.recurrent_inference(latent, torch.tensor([action], dtype=torch.long))
        tr = trajectory["rewards"][i] if i < len(trajectory["rewards"]) else 0.0
        tp = trajectory["policies"][start_index + k] if start_index + k < len(trajectory["policies"]) else uniform
        tv = compute_target_value(net, trajectory, start_index + k, unroll_steps, discount)
        r = F.mse_loss(reward.squeeze(), torch.tensor(tr, dtype=torch.float32))
        v, p = vloss(value, tv), ploss(logits, tp)
        loss = loss + reward_w * r + value_w * v + policy_w * p
        parts["reward"] += reward_w * r.item()
        parts["value"] += value_w * v.item()
        parts["policy"] += policy_w * p.item()
    return loss, parts
