"""Replay buffers and the trajectory record -- SURVEY §8(f) rank 2.

`PrioritizedReplayBuffer` / `MultiVersionReplayBuffer` keep the reference's API and sampling semantics
(/root/reference/main.py:158-244): FIFO eviction at capacity, new items at priority 1.0, sampling without
replacement with probability proportional to priority from ONE `np.random.choice` call (so a seeded numpy RNG draws
the same indices as the reference: tests/golden/replay_sampling.npz), the (buffer, local index) mapping for
priority write-back, `deque(maxlen=num_versions)` of per-model-version buffers.

`episodes_from_rows` turns the batched engine's trajectory rows (`SelfPlayEngine.trajectory_rows`, or the output
of `dist.gather_trajectories`) into the reference's per-game record (main.py:636-711):
    {'observations': [T+2 states, the terminal one twice], 'actions': [T], 'rewards': [T+1, winner last],
     'policies': [T]}
and `save_episodes` / `load_episodes` write them as the pickled list of dicts `self_play.py --output_dir` is meant
to produce (self_play.py:561-583), with the extra keys of that schema (`values`, `returns`, `final_reward`).
"""
import pickle
import random
from collections import deque

import numpy as np
import torch


class PrioritizedReplayBuffer:
    """main.py:158-180."""

    def __init__(self, capacity):
        self.capacity = capacity
        self.buffer = []
        self.priorities = []

    def add(self, trajectory):
        if len(self.buffer) >= self.capacity:
            self.buffer.pop(0)
            self.priorities.pop(0)
        self.buffer.append(trajectory)
        self.priorities.append(1.0)

    def sample(self, batch_size):
        priorities = np.array(self.priorities)
        probs = priorities / priorities.sum()
        indices = np.random.choice(len(self.buffer), batch_size, replace=False, p=probs)
        return [self.buffer[i] for i in indices], indices

    def update_priorities(self, indices, new_priorities):
        for i, p in zip(indices, new_priorities):
            self.priorities[i] = p

    def __len__(self):
        return len(self.buffer)


class MultiVersionReplayBuffer:
    """main.py:183-244: one buffer per model version, the oldest version dropped beyond `num_versions`."""

    def __init__(self, capacity, num_versions, prioritized):
        self.capacity = capacity
        self.num_versions = num_versions
        self.prioritized = prioritized
        self.buffers = deque(maxlen=num_versions)
        self.add_version()

    def add_version(self):
        self.buffers.append(PrioritizedReplayBuffer(self.capacity) if self.prioritized else deque(maxlen=self.capacity))

    def add(self, trajectory):
        if self.prioritized:
            self.buffers[-1].add(trajectory)
        else:
            self.buffers[-1].append(trajectory)

    def __len__(self):
        return sum(len(b) for b in self.buffers)

    def sample(self, batch_size):
        """([], []) while the buffers hold fewer than batch_size trajectories (main.py:215-216, :237-238)."""
        if not self.prioritized:
            combined = [t for buf in self.buffers for t in buf]
            if len(combined) < batch_size:
                return [], []
            return random.sample(combined, batch_size), None
        combined = [t for buf in self.buffers for t in buf.buffer]
        combined_p = [p for buf in self.buffers for p in buf.priorities]
        if len(combined) < batch_size:
            return [], []
        probs = np.array(combined_p) / sum(combined_p)
        indices = np.random.choice(len(combined), batch_size, replace=False, p=probs)
        bounds = np.cumsum([0] + [len(buf.buffer) for buf in self.buffers])
        mapping = []
        for i in indices:
            b = int(np.searchsorted(bounds, i, side="right")) - 1
            mapping.append((b, int(i - bounds[b])))
        return [combined[i] for i in indices], mapping

    def update_priorities(self, mapping, new_priorities):
        if not self.prioritized or mapping is None:
            return
        for (b, local), p in zip(mapping, new_priorities):
            if b < len(self.buffers):
                self.buffers[b].priorities[local] = p


def episodes_from_rows(rows, board_size, carry=None):
    """Trajectory rows (dict of [T,G,...] tensors) -> (finished episodes in the reference's schema, carry).

    A game's record runs from its first row to the row whose `done` flag is set; games still in progress at the
    last row come back in `carry` (pass it to the next call).  Observations are GymGo float64 [6,N,N] states
    expanded from the packed positions by the rules kernels; the terminal state is the finishing row's position
    with its move applied (the engine restarts a finished game in place, so it is never a row of its own)."""
    from .rules import GoBoards
    T, G = rows["action"].shape[:2]
    dev = rows["boards"].device
    n = int(board_size)
    cur = GoBoards(T * G, n, device=dev)
    cur.records.copy_(rows["boards"].reshape(T * G, -1))
    planes = cur.planes(torch.float32).reshape(T, G, 6, n, n).double().cpu().numpy()
    done = rows["done"].bool()
    term = {}
    if bool(done.any()):
        idx = done.reshape(-1).nonzero().flatten()
        last = GoBoards(int(idx.numel()), n, device=dev)
        last.records.copy_(rows["boards"].reshape(T * G, -1)[idx])
        last.step(rows["action"].reshape(-1)[idx].to(torch.int32).contiguous(), return_status=False)
        lp = last.planes(torch.float32).double().cpu().numpy()
        term = {int(i): lp[j] for j, i in enumerate(idx.tolist())}
    action = rows["action"].cpu().numpy()
    policy = rows["policy"].double().cpu().numpy()
    reward = rows["reward"].double().cpu().numpy()
    value = rows["root_value"].double().cpu().numpy()
    winner = rows["winner"].cpu().numpy()
    done_h = done.cpu().numpy()
    carry = carry if carry is not None else [None] * G
    finished = []
    for g in range(G):
        ep = carry[g] or {"observations": [], "actions": [], "rewards": [], "policies": [], "values": []}
        for t in range(T):
            ep["observations"].append(planes[t, g])
            ep["actions"].append(int(action[t, g]))
            ep["policies"].append(policy[t, g])
            ep["rewards"].append(float(reward[t, g]))
            ep["values"].append(float(value[t, g]))
            if done_h[t, g]:
                final = term[t * G + g]
                ep["observations"].append(final)                 # main.py:690
                ep["rewards"].append(float(winner[t, g]))        # main.py:709
                ep["observations"].append(final)                 # main.py:710
                ep["final_reward"] = float(winner[t, g])
                finished.append(ep)
                ep = {"observations": [], "actions": [], "rewards": [], "policies": [], "values": []}
        carry[g] = ep if ep["actions"] else None
    return finished, carry


def calculate_returns(rewards, final_reward, discount):
    """self_play.py:435-447: G_t = r_t + discount * G_{t+1}, backwards from G_T = final_reward."""
    out, run = [0.0] * len(rewards), final_reward
    for i in range(len(rewards) - 1, -1, -1):
        run = rewards[i] + discount * run
        out[i] = run
    return out


def save_episodes(path, episodes, discount=0.99):
    """The pickled list of dicts self_play.py:561-583 means to write (at the reference's HEAD the dump runs after
    the file is closed and nothing is saved): observations, actions, policies, values (search root values),
    rewards (per step, without the outcome main.py appends), returns, final_reward."""
    out = []
    for ep in episodes:
        T = len(ep["actions"])
        final = float(ep.get("final_reward", ep["rewards"][-1] if len(ep["rewards"]) > T else 0.0))
        step_rewards = list(ep["rewards"][:T])
        out.append({"observations": ep["observations"], "actions": ep["actions"], "policies": ep["policies"],
                    "values": ep.get("values", []), "rewards": step_rewards,
                    "returns": calculate_returns(step_rewards, final, discount), "final_reward": final})
    with open(path, "wb") as f:
        pickle.dump(out, f)
    return len(out)


def load_episodes(path):
    with open(path, "rb") as f:
        return pickle.load(f)
