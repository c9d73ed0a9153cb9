"""Replay on the device -- the "replay sampling on GPU" half of SURVEY §8(f) rank 1.

`replay.py` keeps the reference's host-side buffers (lists of per-game dicts).  This is the same content kept where
the trainer needs it: finished games as rows of padded device tensors (packed positions of 256 B instead of
[6,N,N] float64 planes: 19x19 positions are 68x smaller), priorities on the device, and a training batch -- start
indices, unrolled actions, reward / policy / n-step value targets (main.py:396-460, :503-522) -- assembled with
gathers, the bootstrap positions expanded to planes by the rules kernels and evaluated in one batched forward.
`tests/test_train_gpu.py` checks a batch against `trainer.build_batch` on the same games and start indices.

Sampling is proportional to priority without replacement like `PrioritizedReplayBuffer.sample` (main.py:172-178),
drawn by `torch.multinomial` on the device instead of `np.random.choice`: the same distribution, another RNG.
"""
import torch

from .rules import GoBoards
from .trainer import Batch, prediction, representation


class DeviceReplay:
    def __init__(self, capacity, board_size, max_moves=None, device="cuda"):
        self.E, self.n = int(capacity), int(board_size)
        self.A = self.n * self.n + 1
        self.L = int(max_moves if max_moves is not None else int(self.n * self.n * 1.5))     # main.py:53
        dev = self.device = torch.device(device)
        E, L, A = self.E, self.L, self.A
        self.boards = torch.zeros((E, L + 1, 64), dtype=torch.int32, device=dev)     # positions 0..T (T = terminal)
        self.actions = torch.zeros((E, L), dtype=torch.int64, device=dev)
        self.rewards = torch.zeros((E, L + 1), dtype=torch.float64, device=dev)      # T step rewards + the outcome
        self.policies = torch.zeros((E, L, A), dtype=torch.float32, device=dev)
        self.length = torch.zeros((E,), dtype=torch.int64, device=dev)               # T = number of moves
        self.priority = torch.zeros((E,), dtype=torch.float64, device=dev)           # 0 = empty slot
        self.count, self.head = 0, 0
        self._open = None                                                           # per game: rows of unfinished games

    def __len__(self):
        return min(self.count, self.E)

    # ------------------------------------------------------------------ filling
    def add_rows(self, rows):
        """Trajectory rows of the engine (dict of [T,G,...] device tensors).  Finished games go into the ring
        (FIFO eviction, priority 1.0: main.py:163-171); the open tail of every game is kept for the next call."""
        T, G = rows["action"].shape[:2]
        done = rows["done"].bool().cpu().numpy()
        if self._open is None:
            self._open = [None] * G
        for g in range(G):
            lo = 0
            for t in range(T):
                if done[t, g]:
                    self._store(rows, g, lo, t + 1, self._open[g])
                    self._open[g], lo = None, t + 1
            if lo < T:
                part = {k: rows[k][lo:T, g].clone() for k in ("boards", "action", "policy", "reward", "winner")}
                prev = self._open[g]
                self._open[g] = part if prev is None else {k: torch.cat([prev[k], part[k]]) for k in part}

    def _store(self, rows, g, lo, hi, prev):
        seg = {k: rows[k][lo:hi, g] for k in ("boards", "action", "policy", "reward", "winner")}
        if prev is not None:
            seg = {k: torch.cat([prev[k], seg[k]]) for k in seg}
        T = int(seg["action"].shape[0])
        if T > self.L:
            raise ValueError(f"game of {T} moves exceeds max_moves={self.L}")
        e = self.head
        self.head, self.count = (self.head + 1) % self.E, self.count + 1
        last = GoBoards(1, self.n, device=self.device)
        last.records.copy_(seg["boards"][T - 1:T])
        last.step(seg["action"][T - 1:T].to(torch.int32).contiguous(), return_status=False)    # the terminal position
        self.boards[e].zero_()
        self.boards[e, :T] = seg["boards"]
        self.boards[e, T] = last.records[0]
        self.actions[e].zero_()
        self.actions[e, :T] = seg["action"].long()
        self.rewards[e].zero_()
        self.rewards[e, :T] = seg["reward"].double()
        self.rewards[e, T] = seg["winner"][T - 1].double()                                      # main.py:709
        self.policies[e].zero_()
        self.policies[e, :T] = seg["policy"]
        self.length[e] = T
        self.priority[e] = 1.0

    # ------------------------------------------------------------------ sampling
    def sample(self, batch_size, generator=None):
        """-> (episode slots int64 [B], start indices int64 [B]) or None while fewer than batch_size games are
        stored (main.py:383-386); start = random.randint(0, T - 1) per game (main.py:397)."""
        if len(self) < batch_size:
            return None
        slots = torch.multinomial(self.priority, batch_size, replacement=False, generator=generator)
        u = torch.rand(batch_size, device=self.device, dtype=torch.float64, generator=generator)
        starts = torch.minimum((u * self.length[slots]).long(), self.length[slots] - 1)
        return slots, starts

    def update_priorities(self, slots, value):
        self.priority[slots] = float(value)                                                     # main.py:495-497

    # ------------------------------------------------------------------ batch assembly
    @torch.no_grad()
    def build_batch(self, net, slots, starts, cfg):
        """trainer.build_batch's targets, by gathers on the device."""
        K, A, L, dev = cfg.unroll_steps, self.A, self.L, self.device
        B = slots.shape[0]
        T = self.length[slots]                                           # [B] moves; len(rewards) = T + 1
        idx = starts[:, None] + torch.arange(K + 1, device=dev)[None]    # [B, K+1] trajectory index of unroll step k
        rew = torch.cat([self.rewards[slots], torch.zeros((B, 2 * K + 1), dtype=torch.float64, device=dev)], dim=1)
        # n-step value target (main.py:503-522): up to K recorded rewards from idx, rewards[j] for j < T + 1
        table, f = [], 1.0
        for _ in range(K + 1):                                           # factor *= discount, as the reference does
            table.append(f)
            f *= cfg.discount
        disc = torch.tensor(table, dtype=torch.float64, device=dev)
        win = idx[:, :, None] + torch.arange(K, device=dev)[None, None]              # [B, K+1, K]
        terms = torch.gather(rew, 1, win.reshape(B, -1).clamp(max=rew.shape[1] - 1)).reshape(B, K + 1, K)
        target = torch.zeros((B, K + 1), dtype=torch.float64, device=dev)
        for j in range(K):                                               # the reference's left-to-right float64 sum
            target = target + disc[j] * terms[:, :, j]
        n_terms = (T[:, None] + 1 - idx).clamp(min=0, max=K)             # recorded rewards inside the window
        factor = disc[n_terms]
        boot = idx + K < (T[:, None] + 1)                                # main.py:516
        if bool(boot.any()):
            b_i, k_i = boot.nonzero(as_tuple=True)
            pos = GoBoards(int(b_i.numel()), self.n, device=dev)
            pos.records.copy_(self.boards[slots[b_i], idx[b_i, k_i] + K])
            value, _ = prediction(net, representation(net, pos.planes(torch.float32)))
            target[b_i, k_i] += factor[b_i, k_i] * value.flatten().double()
        # policies: recorded for index < T, uniform beyond (main.py:407-410, :447-451)
        pidx = idx.clamp(max=L - 1)
        pol = self.policies[slots[:, None], pidx]                        # [B, K+1, A]
        pol = torch.where((idx < T[:, None])[:, :, None], pol.double(), torch.full_like(pol, 1.0 / A, dtype=torch.float64))
        # actions / rewards of unroll step k >= 1 sit at index - 1 (main.py:433-446)
        prev = idx[:, 1:] - 1
        act = torch.where(prev < T[:, None], self.actions[slots[:, None], prev.clamp(max=L - 1)],
                          torch.full_like(prev, A - 1))
        tr = torch.gather(rew, 1, prev)                                  # zero beyond the T + 1 recorded rewards
        first = GoBoards(B, self.n, device=dev)
        first.records.copy_(self.boards[slots, starts])
        return Batch(first.planes(torch.float32), act, tr.float(), target.float(), pol.float())


def train_on_device(trainer, replay, batch_size, generator=None):
    """`BatchedTrainer.train` (main.py:381-501) with the buffer, the sampling and the batch assembly on the
    device.  Returns the batch-average loss, or None while the buffer cannot fill a batch."""
    drawn = replay.sample(batch_size, generator)
    if drawn is None:
        return None
    slots, starts = drawn
    batch = replay.build_batch(trainer.net, slots, starts, trainer.cfg)
    per_traj, parts = trainer.train_step(batch)
    avg = float(per_traj.sum().item()) / batch_size
    replay.update_priorities(slots, avg)
    trainer._schedule()
    trainer.last = {"training_loss": avg, "lr": trainer.lr,
                    **{k + "_loss": float(v.sum().item()) / batch_size for k, v in parts.items()}}
    return avg
