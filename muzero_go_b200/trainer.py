"""Batched trainer -- SURVEY §8(f) rank 1, first slice.

The reference (`MuZeroAgent.train`, /root/reference/main.py:381-501) takes `batch_size` trajectories and runs
forward, backward, gradient clipping and an Adam step for EACH of them in turn at batch size 1.  Here the K-step
unroll runs over the whole batch and the optimizer steps once on the mean of the per-trajectory losses; every
per-trajectory quantity (targets, the three losses, their weights) is the reference's, so with `batch_size == 1`
a step is the reference's step (tests/test_trainer.py pins the loss and the updated weights against the real
`MuZeroAgent.train`, tests/golden/train_*.npz).

Kernels: on a GPU the 3x3 convolutions of the representation and dynamics networks -- more than 99.9 % of a step's
FLOPs -- run forward, data gradient and weight gradient in libmzgo.so's own float32 kernels (csrc/train.cu through
muzero_go_b200/train_ops.py, one torch.autograd.Function per layer); torch's autograd keeps the bookkeeping around
them (the 1x1 heads, the spatial means, the losses, clipping, Adam).  On CPU tensors (the host-logic parity tests)
the same functions fall through to torch's own convolution.

  reward_value_transform      main.py:67-69
  n-step value target         main.py:503-522 (`compute_target_value`)
  losses                      main.py:396-476
  clip 1.0, Adam 1e-4, StepLR(1000, 0.9)      main.py:377-379, :478-480, :498-499
"""
import dataclasses
import random

import numpy as np
import torch
import torch.nn.functional as F


@dataclasses.dataclass
class TrainConfig:
    """main.py:27-56 (the values the trainer reads)."""
    unroll_steps: int = 10
    discount: float = 0.99
    value_loss_weight: float = 1.0
    policy_loss_weight: float = 1.0
    reward_loss_weight: float = 1.0
    learning_rate: float = 1e-4
    use_value_transform: bool = True
    max_grad_norm: float = 1.0
    lr_step_size: int = 1000
    lr_gamma: float = 0.9


def reward_value_transform(x, epsilon=0.001):
    """h(x) = sign(x) (sqrt(|x| + 1) - 1) + eps x   (main.py:67-69)"""
    return torch.sign(x) * (torch.sqrt(torch.abs(x) + 1) - 1) + epsilon * x


# ---- differentiable forward on the parameter containers of MuZeroNet (main.py:72-142) ---------------------------
def conv3x3_relu(x, weight, bias):
    """relu(conv2d(x, weight, bias, padding=1)): libmzgo.so's kernels on CUDA tensors, torch's on CPU tensors."""
    if x.is_cuda:            # the product path: no library convolution, and a loud error rather than a silent one
        from .train_ops import conv3x3_relu as native
        return native(x, weight, bias)
    return F.relu(F.conv2d(x, weight, bias, padding=1))


def representation(net, obs):
    r = net.representation
    x = conv3x3_relu(obs, r.conv1.weight, r.conv1.bias)
    x = conv3x3_relu(x, r.conv2.weight, r.conv2.bias)
    return conv3x3_relu(x, r.conv3.weight, r.conv3.bias)


def prediction(net, latent):
    p = net.prediction
    value = F.conv2d(latent, p.value_conv.weight, p.value_conv.bias).mean(dim=(2, 3))
    value = F.linear(value, p.value_fc.weight, p.value_fc.bias)
    board = F.conv2d(latent, p.policy_conv.weight, p.policy_conv.bias).flatten(1)
    logits = torch.cat([board, p.pass_logit.expand(board.shape[0], 1)], dim=1)
    return value, logits


def dynamics(net, latent, action):
    d = net.dynamics
    emb = F.embedding(action, d.action_embedding.weight)[:, :, None, None]
    x = conv3x3_relu(latent + emb, d.conv.weight, d.conv.bias)
    r = F.conv2d(x, d.reward_conv.weight, d.reward_conv.bias).mean(dim=(2, 3))
    r = F.linear(F.relu(F.linear(r, d.fc_reward_hidden.weight, d.fc_reward_hidden.bias)), d.fc_reward_output.weight,
                 d.fc_reward_output.bias)
    return x, r


@dataclasses.dataclass
class Batch:
    observations: torch.Tensor     # [B, 6, N, N] float32: the state at the start index
    actions: torch.Tensor          # [B, K] int64 (pass once the recorded actions run out, main.py:433-438)
    target_reward: torch.Tensor    # [B, K] float32
    target_value: torch.Tensor     # [B, K+1] float32
    target_policy: torch.Tensor    # [B, K+1, A] float32 (uniform past the end of the game, main.py:407-410)


def build_batch(net, trajectories, start_indices, cfg, action_size, device):
    """The targets of main.py:396-460 for every trajectory of the batch.  The bootstrap values of the n-step
    targets (a fresh initial inference of observation index + unroll_steps, main.py:516-521) are ONE batched
    no-grad forward over all (trajectory, step) pairs that need one."""
    B, K = len(trajectories), cfg.unroll_steps
    uniform = np.ones(action_size) / action_size
    boot_obs, boot_where = [], []
    tv = np.zeros((B, K + 1), dtype=np.float64)
    factors = np.zeros((B, K + 1), dtype=np.float64)
    tr = np.zeros((B, K), dtype=np.float64)
    tp = np.zeros((B, K + 1, action_size), dtype=np.float64)
    acts = np.zeros((B, K), dtype=np.int64)
    for b, (traj, s) in enumerate(zip(trajectories, start_indices)):
        rewards, T = traj["rewards"], len(traj["rewards"])
        for k in range(K + 1):
            index = s + k
            target, factor = 0.0, 1.0
            for j in range(index, min(index + K, T)):                    # main.py:508-514
                target += factor * rewards[j]
                factor *= cfg.discount
            tv[b, k], factors[b, k] = target, factor
            if index + K < T:                                            # main.py:516
                boot_obs.append(np.asarray(traj["observations"][index + K], dtype=np.float32))
                boot_where.append((b, k))
            tp[b, k] = traj["policies"][index] if index < len(traj["policies"]) else uniform
            if k >= 1:
                i = index - 1
                acts[b, k - 1] = traj["actions"][i] if i < len(traj["actions"]) else action_size - 1
                tr[b, k - 1] = rewards[i] if i < T else 0.0
    if boot_obs:
        with torch.no_grad():
            obs = torch.from_numpy(np.stack(boot_obs)).to(device)
            value, _ = prediction(net, representation(net, obs))
            value = value.flatten().double().cpu().numpy()
        for (b, k), v in zip(boot_where, value):
            tv[b, k] += factors[b, k] * v
    first = np.stack([np.asarray(t["observations"][s], dtype=np.float32) for t, s in zip(trajectories, start_indices)])
    f32 = dict(dtype=torch.float32, device=device)
    return Batch(torch.from_numpy(first).to(device), torch.from_numpy(acts).to(device), torch.tensor(tr, **f32),
                 torch.tensor(tv, **f32), torch.tensor(tp, **f32))


class BatchedTrainer:
    """`MuZeroAgent`'s optimizer side (main.py:377-379, :381-501) on a batch."""

    def __init__(self, net, action_size, cfg=None, cuda_graph=None):
        """cuda_graph (default: on when the network lives on a GPU): a training step is ~1,300 small kernels
        (11 network evaluations forward and backward on a 6x6 .. 19x19 board), i.e. launch bound; the whole step
        -- losses, backward, clipping, Adam -- is captured once per batch shape and replayed."""
        self.net, self.action_size, self.cfg = net, int(action_size), cfg or TrainConfig()
        on_gpu = next(net.parameters()).is_cuda
        self.cuda_graph = on_gpu if cuda_graph is None else bool(cuda_graph and on_gpu)
        if self.cuda_graph:
            # the captured Adam reads its step size from device memory: StepLR below updates it in place
            self._lr = torch.tensor(self.cfg.learning_rate, dtype=torch.float64, device=next(net.parameters()).device)
            self.optimizer = torch.optim.Adam(net.parameters(), lr=self._lr, capturable=True)
            self.scheduler = None
        else:
            self.optimizer = torch.optim.Adam(net.parameters(), lr=self.cfg.learning_rate)
            self.scheduler = torch.optim.lr_scheduler.StepLR(self.optimizer, step_size=self.cfg.lr_step_size,
                                                             gamma=self.cfg.lr_gamma)
        self._sched_calls = 0
        self._graphs = {}

    @property
    def device(self):
        return next(self.net.parameters()).device

    def _weights_changed(self):
        # a replayed graph rewrites the parameters without bumping tensor._version: tell the network explicitly, so the
        # next self-play / arena forward re-derives its kernel-side weight layouts (MuZeroNet.descriptor)
        mark = getattr(self.net, "mark_weights_changed", None)
        if mark is not None:
            mark()

    def losses(self, batch):
        """Per-trajectory step_loss [B] (main.py:396-476) and the three weighted components [B] each."""
        cfg = self.cfg

        def vloss(value, target):
            if cfg.use_value_transform:
                return (reward_value_transform(value) - reward_value_transform(target)) ** 2
            return (value - target) ** 2

        def ploss(logits, target):                     # F.kl_div(..., reduction='batchmean') at batch size 1
            return F.kl_div(F.log_softmax(logits, dim=1), target, reduction="none").sum(dim=1)

        latent = representation(self.net, batch.observations)
        value, logits = prediction(self.net, latent)
        v = cfg.value_loss_weight * vloss(value.flatten(), batch.target_value[:, 0])
        p = cfg.policy_loss_weight * ploss(logits, batch.target_policy[:, 0])
        r = torch.zeros_like(v)
        for k in range(1, cfg.unroll_steps + 1):
            latent, reward = dynamics(self.net, latent, batch.actions[:, k - 1])
            value, logits = prediction(self.net, latent)
            r = r + cfg.reward_loss_weight * (reward.flatten() - batch.target_reward[:, k - 1]) ** 2
            v = v + cfg.value_loss_weight * vloss(value.flatten(), batch.target_value[:, k])
            p = p + cfg.policy_loss_weight * ploss(logits, batch.target_policy[:, k])
        return v + p + r, {"value": v, "policy": p, "reward": r}

    def _step_eager(self, batch):
        self.optimizer.zero_grad(set_to_none=True)
        per_traj, parts = self.losses(batch)
        per_traj.mean().backward()
        torch.nn.utils.clip_grad_norm_(self.net.parameters(), max_norm=self.cfg.max_grad_norm)
        self.optimizer.step()
        self._weights_changed()
        return per_traj.detach(), {k: v.detach() for k, v in parts.items()}

    def _capture(self, batch):
        """One CUDA graph per batch shape.  The warm-up iterations torch asks for before a capture run on a side
        stream and are undone afterwards (weights and Adam state restored), so capturing changes nothing."""
        import copy
        fields = [f.name for f in dataclasses.fields(Batch)]
        static = Batch(**{k: getattr(batch, k).clone() for k in fields})
        net_state = copy.deepcopy(self.net.state_dict())
        # Adam's moments are restored IN PLACE: graphs captured earlier for other batch shapes hold the addresses of
        # these very tensors (optimizer.load_state_dict would swap in new objects and leave them updating freed memory)
        opt_saved = {p: {k: (v.clone() if torch.is_tensor(v) else v) for k, v in st.items()}
                     for p, st in self.optimizer.state.items()}
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                self._step_eager(static)
        torch.cuda.current_stream().wait_stream(side)
        self.net.load_state_dict(net_state)
        for p, st in self.optimizer.state.items():
            saved = opt_saved.get(p)
            for k, v in st.items():
                if not torch.is_tensor(v):
                    continue
                if saved is not None and k in saved:
                    v.copy_(saved[k])
                else:
                    # a fresh optimizer: keep the moment buffers the warm-up created and zero them.  They must exist
                    # BEFORE the capture -- created inside it, their zero-fill would be replayed with every step.
                    v.zero_()
        self.optimizer.zero_grad(set_to_none=True)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out = self._step_eager(static)
        return graph, static, out

    def train_step(self, batch):
        """zero_grad, backward of the mean per-trajectory loss, clip_grad_norm_(1.0), Adam step (main.py:394,
        :478-480).  Returns the per-trajectory losses (detached)."""
        if not self.cuda_graph:
            return self._step_eager(batch)
        key = tuple(tuple(getattr(batch, f.name).shape) for f in dataclasses.fields(Batch))
        if key not in self._graphs:
            self._graphs[key] = self._capture(batch)
        graph, static, (per_traj, parts) = self._graphs[key]
        for f in dataclasses.fields(Batch):
            getattr(static, f.name).copy_(getattr(batch, f.name))
        graph.replay()
        self._weights_changed()
        return per_traj.clone(), {k: v.clone() for k, v in parts.items()}

    def train_reference_order(self, trajectories, start_indices):
        """The reference's update order for a batch of more than one trajectory (main.py:391-481), for parity runs:
        zero_grad ONCE, then for every trajectory in turn -- targets bootstrapped with the weights as updated so far,
        backward on gradients that keep accumulating, clip_grad_norm_ in place, an Adam step.  len(trajectories)
        optimizer steps per call (where train_step takes one on the mean loss).  Eager, one trajectory at a time, like
        the reference.  -> per-trajectory losses [B]."""
        self.optimizer.zero_grad(set_to_none=True)
        out = []
        for traj, s in zip(trajectories, start_indices):
            b = build_batch(self.net, [traj], [s], self.cfg, self.action_size, self.device)
            per_traj, _ = self.losses(b)
            per_traj.sum().backward()
            torch.nn.utils.clip_grad_norm_(self.net.parameters(), max_norm=self.cfg.max_grad_norm)
            self.optimizer.step()
            self._weights_changed()
            out.append(per_traj.detach())
        return torch.cat(out)

    def _schedule(self):
        """StepLR(step_size, gamma) (main.py:379, :498)."""
        self._sched_calls += 1
        if self.scheduler is not None:
            self.scheduler.step()
        elif self._sched_calls % self.cfg.lr_step_size == 0:
            self._lr.mul_(self.cfg.lr_gamma)

    @property
    def lr(self):
        lr = self.optimizer.param_groups[0]["lr"]
        return float(lr.item()) if torch.is_tensor(lr) else float(lr)

    def train(self, replay_buffer, batch_size, rng=random, reference_order=False):
        """main.py:381-501: sample, one random start index per trajectory, train, write the batch-average loss
        back as every sampled trajectory's priority, step the learning-rate schedule.  None if the buffer
        cannot fill a batch.

        reference_order=False (default): ONE optimizer step on the mean of the per-trajectory losses.
        reference_order=True: the reference's own order -- batch_size steps on accumulating gradients
        (train_reference_order); slower, for runs that must be comparable with the reference update for update."""
        batch, indices = replay_buffer.sample(batch_size)
        if not batch or len(batch) < batch_size:
            return None
        starts = [rng.randint(0, len(t["actions"]) - 1) for t in batch]
        if reference_order:
            per_traj = self.train_reference_order(batch, starts)
            parts = {}
        else:
            b = build_batch(self.net, batch, starts, self.cfg, self.action_size, self.device)
            per_traj, parts = self.train_step(b)
        avg_loss = float(per_traj.sum().item()) / batch_size
        if indices is not None:
            replay_buffer.update_priorities(indices, [avg_loss] * len(indices))
        self._schedule()
        self.last = {"training_loss": avg_loss, "lr": self.lr,
                     **{k + "_loss": float(v.sum().item()) / batch_size for k, v in parts.items()}}
        return avg_loss
