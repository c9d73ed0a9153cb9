"""MuZeroNet with the reference's module tree, parameter names and call signatures
(/root/reference/main.py:72-142; copies at self_play.py:63-128, play.py:34-101), evaluated by libmzgo.so.

The nn.Conv2d / nn.Linear / nn.Embedding objects below are parameter containers only: they give the same
state_dict keys, shapes and random initialisation order as the reference, so reference .pth files load and
torch.manual_seed(s) reproduces the reference's weights.  Every forward goes through the C ABI
(mzgo_net_initial / mzgo_net_recurrent / mzgo_net_predict); there is no torch arithmetic here.
"""
import math

import torch
import torch.nn as nn

from . import _lib

PRECISIONS = {"fp32": 0, "bf16": 1}


class RepresentationNetwork(nn.Module):
    """main.py:72-84"""

    def __init__(self, latent_dim):
        super().__init__()
        self.conv1 = nn.Conv2d(6, 64, kernel_size=3, padding=1)
        self.conv2 = nn.Conv2d(64, 64, kernel_size=3, padding=1)
        self.conv3 = nn.Conv2d(64, latent_dim, kernel_size=3, padding=1)
        self._owner = None

    def forward(self, x):
        return self._owner[0]._run_initial(x)[0]


class DynamicsNetwork(nn.Module):
    """main.py:86-106"""

    def __init__(self, latent_dim, max_action_size):
        super().__init__()
        self.action_embedding = nn.Embedding(max_action_size, latent_dim)
        self.conv = nn.Conv2d(latent_dim, latent_dim, kernel_size=3, padding=1)
        self.reward_conv = nn.Conv2d(latent_dim, 1, kernel_size=1)
        self.fc_reward_hidden = nn.Linear(1, 16)
        self.fc_reward_output = nn.Linear(16, 1)
        self._owner = None

    def forward(self, latent, action):
        nxt, reward, _, _ = self._owner[0]._run_recurrent(latent, action)
        return nxt, reward


class PredictionNetwork(nn.Module):
    """main.py:108-125"""

    def __init__(self, latent_dim):
        super().__init__()
        self.value_conv = nn.Conv2d(latent_dim, 1, kernel_size=1)
        self.value_fc = nn.Linear(1, 1)
        self.policy_conv = nn.Conv2d(latent_dim, 1, kernel_size=1)
        self.pass_logit = nn.Parameter(torch.zeros(1))
        self._owner = None

    def forward(self, latent):
        return self._owner[0]._run_predict(latent)


class MuZeroNet(nn.Module):
    """main.py:127-142.  `board_size` defaults to sqrt(max_action_size - 1) (main.py:30);
    play.py:223 sizes the embedding differently, so pass board_size there."""

    def __init__(self, latent_dim, max_action_size, board_size=None, precision="fp32"):
        super().__init__()
        if board_size is None:
            board_size = int(round(math.sqrt(max_action_size - 1)))
            if board_size * board_size + 1 != max_action_size:
                raise ValueError("pass board_size when max_action_size != board_size**2 + 1")
        if max_action_size < board_size * board_size + 1:
            raise ValueError("action embedding smaller than the action space")
        self.latent_dim, self.max_action_size, self.board_size = int(latent_dim), int(max_action_size), int(board_size)
        self.precision = precision
        if precision not in PRECISIONS:
            raise ValueError(f"precision {precision!r} not in {sorted(PRECISIONS)}")
        self.representation = RepresentationNetwork(latent_dim)
        self.dynamics = DynamicsNetwork(latent_dim, max_action_size)
        self.prediction = PredictionNetwork(latent_dim)
        for m in (self.representation, self.dynamics, self.prediction):
            m._owner = [self]          # a list keeps the owner out of the module tree
        self._ws = None            # {ws_key: {"ws": tensor, "key": (max_batch, precision, device), "prepared": key}}
        self._desc = None
        self._weights_version = 0

    # ------------------------------------------------------------------ descriptor / workspace
    @property
    def action_size(self):
        return self.board_size ** 2 + 1

    def _params_in_abi_order(self):
        r, d, p = self.representation, self.dynamics, self.prediction
        return [r.conv1.weight, r.conv1.bias, r.conv2.weight, r.conv2.bias, r.conv3.weight, r.conv3.bias,
                d.action_embedding.weight, d.conv.weight, d.conv.bias, d.reward_conv.weight, d.reward_conv.bias,
                d.fc_reward_hidden.weight, d.fc_reward_hidden.bias, d.fc_reward_output.weight,
                d.fc_reward_output.bias, p.value_conv.weight, p.value_conv.bias, p.value_fc.weight, p.value_fc.bias,
                p.policy_conv.weight, p.policy_conv.bias, p.pass_logit]

    def slot_bytes(self):
        return int(_lib.load().mzgo_net_slot_bytes(self.board_size, self.latent_dim, PRECISIONS[self.precision]))

    def descriptor(self, max_batch, ws_key=0):
        """mzgo_net_t for the current parameters; (re)derives the kernel-side weight layouts when the
        parameters changed (load_state_dict, optimizer step, broadcast).

        ws_key names the workspace: the workspace holds the prepared weight layouts AND the representation network's
        scratch activations, so callers that evaluate the network concurrently on different CUDA streams
        (WaveGroup(streams=2)) must use different keys."""
        _lib.require_device()
        params = self._params_in_abi_order()
        dev = params[0].device
        if dev.type != "cuda":
            raise _lib.MzgoError("MuZeroNet parameters must live on a CUDA device (no CPU fallback): call .cuda()")
        for t in params:
            if t.dtype != torch.float32 or not t.is_contiguous() or t.device != dev:
                raise _lib.MzgoError("MuZeroNet parameters must be contiguous float32 tensors on one device")
        prec = PRECISIONS[self.precision]
        if self._ws is None:
            self._ws = {}
        slot = self._ws.get(ws_key)
        want = (int(max_batch), prec, dev)
        if slot is None or slot["key"][0] < max_batch or slot["key"][1:] != want[1:]:
            nbytes = int(_lib.load().mzgo_net_workspace_bytes(self.board_size, self.latent_dim, prec, int(max_batch)))
            slot = {"ws": torch.zeros(nbytes, dtype=torch.uint8, device=dev), "key": want, "prepared": None}
            self._ws[ws_key] = slot
        ws = slot["ws"]
        d = _lib.NetDesc()
        d.n, d.C, d.A, d.precision = self.board_size, self.latent_dim, self.action_size, prec
        d.max_batch = slot["key"][0]
        for name, t in zip(_lib.NET_PARAM_FIELDS, params):
            setattr(d, name, t.data_ptr())
        d.workspace, d.workspace_bytes = ws.data_ptr(), ws.numel()
        # tensor._version catches eager in-place updates (optimizer.step, load_state_dict, broadcast); a replayed
        # CUDA graph updates the parameters WITHOUT touching it, hence the explicit counter (mark_weights_changed)
        key = tuple((t.data_ptr(), t._version) for t in params) + (ws.data_ptr(), self._weights_version)
        if key != slot["prepared"]:
            _lib.check(_lib.load().mzgo_net_prepare(d, _lib.stream_ptr()))
            slot["prepared"] = key
        self._desc = d
        return d

    def mark_weights_changed(self):
        """Tell the network that its parameters were rewritten behind torch's version counters (a CUDA-graph
        replay of an optimizer step, a raw-pointer copy): the next forward re-derives the kernel-side layouts."""
        self._weights_version += 1

    def new_arena(self, slots):
        """Zero-initialised latent arena (the zero padding of bf16 slots is never rewritten)."""
        dev = self.prediction.pass_logit.device
        return torch.zeros(int(slots) * self.slot_bytes(), dtype=torch.uint8, device=dev)

    # ------------------------------------------------------------------ reference API
    def _check_obs(self, obs):
        n = self.board_size
        obs = obs.to(self.prediction.pass_logit.device, torch.float32).contiguous()
        if obs.dim() != 4 or tuple(obs.shape[1:]) != (6, n, n):
            raise ValueError(f"observation must be [B,6,{n},{n}], got {tuple(obs.shape)}")
        return obs

    @torch.no_grad()
    def _run_initial(self, obs, want_softmax=False):
        obs = self._check_obs(obs)
        B, n, Cc, A = obs.shape[0], self.board_size, self.latent_dim, self.action_size
        d = self.descriptor(B)
        dev = obs.device
        arena = self.new_arena(B)
        slots = torch.arange(B, dtype=torch.int32, device=dev)
        value = torch.empty((B, 1), dtype=torch.float32, device=dev)
        policy = torch.empty((B, A), dtype=torch.float32, device=dev)
        latent = torch.empty((B, Cc, n, n), dtype=torch.float32, device=dev)
        L, s = _lib.load(), _lib.stream_ptr()
        _lib.check(L.mzgo_net_initial(d, _lib.ptr(obs), None, _lib.ptr(arena), _lib.ptr(slots), _lib.ptr(value),
                                      _lib.ptr(policy), int(want_softmax), B, s))
        _lib.check(L.mzgo_net_export_latent(d, _lib.ptr(arena), _lib.ptr(slots), _lib.ptr(latent), B, s))
        return latent, value, policy

    def _import(self, latent, extra_slots):
        n, Cc = self.board_size, self.latent_dim
        latent = latent.to(self.prediction.pass_logit.device, torch.float32).contiguous()
        if latent.dim() != 4 or tuple(latent.shape[1:]) != (Cc, n, n):
            raise ValueError(f"latent must be [B,{Cc},{n},{n}], got {tuple(latent.shape)}")
        B = latent.shape[0]
        d = self.descriptor(B)
        arena = self.new_arena(B + extra_slots)
        slots = torch.arange(B, dtype=torch.int32, device=latent.device)
        _lib.check(_lib.load().mzgo_net_import_latent(d, _lib.ptr(arena), _lib.ptr(slots), _lib.ptr(latent), B,
                                                      _lib.stream_ptr()))
        return d, arena, slots, B

    @torch.no_grad()
    def _run_recurrent(self, latent, action, want_softmax=False):
        d, arena, slots, B = self._import(latent, latent.shape[0])
        dev = arena.device
        n, Cc, A = self.board_size, self.latent_dim, self.action_size
        action = action.to(dev).reshape(-1).to(torch.int32).contiguous()
        if action.numel() != B:
            raise ValueError("one action per latent expected")
        if int(action.min()) < 0 or int(action.max()) >= self.max_action_size:
            raise IndexError("action index out of range of the action embedding")   # nn.Embedding's behaviour
        out_slots = slots + B
        reward = torch.empty((B, 1), dtype=torch.float32, device=dev)
        value = torch.empty((B, 1), dtype=torch.float32, device=dev)
        policy = torch.empty((B, A), dtype=torch.float32, device=dev)
        nxt = torch.empty((B, Cc, n, n), dtype=torch.float32, device=dev)
        L, s = _lib.load(), _lib.stream_ptr()
        _lib.check(L.mzgo_net_recurrent(d, _lib.ptr(arena), _lib.ptr(slots), _lib.ptr(action), _lib.ptr(out_slots),
                                        _lib.ptr(reward), _lib.ptr(value), _lib.ptr(policy), int(want_softmax), B, s))
        _lib.check(L.mzgo_net_export_latent(d, _lib.ptr(arena), _lib.ptr(out_slots), _lib.ptr(nxt), B, s))
        return nxt, reward, value, policy

    @torch.no_grad()
    def _run_predict(self, latent):
        d, arena, slots, B = self._import(latent, 0)
        dev = arena.device
        value = torch.empty((B, 1), dtype=torch.float32, device=dev)
        policy = torch.empty((B, self.action_size), dtype=torch.float32, device=dev)
        _lib.check(_lib.load().mzgo_net_predict(d, _lib.ptr(arena), _lib.ptr(slots), _lib.ptr(value),
                                                _lib.ptr(policy), 0, B, _lib.stream_ptr()))
        return value, policy

    def initial_inference(self, observation):
        """main.py:134-137 -> (latent [B,C,N,N], value [B,1], policy_logits [B,A])"""
        return self._run_initial(observation)

    def recurrent_inference(self, latent, action):
        """main.py:139-142 -> (next_latent, reward [B,1], value [B,1], policy_logits [B,A])"""
        return self._run_recurrent(latent, action)
