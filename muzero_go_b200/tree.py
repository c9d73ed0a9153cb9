"""Structure-of-arrays MCTS trees on the GPU: host-side owner of the mzgo_tree_t arrays.

The arithmetic (root preparation, select, expand, backup; /root/reference/main.py:255-368) runs in
csrc/tree.cu.  This class only allocates the torch tensors behind the descriptor and forwards calls.
"""
from dataclasses import dataclass

import numpy as np
import torch

from . import _lib


@dataclass
class SearchConfig:
    """The search hyper-parameters the reference reads from its module-global `config`
    (main.py:37,44-45,52) and from MCTS.__init__ (main.py:248)."""
    c_puct: float = 2
    discount: float = 0.99
    dirichlet_epsilon: float = 0.25
    dirichlet_alpha: float = 0.03
    pass_epsilon: float = 0.05
    seed: int = 0
    game_base: int = 0        # global id of this engine's game 0 (RNG key), so sharding does not change the games
    variant: str = "main"     # "main" (main.py / play.py) or "self_play" (self_play.py:142-343), SURVEY 2.2

    VARIANTS = {"main": 0, "play": 0, "self_play": 1}

    @classmethod
    def for_variant(cls, variant, **kw):
        """The reference's per-file defaults."""
        defaults = {"main": dict(c_puct=2, dirichlet_alpha=0.03, dirichlet_epsilon=0.25, pass_epsilon=0.05),
                    "play": dict(c_puct=10.0, dirichlet_alpha=0.01, dirichlet_epsilon=0.02, pass_epsilon=0.01),
                    "self_play": dict(c_puct=2.5, dirichlet_alpha=0.15, dirichlet_epsilon=0.02, pass_epsilon=0.01)}[variant]
        defaults.update(kw)
        return cls(variant=variant, **defaults)

    def to_struct(self):
        p = _lib.SearchParams()
        p.c_puct = float(self.c_puct)
        p.discount = float(self.discount)
        p.dirichlet_epsilon = float(self.dirichlet_epsilon)
        p.pass_epsilon = float(self.pass_epsilon)
        p.dirichlet_alpha = float(self.dirichlet_alpha)
        p.c_puct_is_int = int(isinstance(self.c_puct, int))
        p.seed = int(self.seed) & (2 ** 64 - 1)
        if self.variant not in self.VARIANTS:
            raise ValueError(f"unknown search variant {self.variant!r}")
        p.variant = self.VARIANTS[self.variant]
        p.game_base = int(self.game_base)
        return p


class Tree:
    """G search trees with node capacity M over an action space of size A."""

    def __init__(self, num_games, action_size, capacity, device="cuda"):
        _lib.require_device()
        G, A, M = int(num_games), int(action_size), int(capacity)
        if not 2 <= A <= 384:
            raise ValueError("action_size outside [2,384]")
        if not 2 <= M <= 32767:
            raise ValueError("capacity outside [2,32767]")
        self.G, self.A, self.M = G, A, M
        dev = self.device = _lib.norm_device(device)

        def z(shape, dtype):
            return torch.zeros(shape, dtype=dtype, device=dev)
        self.root_prior = z((G, A), torch.float64)
        self.root_mask = z((G, A), torch.float32)
        self.prior = z((G, M, A), torch.float32)
        self.edge_visit = z((G, M, A), torch.int32)
        self.edge_vsum = z((G, M, A), torch.float64)
        self.child = torch.full((G, M, A), -1, dtype=torch.int16, device=dev)
        self.parent = torch.full((G, M), -1, dtype=torch.int32, device=dev)
        self.root_visit = z((G,), torch.int32)
        self.root_vsum = z((G,), torch.float64)
        self.n_nodes = z((G,), torch.int32)
        self.leaf_parent = torch.full((G,), -1, dtype=torch.int32, device=dev)
        self.leaf_action = torch.full((G,), -1, dtype=torch.int32, device=dev)
        self.leaf_node = z((G,), torch.int32)
        self.leaf_depth = z((G,), torch.int32)
        self.leaf_in_slot = torch.full((G,), -1, dtype=torch.int32, device=dev)
        self.leaf_out_slot = torch.full((G,), -1, dtype=torch.int32, device=dev)
        self.choice_ctr = z((G,), torch.int32)
        self.depth_sum = z((G,), torch.int64)
        self.sims_done = z((G,), torch.int32)
        self.dropped = z((G,), torch.int32)
        self.active = torch.ones((G,), dtype=torch.uint8, device=dev)
        self.move_index_dev = z((1,), torch.int32)
        self.desc = self._make_desc()

    def _make_desc(self):
        d = _lib.TreeDesc()
        d.G, d.A, d.M, d.move_index = self.G, self.A, self.M, 0
        for name, _ in _lib.TreeDesc._fields_[4:]:
            setattr(d, name, getattr(self, name).data_ptr())
        return d

    def nbytes(self):
        return sum(getattr(self, n).numel() * getattr(self, n).element_size() for n, _ in _lib.TreeDesc._fields_[4:])

    # ---- main.py:259-288
    def reset_roots(self, cfg, root_soft, boards, board_size, noise=None, move_index=0):
        root_soft = root_soft.to(self.device, torch.float32).contiguous()
        assert tuple(root_soft.shape) == (self.G, self.A)
        if noise is not None:
            noise = torch.as_tensor(np.asarray(noise) if not torch.is_tensor(noise) else noise)
            noise = noise.to(self.device, torch.float64).contiguous()
            assert tuple(noise.shape) == (self.G, self.A)
        self.desc.move_index = int(move_index)
        p = cfg.to_struct()
        _lib.check(_lib.load().mzgo_tree_reset_roots(self.desc, p, _lib.ptr(root_soft), _lib.ptr(boards.records),
                                                     int(board_size), _lib.ptr(noise), int(move_index),
                                                     _lib.stream_ptr()))

    # ---- main.py:318-363
    def select(self, cfg, choice_stream=None):
        stride = 0
        if choice_stream is not None:
            assert choice_stream.dtype == torch.int32 and choice_stream.is_contiguous()
            stride = choice_stream.shape[1]
        _lib.check(_lib.load().mzgo_tree_select(self.desc, cfg.to_struct(), _lib.ptr(choice_stream), stride,
                                                _lib.stream_ptr()))

    # ---- main.py:300-313, :365-368
    def expand_backup(self, cfg, soft, reward, value):
        for t in (soft, reward, value):
            assert t.dtype == torch.float32 and t.is_contiguous() and t.device == self.device
        _lib.check(_lib.load().mzgo_tree_expand_backup(self.desc, cfg.to_struct(), _lib.ptr(soft), _lib.ptr(reward),
                                                       _lib.ptr(value), _lib.stream_ptr()))

    # ---- the fused per-simulation step the engine launches (mzgo_tree_step): expansion + backup of the leaf the
    # previous step selected, then the next selection
    def step(self, cfg, soft, reward, value, choice_stream=None, do_backup=True, do_select=True):
        for t in (soft, reward, value):
            assert t.dtype == torch.float32 and t.is_contiguous() and t.device == self.device
        stride = 0
        if choice_stream is not None:
            assert choice_stream.dtype == torch.int32 and choice_stream.is_contiguous()
            stride = choice_stream.shape[1]
        _lib.check(_lib.load().mzgo_tree_step(self.desc, cfg.to_struct(), _lib.ptr(soft), _lib.ptr(reward),
                                              _lib.ptr(value), _lib.ptr(choice_stream), stride, int(bool(do_backup)),
                                              int(bool(do_select)), _lib.stream_ptr()))

    # ---- main.py:666-669
    def visit_counts(self):
        out = torch.empty((self.G, self.A), dtype=torch.int32, device=self.device)
        _lib.check(_lib.load().mzgo_tree_visit_counts(self.desc, _lib.ptr(out), _lib.stream_ptr()))
        return out

    # ---- main.py:670-677 / self_play.py:389-400
    def choose_action(self, cfg, temperature=0.0, move_index=0):
        action = torch.empty((self.G,), dtype=torch.int32, device=self.device)
        policy = torch.empty((self.G, self.A), dtype=torch.float32, device=self.device)
        root_value = torch.empty((self.G,), dtype=torch.float32, device=self.device)
        _lib.check(_lib.load().mzgo_tree_choose_action(self.desc, cfg.to_struct(), float(temperature), None,
                                                       int(move_index), _lib.ptr(action), _lib.ptr(policy),
                                                       _lib.ptr(root_value), _lib.stream_ptr()))
        return action, policy, root_value

    @staticmethod
    def stream_tensor(stream_u32, device="cuda"):
        """uint32 numpy [G,stride] -> int32 device tensor with the same bits."""
        a = np.ascontiguousarray(stream_u32, dtype=np.uint32).view(np.int32)
        return torch.from_numpy(a).to(device).contiguous()
