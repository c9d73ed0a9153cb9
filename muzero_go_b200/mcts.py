"""Drop-in for the reference's per-move search: MCTS(net, action_size, num_simulations, c_puct).run(obs)
(/root/reference/main.py:247-368, play.py:104-216; variant="self_play": self_play.py:142-343).  The tree lives on the GPU (csrc/tree.cu); run()
returns a read-only MCTSNode view exposing what the callers walk (main.py:666-669, play.py:234-239):
root.visit_count, root.value(), root.children[a]['node'].visit_count, root.children[a]['prior'].
"""
import numpy as np
import torch

from . import _lib
from .rules import GoBoards, govars
from .tree import SearchConfig, Tree


class MCTSNode:
    """Lazy view of one node of a finished search (fields of main.py:145-155)."""

    def __init__(self, snap, index, prior=0):
        self._snap, self._index = snap, index
        self.prior = prior
        self.terminal = False
        self._children = None

    @property
    def visit_count(self):
        s = self._snap
        if self._index == 0:
            return int(s["root_visit"])
        pk = int(s["parent"][self._index])
        return int(s["edge_visit"][pk >> 9, pk & 511])

    @property
    def value_sum(self):
        s = self._snap
        if self._index == 0:
            return float(s["root_vsum"])
        pk = int(s["parent"][self._index])
        return float(s["edge_vsum"][pk >> 9, pk & 511])

    def value(self):
        return self.value_sum / self.visit_count if self.visit_count > 0 else 0

    @property
    def latent(self):
        return self._snap["latent"](self._index)

    @property
    def children(self):
        if self._children is None:
            s, i = self._snap, self._index
            priors = s["root_prior"] if i == 0 else s["prior"][i]
            kids = {}
            for a in range(s["A"]):
                c = int(s["child"][i, a])
                pr = priors[a]
                kids[a] = {"node": MCTSNode(s, c, prior=0) if c >= 0 else None,
                           "prior": pr if pr > 0 else 0, "visit_count": 0, "value_sum": 0, "action": a}
            self._children = kids
        return self._children


class MCTS:
    """main.py:247-316.  `config` supplies discount / dirichlet_epsilon / dirichlet_alpha / pass_epsilon
    (the reference reads them from its module-global config, main.py:264,276,311)."""

    def __init__(self, muzero_net, action_size, num_simulations, c_puct=None, config=None, seed=0, variant="main"):
        self.net = muzero_net
        self.net.eval()
        self.action_size = int(action_size)
        self.num_simulations = int(num_simulations)
        self.variant = variant
        if c_puct is None:                  # main.py:248 (2), play.py:116 (10.0), self_play.py:143 (2.5)
            c_puct = SearchConfig.for_variant(variant).c_puct
        self.c_puct = c_puct
        cfg = SearchConfig.for_variant(variant, c_puct=c_puct, seed=seed)
        for k in ("discount", "dirichlet_epsilon", "dirichlet_alpha", "pass_epsilon"):
            if config is not None and hasattr(config, k):
                setattr(cfg, k, getattr(config, k))
        self.search = cfg
        self._calls = 0
        self._engine = None

    def _get_engine(self):
        if self._engine is None:
            from .selfplay import SelfPlayEngine
            n = self.net.board_size
            if n * n + 1 != self.action_size:
                raise ValueError("action_size must be board_size**2 + 1")
            self._engine = SelfPlayEngine(self.net, n, 1, self.num_simulations, search=self.search,
                                          restart_finished=False, traj_capacity=1)
        return self._engine

    @torch.no_grad()
    def run(self, observation, noise=None, choice_stream=None):
        """observation: GymGo planes [6,N,N].  noise (float64 [A]) / choice_stream (uint32 [k]) replace the
        reference's np.random.dirichlet / random.choice draws (parity tests); None uses the device RNG."""
        eng = self._get_engine()
        n, A = eng.n, eng.A
        dev = eng.device
        obs = torch.as_tensor(np.asarray(observation, dtype=np.float32)).reshape(1, 6, n, n).to(dev)
        eng.boards.load_planes(obs)
        d = eng.descriptor()
        L, s = _lib.load(), _lib.stream_ptr()
        slot0 = torch.zeros(1, dtype=torch.int32, device=dev)
        eng.tree.active.fill_(1)
        _lib.check(L.mzgo_net_initial(d.net, _lib.ptr(obs), None, _lib.ptr(eng.arena), _lib.ptr(slot0),
                                      _lib.ptr(eng.root_value_net), _lib.ptr(eng.root_policy), 1, 1, s))
        nz = None if noise is None else torch.as_tensor(np.asarray(noise, dtype=np.float64)).reshape(1, A)
        eng.tree.reset_roots(self.search, eng.root_policy, eng.boards, n, noise=nz, move_index=self._calls)
        d.tree.move_index = self._calls
        stream_t, stride = None, 0
        if choice_stream is not None:
            stream_t = Tree.stream_tensor(np.asarray(choice_stream).reshape(1, -1), dev)
            stride = stream_t.shape[1]
        _lib.check(L.mzgo_search_run(d, _lib.ptr(stream_t), stride, s))
        self._calls += 1
        root = MCTSNode(self._snapshot(eng), 0)
        if self.variant == "self_play":
            # self_play.py:233-237 returns (root, visit_counts, root_value).  The reference's vector reads the
            # edge dicts' dead 'visit_count' field and is always zero (defect 1); the counts
            # returned here are the children's, i.e. what the agent was meant to receive.
            visit_counts = np.array([c["node"].visit_count if c["node"] is not None else 0
                                     for c in root.children.values()])
            return root, visit_counts, root.value()
        return root

    def _snapshot(self, eng):
        t = eng.tree
        net, arena, dev = self.net, eng.arena, eng.device

        def latent(index):
            desc = net.descriptor(1)
            out = torch.empty((1, net.latent_dim, net.board_size, net.board_size), dtype=torch.float32, device=dev)
            slot = torch.tensor([index], dtype=torch.int32, device=dev)
            _lib.check(_lib.load().mzgo_net_export_latent(desc, _lib.ptr(arena), _lib.ptr(slot), _lib.ptr(out), 1,
                                                          _lib.stream_ptr()))
            return out[0]
        return {"A": t.A, "root_visit": t.root_visit[0].item(), "root_vsum": t.root_vsum[0].item(),
                "root_prior": t.root_prior[0].cpu().numpy(), "prior": t.prior[0].cpu().numpy(),
                "edge_visit": t.edge_visit[0].cpu().numpy(), "edge_vsum": t.edge_vsum[0].cpu().numpy(),
                "child": t.child[0].cpu().numpy(), "parent": t.parent[0].cpu().numpy(),
                "root_mask": t.root_mask[0].cpu().numpy(), "dropped": int(t.dropped[0].item()),
                "depth_sum": int(t.depth_sum[0].item()), "sims_done": int(t.sims_done[0].item()), "latent": latent}


def valid_mask_from_observation(observation, pass_epsilon):
    """main.py:659-665 (host helper for callers that rebuild the mask, e.g. agents)."""
    valid_board = (np.asarray(observation)[govars.INVD_CHNL].flatten() == 0).astype(np.float32)
    pass_prior = pass_epsilon if valid_board.sum() > 0 else 1.0
    return np.concatenate([valid_board, np.array([pass_prior])])
