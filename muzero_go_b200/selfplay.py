"""Batched self-play: G concurrent games in lock step on one GPU.

Host owner of the buffers behind mzgo_selfplay_t.  One call of run(k) enqueues k moves for every game:
MCTS.run + visit-count action + env.step + trajectory append, i.e. the loop body of
/root/reference/main.py:656-702 (self_play.py:465-497), with termination by two passes / env done /
move cap (main.py:696-706) and immediate restart of finished games.  Nothing synchronises with the host
until results are read.
"""
import dataclasses

import numpy as np
import torch

from . import _lib
from .rules import GoBoards
from .tree import SearchConfig, Tree

COUNTER_NAMES = ["moves", "simulations", "games_finished", "dropped_simulations", "leaf_depth_sum", "black_wins",
                 "white_wins", "move_cap_endings"]


class SelfPlayEngine:
    def __init__(self, net, board_size, num_games, num_simulations, search=None, max_game_moves=None,
                 temperature=0.0, temperature_moves=0, komi=0.0, restart_finished=True, traj_capacity=8,
                 device=None, share_search_buffers_with=None, ws_key=0):
        _lib.require_device()
        self.net = net
        self.n = int(board_size)
        if net.board_size != self.n:
            raise ValueError("network board size differs from the engine's")
        self.G, self.S = int(num_games), int(num_simulations)
        self.A = self.n * self.n + 1
        self.M = self.S + 1
        self.search = search or SearchConfig()
        self.ws_key = ws_key          # which network workspace (weight layouts + scratch) this engine evaluates with
        dev = self.device = _lib.norm_device(device if device is not None else net.prediction.pass_logit.device)
        self.max_game_moves = int(self.n * self.n * 1.5) if max_game_moves is None else int(max_game_moves)  # main.py:53
        self.temperature, self.temperature_moves = float(temperature), int(temperature_moves)
        self.komi, self.restart_finished = float(komi), bool(restart_finished)
        self.T = int(traj_capacity)
        G, A, T = self.G, self.A, self.T

        self.boards = GoBoards(G, self.n, dev).reset()

        def z(shape, dtype):
            return torch.zeros(shape, dtype=dtype, device=dev)
        other = share_search_buffers_with
        if other is None:
            self.tree = Tree(G, A, self.M, dev)
            self.arena = net.new_arena(G * self.M)
            self.root_policy = z((G, A), torch.float32)
            self.root_value_net = z((G,), torch.float32)
            self.leaf_policy = z((G, A), torch.float32)
            self.leaf_reward = z((G,), torch.float32)
            self.leaf_value = z((G,), torch.float32)
            self.in_slot = z((G,), torch.int32)
            self.out_slot = z((G,), torch.int32)
        else:
            # A search only lives for one move: engines that run their moves one after the other on one stream
            # (the waves of a WaveGroup) can share the trees, the latent arena and the per-simulation scratch.
            if (other.G, other.S, other.n, other.net) != (self.G, self.S, self.n, net) or other.device != dev:
                raise ValueError("engines sharing search buffers must have the same shape, network and device")
            for name in ("tree", "arena", "root_policy", "root_value_net", "leaf_policy", "leaf_reward", "leaf_value",
                         "in_slot", "out_slot"):
                setattr(self, name, getattr(other, name))
        self.action = z((G,), torch.int32)
        self.move_count = z((G,), torch.int32)
        self.pass_streak = z((G,), torch.int32)
        self.game_temperature = z((G,), torch.float32)
        self.score_winner = z((G,), torch.int32)
        self.traj_boards = z((T, G, _lib.BOARD_WORDS), torch.int32)
        self.traj_action = torch.full((T, G), -1, dtype=torch.int32, device=dev)
        self.traj_policy = z((T, G, A), torch.float32)
        self.traj_root_value = z((T, G), torch.float32)
        self.traj_reward = z((T, G), torch.float32)
        self.traj_done = z((T, G), torch.uint8)
        self.traj_winner = z((T, G), torch.int8)
        self.counters = z((8,), torch.int64)
        self.moves_enqueued = 0
        self._desc = None

    def memory_bytes(self):
        return self.arena.numel() + self.tree.nbytes()

    def descriptor(self):
        d = _lib.SelfPlayDesc()
        d.tree = self.tree.desc
        d.params = self.search.to_struct()
        d.net = self.net.descriptor(self.G, self.ws_key)
        d.n, d.num_simulations, d.max_game_moves = self.n, self.S, self.max_game_moves
        d.temperature_moves, d.temperature, d.komi = self.temperature_moves, self.temperature, self.komi
        d.restart_finished, d.traj_capacity = int(self.restart_finished), self.T
        d.boards = self.boards.records.data_ptr()
        d.arena = self.arena.data_ptr()
        for name in ("root_policy", "root_value_net", "leaf_policy", "leaf_reward", "leaf_value", "in_slot",
                     "out_slot", "action", "move_count", "pass_streak", "game_temperature", "score_winner",
                     "traj_boards", "traj_action", "traj_policy", "traj_root_value", "traj_reward", "traj_done",
                     "traj_winner", "counters"):
            setattr(d, name, getattr(self, name).data_ptr())
        self._desc = d
        return d

    def run(self, num_moves, traj_offset=0):
        """Enqueue `num_moves` lock-step moves; trajectory rows [traj_offset, traj_offset+num_moves)."""
        d = self.descriptor()
        _lib.check(_lib.load().mzgo_selfplay_run(d, int(num_moves), int(traj_offset), self.moves_enqueued,
                                                 _lib.stream_ptr()))
        self.moves_enqueued += int(num_moves)

    def pack_rows(self, lo, hi, out, g_offset=0):
        """Trajectory rows [lo,hi) in the wire layout of the trajectory gather (dist.ROW_FIELDS) into
        out[0:hi-lo, g_offset:g_offset+G, :], a uint8 [rows, G_total, stride] device tensor: ONE launch
        (mzgo_selfplay_pack_rows), no torch.cat."""
        assert out.dtype == torch.uint8 and out.is_contiguous() and out.dim() == 3 and out.shape[0] >= hi - lo
        assert out.shape[2] == 270 + 4 * self.A
        d = self._desc if self._desc is not None else self.descriptor()
        _lib.check(_lib.load().mzgo_selfplay_pack_rows(d, int(lo), int(hi - lo), _lib.ptr(out), int(g_offset),
                                                       int(out.shape[1]), _lib.stream_ptr()))

    def read_counters(self):
        """Synchronises.  -> dict of totals since construction."""
        return dict(zip(COUNTER_NAMES, self.counters.cpu().tolist()))

    def trajectory_rows(self, lo, hi):
        """Rows [lo,hi) of the trajectory buffers as a dict of device tensors (views)."""
        return {"boards": self.traj_boards[lo:hi], "action": self.traj_action[lo:hi],
                "policy": self.traj_policy[lo:hi], "root_value": self.traj_root_value[lo:hi],
                "reward": self.traj_reward[lo:hi], "done": self.traj_done[lo:hi], "winner": self.traj_winner[lo:hi]}


class WaveGroup:
    """`waves` engines of num_games / waves games each.  Every call of run(k) plays k moves of every game.

    Waves that run on the same stream take turns on ONE set of search buffers (trees, latent arena, scratch): a search
    only lives for one move.  This is how a game count whose latent arena would not fit the HBM runs on one GPU
    (19x19, 800 simulations: 2048 games need 177 GB of latents at once).

    `streams` > 1 runs the waves round-robin on that many CUDA streams, each with its own set of search buffers, so
    that the kernels of different waves overlap: the per-simulation loop of a wave is a chain [tree step -> network
    kernel -> tree step ...] in which the latency-bound tree step (three dependent memory phases, ~15 % of a 19x19
    simulation) leaves the SMs idle; with two waves in flight one wave's tree step runs under the other wave's
    network kernel (SURVEY section 7 hard part 4).  The games are independent and the RNG is keyed
    by the global game id, so the moves do not depend on how the games are cut into waves or streams."""

    def __init__(self, net, board_size, num_games, num_simulations, waves=1, search=None, streams=1, **kw):
        if num_games % waves:
            raise ValueError("num_games must be a multiple of waves")
        streams = max(1, min(int(streams), waves))
        if waves % streams:
            raise ValueError("waves must be a multiple of streams")
        per = num_games // waves
        self.engines = []
        for w in range(waves):
            # one RNG key space for the whole group: wave w owns the global game ids [base + w*per, base + (w+1)*per)
            base_cfg = search if search is not None else SearchConfig()
            cfg = dataclasses.replace(base_cfg, game_base=base_cfg.game_base + w * per)
            owner = self.engines[w % streams] if w >= streams else None      # the first wave of this wave's stream
            self.engines.append(SelfPlayEngine(net, board_size, per, num_simulations, search=cfg,
                                               share_search_buffers_with=owner, ws_key=w % streams, **kw))
        e0 = self.engines[0]
        self.G, self.S, self.A, self.n, self.M = num_games, e0.S, e0.A, e0.n, e0.M
        self.arena, self.tree, self.device = e0.arena, e0.tree, e0.device
        self.streams = [torch.cuda.Stream(device=self.device) for _ in range(streams)] if streams > 1 else []

    def memory_bytes(self):
        return sum(e.memory_bytes() for e in self.engines[:max(1, len(self.streams))])

    def run(self, num_moves, traj_offset=0):
        if not self.streams:
            for m in range(int(num_moves)):
                for e in self.engines:
                    e.run(1, traj_offset=traj_offset + m)
            return
        # fork: every side stream starts behind the caller's stream; join: the caller's stream continues behind all
        main = torch.cuda.current_stream(self.device)
        start = torch.cuda.Event()
        start.record(main)
        ns = len(self.streams)
        for k, st in enumerate(self.streams):
            st.wait_event(start)
            with torch.cuda.stream(st):
                for m in range(int(num_moves)):
                    for e in self.engines[k::ns]:
                        e.run(1, traj_offset=traj_offset + m)
            main.wait_stream(st)

    def read_counters(self):
        tot = dict.fromkeys(COUNTER_NAMES, 0)
        for e in self.engines:
            for k, v in e.read_counters().items():
                tot[k] += v
        return tot

    def trajectory_rows(self, lo, hi):
        rows = [e.trajectory_rows(lo, hi) for e in self.engines]
        return {k: torch.cat([r[k] for r in rows], dim=1) for k in rows[0]}

    def pack_rows(self, lo, hi, out):
        """All waves' rows [lo,hi) into out (uint8 [rows, num_games, stride], wire layout): one launch per wave."""
        g0 = 0
        for e in self.engines:
            e.pack_rows(lo, hi, out, g_offset=g0)
            g0 += e.G
