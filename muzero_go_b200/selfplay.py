"""Batched self-play: G concurrent games in lock step on one GPU.

Host owner of the buffers behind mzgo_selfplay_t.  One call of run(k) enqueues k moves for every game:
MCTS.run + visit-count action + env.step + trajectory append, i.e. the loop body of
/root/reference/main.py:656-702 (self_play.py:465-497), with termination by two passes / env done /
move cap (main.py:696-706) and immediate restart of finished games.  Nothing synchronises with the host
until results are read.
"""
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
                 device=None):
        _lib.require_device()
        self.net = net
        self.n = int(board_size)
        if net.board_size != self.n:
            raise ValueError("network board size differs from the engine's")
        self.G, self.S = int(num_games), int(num_simulations)
        self.A = self.n * self.n + 1
        self.M = self.S + 1
        self.search = search or SearchConfig()
        dev = self.device = _lib.norm_device(device if device is not None else net.prediction.pass_logit.device)
        self.max_game_moves = int(self.n * self.n * 1.5) if max_game_moves is None else int(max_game_moves)  # main.py:53
        self.temperature, self.temperature_moves = float(temperature), int(temperature_moves)
        self.komi, self.restart_finished = float(komi), bool(restart_finished)
        self.T = int(traj_capacity)
        G, A, T = self.G, self.A, self.T

        self.boards = GoBoards(G, self.n, dev).reset()
        self.tree = Tree(G, A, self.M, dev)
        self.arena = net.new_arena(G * self.M)

        def z(shape, dtype):
            return torch.zeros(shape, dtype=dtype, device=dev)
        self.root_policy = z((G, A), torch.float32)
        self.root_value_net = z((G,), torch.float32)
        self.leaf_policy = z((G, A), torch.float32)
        self.leaf_reward = z((G,), torch.float32)
        self.leaf_value = z((G,), torch.float32)
        self.in_slot = z((G,), torch.int32)
        self.out_slot = z((G,), torch.int32)
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
        d.net = self.net.descriptor(self.G)
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

    def read_counters(self):
        """Synchronises.  -> dict of totals since construction."""
        return dict(zip(COUNTER_NAMES, self.counters.cpu().tolist()))

    def trajectory_rows(self, lo, hi):
        """Rows [lo,hi) of the trajectory buffers as a dict of device tensors (views)."""
        return {"boards": self.traj_boards[lo:hi], "action": self.traj_action[lo:hi],
                "policy": self.traj_policy[lo:hi], "root_value": self.traj_root_value[lo:hi],
                "reward": self.traj_reward[lo:hi], "done": self.traj_done[lo:hi], "winner": self.traj_winner[lo:hi]}
