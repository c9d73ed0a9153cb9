"""Go rules: batched packed boards on the GPU and a GymGo-compatible single-game env.

Host mirror of the reference's use of GymGo (gym.make("gym_go:go-v0", size, komi=0, reward_method='real'),
/root/reference/main.py:628; reset/step/winner at main.py:637,681,709; planes at main.py:261).  All rule
arithmetic runs in libmzgo.so (csrc/rules.cu).
"""
import numpy as np
import torch

from . import _lib


class govars:
    """gym_go.govars (channel indices of the [6,N,N] state)."""
    ANYONE = None
    NOONE = -1
    BLACK = 0
    WHITE = 1
    TURN_CHNL = 2
    INVD_CHNL = 3
    PASS_CHNL = 4
    DONE_CHNL = 5
    NUM_CHNLS = 6


_DTYPES = {torch.float32: 0, torch.float64: 1, torch.uint8: 2}


class GoBoards:
    """`num` packed positions (uint32 [num,64] records, see include/mzgo.h) resident on one GPU."""

    def __init__(self, num, size, device="cuda"):
        _lib.require_device()
        if not 2 <= size <= _lib.MAX_N:
            raise ValueError(f"board size {size} outside [2,{_lib.MAX_N}]")
        self.num, self.size = int(num), int(size)
        self.device = _lib.norm_device(device)
        self.records = torch.zeros((self.num, _lib.BOARD_WORDS), dtype=torch.int32, device=self.device)

    @property
    def action_size(self):
        return self.size * self.size + 1

    def reset(self):
        _lib.check(_lib.load().mzgo_rules_reset(_lib.ptr(self.records), self.num, self.size, _lib.stream_ptr()))
        return self

    def step(self, actions, return_status=True):
        """actions: int32 [num] on the device (negative = leave the board alone)."""
        actions = torch.as_tensor(actions, dtype=torch.int32, device=self.device).contiguous()
        assert actions.numel() == self.num
        status = torch.empty(self.num, dtype=torch.int32, device=self.device) if return_status else None
        _lib.check(_lib.load().mzgo_rules_step(_lib.ptr(self.records), _lib.ptr(actions), _lib.ptr(status),
                                               self.num, self.size, _lib.stream_ptr()))
        return status

    def score(self, komi=0.0):
        """-> (areas int32 [num,2] (black, white), winner int32 [num] from black's point of view)."""
        areas = torch.empty((self.num, 2), dtype=torch.int32, device=self.device)
        winner = torch.empty(self.num, dtype=torch.int32, device=self.device)
        _lib.check(_lib.load().mzgo_rules_score(_lib.ptr(self.records), _lib.ptr(areas), _lib.ptr(winner),
                                                float(komi), self.num, self.size, _lib.stream_ptr()))
        return areas, winner

    def planes(self, dtype=torch.float32):
        """GymGo state planes [num,6,N,N]."""
        out = torch.empty((self.num, 6, self.size, self.size), dtype=dtype, device=self.device)
        _lib.check(_lib.load().mzgo_rules_expand_planes(_lib.ptr(self.records), _lib.ptr(out), _DTYPES[dtype],
                                                        self.num, self.size, _lib.stream_ptr()))
        return out

    def load_planes(self, planes):
        """Inverse of planes(): planes float32 [num,6,N,N] (device or host)."""
        planes = torch.as_tensor(np.asarray(planes) if not torch.is_tensor(planes) else planes)
        planes = planes.to(device=self.device, dtype=torch.float32).contiguous()
        assert tuple(planes.shape) == (self.num, 6, self.size, self.size), planes.shape
        _lib.check(_lib.load().mzgo_rules_pack_planes(_lib.ptr(planes), _lib.ptr(self.records), self.num, self.size,
                                                      _lib.stream_ptr()))
        return self

    def playout(self, R, want_actions=True, want_hash=False, want_trace=False):
        """BASELINE.json configs[2] random playout; R uint32 [num,T] (as a numpy array or int64/int32 tensor)."""
        if not torch.is_tensor(R):
            R = torch.from_numpy(np.ascontiguousarray(R, dtype=np.uint32).view(np.int32))
        R = R.to(self.device).contiguous()
        T = R.shape[1]
        acts = torch.empty((self.num, T), dtype=torch.int32, device=self.device) if want_actions else None
        hsh = torch.empty((self.num, T), dtype=torch.int64, device=self.device) if want_hash else None
        trace = (torch.empty((self.num, T, 6, self.size, self.size), dtype=torch.uint8, device=self.device)
                 if want_trace else None)
        _lib.check(_lib.load().mzgo_rules_playout(_lib.ptr(self.records), _lib.ptr(R), T, _lib.ptr(acts),
                                                  _lib.ptr(hsh), _lib.ptr(trace), self.num, self.size,
                                                  _lib.stream_ptr()))
        return acts, hsh, trace

    def done(self):
        return ((self.records[:, 61] >> 1) & 1).bool()

    def turn(self):
        return self.records[:, 60] & 1


class GoEnv:
    """Duck-typed gym_go GoEnv (reward_method='real'), one game, state on the GPU.

    reset() / step() return numpy float64 [6,N,N] like GymGo, so /root/reference/main.py:637-711,
    self_play.py:453-526 and play.py:283-348 run against it unchanged."""

    govars = govars

    def __init__(self, size, komi=0, reward_method="real", device="cuda"):
        if reward_method != "real":
            raise ValueError("only reward_method='real' (the reference's setting, main.py:628) is supported")
        self.size, self.komi = int(size), komi
        self.boards = GoBoards(1, size, device)
        self.done = False
        self._state = None

    def _sync(self):
        self._state = self.boards.planes(torch.float64)[0].cpu().numpy()
        self.done = bool(self._state[govars.DONE_CHNL, 0, 0] == 1)
        return np.copy(self._state)

    def reset(self):
        self.boards.reset()
        return self._sync()

    def step(self, action):
        assert not self.done
        if isinstance(action, (tuple, list, np.ndarray)):
            assert 0 <= action[0] < self.size and 0 <= action[1] < self.size
            action = self.size * int(action[0]) + int(action[1])
        elif action is None:
            action = self.size ** 2
        status = self.boards.step(torch.tensor([int(action)], dtype=torch.int32))
        st = int(status.item())
        assert st == 0, ("Invalid move", action, st)
        state = self._sync()
        return state, self.reward(), self.done, self.info()

    def game_ended(self):
        return self.done

    def turn(self):
        return int(self._state[govars.TURN_CHNL, 0, 0])

    def state(self):
        return np.copy(self._state)

    def invalid_moves(self):
        if self.done:
            return np.zeros(self.size ** 2 + 1)
        return np.append(self._state[govars.INVD_CHNL].flatten(), 0)

    def valid_moves(self):
        return 1 - self.invalid_moves()

    def winning(self):
        _, winner = self.boards.score(self.komi)
        return int(winner.item())

    def winner(self):
        return self.winning() if self.done else 0

    def reward(self):
        return self.winner()

    def info(self):
        return {"turn": self.turn(), "invalid_moves": self.invalid_moves(),
                "prev_player_passed": bool(self._state[govars.PASS_CHNL, 0, 0] == 1)}


def make(env_id="gym_go:go-v0", size=19, komi=0, reward_method="real", **_):
    """Stand-in for gym.make("gym_go:go-v0", ...) (main.py:628, self_play.py:544, play.py:283)."""
    if not str(env_id).endswith("go-v0"):
        raise ValueError(f"unknown env id {env_id!r}")
    return GoEnv(size, komi=komi, reward_method=reward_method)
