"""Oracle: CPU restatement of the GymGo rules engine (TEST INFRASTRUCTURE).

PARITY UNPINNED.  The reference calls an external package
(huangeddie/GymGo, submodule pinned at 3a3d97bea329ac263bd43990dce73dea8d57534b,
/root/reference/.SUBMODULES.json:9-15, status "missing").  Its source is not
in /root/reference and not installed, so this file restates the *published*
algorithm of gym_go/gogame.py, gym_go/state_utils.py, gym_go/govars.py and
gym_go/envs/go_env.py.  It is anchored on the reference's call sites:

  gym.make("gym_go:go-v0", size, komi=0, reward_method='real')  main.py:628
  env.reset() -> [6,N,N] float64                                 main.py:637
  env.step(a) -> (state, reward, done, info)                     main.py:681
  env.winner()                                                   main.py:709
  observation[govars.INVD_CHNL]                                  main.py:261

and on hand-written known-answer Go positions in tests/test_rules_oracle.py.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
import this module.
"""
import numpy as np
from scipy import ndimage


# ---- gym_go/govars.py ------------------------------------------------------
class govars:
    ANYONE = None
    NOONE = -1
    BLACK = 0
    WHITE = 1
    TURN_CHNL = 2
    INVD_CHNL = 3
    PASS_CHNL = 4
    DONE_CHNL = 5
    NUM_CHNLS = 6


# 4-neighbourhood without the centre (state_utils.surround_struct)
_SURROUND = np.array([[0, 1, 0],
                      [1, 0, 1],
                      [0, 1, 0]])
_NEIGHBOR_DELTAS = np.array([[-1, 0], [1, 0], [0, -1], [0, 1]])


# ---- gym_go/state_utils.py ---------------------------------------------------
def _adj_data(state, action2d, player):
    """In-board 4-neighbours of the new stone, and whether all of them hold
    opponent stones (evaluated BEFORE captures)."""
    size = state.shape[1]
    nb = _NEIGHBOR_DELTAS + np.asarray(action2d)
    ok = np.prod((nb >= 0) & (nb < size), axis=1)
    nb = nb[np.nonzero(ok)]
    opp = state[1 - player]
    surrounded = bool((opp[nb[:, 0], nb[:, 1]] > 0).all())
    return nb, surrounded


def _update_pieces(state, adj_locs, player):
    """Remove every opponent group that touches adj_locs and has no liberty.
    Liberties are measured once, on the board after the stone was placed."""
    opponent = 1 - player
    killed = []
    all_pieces = state[govars.BLACK] + state[govars.WHITE]
    empties = 1 - all_pieces
    labels, _ = ndimage.label(state[opponent])
    adj_labels = np.unique(labels[adj_locs[:, 0], adj_locs[:, 1]])
    for idx in adj_labels[np.nonzero(adj_labels)]:
        group = labels == idx
        libs = empties * ndimage.binary_dilation(group)
        if np.sum(libs) <= 0:
            locs = np.argwhere(group)
            state[opponent, locs[:, 0], locs[:, 1]] = 0
            killed.append(locs)
    return killed


def compute_invalid_moves(state, player, ko_protect=None):
    """Invalid-move plane for the NEXT mover (the opponent of `player`).

    invalid = occupied
            | (surrounded & possibly_invalid & ~definitely_valid)
            | ko point
    possibly_invalid : liberty of a `player` group with >1 liberties, or of a
                       next-mover group with exactly 1 liberty
    definitely_valid : liberty of a `player` group with exactly 1 liberty
                       (capture), or of a next-mover group with >1 liberties
    surrounded       : all four neighbours occupied, off-board counts as
                       occupied
    """
    all_pieces = state[govars.BLACK] + state[govars.WHITE]
    empties = 1 - all_pieces
    shape = state.shape[1:]
    possible = np.zeros(shape)
    definite = np.zeros(shape)

    own_labels, n_own = ndimage.label(state[player])
    opp_labels, n_opp = ndimage.label(state[1 - player])
    for i in range(1, n_own + 1):
        libs = empties * ndimage.binary_dilation(own_labels == i, _SURROUND)
        cnt = np.sum(libs)
        if cnt > 1:
            possible += libs
        elif cnt == 1:
            definite += libs
    for i in range(1, n_opp + 1):
        libs = empties * ndimage.binary_dilation(opp_labels == i, _SURROUND)
        cnt = np.sum(libs)
        if cnt == 1:
            possible += libs
        elif cnt > 1:
            definite += libs

    surrounded = ndimage.convolve(all_pieces, _SURROUND, mode='constant', cval=1) == 4
    invalid = all_pieces + possible * (definite == 0) * surrounded
    if ko_protect is not None:
        invalid[ko_protect[0], ko_protect[1]] = 1
    return invalid > 0


# ---- gym_go/gogame.py ----------------------------------------------------------
def init_state(size):
    return np.zeros((govars.NUM_CHNLS, size, size))


def turn(state):
    return int(np.max(state[govars.TURN_CHNL]))


def action_size(state=None, board_size=None):
    if state is not None:
        return state.shape[1] * state.shape[2] + 1
    return board_size ** 2 + 1


def game_ended(state):
    m, n = state.shape[1:]
    return int(np.count_nonzero(state[govars.DONE_CHNL] == 1) == m * n)


def next_state(state, action1d):
    state = np.copy(state)
    size = state.shape[1]
    pass_idx = size * size
    passed = action1d == pass_idx
    action2d = (action1d // size, action1d % size)
    player = turn(state)
    previously_passed = np.max(state[govars.PASS_CHNL]) == 1
    ko_protect = None

    if passed:
        state[govars.PASS_CHNL] = 1
        if previously_passed:
            state[govars.DONE_CHNL] = 1
    else:
        state[govars.PASS_CHNL] = 0
        assert state[govars.INVD_CHNL, action2d[0], action2d[1]] == 0, ("Invalid move", action2d)
        state[player, action2d[0], action2d[1]] = 1
        adj_locs, surrounded = _adj_data(state, action2d, player)
        killed = _update_pieces(state, adj_locs, player)
        if len(killed) == 1 and surrounded:
            if len(killed[0]) == 1:
                ko_protect = killed[0][0]

    state[govars.INVD_CHNL] = compute_invalid_moves(state, player, ko_protect)
    state[govars.TURN_CHNL] = 1 - state[govars.TURN_CHNL]
    return state


def invalid_moves(state):
    if game_ended(state):
        return np.zeros(action_size(state))
    return np.append(state[govars.INVD_CHNL].flatten(), 0)


def valid_moves(state):
    return 1 - invalid_moves(state)


def areas(state):
    """Area score: stones plus empty regions bordered by one colour only."""
    all_pieces = state[govars.BLACK] + state[govars.WHITE]
    empties = 1 - all_pieces
    labels, n_regions = ndimage.label(empties)
    black_area = np.sum(state[govars.BLACK])
    white_area = np.sum(state[govars.WHITE])
    for lab in range(1, n_regions + 1):
        region = labels == lab
        nb = ndimage.binary_dilation(region)
        black_claim = bool((state[govars.BLACK] * nb > 0).any())
        white_claim = bool((state[govars.WHITE] * nb > 0).any())
        if black_claim and not white_claim:
            black_area += np.sum(region)
        elif white_claim and not black_claim:
            white_area += np.sum(region)
    return black_area, white_area


def winning(state, komi=0):
    b, w = areas(state)
    return np.sign(b - w - komi)


# ---- gym_go/envs/go_env.py -------------------------------------------------------
class GoEnv:
    """Duck-typed stand-in for gym.make("gym_go:go-v0", size, komi, reward_method)."""

    govars = govars

    def __init__(self, size, komi=0, reward_method='real'):
        assert reward_method == 'real', "only the reference's reward_method='real' is restated"
        self.size = size
        self.komi = komi
        self.state_ = init_state(size)
        self.done = False

    def reset(self):
        self.state_ = init_state(self.size)
        self.done = False
        return np.copy(self.state_)

    def step(self, action):
        assert not self.done
        if isinstance(action, (tuple, list, np.ndarray)):
            assert 0 <= action[0] < self.size and 0 <= action[1] < self.size
            action = self.size * action[0] + action[1]
        elif action is None:
            action = self.size ** 2
        self.state_ = next_state(self.state_, int(action))
        self.done = bool(game_ended(self.state_))
        return np.copy(self.state_), self.reward(), self.done, self.info()

    def game_ended(self):
        return self.done

    def turn(self):
        return turn(self.state_)

    def state(self):
        return np.copy(self.state_)

    def valid_moves(self):
        return valid_moves(self.state_)

    def winning(self):
        return winning(self.state_, self.komi)

    def winner(self):
        return self.winning() if self.game_ended() else 0

    def reward(self):
        return self.winner()

    def info(self):
        return {'turn': turn(self.state_),
                'invalid_moves': invalid_moves(self.state_),
                'prev_player_passed': bool(np.max(self.state_[govars.PASS_CHNL]) == 1)}
