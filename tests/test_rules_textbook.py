"""CPU: the rules of Go as a THIRD checker (oracle/go_textbook.py, go_textbook.c; VERDICT r1 missing 1).

oracle/gymgo.py and oracle/gymgo_c.c restate GymGo's liberty-count formula; the textbook engine knows nothing of
it (place the stone, remove dead opponent groups, illegal iff the mover's group then has no liberty, simple ko =
one stone captured by a stone that is then a lone stone in atari, two passes end the game, area scoring by region
flood).  Here:
  * hand-derived known answers (ko in the corner / on the edge, two stones at once, recapture that is not ko,
    multi-stone suicide, snapback, eyes, passes, tiny boards, komi) driven through ALL FOUR engines at once, every
    plane compared after every move;
  * random playouts: identical planes, actions, areas on every position -- pure Python engine on small boards,
    the C engine on BASELINE configs[2] at FULL size (4096 boards x 722 plies of 19x19) against oracle/gymgo_c.c.
So the restated formula equals textbook legality + GymGo's plane conventions on every position reached; GymGo's own
source is absent (also from the GPU box, profiles/r02/gymgo_probe.txt), so "unpinned by the reference" stands."""
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

from oracle import go_textbook as tb
from oracle import gymgo, gymgo_fast
from oracle.gymgo import govars


class Quad:
    """One position in four engines: gymgo.py (formula, scipy), gymgo_c.c (formula, BFS), TextbookGo (Python),
    go_textbook.c.  Every operation asserts that all of them agree."""

    def __init__(self, rows, to_move="b"):
        n = self.n = len(rows)
        mover = 0 if to_move == "b" else 1
        self.py = tb.TextbookGo(n)
        self.py.to_move = mover
        st = gymgo.init_state(n)
        for r, line in enumerate(rows):
            assert len(line) == n
            for c, ch in enumerate(line):
                if ch in "XO":
                    col = 0 if ch == "X" else 1
                    st[col, r, c] = 1
                    self.py.stones[(r, c)] = col
        st[govars.TURN_CHNL] = mover
        st[govars.INVD_CHNL] = gymgo.compute_invalid_moves(st, 1 - mover, None)
        for (r, c) in self.py.stones:                 # the position itself must be legal: no group without liberties
            assert self.py.group(self.py.stones, (r, c))[1], ("dead group in the test position", r, c)
        self.np_state = st
        self.c_state = st.astype(np.uint8)
        self.t_state, self.t_ko = st.astype(np.uint8), -1
        self.check()

    def check(self):
        want = self.py.planes()
        assert np.array_equal(self.np_state.astype(np.uint8), want), "gymgo.py differs from the textbook engine"
        assert np.array_equal(self.c_state, want), "gymgo_c.c differs from the textbook engine"
        assert np.array_equal(self.t_state, want), "go_textbook.c differs from go_textbook.py"

    def legal(self, r, c):
        return self.np_state[govars.INVD_CHNL, r, c] == 0

    def play(self, r=None, c=None):
        a = self.n * self.n if r is None else r * self.n + c
        self.np_state = gymgo.next_state(self.np_state, a)
        assert gymgo_fast.next_state(self.c_state, a) == 0
        rc, self.t_state, self.t_ko = tb.next_state(self.t_state, self.t_ko, a, literal=True)
        assert rc == 0
        self.py.play(a)
        self.check()
        return self

    def refuses(self, r, c):
        """All engines reject the move."""
        a = r * self.n + c
        with pytest.raises(AssertionError):
            gymgo.next_state(self.np_state, a)
        assert gymgo_fast.next_state(self.c_state.copy(), a) == -1
        assert tb.next_state(self.t_state, self.t_ko, a)[0] == -1
        assert self.py.try_move((r, c)) is None
        return True

    def stones(self):
        return int(self.np_state[0].sum()), int(self.np_state[1].sum())

    def areas(self):
        a = tuple(int(x) for x in gymgo.areas(self.np_state))
        assert a == gymgo_fast.areas(self.c_state) == tb.areas(self.t_state) == self.py.areas()
        return a


E5 = ["....."] * 5


def test_corner_ko():
    q = Quad(["O.O..", "XO...", ".....", ".....", "....."], "b")
    assert q.legal(0, 1)
    q.play(0, 1)                                       # captures the corner stone; the capturing stone is alone in atari
    assert q.stones() == (2, 2) and q.py.ko == (0, 0)
    assert not q.legal(0, 0) and q.refuses(0, 0)       # white may not retake at once
    q.play(4, 4).play(4, 0)
    assert q.legal(0, 0)                               # one exchange later it may
    q.play(0, 0)
    assert q.stones() == (2, 4) and q.py.ko == (0, 1) and not q.legal(0, 1)


def test_edge_ko():
    q = Quad(["XO.O.", ".XO..", ".....", ".....", "....."], "b")
    q.play(0, 2)
    assert q.stones() == (3, 2) and q.py.ko == (0, 1) and not q.legal(0, 1)
    q.play(3, 3)                                       # white elsewhere
    assert q.legal(0, 1)                               # black may fill (connect) the ko
    q.play(0, 1)
    assert q.py.ko is None


def test_ko_ban_is_lifted_by_a_pass():
    q = Quad([".XO..", "XO.O.", ".XO..", ".....", "....."], "b")
    q.play(1, 2)
    assert q.py.ko == (1, 1) and not q.legal(1, 1)
    q.play()                                           # white passes: PASS plane set, ko ban gone for everybody
    assert q.np_state[govars.PASS_CHNL].all() and q.legal(1, 1)
    q.play(4, 4)                                       # black elsewhere: PASS plane cleared again
    assert not q.np_state[govars.PASS_CHNL].any()
    assert q.legal(1, 1)                               # white retakes now: not "immediately"
    q.play(1, 1)
    assert q.py.ko == (1, 2)


def test_two_single_stones_captured_at_once_is_not_ko():
    q = Quad(["O.OX.", "XOX..", ".....", ".....", "....."], "b")
    q.play(0, 1)
    assert q.stones() == (4, 1) and q.py.ko is None
    # both vacated points are plain suicide for white (no capture possible there), not ko
    assert not q.legal(0, 0) and not q.legal(0, 2)
    q.play().play(4, 4)                                # still illegal a move later: it never was a ko ban
    assert not q.legal(0, 0) and not q.legal(0, 2)


def test_one_stone_captured_by_a_group_is_not_ko_and_fills_own_last_liberty():
    q = Quad(["X.OX.", "OOX..", ".....", ".....", "....."], "b")
    assert q.legal(0, 1)                               # fills the last liberty of the corner stone, but captures
    q.play(0, 1)
    assert q.stones() == (4, 2) and q.py.ko is None
    assert q.legal(0, 2)                               # white recaptures TWO stones at once: allowed
    q.play(0, 2)
    assert q.stones() == (2, 3) and q.py.ko is None and q.legal(0, 1) and q.legal(0, 0)


def test_multi_stone_suicide():
    rows = ["X.XO.", "OOO..", ".....", ".....", "....."]
    q = Quad(rows, "b")
    assert not q.legal(0, 1) and q.refuses(0, 1)       # would join two stones into a group without liberties
    w = Quad(rows, "w")
    assert w.legal(0, 1)
    w.play(0, 1)                                       # for white the same point captures both
    assert w.stones() == (0, 5) and w.py.ko is None


def test_snapback():
    q = Quad(["..OX.", "OOOX.", "XXX..", ".....", "....."], "b")
    q.play(0, 1)                                       # throw-in: one stone in atari, white group in atari too
    assert q.stones() == (6, 4)
    q.play(0, 0)                                       # white captures the throw-in ...
    assert q.stones() == (5, 5) and q.py.ko is None    # ... with a stone that joins a group: no ko
    assert q.legal(0, 1)
    q.play(0, 1)                                       # black recaptures five stones
    assert q.stones() == (6, 0)


def test_centre_point_suicide_capture_and_centre_ko():
    q = Quad([".....", "..O..", ".O.O.", "..O..", "....."], "b")
    assert not q.legal(2, 2) and q.refuses(2, 2)       # four white stones with outside liberties: suicide
    q = Quad([".....", ".XO..", ".O.O.", ".XO..", "....."], "b")
    assert not q.legal(2, 2)                           # (2,1) still has the liberty (2,0)
    q = Quad([".....", ".XO..", "XO.O.", ".XO..", "....."], "b")
    assert q.legal(2, 2)                               # now (2,2) is the last liberty of (2,1): capture, legal
    q.play(2, 2)
    assert q.stones() == (4, 3) and q.py.ko == (2, 1) and not q.legal(2, 1)      # ... and a ko in the centre


def test_two_eyes_live_and_single_eye_dies():
    live = ["X.X.X", "XXXXX", "OOOOO", ".....", "....."]
    q = Quad(live, "w")
    assert not q.legal(0, 1) and not q.legal(0, 3)     # neither eye can be entered
    q = Quad(live, "b")
    assert q.legal(0, 1) and q.legal(0, 3)             # the owner may fill one (the other remains)
    q.play(0, 1)
    assert q.legal(0, 3)                               # ... which leaves one eye: white captures all nine stones
    q.play(0, 3)
    assert q.stones() == (0, 6)
    w = Quad(["XXX.X", "XXXXX", "OOOOO", ".....", "....."], "w")
    assert w.legal(0, 3)                               # one eye only: white captures everything
    w.play(0, 3)
    assert w.stones() == (0, 6)
    b = Quad(["XXX.X", "XXXXX", "OOOOO", ".....", "....."], "b")
    assert not b.legal(0, 3)                           # and black filling it is suicide


def test_self_atari_is_legal_suicide_of_a_group_is_not():
    q = Quad(["XO...", "XO...", ".O...", "O....", "....."], "b")
    assert not q.legal(2, 0)                           # the three stones would have no liberty and capture nothing
    q2 = Quad(["XO...", "XO...", ".O...", ".....", "....."], "b")
    assert q2.legal(2, 0)                              # one liberty left at (3,0): self-atari, legal
    q2.play(2, 0)
    assert q2.stones() == (3, 3)
    q2.play(3, 0)                                      # white takes the last liberty
    assert q2.stones() == (0, 4)


def test_capture_of_two_groups_of_different_size():
    q = Quad(["OO.OX", "XX.X.", ".....", ".....", "....."], "b")
    q.play(0, 2)
    assert q.stones() == (5, 0) and q.py.ko is None
    assert q.legal(0, 0) and q.legal(0, 1)             # vacated points with an empty neighbour: playable again
    assert not q.legal(0, 3)                           # the third is walled in by black stones that have liberties
    q.play(0, 0)
    assert q.stones() == (5, 1)


def test_large_capture_frees_the_points():
    q = Quad(["OOOO.", "OOOX.", "XXX..", ".....", "....."], "b")
    q.play(0, 4)
    assert q.stones() == (5, 0)
    assert all(q.legal(r, c) for r, c in [(0, 0), (0, 1), (0, 2), (0, 3), (1, 0), (1, 1), (1, 2)])
    assert q.areas() == (25, 0)


def test_tiny_boards():
    q = Quad(["XO", ".."], "b")
    q.play(1, 1)                                       # captures (0,1): its neighbours are both black now
    assert q.stones() == (2, 0) and q.py.ko is None
    q = Quad(["X.", ".O"], "w")
    assert q.legal(0, 1) and q.legal(1, 0)
    q.play(0, 1)                                       # no capture: the corner stone keeps the liberty (1,0)
    assert q.stones() == (1, 2)
    assert q.legal(1, 0)                               # black fills its own last liberty -- and captures both white stones
    q.play(1, 0)
    assert q.stones() == (2, 0) and q.py.ko is None
    q3 = Quad(["X.X", ".X.", "X.X"], "w")
    assert not any(q3.legal(r, c) for r, c in [(0, 1), (1, 0), (1, 2), (2, 1)])   # every empty point is a black eye


def test_first_moves_on_9x9_and_19x19():
    for n in (9, 19):
        q = Quad(["." * n] * n, "b")
        assert not q.np_state[govars.INVD_CHNL].any()
        q.play(n // 2, n // 2)
        assert q.np_state[govars.INVD_CHNL].sum() == 1 and q.np_state[govars.TURN_CHNL].all()
        q.play(0, 0).play(0, 1)
        assert q.legal(1, 0)
        q.play(n - 1, n - 1).play(1, 0)                # the corner stone is captured on a big board too
        assert q.stones() == (3, 1) and q.py.ko is None


def test_corner_ko_on_19x19():
    rows = ["." * 19 for _ in range(19)]
    rows[0] = "O.O" + "." * 16
    rows[1] = "XO" + "." * 17
    q = Quad(rows, "b")
    q.play(0, 1)
    assert q.py.ko == (0, 0) and not q.legal(0, 0)
    q.play(18, 18).play(18, 0)
    assert q.legal(0, 0)


def test_two_passes_planes_and_invalid_plane_after_the_end():
    q = Quad([".O...", "O....", ".....", "...X.", "....."], "b")
    assert not q.legal(0, 0)
    q.play()
    assert q.np_state[govars.PASS_CHNL].all() and not q.np_state[govars.DONE_CHNL].any()
    assert q.np_state[govars.TURN_CHNL].all()
    q.play()
    assert q.np_state[govars.DONE_CHNL].all() and q.np_state[govars.PASS_CHNL].all()
    assert not q.np_state[govars.TURN_CHNL].any()      # the turn plane keeps toggling
    # the invalid plane of the finished game is still "occupied | illegal for the side to move" (black: the corner)
    assert q.np_state[govars.INVD_CHNL, 0, 0] == 1 and q.np_state[govars.INVD_CHNL].sum() == 4
    assert gymgo.valid_moves(q.np_state).sum() == 26   # ... while GymGo's valid_moves() opens everything once done
    assert q.py.over and gymgo_fast.next_state(q.c_state.copy(), 3) == -2


def test_area_scoring_known_answers():
    assert Quad(E5).areas() == (0, 0)
    assert Quad([".....", ".....", "..X..", ".....", "....."]).areas() == (25, 0)
    assert Quad([".....", ".....", "..X..", ".....", "....O"]).areas() == (1, 1)        # one region touching both
    assert Quad(["..X..", "..X..", "..X..", "..X..", "..X.."]).areas() == (25, 0)
    assert Quad(["..XO.", "..XO.", "..XO.", "..XO.", "..XO."]).areas() == (15, 10)
    assert Quad([".X.O.", "XX.OO", ".....", "OO.XX", ".O.X."]).areas() == (6 + 2, 6 + 2)  # four corners, centre neutral
    st = Quad(["..XO.", "..XO.", "..XO.", "..XO.", "..XO."]).np_state
    assert gymgo.winning(st, komi=0) == 1 and gymgo.winning(st, komi=5) == 0 and gymgo.winning(st, komi=5.5) == -1


@pytest.mark.parametrize("n,games,seed", [(2, 20, 0), (3, 20, 1), (5, 12, 2), (7, 4, 3), (9, 2, 4)])
def test_python_textbook_engine_equals_restatement_on_every_position(n, games, seed):
    """Full random games, literal play-and-check of every empty point in pure Python, against oracle/gymgo.py."""
    rng = np.random.default_rng(seed)
    for _ in range(games):
        q = Quad(["." * n] * n, "b")
        for _ply in range(3 * n * n):
            if q.py.over:
                break
            valid = np.nonzero(gymgo.valid_moves(q.np_state))[0]
            a = int(valid[-1]) if len(valid) == 1 or rng.random() < 0.06 else int(valid[rng.integers(len(valid) - 1)])
            if a == n * n:
                q.play()
            else:
                q.play(a // n, a % n)
        q.areas()


@pytest.mark.parametrize("n,B,T", [(2, 512, 12), (5, 512, 75), (6, 256, 108), (9, 256, 243), (13, 64, 400)])
def test_c_textbook_engine_equals_restatement_on_playouts(n, B, T):
    """configs[2]-style playouts: the textbook engine (literal mode and with the empty-neighbour shortcut) and
    oracle/gymgo_c.c choose the same actions and reach the same planes after EVERY ply (per-ply hashes), same areas."""
    R = np.random.Generator(np.random.Philox(n)).integers(0, 2 ** 32, size=(B, T), dtype=np.uint64).astype(np.uint32)
    lit = tb.playout(n, R, literal=True)
    fast = tb.playout(n, R)
    states = gymgo_fast.init_states(B, n)
    plies, actions, _, hashes = gymgo_fast.playout(states, R, hashes=True)
    for got in (lit, fast):
        assert got["plies"] == plies
        assert np.array_equal(got["actions"], actions)
        assert np.array_equal(got["hashes"], hashes)
        assert np.array_equal(got["planes"], states)
        assert np.array_equal(got["areas"], np.array([gymgo_fast.areas(s) for s in states]))


def test_config3_full_size_textbook_equals_restatement():
    """BASELINE.json configs[2] at FULL size -- 4096 boards x 722 plies of 19x19, R from Philox(0) exactly as the GPU
    test draws it -- textbook engine vs oracle/gymgo_c.c: same actions and same planes after every one of the
    ~2.4 M plies (per-ply state hashes), same final areas.  (The GPU kernels are compared with gymgo_c.c on the same
    table in tests/test_rules_gpu.py, so this closes the chain kernels == restatement == textbook rules.)"""
    n, B, T = 19, 4096, 722
    R = np.random.Generator(np.random.Philox(0)).integers(0, 2 ** 32, size=(B, T), dtype=np.uint64).astype(np.uint32)
    chunks = [slice(i, i + 128) for i in range(0, B, 128)]

    def one(sl):
        got = tb.playout(n, R[sl])
        states = gymgo_fast.init_states(sl.stop - sl.start, n)
        plies, actions, _, hashes = gymgo_fast.playout(states, R[sl], hashes=True)
        ok = (got["plies"] == plies and np.array_equal(got["actions"], actions) and np.array_equal(got["hashes"], hashes)
              and np.array_equal(got["planes"], states)
              and np.array_equal(got["areas"], np.array([gymgo_fast.areas(s) for s in states])))
        return ok, plies

    tb.lib(), gymgo_fast.lib()
    with ThreadPoolExecutor(8) as pool:                # ctypes releases the GIL inside the C calls
        res = list(pool.map(one, chunks))
    assert all(ok for ok, _ in res)
    assert sum(p for _, p in res) > 2_000_000
