"""CPU: the GymGo restatements (oracle/gymgo.py literal numpy/scipy, oracle/gymgo_c.c
BFS) against hand-written known-answer Go positions and against each other.
GymGo's source is absent from /root/reference, so these known answers are the
only pins of the rules oracle ("parity unpinned by the reference")."""
import numpy as np
import pytest

from oracle import gymgo, gymgo_fast
from oracle.gymgo import govars


def board(rows, to_move="b"):
    """rows: list of strings with 'X' black, 'O' white, '.' empty."""
    n = len(rows)
    st = gymgo.init_state(n)
    for r, line in enumerate(rows):
        for c, ch in enumerate(line):
            if ch == "X":
                st[govars.BLACK, r, c] = 1
            elif ch == "O":
                st[govars.WHITE, r, c] = 1
    mover = 0 if to_move == "b" else 1
    st[govars.TURN_CHNL] = mover
    # invalid plane for `mover` = compute_invalid_moves(state, player=opponent-of-mover)
    st[govars.INVD_CHNL] = gymgo.compute_invalid_moves(st, 1 - mover, None)
    return st


def play(st, r, c):
    n = st.shape[1]
    nxt = gymgo.next_state(st, r * n + c)
    su = st.astype(np.uint8)
    assert gymgo_fast.next_state(su, r * n + c) == 0
    assert np.array_equal(nxt.astype(np.uint8), su), "python and C restatements disagree"
    return nxt


def test_single_stone_capture():
    st = board([".X...",
                "XO...",
                ".....",
                ".....",
                "....."], "b")
    st = play(st, 1, 2)
    assert st[govars.WHITE, 1, 1] == 1          # one liberty left at (2,1)
    st = play(st, 4, 4)                          # white elsewhere
    st = play(st, 2, 1)                          # black captures
    assert st[govars.WHITE, 1, 1] == 0
    assert st[govars.BLACK].sum() == 4 and st[govars.WHITE].sum() == 1


def test_multi_stone_capture_and_no_ko():
    st = board([".XX..",
                "XOO..",
                ".XX..",
                ".....",
                "....."], "b")
    st = play(st, 1, 3)
    assert st[govars.WHITE].sum() == 0
    # two stones died: no ko point, both points are playable for white again
    assert st[govars.INVD_CHNL, 1, 1] == 0 and st[govars.INVD_CHNL, 1, 2] == 0


def test_ko_and_release():
    st = board([".XO..",
                "XO.O.",
                ".XO..",
                ".....",
                "....."], "b")
    st = play(st, 1, 2)                          # black takes the ko
    assert st[govars.WHITE, 1, 1] == 0
    assert st[govars.INVD_CHNL, 1, 1] == 1       # white may not retake at once
    assert gymgo.turn(st) == 1
    st = play(st, 4, 4)                          # white plays elsewhere
    st = play(st, 4, 0)                          # black plays elsewhere
    assert st[govars.INVD_CHNL, 1, 1] == 0       # ko released: white may retake now
    st = play(st, 1, 1)
    assert st[govars.BLACK, 1, 2] == 0
    assert st[govars.INVD_CHNL, 1, 2] == 1       # and now black is ko-barred


def test_suicide_illegal_capture_legal():
    st = board([".O...",
                "O....",
                ".....",
                ".....",
                "....."], "b")
    assert st[govars.INVD_CHNL, 0, 0] == 1       # corner suicide for black
    st2 = board([".OX..",
                 "OX...",
                 "X....",
                 ".....",
                 "....."], "b")
    # (0,0) is surrounded by white stones that are each in atari -> capture, legal
    assert st2[govars.INVD_CHNL, 0, 0] == 0
    st2 = play(st2, 0, 0)
    assert st2[govars.WHITE].sum() == 0
    with pytest.raises(AssertionError):
        gymgo.next_state(st, 0)
    assert gymgo_fast.next_state(st.astype(np.uint8), 0) == -1


def test_own_eye_fill_rules():
    # black group with a single eye and outside liberties: filling own eye is legal (connects, >1 libs)
    st = board(["XXX..",
                "X.X..",
                "XXX..",
                ".....",
                "....."], "b")
    assert st[govars.INVD_CHNL, 1, 1] == 0
    # same shape but no outside liberties: filling the last liberty is suicide
    st = board(["XXXO.",
                "X.XO.",
                "XXXO.",
                "OOO..",
                "....."], "b")
    assert st[govars.INVD_CHNL, 1, 1] == 1
    # ... and for white it is a capture
    st = board(["XXXO.",
                "X.XO.",
                "XXXO.",
                "OOO..",
                "....."], "w")
    assert st[govars.INVD_CHNL, 1, 1] == 0
    st = play(st, 1, 1)
    assert st[govars.BLACK].sum() == 0


def test_snapback():
    st = board([".XO..",
                "XO.O.",
                "XOOO.",
                "XXXX.",
                "....."], "b")
    # black throws in at (1,2)?  no: classic snapback: white captures one stone then is recaptured
    st = board(["OX.X.",
                "OXXX.",
                "OOOO.",
                ".....",
                "....."], "w")
    # white plays (0,2): stone has no liberties but captures nothing -> must be invalid unless capture
    assert st[govars.INVD_CHNL, 0, 2] == 1
    st = board(["XO.O.",
                "XOOOX",
                "XXXXX",
                ".....",
                "....."], "b")
    # black throw-in at (0,2) is suicide-like (no capture: white group has libs (0,2),(0,4))
    assert st[govars.INVD_CHNL, 0, 2] == 1
    st = play(st, 0, 4)                          # black takes the outside liberty: legal, white now in atari
    assert st[govars.INVD_CHNL, 0, 2] == 1       # white filling its own last liberty is suicide
    st = play(st, 4, 4)                          # white elsewhere
    assert st[govars.INVD_CHNL, 0, 2] == 0       # black may capture five stones
    st = play(st, 0, 2)
    assert st[govars.WHITE].sum() == 1           # only the (4,4) stone remains
    assert st[govars.INVD_CHNL].sum() == st[govars.BLACK].sum() + 1   # multi-stone capture: no ko


def test_two_pass_end_and_winner():
    env = gymgo.GoEnv(5)
    env.reset()
    s, r, done, _ = env.step(12)
    assert not done and r == 0 and env.winner() == 0
    s, r, done, _ = env.step(None)               # white passes
    assert not done and s[govars.PASS_CHNL].all()
    s, r, done, _ = env.step(25)                 # black passes -> end
    assert done and s[govars.DONE_CHNL].all()
    assert env.winner() == 1 and r == 1          # a lone black stone owns the board
    with pytest.raises(AssertionError):
        env.step(0)
    env.reset()
    env.step(None)
    s, _, done, _ = env.step(3)                  # a move clears the pass plane
    assert not done and not s[govars.PASS_CHNL].any()


def test_area_score_dame_and_seki():
    st = board(["X.XO.",
                "XXXO.",
                "..XOO",
                "XXXO.",
                "X.XOO"], "b")
    b, w = gymgo.areas(st)
    # black: 11 stones + (0,1) + (4,1) eyes + (2,0),(2,1) pocket = 15 ; white: 7 stones + 3 empties = 10
    assert (int(b), int(w)) == (15, 10)
    assert gymgo_fast.areas(st.astype(np.uint8)) == (15, 10)
    dame = board(["X.O..",
                  "X.O..",
                  "X.O..",
                  "X.O..",
                  "X.O.."], "b")
    b, w = gymgo.areas(dame)
    assert (int(b), int(w)) == (5, 15)           # the middle column is neutral
    assert gymgo.winning(dame) == -1
    seki = board([".XO.O",
                  "XXOOO",
                  "OOXXX",
                  "O.OX.",
                  "OOOX."], "b")
    b, w = gymgo.areas(seki)
    assert gymgo_fast.areas(seki.astype(np.uint8)) == (int(b), int(w))
    assert gymgo.winning(board(["....."] * 5)) == 0


@pytest.mark.parametrize("n,games", [(5, 6), (9, 4), (19, 1)])
def test_random_playout_invariants_and_differential(n, games):
    rng = np.random.default_rng(100 + n)
    for g in range(games):
        st = gymgo.init_state(n)
        su = np.zeros((6, n, n), np.uint8)
        ko_point = None
        for t in range(3 * n * n):
            if gymgo.game_ended(st):
                break
            v = np.nonzero(gymgo.valid_moves(st))[0]
            a = int(v[-1]) if (len(v) == 1 or rng.random() < 0.04) else int(v[rng.integers(len(v) - 1)])
            before = st[0].sum() + st[1].sum()
            st = gymgo.next_state(st, a)
            assert gymgo_fast.next_state(su, a) == 0
            assert np.array_equal(st.astype(np.uint8), su)
            occ = st[0] + st[1]
            assert occ.max() <= 1
            assert (st[govars.INVD_CHNL] >= occ).all()                     # invalid contains occupied
            if a != n * n:
                assert occ.sum() <= before + 1
            # no group without liberties survives a move
            for colour in (0, 1):
                lab, k = gymgo.ndimage.label(st[colour])
                for i in range(1, k + 1):
                    assert ((1 - occ) * gymgo.ndimage.binary_dilation(lab == i)).sum() > 0
            # at most one empty point is invalid purely because of ko; it is released one ply later
            b, w = gymgo.areas(st)
            assert b + w <= n * n
        assert gymgo_fast.areas(su) == tuple(int(x) for x in gymgo.areas(st))


def test_c_playout_matches_stepwise():
    n, B, T = 9, 8, 200
    R = np.random.Generator(np.random.Philox(0)).integers(0, 2 ** 32, size=(B, T), dtype=np.uint64).astype(np.uint32)
    states = gymgo_fast.init_states(B, n)
    plies, actions, trace = gymgo_fast.playout(states, R, trace=True)
    assert plies > 0
    for b in range(2):
        st = gymgo.init_state(n)
        for t in range(T):
            if gymgo.game_ended(st):
                assert actions[b, t] == -1
                continue
            v = np.nonzero(gymgo.valid_moves(st))[0]
            a = int(v[R[b, t] % len(v)])
            assert a == actions[b, t]
            st = gymgo.next_state(st, a)
            assert np.array_equal(st.astype(np.uint8), trace[b, t])
