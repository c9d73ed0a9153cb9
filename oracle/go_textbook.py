"""Oracle: the rules of Go as a rule book states them (TEST INFRASTRUCTURE, see oracle/__init__.py).

A third, independent checker of the rules engine (missing 1): oracle/gymgo.py and oracle/gymgo_c.c
both restate GymGo's liberty-count formula for the invalid-move plane; nothing here knows that formula.

  `TextbookGo`    pure Python (sets, flood fill), literal play-and-check of EVERY empty point: place the stone,
                  remove the opponent's groups without liberties, illegal iff the mover's own group then has none;
                  simple ko = exactly one stone captured by a stone that is then a lone stone in atari; two passes
                  end the game; area scoring by region flood.  Small cases.
  go_textbook.c   the same rules in C for full-size runs (4096 boards x 722 plies), through `playout()` below.

Both derive the six GymGo planes the reference reads (main.py:261, :637, :681): invalid = occupied | illegal |
ko point, for the player to move.

Only tests/ may import this module.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libgo_textbook.so")
_lib = None


class TextbookGo:
    def __init__(self, n):
        self.n = n
        self.stones = {}            # (r, c) -> 0 black / 1 white
        self.to_move = 0
        self.ko = None              # point the player to move may not play
        self.last_pass = False
        self.over = False

    def neighbours(self, p):
        r, c = p
        return [(r + dr, c + dc) for dr, dc in ((-1, 0), (1, 0), (0, -1), (0, 1))
                if 0 <= r + dr < self.n and 0 <= c + dc < self.n]

    def group(self, stones, p):
        """(members, liberties) of the group of the stone at p."""
        colour, members, libs, todo = stones[p], {p}, set(), [p]
        while todo:
            for q in self.neighbours(todo.pop()):
                if q not in stones:
                    libs.add(q)
                elif stones[q] == colour and q not in members:
                    members.add(q)
                    todo.append(q)
        return members, libs

    def try_move(self, p):
        """-> (stones after the move, captured points) or None when the move is illegal."""
        if p in self.stones or p == self.ko:
            return None
        stones = dict(self.stones)
        stones[p] = self.to_move
        captured = []
        for q in self.neighbours(p):
            if q in stones and stones[q] != self.to_move:
                members, libs = self.group(stones, q)
                if not libs:
                    for m in members:
                        del stones[m]
                    captured.extend(members)
        if not self.group(stones, p)[1]:
            return None                                    # suicide
        return stones, captured

    def legal_points(self):
        return [(r, c) for r in range(self.n) for c in range(self.n) if self.try_move((r, c)) is not None]

    def play(self, action):
        assert not self.over
        n = self.n
        if action == n * n:
            self.over = self.last_pass
            self.last_pass, self.ko = True, None
        else:
            p = divmod(action, n)
            res = self.try_move(p)
            assert res is not None, ("illegal move", p)
            self.stones, captured = res
            members, libs = self.group(self.stones, p)
            self.ko = captured[0] if len(captured) == 1 and len(members) == 1 and len(libs) == 1 else None
            self.last_pass = False
        self.to_move ^= 1

    def planes(self):
        n = self.n
        st = np.zeros((6, n, n), dtype=np.uint8)
        for (r, c), col in self.stones.items():
            st[col, r, c] = 1
        st[2] = self.to_move
        st[3] = 1
        for r, c in self.legal_points():
            st[3, r, c] = 0
        st[4] = int(self.last_pass)
        st[5] = int(self.over)
        return st

    def areas(self):
        n, seen, area = self.n, set(), [0, 0]
        for col in self.stones.values():
            area[col] += 1
        for r in range(n):
            for c in range(n):
                if (r, c) in self.stones or (r, c) in seen:
                    continue
                region, touch, todo = {(r, c)}, set(), [(r, c)]
                while todo:
                    for q in self.neighbours(todo.pop()):
                        if q in self.stones:
                            touch.add(self.stones[q])
                        elif q not in region:
                            region.add(q)
                            todo.append(q)
                seen |= region
                if len(touch) == 1:
                    area[touch.pop()] += len(region)
        return tuple(area)

    def winner(self, komi=0):
        b, w = self.areas()
        return int(np.sign(b - w - komi))


# ---------------------------------------------------------------------------------------------- C engine
def build(force=False):
    src = os.path.join(_HERE, "go_textbook.c")
    if not force and os.path.exists(_SO) and os.path.getmtime(_SO) >= os.path.getmtime(src):
        return _SO
    os.makedirs(os.path.dirname(_SO), exist_ok=True)
    subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-o", _SO, src])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_SO)
        vp, ci = ctypes.c_void_p, ctypes.c_int
        L.tb_next_state.argtypes = [vp, ci, ci, ci, vp, ctypes.POINTER(ci), ci]
        L.tb_next_state.restype = ci
        L.tb_areas.argtypes = [vp, ci, ctypes.POINTER(ci), ctypes.POINTER(ci)]
        L.tb_areas.restype = None
        L.tb_playout.argtypes = [ci, ci, vp, ci, vp, vp, vp, vp, ci]
        L.tb_playout.restype = ctypes.c_long
        _lib = L
    return _lib


def next_state(planes_u8, ko, action, literal=False):
    """-> (status, planes after the move, new ko point or -1)."""
    st = np.ascontiguousarray(planes_u8, dtype=np.uint8)
    out = np.zeros_like(st)
    ko_out = ctypes.c_int(-1)
    rc = lib().tb_next_state(st.ctypes.data, st.shape[1], int(ko), int(action), out.ctypes.data, ctypes.byref(ko_out),
                             int(literal))
    return rc, out, ko_out.value


def areas(planes_u8):
    st = np.ascontiguousarray(planes_u8, dtype=np.uint8)
    b, w = ctypes.c_int(0), ctypes.c_int(0)
    lib().tb_areas(st.ctypes.data, st.shape[1], ctypes.byref(b), ctypes.byref(w))
    return b.value, w.value


def playout(n, R, literal=False):
    """configs[2] playout by the textbook engine.  R uint32 [B,T] -> dict(plies, actions [B,T], hashes [B,T] uint64,
    planes [B,6,n,n] uint8 (final), areas [B,2])."""
    R = np.ascontiguousarray(R, dtype=np.uint32)
    B, T = R.shape
    actions = np.full((B, T), -1, dtype=np.int32)
    hashes = np.zeros((B, T), dtype=np.uint64)
    planes = np.zeros((B, 6, n, n), dtype=np.uint8)
    ar = np.zeros((B, 2), dtype=np.int32)
    plies = lib().tb_playout(n, B, R.ctypes.data, T, actions.ctypes.data, hashes.ctypes.data, planes.ctypes.data,
                             ar.ctypes.data, int(literal))
    return dict(plies=plies, actions=actions, hashes=hashes, planes=planes, areas=ar)
