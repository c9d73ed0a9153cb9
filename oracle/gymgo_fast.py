"""ctypes wrapper over oracle/gymgo_c.c (TEST INFRASTRUCTURE, see oracle/__init__.py).

build() compiles the C restatement with gcc into oracle/_build/libgymgo_c.so.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libgymgo_c.so")
_lib = None


def build(force=False):
    src = os.path.join(_HERE, "gymgo_c.c")
    if (not force and os.path.exists(_SO)
            and os.path.getmtime(_SO) >= os.path.getmtime(src)):
        return _SO
    os.makedirs(os.path.dirname(_SO), exist_ok=True)
    subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-o", _SO, src])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_SO)
        L.gymgo_c_next_state.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
        L.gymgo_c_next_state.restype = ctypes.c_int
        L.gymgo_c_areas.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        L.gymgo_c_areas.restype = None
        L.gymgo_c_playout.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                      ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.gymgo_c_hash.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.gymgo_c_hash.restype = ctypes.c_uint64
        L.gymgo_c_playout.restype = ctypes.c_long
        _lib = L
    return _lib


def init_states(nboards, n):
    return np.zeros((nboards, 6, n, n), dtype=np.uint8)


def next_state(state_u8, action):
    """state_u8: [6,n,n] uint8 (modified in place). Returns status (0 ok)."""
    assert state_u8.dtype == np.uint8 and state_u8.flags.c_contiguous
    return lib().gymgo_c_next_state(state_u8.ctypes.data, state_u8.shape[1], int(action))


def areas(state_u8):
    b = ctypes.c_int(0)
    w = ctypes.c_int(0)
    st = np.ascontiguousarray(state_u8, dtype=np.uint8)
    lib().gymgo_c_areas(st.ctypes.data, st.shape[1], ctypes.byref(b), ctypes.byref(w))
    return b.value, w.value


def board_hash(state_u8):
    st = np.ascontiguousarray(state_u8, dtype=np.uint8)
    return int(lib().gymgo_c_hash(st.ctypes.data, st.shape[1]))


def playout(states_u8, R, trace=False, hashes=False):
    """states_u8 [B,6,n,n] uint8 in place; R [B,T] uint32.
    Returns (plies, actions[B,T] int32, trace[B,T,6,n,n] or None[, hashes[B,T] uint64])."""
    assert states_u8.dtype == np.uint8 and states_u8.flags.c_contiguous
    B, _, n, _ = states_u8.shape
    R = np.ascontiguousarray(R, dtype=np.uint32)
    T = R.shape[1]
    actions = np.full((B, T), -1, dtype=np.int32)
    tr = np.zeros((B, T, 6, n, n), dtype=np.uint8) if trace else None
    hs = np.zeros((B, T), dtype=np.uint64) if hashes else None
    plies = lib().gymgo_c_playout(states_u8.ctypes.data, n, B, R.ctypes.data, T,
                                  tr.ctypes.data if trace else None, actions.ctypes.data,
                                  hs.ctypes.data if hashes else None)
    if hashes:
        return plies, actions, tr, hs
    return plies, actions, tr
