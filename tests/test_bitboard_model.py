"""CPU: the bitboard algorithm of csrc/rules.cu (as modelled lane-by-lane in tests/bitboard_model.py)
against the C oracle on seeded random playouts, every plane after every ply."""
import numpy as np
import pytest

from oracle import gymgo_fast
from tests.bitboard_model import Board


@pytest.mark.parametrize("n,games,seed", [(5, 40, 1), (6, 20, 2), (9, 12, 3), (19, 2, 4)])
def test_bitboard_model_matches_oracle(n, games, seed):
    rng = np.random.default_rng(seed)
    T = 2 * n * n
    for g in range(games):
        st = gymgo_fast.init_states(1, n)[0]
        b = Board(n)
        for t in range(T):
            if st[5, 0, 0]:
                break
            valid = [p for p in range(n * n) if not st[3].flat[p]] + [n * n]
            # mostly board moves: passes only now and then so games run long and fill up
            a = valid[rng.integers(len(valid) - 1)] if len(valid) > 1 and rng.random() < 0.97 else valid[-1]
            assert gymgo_fast.next_state(st, a) == 0
            assert b.step(a) == 0
            assert np.array_equal(b.planes(), st), (n, g, t, a)
        assert b.score() == gymgo_fast.areas(st)
    # illegal / finished handling
    b = Board(n)
    assert b.step(0) == 0 and b.step(0) == -1
    assert b.step(n * n) == 0 and b.step(n * n) == 0 and b.done and b.step(1) == -2
