"""GPU: the CUDA rules engine (through the C ABI) against the oracle restatements of GymGo.
Bit-exact: every plane after every ply, actions, areas, winner."""
import numpy as np
import pytest
import torch

from oracle import gymgo, gymgo_fast

pytestmark = pytest.mark.gpu


def philox_table(seed, shape):
    return np.random.Generator(np.random.Philox(seed)).integers(0, 2 ** 32, size=shape, dtype=np.uint64).astype(np.uint32)


@pytest.mark.parametrize("n,boards,T", [(2, 16, 24), (5, 64, 80), (6, 64, 100), (9, 64, 200), (13, 16, 400), (19, 16, 722)])
def test_playout_trace_bit_exact(n, boards, T):
    import muzero_go_b200 as mz
    R = philox_table(100 + n, (boards, T))
    ref_states = gymgo_fast.init_states(boards, n)
    plies, ref_actions, ref_trace = gymgo_fast.playout(ref_states, R, trace=True)
    gb = mz.GoBoards(boards, n).reset()
    acts, hsh, trace = gb.playout(R, want_hash=True, want_trace=True)
    assert np.array_equal(acts.cpu().numpy(), ref_actions)
    got = trace.cpu().numpy()
    if not np.array_equal(got, ref_trace):
        bad = np.argwhere((got != ref_trace).reshape(boards, T, -1).any(axis=2))[0]
        raise AssertionError(f"first mismatch at board {bad[0]} ply {bad[1]} action {ref_actions[bad[0], bad[1]]}")
    assert np.array_equal(gb.planes(torch.uint8).cpu().numpy(), ref_states)
    # the state hash agrees with the oracle's implementation of the same function
    ref_hash = np.array([[gymgo_fast.board_hash(ref_trace[b, t]) for t in range(0, T, 7)] for b in range(boards)], dtype=np.uint64)
    assert np.array_equal(hsh.cpu().numpy().view(np.uint64)[:, ::7], ref_hash)
    # areas / winner on the final positions
    areas, winner = gb.score(0.0)
    ref_areas = np.array([gymgo_fast.areas(s) for s in ref_states])
    assert np.array_equal(areas.cpu().numpy(), ref_areas)
    assert np.array_equal(winner.cpu().numpy(), np.sign(ref_areas[:, 0] - ref_areas[:, 1]))
    assert plies == int((ref_actions >= 0).sum())


def test_config3_full_size_4096_boards():
    """BASELINE.json configs[2]: 19x19, 4096 boards, up to 722 plies; per-ply state hashes, actions,
    final planes and areas against the C oracle."""
    import muzero_go_b200 as mz
    n, boards, T = 19, 4096, 722
    R = philox_table(0, (boards, T))
    ref_states = gymgo_fast.init_states(boards, n)
    plies, ref_actions, _, ref_hash = gymgo_fast.playout(ref_states, R, hashes=True)
    gb = mz.GoBoards(boards, n).reset()
    acts, hsh, _ = gb.playout(R, want_hash=True)
    assert np.array_equal(acts.cpu().numpy(), ref_actions)
    assert np.array_equal(hsh.cpu().numpy().view(np.uint64), ref_hash)
    assert np.array_equal(gb.planes(torch.uint8).cpu().numpy(), ref_states)
    areas, _ = gb.score(0.0)
    assert np.array_equal(areas.cpu().numpy(), np.array([gymgo_fast.areas(s) for s in ref_states]))
    assert plies > boards * 300


@pytest.fixture
def thread_per_board():
    """Route step / playout through the thread-per-board kernels (rules_tpb.cuh) whatever the batch size."""
    from muzero_go_b200 import _lib
    lib = _lib.load()
    old = lib.mzgo_debug_rules_tpb_min(0)
    yield
    lib.mzgo_debug_rules_tpb_min(old)


@pytest.mark.parametrize("n,boards,T", [(5, 70, 80), (6, 64, 100), (9, 200, 200), (13, 33, 400), (19, 130, 722)])
def test_thread_per_board_playout_bit_exact(thread_per_board, n, boards, T):
    """The large-batch kernels: actions, per-ply state hashes, final planes and areas against the C oracle
    (ragged last block, every supported board size)."""
    import muzero_go_b200 as mz
    R = philox_table(300 + n, (boards, T))
    ref_states = gymgo_fast.init_states(boards, n)
    plies, ref_actions, _, ref_hash = gymgo_fast.playout(ref_states, R, hashes=True)
    gb = mz.GoBoards(boards, n).reset()
    acts, hsh, _ = gb.playout(R, want_hash=True)
    assert np.array_equal(acts.cpu().numpy(), ref_actions)
    assert np.array_equal(hsh.cpu().numpy().view(np.uint64), ref_hash)
    assert np.array_equal(gb.planes(torch.uint8).cpu().numpy(), ref_states)
    assert plies == int((ref_actions >= 0).sum())


def test_thread_per_board_equals_warp_per_board_at_scale(thread_per_board):
    """65,536 boards of 19x19, 40 plies: the two kernel families leave identical records and actions."""
    import muzero_go_b200 as mz
    from muzero_go_b200 import _lib
    n, boards, T = 19, 65536, 40
    R = philox_table(9, (boards, T))
    a = mz.GoBoards(boards, n).reset()
    acts_a, hash_a, _ = a.playout(R, want_hash=True)
    _lib.load().mzgo_debug_rules_tpb_min(1 << 30)
    b = mz.GoBoards(boards, n).reset()
    acts_b, hash_b, _ = b.playout(R, want_hash=True)
    assert torch.equal(acts_a, acts_b) and torch.equal(hash_a, hash_b) and torch.equal(a.records, b.records)
    # one more ply through the step entry point, with some boards left alone and some illegal moves
    _lib.load().mzgo_debug_rules_tpb_min(0)
    move = torch.randint(-1, n * n + 1, (boards,), dtype=torch.int32, device="cuda")
    st_a = a.step(move)
    _lib.load().mzgo_debug_rules_tpb_min(1 << 30)
    st_b = b.step(move)
    assert torch.equal(st_a, st_b) and torch.equal(a.records, b.records) and int((st_a != 0).sum()) > 0


def test_step_api_matches_python_oracle_and_reports_errors():
    import muzero_go_b200 as mz
    n, boards = 7, 32
    rng = np.random.default_rng(5)
    gb = mz.GoBoards(boards, n).reset()
    states = [gymgo.init_state(n) for _ in range(boards)]
    for t in range(90):
        actions = np.full(boards, -1, dtype=np.int32)
        for b, st in enumerate(states):
            if gymgo.game_ended(st) or (b + t) % 11 == 0:
                continue                      # -1: leave this board alone
            v = np.nonzero(gymgo.valid_moves(st))[0]
            actions[b] = v[rng.integers(len(v))] if rng.random() < 0.9 else n * n
            states[b] = gymgo.next_state(st, int(actions[b]))
        status = gb.step(torch.from_numpy(actions))
        assert int(status.abs().sum()) == 0
        assert np.array_equal(gb.planes(torch.float64).cpu().numpy(), np.stack(states))
    # errors leave the board untouched
    before = gb.records.clone()
    bad = np.zeros(boards, dtype=np.int32)
    for b, st in enumerate(states):
        occupied = np.nonzero((st[0] + st[1]).flatten())[0]
        bad[b] = occupied[0] if len(occupied) else n * n + 5
    status = gb.step(torch.from_numpy(bad)).cpu().numpy()
    for b, st in enumerate(states):
        assert status[b] == (-2 if gymgo.game_ended(st) else -1)
    assert torch.equal(before, gb.records)


def test_pack_expand_round_trip_and_env():
    import muzero_go_b200 as mz
    n = 9
    R = philox_table(9, (8, 60))
    gb = mz.GoBoards(8, n).reset()
    gb.playout(R)
    planes = gb.planes(torch.float32)
    gb2 = mz.GoBoards(8, n).load_planes(planes)
    assert torch.equal(gb2.planes(torch.float32), planes)
    # GoEnv: same trajectory as the oracle env, including reward / done / winner
    env, ref = mz.make("gym_go:go-v0", size=5, komi=0, reward_method="real"), gymgo.GoEnv(5)
    assert np.array_equal(env.reset(), ref.reset())
    rng = np.random.default_rng(0)
    done = False
    while not done:
        v = np.nonzero(ref.valid_moves())[0]
        a = int(v[rng.integers(len(v))])
        s1, r1, d1, i1 = env.step(a)
        s2, r2, d2, i2 = ref.step(a)
        assert np.array_equal(s1, s2) and r1 == r2 and d1 == d2 and i1["turn"] == i2["turn"]
        assert np.array_equal(i1["invalid_moves"], i2["invalid_moves"])
        done = d1
    assert env.winner() == ref.winner()


@pytest.mark.parametrize("n,boards,T", [(5, 512, 75), (9, 512, 243), (19, 192, 722)])
def test_playout_equals_textbook_rules_engine(n, boards, T):
    """The kernels against the THIRD checker directly (oracle/go_textbook.c: place the stone, remove dead opponent
    groups, suicide test, simple ko, two passes, region-flood area score -- no liberty-count formula in it): same
    actions, same six planes after every ply (state hashes), same final planes and areas."""
    import muzero_go_b200 as mz
    from oracle import go_textbook
    R = philox_table(700 + n, (boards, T))
    want = go_textbook.playout(n, R)
    gb = mz.GoBoards(boards, n).reset()
    acts, hsh, _ = gb.playout(R, want_hash=True)
    assert np.array_equal(acts.cpu().numpy(), want["actions"])
    assert np.array_equal(hsh.cpu().numpy().view(np.uint64), want["hashes"])
    assert np.array_equal(gb.planes(torch.uint8).cpu().numpy(), want["planes"])
    areas, winner = gb.score(0.0)
    assert np.array_equal(areas.cpu().numpy(), want["areas"])
    assert np.array_equal(winner.cpu().numpy(), np.sign(want["areas"][:, 0] - want["areas"][:, 1]))


def test_known_answer_positions_on_the_kernels():
    """The hand-derived positions of tests/test_rules_textbook.py (ko in the corner / on the edge / in the centre,
    two stones at once, recapture that is not ko, multi-stone suicide, snapback, eyes, tiny boards) played on the
    GPU: planes after every move equal the textbook engine's, illegal moves are refused with status -1."""
    import muzero_go_b200 as mz
    from oracle.go_textbook import TextbookGo
    cases = [
        (["O.O..", "XO...", ".....", ".....", "....."], "b", [(0, 1), "x00", (4, 4), (4, 0), (0, 0), "x01"]),
        (["XO.O.", ".XO..", ".....", ".....", "....."], "b", [(0, 2), "x01", (3, 3), (0, 1)]),
        ([".XO..", "XO.O.", ".XO..", ".....", "....."], "b", [(1, 2), "x11", None, (4, 4), (1, 1), "x12"]),
        (["O.OX.", "XOX..", ".....", ".....", "....."], "b", [(0, 1), "x00", "x02", None, (4, 4), "x00"]),
        (["X.OX.", "OOX..", ".....", ".....", "....."], "b", [(0, 1), (0, 2), (0, 1)]),
        (["X.XO.", "OOO..", ".....", ".....", "....."], "b", ["x01", None, (0, 1)]),
        (["..OX.", "OOOX.", "XXX..", ".....", "....."], "b", [(0, 1), (0, 0), (0, 1)]),
        ([".....", ".XO..", "XO.O.", ".XO..", "....."], "b", [(2, 2), "x21", (0, 0), (4, 4), (2, 1)]),
        (["X.X.X", "XXXXX", "OOOOO", ".....", "....."], "w", ["x01", "x03", None, (0, 1), (0, 3)]),
        (["XXX.X", "XXXXX", "OOOOO", ".....", "....."], "b", ["x03", None, (0, 3)]),
        (["XO", ".."], "b", [(1, 1), "x01", "x10", None, (0, 1)]),
        (["X.", ".O"], "w", [(0, 1), (1, 0), None, None]),
        (["X.X", ".X.", "X.X"], "w", ["x01", "x10", "x12", "x21", None, (0, 1)]),
    ]
    for rows, to_move, script in cases:
        n = len(rows)
        ref = TextbookGo(n)
        ref.to_move = 0 if to_move == "b" else 1
        for r, line in enumerate(rows):
            for c, ch in enumerate(line):
                if ch in "XO":
                    ref.stones[(r, c)] = 0 if ch == "X" else 1
        gb = mz.GoBoards(1, n).load_planes(ref.planes()[None].astype(np.float32))
        for step in script:
            if isinstance(step, str):                  # "xRC": the move must be refused, board untouched
                r, c = int(step[1]), int(step[2])
                before = gb.records.clone()
                assert ref.try_move((r, c)) is None, (rows, step)
                status = gb.step(torch.tensor([r * n + c], dtype=torch.int32))
                assert int(status[0]) == -1 and torch.equal(before, gb.records), (rows, step)
                continue
            a = n * n if step is None else step[0] * n + step[1]
            ref.play(a)
            status = gb.step(torch.tensor([a], dtype=torch.int32))
            assert int(status[0]) == 0, (rows, step)
            assert np.array_equal(gb.planes(torch.uint8).cpu().numpy()[0], ref.planes()), (rows, step)
        areas, _ = gb.score(0.0)
        assert tuple(areas[0].tolist()) == ref.areas(), rows
