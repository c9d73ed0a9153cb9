"""arena evaluator.  Scoring and Elo against the REAL SelfPlayEvaluator.evaluate
(tests/golden/arena_evaluate.npz, main.py:595-608); the games themselves on the GPU against the oracle rules."""
import os

import numpy as np
import pytest
import torch

from oracle import gymgo

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def test_scoring_and_elo_match_reference():
    from muzero_go_b200.arena import ArenaConfig, elo_update, game_result, score_games
    z = np.load(os.path.join(GOLDEN, "arena_evaluate.npz"))
    cfg = ArenaConfig(initial_elo=float(z["initial_elo"]), elo_k=float(z["elo_k"]), evaluation_num_games=10,
                      evaluation_win_threshold=float(z["threshold"]))
    cur = best = cfg.initial_elo
    for row in z["rows"]:
        results = [int(v) for v in row[:10]]
        wr = score_games(results)
        cur, best = elo_update(wr, cur, best, cfg)
        assert wr == row[10] and cur == row[11] and best == row[12]          # same float64 arithmetic: exact
    # play_game's mapping (main.py:589-593): starting player 0 = the current network is black
    assert [game_result(0, w) for w in (1, 0, -1)] == [1, 0, 0]
    assert [game_result(1, w) for w in (1, 0, -1)] == [0, 0, 1]


@pytest.mark.gpu
def test_arena_games_are_legal_and_scored():
    import muzero_go_b200 as mz
    from muzero_go_b200.arena import ArenaEvaluator
    n, c, G, S = 5, 16, 10, 16
    A = n * n + 1
    torch.manual_seed(0)
    cur = mz.MuZeroNet(c, A).cuda()
    torch.manual_seed(1)
    best = mz.MuZeroNet(c, A).cuda()
    ev = ArenaEvaluator(cur, best, n, S, num_games=G, search=mz.SearchConfig(seed=9))
    wr, elo = ev.evaluate()
    last = ev.last
    assert 0.0 <= wr <= 1.0 and last["plies"] <= int(n * n * 1.5)
    acts = last["actions"].numpy()
    for g in range(G):
        st, passes, count = gymgo.init_state(n), 0, 0
        for p in range(last["plies"]):
            a = int(acts[p, g])
            over = bool(gymgo.game_ended(st)) or passes >= 2 or count >= int(n * n * 1.5)
            assert (a < 0) == over, (g, p)
            if over:
                continue
            assert gymgo.valid_moves(st)[a] > 0
            st = gymgo.next_state(st, a)
            passes = passes + 1 if a == n * n else 0
            count += 1
        want = int(gymgo.winning(st)) if gymgo.game_ended(st) else 0
        assert int(last["winner"][g]) == want
        assert last["results"][g] == (1 if want == (1 if g % 2 == 0 else -1) else 0)
    assert abs(last["true_win_rate"] - sum(last["results"]) / G) < 1e-12
    # gating (main.py:741-748)
    promoted = ev.gate(cur, best)
    a, b = cur.state_dict(), best.state_dict()
    assert all(torch.equal(a[k], b[k]) for k in a) and promoted == (wr > 0.55)
