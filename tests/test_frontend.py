"""play.py's terminal helpers (tests/golden/frontend.json holds the output of the REAL
print_board / action_to_coord / coord_to_action, play.py:248-278) and checkpoint files."""
import json
import os

import numpy as np
import pytest
import torch

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def test_terminal_helpers_match_reference():
    from muzero_go_b200.frontend import action_to_coord, coord_to_action, format_board, print_board
    z = json.load(open(os.path.join(GOLDEN, "frontend.json")))
    for b in z["boards"]:
        obs = np.zeros((6, b["n"], b["n"]))
        obs[0], obs[1] = np.asarray(b["black"]), np.asarray(b["white"])
        assert format_board(obs, b["n"]) == b["text"]
        lines = []
        print_board(obs, b["n"], print_fn=lambda s, end="\n": lines.append(s + end))
        assert "".join(lines) == b["text"]
    for a, n, want in z["a2c"]:
        assert action_to_coord(a, n) == want
    for c, n, want in z["c2a"]:
        assert coord_to_action(c, n) == want
    for c, kind, msg in z["c2a_errors"]:
        with pytest.raises(ValueError, match=msg):
            coord_to_action(c, 9)


def test_checkpoint_files_round_trip_with_reference_modules(tmp_path):
    """A file written from the oracle's restatement of the reference module (same key names and shapes as
    main.py:72-142, see tests/test_abi.py) loads, and a file written here loads back into it."""
    from oracle import nets_ref
    from muzero_go_b200.frontend import load_checkpoint, save_checkpoint
    from muzero_go_b200.networks import MuZeroNet
    torch.manual_seed(3)
    ref = nets_ref.MuZeroNetRef(32, 37)
    p1, p2 = os.path.join(tmp_path, "ref.pth"), os.path.join(tmp_path, "ours.pth")
    torch.save(ref.state_dict(), p1)                         # main.py:752-753
    net = load_checkpoint(MuZeroNet(32, 37), p1, map_location="cpu")
    for k, v in ref.state_dict().items():
        assert torch.equal(net.state_dict()[k], v)
    save_checkpoint(net, p2)
    ref2 = nets_ref.MuZeroNetRef(32, 37)
    ref2.load_state_dict(torch.load(p2, map_location="cpu"))   # play.py:290-291
    assert all(torch.equal(a, b) for a, b in zip(ref.state_dict().values(), ref2.state_dict().values()))


def test_play_loop_with_scripted_terminal():
    """play.py:305-346 with a scripted human, a scripted agent and a stub environment."""
    from muzero_go_b200.frontend import play

    class Env:
        def __init__(self):
            self.moves = []

        def reset(self):
            return np.zeros((6, 3, 3))

        def step(self, a):
            self.moves.append(a)
            obs = np.zeros((6, 3, 3))
            done = len(self.moves) >= 2 and self.moves[-1] == 9 and self.moves[-2] == 9
            return obs, (-1.0 if done else 0.0), done, {}

    class Agent:
        def select_action(self, obs):
            return 9

    said, script = [], iter(["zz", "B2", "pass"])
    env = Env()
    outcome, reward = play(Agent(), env, 3, True, input_fn=lambda prompt: next(script),
                           print_fn=lambda s, end="\n": said.append(s))
    assert env.moves == [4, 9, 9, 9][:len(env.moves)] and env.moves[:2] == [4, 9]
    assert "Invalid move format. Try again." in said and "Agent move: pass" in said and "Game over!" in said
    assert (outcome, reward) == ("loss", -1.0)
