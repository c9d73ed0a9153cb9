"""Calibrate the CPU arm once (TEST INFRASTRUCTURE; build container only, needs /root/reference).

    python -m oracle.calibrate_cpu [--out profiles/r02/cpu_calibration.json]

bench.py's CPU legs time the oracle PORT (oracle/mcts_ref.py + oracle/nets_ref.py + oracle/gymgo.py), because the
reference is Python and does not travel to the GPU box.  This script records, in the build container, how the port
compares with the REAL reference code on the same inputs, one thread each:

  * main.MCTS(net, A, S).run(obs)   (/root/reference/main.py:255-316, unmodified, via oracle/ref_loader.py)
    vs oracle.mcts_ref.run_mcts on the same network weights and position: 9x9 / 50 sims and 19x19 / 800 sims;
  * BASELINE.json configs[0], literally: self_play.run_self_play_game (self_play.py:453-526) for ONE game at
    50 sims/move on 9x9 against the restated GymGo env, with latent_dim 96 (the script's default, self_play.py:21)
    and 128 (main.py:29).  Its action choice ignores the search (defect 1), so game length, not the
    policy, drives its moves/sec.
"""
import argparse
import contextlib
import io
import json
import os
import random
import time

import numpy as np
import torch

from . import gymgo, mcts_ref, ref_loader
from .make_golden import midgame_obs
from .nets_ref import MuZeroNetRef


def time_mcts(n, sims, latent, moves):
    main = ref_loader.configure(ref_loader.load("main"), n)
    A = n * n + 1
    torch.manual_seed(0)
    real_net = main.MuZeroNet(latent, A).eval()
    port_net = MuZeroNetRef(latent, A).eval()
    port_net.load_state_dict(real_net.state_dict())
    initial, recurrent = mcts_ref.torch_net_callbacks(port_net)
    obs_list = [midgame_obs(n, int(0.2 * n * n * k), seed=k) for k in range(moves)]
    rng = np.random.default_rng(0)
    random.seed(0)
    np.random.seed(0)
    t0 = time.perf_counter()
    done_real = 0
    for obs in obs_list:
        root = main.MCTS(real_net, A, sims).run(obs)
        done_real += root.visit_count
    t_real = time.perf_counter() - t0
    t0 = time.perf_counter()
    done_port = 0
    for obs in obs_list:
        noise = rng.dirichlet([0.03] * A)
        stream = rng.integers(0, 2 ** 32, size=sims + 8, dtype=np.uint64).astype(np.uint32)
        tree, _ = mcts_ref.run_mcts(obs, initial, recurrent, sims, noise, stream)
        done_port += tree.visit[0]
    t_port = time.perf_counter() - t0
    return {"board": n, "sims_per_move": sims, "latent_dim": latent, "moves": moves, "threads": 1,
            "real_main_mcts_sims_per_s": done_real / t_real, "oracle_port_sims_per_s": done_port / t_port,
            "port_over_real": (done_port / t_port) / (done_real / t_real)}


def time_config0(latent, sims=50, n=9):
    sp = ref_loader.configure(ref_loader.load("self_play"), n)
    sp.config.latent_dim = latent
    sp.config.mcts_simulations = sims
    torch.manual_seed(0)
    random.seed(0)
    np.random.seed(0)
    agent = sp.MuZeroAgent(n, latent, n * n + 1, sims)
    calls = [0]
    orig = agent.net.recurrent_inference

    def counted(*a, **k):
        calls[0] += 1
        return orig(*a, **k)
    agent.net.recurrent_inference = counted
    env = gymgo.GoEnv(n)
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):                     # the reference prints the board every move
        hist = sp.run_self_play_game(agent, env, sp.config)
    dt = time.perf_counter() - t0
    return {"board": n, "sims_per_move": sims, "latent_dim": latent, "threads": 1, "moves": len(hist), "seconds": dt,
            "sims_per_s": calls[0] / dt, "moves_per_s": len(hist) / dt,
            "what": "self_play.run_self_play_game, one game, restated GymGo env (configs[0])"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    torch.set_num_threads(1)
    out = {"host": {"cpus": os.cpu_count(), "torch": torch.__version__, "numpy": np.__version__,
                    "where": "build container (not the GPU box)"},
           "mcts": [time_mcts(9, 50, 128, 12), time_mcts(19, 800, 128, 3)],
           "config0": [time_config0(96), time_config0(128)]}
    text = json.dumps(out, indent=1)
    print(text)
    if args.out:
        with open(args.out, "w") as fh:
            fh.write(text + "\n")


if __name__ == "__main__":
    main()
