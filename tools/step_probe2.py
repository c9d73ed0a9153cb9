#!/usr/bin/env python
"""Time of one 9x9 lock-step move (WaveGroup.run(1)) on mixed-phase positions vs openings, with and without the
captured search graph (MZGO_NO_GRAPH=1 from the environment), launches per move, dropped simulations."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import muzero_go_b200 as mz  # noqa: E402
from muzero_go_b200 import _lib  # noqa: E402

n, G, S = 9, 1024, 50
dev = torch.device("cuda")
torch.manual_seed(0)
net = mz.MuZeroNet(128, n * n + 1, precision="bf16").to(dev).eval()
lib = _lib.load()
for mixed in (False, True):
    group = mz.WaveGroup(net, n, G, S, waves=1, search=mz.SearchConfig(seed=1000), traj_capacity=2)
    if mixed:
        bench.advance_games(mz, torch, group, n, seed=77)
    for _ in range(5):
        group.run(1)
    torch.cuda.synchronize()
    c0 = group.read_counters()
    lib.mzgo_launch_count(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    K = 60
    for _ in range(K):
        group.run(1)
    e1.record()
    torch.cuda.synchronize()
    c1 = group.read_counters()
    print(f"mixed={mixed} graph={'off' if os.environ.get('MZGO_NO_GRAPH') else 'on'}: {e0.elapsed_time(e1) / K:.3f} ms per move, "
          f"{lib.mzgo_launch_count(0) / K:.1f} launches per move, sims {c1['simulations'] - c0['simulations']}, dropped "
          f"{c1['dropped_simulations'] - c0['dropped_simulations']}, games finished {c1['games_finished'] - c0['games_finished']}, "
          f"mean depth {(c1['leaf_depth_sum'] - c0['leaf_depth_sum']) / max(1, c1['simulations'] - c0['simulations']):.2f}")
