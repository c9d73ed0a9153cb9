#!/usr/bin/env python
"""Where a fused 9x9 launch (network kernel + tree phase) spends its time on mixed-phase positions: per-CTA stamps of
"roles done" (30) and "tree phase done" (31) of the LAST simulation's launch of a move, with the leaf depths of the
games.  Run with MZGO_NO_GRAPH=1 (the trace pointer is a launch argument).  python tools/step_probe.py [--n 9]"""
import argparse
import os
import sys

import numpy as np
import torch

os.environ["MZGO_NO_GRAPH"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import muzero_go_b200 as mz  # noqa: E402
from muzero_go_b200 import _lib  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=9)
    ap.add_argument("--games", type=int, default=1024)
    ap.add_argument("--sims", type=int, default=50)
    ap.add_argument("--openings", action="store_true")
    args = ap.parse_args()
    n, G, S = args.n, args.games, args.sims
    dev = torch.device("cuda")
    torch.manual_seed(0)
    net = mz.MuZeroNet(128, n * n + 1, precision="bf16").to(dev).eval()
    group = mz.WaveGroup(net, n, G, S, waves=1, search=mz.SearchConfig(seed=1000), traj_capacity=2)
    if not args.openings:
        bench.advance_games(mz, torch, group, n, seed=77)
    lib = _lib.load()
    for _ in range(3):
        group.run(1)
    torch.cuda.synchronize()
    trace = torch.zeros(148 * 64, dtype=torch.int64, device=dev)
    lib.mzgo_debug_tc_trace(trace.data_ptr())
    group.run(1)
    torch.cuda.synchronize()
    lib.mzgo_debug_tc_trace(None)
    t = trace.cpu().numpy().reshape(148, 64).astype(np.float64)
    ok = t[:, 0] > 0
    t0 = t[ok, 0].min()
    roles = (t[ok, 30] - t0) / 1e3
    done = (t[ok, 31] - t0) / 1e3
    print(f"n={n} games={G} sims={S} {'openings' if args.openings else 'mixed'}: CTAs reporting {ok.sum()}")
    for name, v in (("roles done (us)", roles), ("tree phase done (us)", done), ("tree phase duration (us)", done - roles)):
        print(f"  {name:26s} min {v.min():7.2f}  median {np.median(v):7.2f}  p90 {np.percentile(v, 90):7.2f}  max {v.max():7.2f}")
    e = group.engines[0]
    d = e.tree.leaf_depth.cpu().numpy()
    sd = (e.tree.depth_sum.cpu().numpy() / np.maximum(1, e.tree.sims_done.cpu().numpy()))
    print("  last leaf depth per game: mean %.2f  p90 %d  max %d;  mean depth over the search: mean %.2f max %.2f" %
          (d.mean(), np.percentile(d, 90), d.max(), sd.mean(), sd.max()))


if __name__ == "__main__":
    main()
