#!/usr/bin/env python
"""Leaf depth of the last selection of a move, per game, on the bench's mixed-phase positions (how deep the deepest
games of a lock-step tree step are, and how many live root edges they have).  python tools/depth_probe.py [--n 19]"""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import muzero_go_b200 as mz  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=19)
    ap.add_argument("--games", type=int, default=1024)
    ap.add_argument("--sims", type=int, default=800)
    ap.add_argument("--moves", type=int, default=4)
    args = ap.parse_args()
    n, G, S = args.n, args.games, args.sims
    dev = torch.device("cuda")
    torch.manual_seed(0)
    net = mz.MuZeroNet(128, n * n + 1, precision="bf16").to(dev).eval()
    group = mz.WaveGroup(net, n, G, S, waves=1, search=mz.SearchConfig(seed=1000), traj_capacity=max(2, args.moves))
    bench.advance_games(mz, torch, group, n, seed=77)
    e = group.engines[0]
    for m in range(args.moves):
        group.run(1, traj_offset=m)
        torch.cuda.synchronize()
        d = e.tree.leaf_depth.cpu().numpy()
        live = (e.tree.root_mask > 0).sum(dim=1).cpu().numpy()
        order = np.argsort(-d)[:8]
        print(f"move {m}: leaf depth mean {d.mean():.2f} p90 {np.percentile(d, 90):.0f} p99 {np.percentile(d, 99):.0f} max {d.max()}; "
              f"deepest games: depth {d[order].tolist()} live root edges {live[order].tolist()}; games with <= 2 live edges: {(live <= 2).sum()}")


if __name__ == "__main__":
    main()
