#!/usr/bin/env python
"""Phase timeline of k_tree_step inside the real search loop (graph replay, PDL): per-game globaltimer stamps of the
last selecting step of a move.  Needs the trace build:
    python -m muzero_go_b200.build --define MZGO_TRACE_TREE && MZGO_KEEP_SO=1 python tools/tree_trace.py"""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import muzero_go_b200 as mz  # noqa: E402
from muzero_go_b200 import _lib  # noqa: E402

NAMES = {0: "block entered", 1: "previous kernel complete (griddep wait)", 2: "first batch of loads arrived",
         3: "expand + backup issued, path statistics arrived", 4: "barrier before the selection",
         5: "root level selected", 6: "level 2 loads arrived", 7: "level 2 selected", 8: "level 3 loads arrived",
         9: "level 3 selected", 15: "leaf record written (end)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=19)
    ap.add_argument("--games", type=int, default=1024)
    ap.add_argument("--sims", type=int, default=800)
    args = ap.parse_args()
    n, G, S = args.n, args.games, args.sims
    dev = torch.device("cuda")
    torch.manual_seed(0)
    net = mz.MuZeroNet(128, n * n + 1, precision="bf16").to(dev).eval()
    group = mz.WaveGroup(net, n, G, S, waves=1, search=mz.SearchConfig(seed=1000), traj_capacity=2)
    bench.advance_games(mz, torch, group, n, seed=77)
    L = _lib.load()
    for _ in range(3):
        group.run(1)
    torch.cuda.synchronize()
    trace = torch.zeros(G * 16, dtype=torch.int64, device=dev)
    _lib.check(L.mzgo_debug_tree_trace(trace.data_ptr()))
    group.run(1)
    torch.cuda.synchronize()
    _lib.check(L.mzgo_debug_tree_trace(None))
    t = trace.cpu().numpy().reshape(G, 16).astype(np.float64)
    t0 = t[:, 0][t[:, 0] > 0].min()
    print(f"n={n} games={G} sims={S}: the last selecting tree step of a move (us after the first block entered; games reporting)")
    for k in sorted(NAMES):
        col = t[:, k]
        col = col[col >= t0]
        if col.size == 0:
            continue
        r = (col - t0) / 1e3
        print(f"  {NAMES[k]:48s} min {r.min():6.2f}  median {np.median(r):6.2f}  p90 {np.percentile(r, 90):6.2f}  max {r.max():6.2f}   {col.size}")
    dur = (t[:, 15] - t[:, 0]) / 1e3
    print(f"  per game, entry -> end: median {np.median(dur):.2f} us, max {dur.max():.2f} us")


if __name__ == "__main__":
    main()
