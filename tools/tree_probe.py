#!/usr/bin/env python
"""Brings G trees to the late state of an S-simulation search with random network outputs and then runs a few
fused tree steps: a small target for `ncu -k regex:k_tree_step -s <S> -c 2` (the launches after the search)."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import muzero_go_b200 as mz  # noqa: E402
from muzero_go_b200 import _lib  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=19)
    ap.add_argument("--games", type=int, default=1024)
    ap.add_argument("--sims", type=int, default=700)
    ap.add_argument("--extra", type=int, default=8)
    args = ap.parse_args()
    n, G, S = args.n, args.games, args.sims
    A = n * n + 1
    dev = torch.device("cuda")
    torch.manual_seed(0)
    lib = _lib.load()
    cfg = mz.SearchConfig(seed=3)
    boards = mz.GoBoards(G, n).reset()
    tree = mz.Tree(G, A, S + args.extra + 2)
    soft = torch.softmax(torch.randn(G, A, device=dev), dim=1).contiguous()
    tree.reset_roots(cfg, soft, boards, n, noise=None, move_index=0)
    rew = (0.1 * torch.randn(G, device=dev)).contiguous()
    val = (0.5 * torch.randn(G, device=dev)).contiguous()
    p = cfg.to_struct()
    a = (tree.desc, p, _lib.ptr(soft), _lib.ptr(rew), _lib.ptr(val), None, 0)
    _lib.check(lib.mzgo_tree_step(*a, 0, 1, _lib.stream_ptr()))
    for _ in range(S - 1):
        _lib.check(lib.mzgo_tree_step(*a, 1, 1, _lib.stream_ptr()))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            for _ in range(args.extra):
                _lib.check(lib.mzgo_tree_step(*a, 1, 1, _lib.stream_ptr()))
    d0 = int(tree.depth_sum.sum())
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    depth = (int(tree.depth_sum.sum()) - d0) / (G * args.extra)
    print(f"n={n} G={G} after {S} sims: {1e3 * e0.elapsed_time(e1) / args.extra:.2f} us per tree step, mean leaf depth {depth:.2f}")


if __name__ == "__main__":
    main()
