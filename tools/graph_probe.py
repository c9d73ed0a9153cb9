#!/usr/bin/env python
"""Does a CUDA graph of the search loop beat stream launches?  Times mzgo_search_run (S simulations on G games)
launched directly and replayed from a captured graph, plus the tree step and the recurrent kernel alone
(graph replay, so no host launch overhead is in the numbers)."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import muzero_go_b200 as mz  # noqa: E402
from muzero_go_b200 import _lib  # noqa: E402


def timed(fn, reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=9)
    ap.add_argument("--games", type=int, default=1024)
    ap.add_argument("--sims", type=int, default=50)
    args = ap.parse_args()
    n, G, S = args.n, args.games, args.sims
    A = n * n + 1
    dev = torch.device("cuda")
    torch.manual_seed(0)
    net = mz.MuZeroNet(128, A, precision="bf16").to(dev).eval()
    eng = mz.SelfPlayEngine(net, n, G, S, traj_capacity=2)
    lib = _lib.load()
    eng.run(1)
    torch.cuda.synchronize()
    d = eng.descriptor()
    side = torch.cuda.Stream()

    def prep():
        eng.tree.reset_roots(eng.search, eng.root_policy, eng.boards, n, noise=None, move_index=0)

    def search():
        _lib.check(lib.mzgo_search_run(d, None, 0, _lib.stream_ptr()))

    prep(); search()
    t_direct = timed(lambda: (prep(), search()), 5)
    t_prep = timed(prep, 5)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        prep()
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=side):
            search()
    def replay():
        prep(); g.replay()
    replay()
    t_graph = timed(replay, 5)
    print(f"n={n} G={G} S={S}: search direct {1e3 * (t_direct - t_prep) / S:.2f} us/sim, graph {1e3 * (t_graph - t_prep) / S:.2f} us/sim")

    # single kernels from graphs of 16 launches each
    p = eng.search.to_struct()
    args_step = (eng.tree.desc, p, _lib.ptr(eng.leaf_policy), _lib.ptr(eng.leaf_reward), _lib.ptr(eng.leaf_value), None, 0)
    d2 = eng.descriptor(); d2.num_simulations = max(1, S - 40)
    prep(); _lib.check(lib.mzgo_search_run(d2, None, 0, _lib.stream_ptr()))
    _lib.check(lib.mzgo_tree_step(*args_step, 0, 1, _lib.stream_ptr()))
    torch.cuda.synchronize()
    g2 = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g2, stream=side):
            for _ in range(16):
                _lib.check(lib.mzgo_tree_step(*args_step, 1, 1, _lib.stream_ptr()))
    t_tree = timed(g2.replay, 1) / 16
    args_rec = (d.net, _lib.ptr(eng.arena), _lib.ptr(eng.tree.leaf_in_slot), _lib.ptr(eng.tree.leaf_action),
                _lib.ptr(eng.tree.leaf_out_slot), _lib.ptr(eng.leaf_reward), _lib.ptr(eng.leaf_value),
                _lib.ptr(eng.leaf_policy), 1, G, _lib.stream_ptr())
    g3 = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g3, stream=side):
            for _ in range(16):
                a = list(args_rec); a[-1] = _lib.stream_ptr()
                _lib.check(lib.mzgo_net_recurrent(*a))
    g3.replay()
    t_rec = timed(g3.replay, 3) / 16
    print(f"   tree step {1e3 * t_tree:.2f} us, recurrent kernel {1e3 * t_rec:.2f} us (graph replay, 16 launches each)")


if __name__ == "__main__":
    main()
