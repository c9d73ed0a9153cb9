#!/usr/bin/env python
"""How the placement of the latent slots in a 160 GB arena changes the time of one dynamics launch (19x19, 2048 leaves):
sequential slots, game-major slots (slot = g * M + node, the engine's layout) and node-major slots (slot = node * G + g).
    python tools/arena_probe.py [--games 2048] [--nodes 801]"""
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
    ap.add_argument("--games", type=int, default=2048)
    ap.add_argument("--nodes", type=int, default=801)
    args = ap.parse_args()
    n, G, M = args.n, args.games, args.nodes
    A = n * n + 1
    dev = torch.device("cuda")
    torch.manual_seed(0)
    net = mz.MuZeroNet(128, A, precision="bf16").to(dev).eval()
    lib = _lib.load()
    d = net.descriptor(G)
    arena = torch.empty(G * M * net.slot_bytes(), dtype=torch.uint8, device=dev)
    arena[: 64 << 20].zero_()
    g = torch.arange(G, dtype=torch.int64, device=dev)
    gen = torch.Generator(device="cpu").manual_seed(1)
    action = torch.randint(0, A, (G,), dtype=torch.int32, generator=gen).to(dev)
    rew, val, pol = torch.empty(G, device=dev), torch.empty(G, device=dev), torch.empty(G, A, device=dev)

    def time_it(name, in_slot, out_slot):
        in_slot, out_slot = in_slot.to(torch.int32).contiguous(), out_slot.to(torch.int32).contiguous()
        call = (d, _lib.ptr(arena), _lib.ptr(in_slot), _lib.ptr(action), _lib.ptr(out_slot), _lib.ptr(rew), _lib.ptr(val),
                _lib.ptr(pol), 1, G, _lib.stream_ptr())
        for _ in range(3):
            _lib.check(lib.mzgo_net_recurrent(*call))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(20):
            _lib.check(lib.mzgo_net_recurrent(*call))
        e1.record()
        torch.cuda.synchronize()
        print(f"{name:64s} {e0.elapsed_time(e1) / 20 * 1e3:7.1f} us per launch")

    node_in = torch.randint(0, M - 1, (G,), generator=gen).to(dev)
    half_root = torch.where(torch.rand(G, generator=gen).to(dev) < 0.45, torch.zeros_like(node_in), node_in)
    node_out = torch.full((G,), M - 1, dtype=torch.int64, device=dev)
    print(f"n={n} games={G} nodes per game={M}: arena {arena.numel() / 1e9:.1f} GB")
    time_it("sequential slots (in = g, out = G + g)", g, G + g)
    time_it("game-major, random parent (in = g M + rand, out = g M + M-1)", g * M + node_in, g * M + node_out)
    time_it("game-major, 45 % of the parents are the root", g * M + half_root, g * M + node_out)
    time_it("node-major, random parent (in = rand G + g, out = (M-1) G + g)", node_in * G + g, node_out * G + g)
    time_it("node-major, 45 % of the parents are the root", half_root * G + g, node_out * G + g)


if __name__ == "__main__":
    main()
