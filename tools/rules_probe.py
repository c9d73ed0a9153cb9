#!/usr/bin/env python
"""One rules ply on 2^20 mid-game boards, warp-per-board vs thread-per-board kernels (mzgo_debug_rules_tpb_min).
Usage (on a B200): python tools/rules_probe.py [board sizes, default 9 13 19]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import muzero_go_b200 as mz  # noqa: E402
from muzero_go_b200 import _lib  # noqa: E402


def main():
    lib = _lib.load()
    NB = 1 << 20
    sizes = [int(a) for a in sys.argv[1:]] or [9, 13, 19]
    for n, warm in ((9, 40), (13, 90), (19, 180)):
        if n not in sizes:
            continue
        rng = np.random.Generator(np.random.Philox(n))
        for mode, tmin in (("warp per board", 1 << 30), ("thread per board", 0)):
            lib.mzgo_debug_rules_tpb_min(1 << 30)
            gb = mz.GoBoards(NB, n).reset()
            R0 = torch.from_numpy(rng.integers(0, 2 ** 31, size=(NB, warm), dtype=np.int64)).to(torch.int32).cuda()
            gb.playout(R0, want_actions=False)                     # the same mid-game positions for both modes
            lib.mzgo_debug_rules_tpb_min(tmin)
            R = torch.from_numpy(np.random.Generator(np.random.Philox(7)).integers(0, 2 ** 31, size=(NB, 1), dtype=np.int64)).to(torch.int32).cuda()
            times = []
            for _ in range(6):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                gb.playout(R, want_actions=False)
                e1.record()
                torch.cuda.synchronize()
                times.append(e0.elapsed_time(e1))
            ms = float(np.median(times[1:]))
            print(f"n={n:2d} {mode:16s}: {ms:7.3f} ms per ply of {NB} boards = {NB / ms / 1e3:7.1f} M plies/s"
                  f" ({(3 * n * 4 + 16) * 2 * NB / ms / 1e6:6.1f} GB/s of packed state)")
            rng = np.random.Generator(np.random.Philox(n))
    lib.mzgo_debug_rules_tpb_min(32768)


if __name__ == "__main__":
    main()
