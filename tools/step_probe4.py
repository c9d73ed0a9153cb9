#!/usr/bin/env python
"""Where a 9x9 lock-step move spends its time on mixed-phase positions vs openings: the pieces of mzgo_selfplay_run
driven one by one through the C ABI between CUDA events (medians over moves)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import muzero_go_b200 as mz  # noqa: E402
from muzero_go_b200 import _lib  # noqa: E402

n, G, S = 9, 1024, 50
dev = torch.device("cuda")
torch.manual_seed(0)
net = mz.MuZeroNet(128, n * n + 1, precision="bf16").to(dev).eval()
L = _lib.load()
for mixed in (False, True):
    group = mz.WaveGroup(net, n, G, S, waves=1, search=mz.SearchConfig(seed=1000), traj_capacity=2)
    if mixed:
        bench.advance_games(mz, torch, group, n, seed=77)
    e = group.engines[0]
    group.run(2)
    torch.cuda.synchronize()
    d = e.descriptor()
    s = _lib.stream_ptr()
    roots = (torch.arange(G, dtype=torch.int32, device=dev) * e.M).contiguous()
    names = ["initial", "reset_roots", "search", "choose", "rules_step", "score"]
    acc = {k: [] for k in names}
    for mv in range(12):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(len(names) + 1)]
        ev[0].record()
        _lib.check(L.mzgo_net_initial(d.net, None, _lib.ptr(e.boards.records), _lib.ptr(e.arena), _lib.ptr(roots),
                                      _lib.ptr(e.root_value_net), _lib.ptr(e.root_policy), 1, G, s))
        ev[1].record()
        e.tree.active.fill_(1)
        e.tree.reset_roots(e.search, e.root_policy, e.boards, n, noise=None, move_index=100 + mv)
        d.tree.move_index = 100 + mv
        ev[2].record()
        _lib.check(L.mzgo_search_run(d, None, 0, s))
        ev[3].record()
        action, policy, rv = e.tree.choose_action(e.search, 0.0, move_index=100 + mv)
        ev[4].record()
        e.boards.step(action, return_status=False)
        ev[5].record()
        e.boards.score(0.0)
        ev[6].record()
        torch.cuda.synchronize()
        for i, k in enumerate(names):
            acc[k].append(ev[i].elapsed_time(ev[i + 1]) * 1e3)
    print(f"mixed={mixed}: " + "  ".join(f"{k} {np.median(v):.0f} us" for k, v in acc.items()),
          f"  dropped {int(e.tree.dropped.sum())} sims_done {int(e.tree.sims_done.sum())}")
