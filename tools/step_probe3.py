#!/usr/bin/env python
"""Per-CTA timeline of ONE mid-search fused launch (network kernel + tree phase with selection) at 9x9 on mixed-phase
positions vs openings: the search loop driven launch by launch through the C ABI, the trace armed for simulation k."""
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
MOVES = int(os.environ.get("PROBE_MOVES", "2"))      # moves played before the probed search
for mixed in (False, True):
    group = mz.WaveGroup(net, n, G, S, waves=1, search=mz.SearchConfig(seed=1000), traj_capacity=max(2, MOVES))
    if mixed:
        bench.advance_games(mz, torch, group, n, seed=77)
    e = group.engines[0]
    group.run(MOVES)
    torch.cuda.synchronize()
    d = e.descriptor()
    s = _lib.stream_ptr()
    roots = (torch.arange(G, dtype=torch.int32, device=dev) * e.M).contiguous()
    _lib.check(L.mzgo_net_initial(d.net, None, _lib.ptr(e.boards.records), _lib.ptr(e.arena), _lib.ptr(roots),
                                  _lib.ptr(e.root_value_net), _lib.ptr(e.root_policy), 1, G, s))
    e.tree.active.fill_(1)
    e.tree.reset_roots(e.search, e.root_policy, e.boards, n, noise=None, move_index=7)
    p = e.search.to_struct()
    _lib.check(L.mzgo_tree_step(e.tree.desc, p, _lib.ptr(e.leaf_policy), _lib.ptr(e.leaf_reward), _lib.ptr(e.leaf_value), None, 0, 0, 1, s))
    trace = torch.zeros(148 * 64, dtype=torch.int64, device=dev)
    for k in range(S):
        if k in (10, 30, 45):
            torch.cuda.synchronize()
            trace.zero_()
            L.mzgo_debug_tc_trace(trace.data_ptr())
        _lib.check(L.mzgo_net_recurrent_tree(d.net, _lib.ptr(e.arena), e.tree.desc, p, None, 0, _lib.ptr(e.leaf_reward),
                                             _lib.ptr(e.leaf_value), _lib.ptr(e.leaf_policy), 1 if k + 1 < S else 0, s))
        if k in (10, 30, 45):
            torch.cuda.synchronize()
            L.mzgo_debug_tc_trace(None)
            t = trace.cpu().numpy().reshape(148, 64).astype(np.float64)
            t0 = t[:, 0].min()
            roles, done = (t[:, 30] - t0) / 1e3, (t[:, 31] - t0) / 1e3
            depth = e.tree.leaf_depth.cpu().numpy()
            print(f"mixed={mixed} sim {k}: roles done median {np.median(roles):.1f} max {roles.max():.1f} us; tree phase duration "
                  f"median {np.median(done - roles):.1f} p90 {np.percentile(done - roles, 90):.1f} max {(done - roles).max():.1f} us; "
                  f"kernel end max {done.max():.1f} us; next leaf depth mean {depth.mean():.2f} max {depth.max()}")
