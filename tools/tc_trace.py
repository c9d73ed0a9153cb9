#!/usr/bin/env python
"""Role timeline of the fused dynamics kernel (k_tc_conv<128,128,DYN>): globaltimer stamps written by
the kernel when mzgo_debug_tc_trace() is armed.  Usage (on a B200): python tools/tc_trace.py --n 9 --leaves 1024"""
import argparse
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import muzero_go_b200 as mz  # noqa: E402
from muzero_go_b200 import _lib  # noqa: E402

NAMES = {0: "entry", 1: "setup done", 2: "mma g0 start", 3: "mma g0 issued", 4: "mma g1 start", 5: "mma g1 issued",
         6: "mma g2 start", 7: "mma g2 issued", 10: "epi g0 first acc", 11: "epi g0 done", 12: "epi g1 first acc",
         13: "epi g1 done", 14: "epi g2 first acc", 15: "epi g2 done", 27: "fin g1 woke", 28: "fin g1 heads", 29: "fin g1 exps", 24: "fin g2 woke", 25: "fin g2 heads", 26: "fin g2 exps",
         18: "epi g2 acc in regs", 19: "epi g2 stored",
         20: "fin g0", 21: "fin g1", 22: "fin g2", 30: "exit"}
for _g in range(12):            # every group of the launch (19x19: seven)
    NAMES[32 + 2 * _g] = f"MMA group {_g} start"
    NAMES[33 + 2 * _g] = f"MMA group {_g} issued"
NAMES[56] = "epilogue of the last group done"
NAMES[57] = "finaliser of the last group done"
NAMES[58] = "TMEM released (leave)"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=9)
    ap.add_argument("--leaves", type=int, default=1024)
    args = ap.parse_args()
    n, B = args.n, args.leaves
    A = n * n + 1
    dev = torch.device("cuda")
    torch.manual_seed(0)
    net = mz.MuZeroNet(128, A, precision="bf16").to(dev).eval()
    lib = _lib.load()
    lib.mzgo_debug_tc_trace.restype = ctypes.c_int
    lib.mzgo_debug_tc_trace.argtypes = [ctypes.c_void_p]
    slots = 2 * B
    arena = torch.zeros(slots * net.slot_bytes() // 2, dtype=torch.bfloat16, device=dev)
    lat = torch.randn(B, 128, n, n, device=dev).relu()
    d = net.descriptor(B)
    in_slot = torch.arange(B, dtype=torch.int32, device=dev)
    out_slot = in_slot + B
    _lib.check(lib.mzgo_net_import_latent(d, _lib.ptr(arena), _lib.ptr(in_slot), _lib.ptr(lat), B, _lib.stream_ptr()))
    action = torch.randint(0, A, (B,), dtype=torch.int32, device=dev)
    rew = torch.empty(B, device=dev)
    val = torch.empty(B, device=dev)
    pol = torch.empty(B, A, device=dev)
    call = (d, _lib.ptr(arena), _lib.ptr(in_slot), _lib.ptr(action), _lib.ptr(out_slot), _lib.ptr(rew), _lib.ptr(val),
            _lib.ptr(pol), 1, B, _lib.stream_ptr())
    for _ in range(3):
        _lib.check(lib.mzgo_net_recurrent(*call))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(10):
        _lib.check(lib.mzgo_net_recurrent(*call))
    e1.record()
    torch.cuda.synchronize()
    print(f"n={n} leaves={B}: {e0.elapsed_time(e1) / 10 * 1e3:.1f} us per launch (no trace)")
    trace = torch.zeros(148 * 64, dtype=torch.int64, device=dev)
    lib.mzgo_debug_tc_trace(trace.data_ptr())
    _lib.check(lib.mzgo_net_recurrent(*call))
    torch.cuda.synchronize()
    lib.mzgo_debug_tc_trace(None)
    t = trace.cpu().numpy().reshape(148, 64).astype(np.float64)
    t0 = t[:, 0][t[:, 0] > 0].min()
    print(f"{'event':18s} {'min':>8s} {'median':>8s} {'max':>8s}   (us after the first CTA entered; CTAs reporting)")
    for k in sorted(NAMES):
        col = t[:, k]
        col = col[col > 0]
        if col.size == 0:
            continue
        r = (col - t0) / 1e3
        print(f"{NAMES[k]:18s} {r.min():8.2f} {np.median(r):8.2f} {r.max():8.2f}   {col.size}")


if __name__ == "__main__":
    main()
