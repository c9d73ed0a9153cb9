#!/usr/bin/env python
"""Self-play throughput bench (BASELINE.json metric: MCTS simulations/sec; moves/sec beside it).

    python bench.py --gpus N --steps K --warmup W            # our arm: CUDA hot path through the C ABI
    python bench.py --impl reference ...                      # CPU arm: the oracle port of the reference

Default workload: BASELINE.json configs[3] (19x19, 2048 games, 800 sims/move, bf16), the north-star configuration;
the configs[1] figures (9x9, 1024 games, 50 sims) ride in the same line under "secondary".

A step = one lock-step move of every concurrent game: initial inference, root preparation,
`sims` x (select -> recurrent inference -> expand/backup), action choice, rules step, trajectory row.
Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for the definitions used here.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # BASELINE.json configs[1] / configs[3]
    "9x9": dict(n=9, games=1024, sims=50, name="9x9 Go batched self-play, 1024 concurrent games, 50 sims/move"),
    "19x19": dict(n=19, games=2048, sims=800, name="19x19 Go batched self-play, 2048 games, 800 sims/move"),
}
LATENT = 128


def flops_recurrent(n, c=LATENT):
    """dynamics conv3x3 + the three 1x1 heads, per simulation."""
    return 2 * 9 * c * c * n * n + 2 * 3 * c * n * n


def flops_initial(n, c=LATENT):
    return 2 * 9 * n * n * (6 * 64 + 64 * 64 + 64 * c) + 2 * 2 * c * n * n


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            p = json.load(fh)
        return dict(hbm_gbs=p["hbm_gbs"], bf16_burst=p["bf16_tflops"], bf16_sustained=p["bf16_tflops_sustained"],
                    source="measured")
    return dict(hbm_gbs=6650.0, bf16_burst=1590.0, bf16_sustained=1400.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def count(self, since):
        return sum(1 for t, _ in self.rows if t >= since)

    def stop(self, since=0.0):
        """Summary of the samples read after host time `since` (the GPU has been under the bench load from then on)."""
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:                                   # gone for good before anything else is timed (its NVML teardown stalls launches)
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        time.sleep(0.2)
        sm, pw, mx, reasons = [], [], None, set()
        for t, r in self.rows:
            if len(r) < 7 or t < since:
                continue
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                pw.append(float(r[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        pw.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "power_w_median": pw[len(pw) // 2] if pw else None}


# ------------------------------------------------------------------------------------------------ CPU arm
def _cpu_worker(args):
    """Oracle port of the reference's self-play move loop (main.py:656-702): reference networks in torch
    fp32 on the CPU, the oracle MCTS restatement, the oracle GymGo restatement.  Runs for ~`budget` s."""
    n, sims, budget, seed = args
    import numpy as np
    import torch
    from oracle import gymgo, mcts_ref
    from oracle.nets_ref import MuZeroNetRef
    torch.set_num_threads(1)
    torch.manual_seed(0)
    net = MuZeroNetRef(LATENT, n * n + 1).eval()
    initial, recurrent = mcts_ref.torch_net_callbacks(net)
    rng = np.random.default_rng(seed)
    A = n * n + 1
    env = gymgo.GoEnv(n)
    obs = env.reset()
    # mixed game phases like the GPU arm (config "positions"): a random 0-60 % of the move cap, untimed
    for _ in range(int(rng.integers(0, int(0.6 * 1.5 * n * n) + 1))):
        v = np.nonzero(env.valid_moves())[0]
        obs, _, done, _ = env.step(int(v[rng.integers(len(v) - 1)]) if len(v) > 1 else int(v[0]))
        if done:
            obs = env.reset()
            break
    sims_done = moves = 0
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < budget:
        noise = rng.dirichlet([0.03] * A)
        stream = rng.integers(0, 2 ** 32, size=sims + 8, dtype=np.uint64).astype(np.uint32)
        tree, mask = mcts_ref.run_mcts(obs, initial, recurrent, sims, noise, stream)
        vc = tree.root_visit_counts(mask)
        action, _ = mcts_ref.choose_action(vc, mask, int(rng.integers(1 << 30)))
        obs, _, done, _ = env.step(action)
        sims_done += tree.visit[0]
        moves += 1
        if done or moves % int(1.5 * n * n) == 0:
            obs = env.reset()
    return sims_done, moves, time.perf_counter() - t0


def cpu_arm(n, sims, budget, procs):
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    with ctx.Pool(procs) as pool:
        res = pool.map(_cpu_worker, [(n, sims, budget, 100 + i) for i in range(procs)])
    total_s = sum(r[0] for r in res)
    total_m = sum(r[1] for r in res)
    wall = max(r[2] for r in res)
    return total_s / wall, total_m / wall, wall


def run_reference(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    procs = os.cpu_count() or 1
    n, sims = wl["n"], args.sims or wl["sims"]
    per_step = max(4.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_arm(n, sims, min(2.0, per_step), procs)
    vals, mvals, t_total = [], [], 0.0
    for _ in range(args.steps):
        s, m, wall = cpu_arm(n, sims, per_step, procs)
        vals.append(s)
        mvals.append(m)
        t_total += wall
    v = sum(vals) / len(vals)
    sample = (f"{procs} single-thread processes x {per_step:.0f} s of the oracle port of main.py's self-play loop "
              f"({n}x{n}, {sims} sims/move, C={LATENT}, one game per process)")
    line = {"impl": "reference", "metric": "mcts_simulations_per_sec", "value": v, "unit": "sims/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * t_total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(wl, args.games or wl["games"], sims, args.precision),
            "moves_per_sec": sum(mvals) / len(mvals),
            "cpu_baseline": {"value": v, "unit": "sims/s", "cores": procs, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "sims/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# ------------------------------------------------------------------------------------------------ GPU arm
def config_dict(wl, games, sims, precision, positions="mixed"):
    """The same dict in both arms (the CPU arm runs this workload's per-game loop on the host cores)."""
    return {"workload": wl["name"], "board": wl["n"], "games_per_gpu": games, "sims_per_move": sims, "latent_dim": LATENT,
            "weights": "random-init (torch.manual_seed(0))", "precision": precision,
            "positions": ("every game advanced a random 0-60 % of the move cap by random legal moves before the warm-up"
                          if positions == "mixed" else "every game starts from the empty board (openings only)"),
            "cache": "inputs larger than L2 (the latent arena is > 1 GB per GPU and every simulation reads and writes fresh slots)",
            "parallelism": "games sharded over the GPUs, no collective inside the search"}


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per unit from the newest committed ncu --set full capture
    (profiles/<round>/ncu_traffic.json, written by tools/summarise_ncu.py).  None when there is no capture."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*", "ncu_traffic.json")))
    if not files:
        return None
    with open(files[-1]) as fh:
        d = json.load(fh)
    d["file"] = os.path.relpath(files[-1], ROOT)
    return d


def advance_games(mz, torch, group, n, seed):
    """Mixed game phases: every game is advanced by a random 0-60 % of the move cap with random legal moves (the
    configs[2] playout kernel), so the timed region sees openings, middle games and endgames side by side -- the steady
    state of a long self-play job -- instead of 2048 empty boards."""
    gen = torch.Generator(device="cpu").manual_seed(seed)
    buckets = 8
    for e in group.engines:
        cap = e.max_game_moves
        which = torch.randint(0, buckets, (e.G,), generator=gen)
        for b in range(1, buckets):
            idx = torch.nonzero(which == b).flatten().to(e.device)
            plies = int(0.6 * cap * b / (buckets - 1))
            if idx.numel() == 0 or plies == 0:
                continue
            tmp = mz.GoBoards(int(idx.numel()), n, e.device).reset()
            R = torch.randint(0, 2 ** 31 - 1, (int(idx.numel()), plies), dtype=torch.int32, generator=gen)
            tmp.playout(R, want_actions=False)
            e.boards.records[idx] = tmp.records
            e.move_count[idx] = tmp.records[:, 62]
            e.pass_streak[idx] = tmp.records[:, 61] & 1
    torch.cuda.synchronize()


class TrajectoryExchange:
    """The trajectory gather of north_star (4), off the search's critical path: every step's rows are packed with
    slice copies into a ring of `depth` fixed-stride rows (no allocation), and every `depth` steps the ring is
    all-gathered on a side stream ordered behind the compute stream by an event ("batched per K moves so
    messages are >= 1 MB").  With one rank the rows are only packed."""

    def __init__(self, torch, dist, mzdist, group, A, world, depth):
        self.torch, self.mzdist, self.world, self.depth = torch, mzdist, world, depth
        self.ring = mzdist.RowRing(group.G, A, depth, group.device)
        self.side = torch.cuda.Stream()
        self.out = (torch.empty((world,) + tuple(self.ring.buf.shape), dtype=torch.uint8, device=group.device)
                    if world > 1 else None)
        self.filled = 0
        self.bytes_gathered = 0

    def push(self, group, row):
        self.ring.write_from(self.filled, group, row)
        self.filled += 1
        if self.filled == self.depth:
            self.flush()

    def flush(self):
        if self.filled and self.world > 1:
            torch = self.torch
            ev = torch.cuda.Event()
            ev.record()
            with torch.cuda.stream(self.side):
                self.side.wait_event(ev)
                self.mzdist.all_gather_rows(self.out, self.ring.buf)
            # the ring is rewritten by the next step: order that step behind the gather
            torch.cuda.current_stream().wait_stream(self.side)
            self.bytes_gathered += self.ring.buf.numel() * self.world
        self.filled = 0


def measure(args, wl, G, S, K, W, rank, world, dev, primary, positions="mixed"):
    """One workload on this rank: whole-step throughput (device timed), end to end with host buffers, and -- for the
    primary workload -- the three kernel families timed alone."""
    import torch
    import torch.distributed as dist
    import muzero_go_b200 as mz
    from muzero_go_b200 import _lib
    from muzero_go_b200 import dist as mzdist

    n = wl["n"]
    A = n * n + 1
    local = dev.index
    torch.manual_seed(0)                                    # random-init weights of the reference architecture
    net = mz.MuZeroNet(LATENT, A, precision=args.precision).to(dev).eval()
    bytes_bcast = mzdist.broadcast_weights(net, src=0)       # weight version 0 from the learner rank
    # The latent arena of one search is G * (S + 1) slots.  When the named game count does not fit the HBM at
    # once (19x19 / 800 sims / 2048 games: 177 GB of bf16 latents) the games run as `waves` equal groups that
    # take turns on one set of search buffers (muzero_go_b200.WaveGroup): still G games, one move each per step.
    free_b, _ = torch.cuda.mem_get_info()
    per_game = (S + 1) * (net.slot_bytes() + A * 18 + 8) + 64 * 1024
    # `streams` sets of search buffers are alive at once (waves on different streams overlap: one wave's tree step
    # under another wave's network kernel, muzero_go_b200.WaveGroup)
    streams = max(1, args.streams)
    waves = max(streams, (max(args.waves, 1) + streams - 1) // streams * streams)
    while streams * (G // waves) * per_game > 0.92 * free_b or G % waves:
        waves += streams
    group = mz.WaveGroup(net, n, G, S, waves=waves, streams=streams, search=mz.SearchConfig(seed=1000, game_base=rank * G),
                         traj_capacity=2, restart_finished=True)
    eng = group.engines[0]          # the single-kernel timings below run on one wave's buffers
    GW = G // waves
    lib = _lib.load()
    if positions == "mixed":
        advance_games(mz, torch, group, n, seed=77 + rank)
    exch_depth = max(1, min(64, (1 << 20) // max(1, G * mzdist.row_stride(A)) + 1))
    exch = TrajectoryExchange(torch, dist, mzdist, group, A, world, exch_depth)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(row=0):
        group.run(1, traj_offset=row)
        exch.push(group, row)

    sampler = ClockSampler(local) if primary else None
    if sampler:
        sampler.start()             # nvidia-smi takes longer to start than a short timed region lasts: start it first
    t_load = time.perf_counter()
    for _ in range(W):
        step()
    exch.flush()
    barrier()
    lib.mzgo_launch_count(1)
    c0 = group.read_counters()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(K):
        step()
    exch.flush()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    launches = int(lib.mzgo_launch_count(0))
    c1 = group.read_counters()
    clocks = None
    if sampler:
        # a short timed region (9x9: a few ms per step) ends before nvidia-smi's 100 ms sampler has seen it: the same
        # steps keep running, untimed, until 0.6 s of load have been sampled (a function of the max-over-ranks time, so
        # that every rank runs the same number of gathers); the games are put back afterwards
        ms_all = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms_all, op=dist.ReduceOp.MAX)
        ms_max = float(ms_all.item())
        extra = 0 if ms_max >= 600.0 or os.environ.get("MZGO_BENCH_NO_EXTRA") else min(2000, int((600.0 - ms_max) / max(ms_max / K, 1e-3)) + 1)
        saved = [(e.boards.records.clone(), e.move_count.clone(), e.pass_streak.clone(), e.game_temperature.clone(),
                  e.moves_enqueued) for e in group.engines]
        for _ in range(extra):
            step()
        exch.flush()
        torch.cuda.synchronize()
        clocks = sampler.stop(t_load + 0.05)
        for e, sv in zip(group.engines, saved):
            e.boards.records.copy_(sv[0]); e.move_count.copy_(sv[1]); e.pass_streak.copy_(sv[2])
            e.game_temperature.copy_(sv[3]); e.moves_enqueued = sv[4]
        del saved
    barrier()
    sims_done = c1["simulations"] - c0["simulations"]
    moves_done = c1["moves"] - c0["moves"]
    depth = (c1["leaf_depth_sum"] - c0["leaf_depth_sum"]) / max(1, sims_done)

    # ---- end to end through the public API with HOST buffers: positions up, trajectory row down, every step
    host_boards = torch.empty((G, 64), dtype=torch.int32).pin_memory()
    dev_boards = [e.boards.records for e in group.engines]

    def boards_up():
        for w, rec in enumerate(dev_boards):
            rec.copy_(host_boards[w * GW:(w + 1) * GW], non_blocking=True)

    def boards_down(non_blocking=False):
        for w, rec in enumerate(dev_boards):
            host_boards[w * GW:(w + 1) * GW].copy_(rec, non_blocking=non_blocking)

    boards_down()
    row_ring = mzdist.RowRing(G, A, 1, dev)
    host_row = torch.empty(row_ring.buf.shape, dtype=torch.uint8).pin_memory()

    def e2e_step():
        boards_up()
        step()
        row_ring.write_from(0, group, 0)
        host_row.copy_(row_ring.buf, non_blocking=True)
        boards_down(non_blocking=True)                       # next step's input comes from the host again
        torch.cuda.synchronize()                             # both read-backs have landed in pinned memory

    for _ in range(2):                                       # untimed: first use of the pinned buffers and copy paths
        e2e_step()
    barrier()
    c1e = group.read_counters()
    t0 = time.perf_counter()
    for _ in range(K):
        e2e_step()
    exch.flush()
    barrier()
    e2e_s = time.perf_counter() - t0
    c2 = group.read_counters()
    e2e_sims = c2["simulations"] - c1e["simulations"]
    if os.environ.get("MZGO_BENCH_DEBUG"):
        print(f"[bench debug] rank {rank}: timed {ms:.3f} ms / {sims_done} sims, e2e {e2e_s * 1e3:.3f} ms / {e2e_sims} sims",
              file=sys.stderr)
    res = dict(ms=ms, e2e_ms=e2e_s * 1000.0, sims=sims_done, moves=moves_done, e2e_sims=e2e_sims, launches=launches, G=G,
               depth=depth, waves=waves, GW=GW, streams=streams, clocks=clocks, bytes_bcast=bytes_bcast, arena_gb=eng.arena.numel() / 1e9,
               h2d=G * 256, d2h=int(row_ring.buf.numel()) + G * 256, graph_replays=int(lib.mzgo_graph_replays()),
               exchange={"rows_per_gather": exch_depth, "bytes_per_gather_per_rank": int(exch.ring.buf.numel()),
                         "stream": "side stream, event-ordered"})

    # ---- dominant kernel (recurrent inference of one leaf batch), timed alone with CUDA events
    d = eng.descriptor()
    reps = 20
    eng.tree.reset_roots(eng.search, eng.root_policy, eng.boards, n, noise=None, move_index=0)
    eng.tree.select(eng.search, None)
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    args_rec = (d.net, _lib.ptr(eng.arena), _lib.ptr(eng.tree.leaf_in_slot), _lib.ptr(eng.tree.leaf_action),
                _lib.ptr(eng.tree.leaf_out_slot), _lib.ptr(eng.leaf_reward), _lib.ptr(eng.leaf_value),
                _lib.ptr(eng.leaf_policy), 1, GW, _lib.stream_ptr())
    # single kernels are timed as CUDA-graph replays of `reps` launches, so that no host launch overhead sits
    # between the events (a Python-driven loop of 10 us kernels measures the host, not the GPU)
    side = torch.cuda.Stream()

    def graph_time(fn, count, replays=3):
        fn(_lib.stream_ptr())                                # warm (function attributes, tables)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side):
                for _ in range(count):
                    fn(_lib.stream_ptr())
        g.replay()
        torch.cuda.synchronize()
        k0.record()
        for _ in range(replays):
            g.replay()
        k1.record()
        torch.cuda.synchronize()
        return k0.elapsed_time(k1) / (replays * count)

    rec_fn = lambda st: _lib.check(lib.mzgo_net_recurrent(*args_rec[:-1], st))
    # "timed alone" = from an idle GPU: the step above ran for seconds under the power cap, and the clocks take a moment
    # to come back; a kernel timed straight behind it is measured at the capped clocks and held against the BURST peak
    idle_s = 1.5 if primary else 0.0
    torch.cuda.synchronize()
    time.sleep(idle_s)
    kern_ms = graph_time(rec_fn, reps)
    # ... and back to back for about a second (primary workload): the same kernel under the power cap, the way it runs
    # inside a seconds-long step -- this is the figure to hold against the SUSTAINED cuBLAS peak
    kern_ms_sustained = graph_time(rec_fn, reps, replays=max(3, int(1000.0 / (reps * kern_ms)))) if primary else None
    live = int((eng.tree.leaf_action >= 0).sum())
    peaks = measured_peaks()
    achieved = live * flops_recurrent(n) / (kern_ms * 1e-3) / 1e12
    traffic = ncu_traffic()
    tkey = f"rec_bytes_per_leaf_{n}"
    res["roofline"] = {
        "bound": "tensor", "kernel": "mzgo_net_recurrent (dynamics conv3x3 + heads, one leaf batch), timed alone: "
                                     f"CUDA-graph replay of {reps} launches between CUDA events",
        "achieved": achieved, "peak": peaks["bf16_burst"], "unit": "TFLOP/s", "frac": achieved / peaks["bf16_burst"],
        "peak_source": peaks["source"] + " bf16 burst (the kernel is timed in isolation)",
        "frac_of_sustained": achieved / peaks["bf16_sustained"],
        "traffic": (traffic[tkey] * live if traffic and tkey in traffic else None),
        "traffic_unit": "bytes per launch (ncu dram read+write per leaf x leaves)",
        "traffic_source": traffic["file"] if traffic and tkey in traffic else None,
        "algorithmic_bytes": 2 * LATENT * n * n * 2 * live, "launch_ms": kern_ms, "leaves_per_launch": live,
        "idle_before_s": idle_s}
    if kern_ms_sustained:
        sus = live * flops_recurrent(n) / (kern_ms_sustained * 1e-3) / 1e12
        res["roofline"]["sustained"] = {"launch_ms": kern_ms_sustained, "achieved": sus, "peak": peaks["bf16_sustained"],
                                        "frac": sus / peaks["bf16_sustained"],
                                        "what": "the same launches back to back for ~1 s (power cap reached), against the "
                                                "measured sustained bf16 peak"}
    if not primary:
        return res

    # ---- tree step (HBM bound): expand + backup + next selection of all G trees, late in a search
    def bytes_tree(depth_mean):
        return 20 * A * depth_mean + 8 * A + 24 * (depth_mean + 1)

    tree_reps = min(16, max(1, S // 4))
    d_part = eng.descriptor()
    d_part.num_simulations = S - tree_reps - 2
    p_struct = eng.search.to_struct()
    eng.tree.reset_roots(eng.search, eng.root_policy, eng.boards, n, noise=None, move_index=0)
    _lib.check(lib.mzgo_search_run(d_part, None, 0, _lib.stream_ptr()))
    step_args = (eng.tree.desc, p_struct, _lib.ptr(eng.leaf_policy), _lib.ptr(eng.leaf_reward), _lib.ptr(eng.leaf_value),
                 None, 0)
    # mzgo_search_run ends with a backup-only step: select the next leaf first, so that every step below backs up a
    # leaf that a selection really produced (with whatever network outputs are in the buffers) and selects the next
    _lib.check(lib.mzgo_tree_step(*step_args, 0, 1, _lib.stream_ptr()))
    _lib.check(lib.mzgo_tree_step(*step_args, 1, 1, _lib.stream_ptr()))    # warm
    torch.cuda.synchronize()
    g_tree = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g_tree, stream=side):
            for _ in range(tree_reps):
                _lib.check(lib.mzgo_tree_step(*step_args, 1, 1, _lib.stream_ptr()))
    torch.cuda.synchronize()
    dsum0, sims0 = int(eng.tree.depth_sum.sum()), int(eng.tree.sims_done.sum())
    k0.record()
    g_tree.replay()                                          # one replay: every launch expands a node
    k1.record()
    torch.cuda.synchronize()
    tree_ms = k0.elapsed_time(k1) / tree_reps
    tree_sims = int(eng.tree.sims_done.sum()) - sims0
    tree_depth = (int(eng.tree.depth_sum.sum()) - dsum0) / max(1, tree_sims)
    tree_bytes = tree_sims / tree_reps * bytes_tree(tree_depth)
    tree_gbs = tree_bytes / (tree_ms * 1e-3) / 1e9
    tkey = f"tree_bytes_per_game_{n}"
    res["roofline_tree"] = {
        "bound": "hbm", "kernel": "k_tree_step (expand + backup + select, all trees of a wave), timed alone",
        "achieved": tree_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": tree_gbs / peaks["hbm_gbs"],
        "traffic": (traffic[tkey] * tree_sims / tree_reps if traffic and tkey in traffic else None),
        "traffic_source": traffic["file"] if traffic and tkey in traffic else None, "launch_ms": tree_ms,
        "bytes_per_sim": bytes_tree(tree_depth), "mean_leaf_depth": tree_depth, "sims_per_launch": tree_sims / tree_reps,
        "state": f"simulations {S - tree_reps}..{S} of a search"}

    # ---- rules step (HBM bound nominally, 324 B per ply at 19x19 / 108 B at 9x9): 2^20 mid-game boards, one
    # random legal move each per launch (the configs[2] playout kernel with T = 1)
    NB = 1 << 20
    rb = mz.GoBoards(NB, n).reset()
    Rnd = torch.randint(0, 2 ** 31 - 1, (NB, n * n // 2), dtype=torch.int32, device=dev)
    rb.playout(Rnd, want_actions=False)
    rules_reps = 6
    acts = [torch.randint(0, 2 ** 31 - 1, (NB, 1), dtype=torch.int32, device=dev) for _ in range(rules_reps + 2)]
    for i in range(2):
        rb.playout(acts[i], want_actions=False)
    torch.cuda.synchronize()
    g_rules = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g_rules, stream=side):
            for i in range(rules_reps):
                rb.playout(acts[2 + i], want_actions=False)
    torch.cuda.synchronize()
    k0.record()
    g_rules.replay()
    k1.record()
    torch.cuda.synchronize()
    rules_ms = k0.elapsed_time(k1) / rules_reps
    rules_bytes = NB * (2 * (3 * 4 * ((n * n + 31) // 32) + 16) + 4)
    rules_gbs = rules_bytes / (rules_ms * 1e-3) / 1e9
    del rb, Rnd, acts
    tkey = f"rules_bytes_per_ply_{n}"
    res["roofline_rules"] = {
        "bound": "hbm", "kernel": "k_rules_playout, T=1 (one random legal move on each of 2^20 mid-game boards)",
        "achieved": rules_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": rules_gbs / peaks["hbm_gbs"],
        "traffic": (traffic[tkey] * NB if traffic and tkey in traffic else None),
        "traffic_source": traffic["file"] if traffic and tkey in traffic else None,
        "launch_ms": rules_ms, "bytes_per_ply": rules_bytes / NB, "plies_per_sec": NB / (rules_ms * 1e-3)}
    return res


def rules_cpu_baseline(n):
    """BASELINE.md section 3, rules-only CPU point: the C restatement of GymGo (oracle/gymgo_c.c, one host thread) on
    configs[2]'s move table, a bounded sample of boards."""
    import numpy as np
    from oracle import gymgo_fast
    boards, T = (64, 2 * n * n) if n > 9 else (1024, 2 * n * n)
    R = np.random.Generator(np.random.Philox(0)).integers(0, 2 ** 32, size=(boards, T), dtype=np.uint64).astype(np.uint32)
    st = gymgo_fast.init_states(boards, n)
    gymgo_fast.lib()
    t0 = time.perf_counter()
    plies = gymgo_fast.playout(st, R)[0]
    dt = time.perf_counter() - t0
    return {"value": plies / dt, "unit": "plies/s", "cores": 1, "kind": "port",
            "sample": f"oracle/gymgo_c.c, {boards} boards x up to {T} plies of the configs[2] playout ({n}x{n}), one thread, {dt:.1f} s"}


def run_ours(args, wl):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    G, S = args.games or wl["games"], args.sims or wl["sims"]
    K, W = args.steps, args.warmup
    r = measure(args, wl, G, S, K, W, rank, world, dev, primary=True)
    second = None
    if args.workload == "19x19" and not args.no_secondary and not args.games and not args.sims:
        # BASELINE configs[1] beside it (extra key, same line): the 9x9 / 1024 games / 50 sims workload, whose steps
        # are ~2 ms -- timed over enough steps for half a second of device time
        torch.cuda.empty_cache()
        wl9 = WORKLOADS["9x9"]
        r9 = measure(args, wl9, wl9["games"], wl9["sims"], max(K, 250), max(W, 5), rank, world, dev, primary=False)
        # ... and started from empty boards and played on: over these hundreds of moves the games run through whole
        # self-play games (cap 121 moves, restarted at once), i.e. the phases self-play itself produces with this
        # random-init network (round 1 timed 20 moves from the empty board: openings only, 25 M sims/s)
        r9o = measure(args, wl9, wl9["games"], wl9["sims"], max(K, 250), max(W, 5), rank, world, dev, primary=False,
                      positions="openings")
        second = (wl9, r9, r9o)

    def reduce(res):
        t = torch.tensor([res["ms"], res["e2e_ms"]], dtype=torch.float64, device=dev)
        tot = torch.tensor([res["sims"], res["moves"], res["e2e_sims"], res["launches"], res["G"]], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        return t.tolist(), tot.tolist()

    (ms, e2e_ms), (sims_all, moves_all, e2e_sims_all, launches_all, G_all) = reduce(r)
    red9 = reduce(second[1]) if second else None
    red9o = reduce(second[2]) if second else None
    if rank == 0:
        n = wl["n"]
        peaks = measured_peaks()
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            procs = os.cpu_count() or 1
            s, m, wall = cpu_arm(n, S, 12.0, procs)
            cpu = {"value": s, "unit": "sims/s", "cores": procs, "kind": "port", "moves_per_sec": m,
                   "sample": f"{procs} single-thread processes x 12 s of the oracle port of main.py's self-play loop "
                             f"({n}x{n}, {S} sims/move, C={LATENT}); calibration of the port against the real "
                             "main.MCTS.run: BASELINE.md section 3"}
            r["roofline_rules"]["cpu_baseline"] = rules_cpu_baseline(n)
        value = sims_all / (ms * 1e-3)
        cfg = config_dict(wl, G, S, args.precision)
        line = {"metric": "mcts_simulations_per_sec", "value": value, "unit": "sims/s", "n_gpus": world,
                "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
                "config": cfg,
                "run": {"games_total": int(G_all), "waves": r["waves"], "games_per_wave": r["GW"], "streams": r["streams"],
                        "latent_arena_gb_per_gpu": r["arena_gb"], "trajectory_exchange": r["exchange"]},
                "moves_per_sec": moves_all / (ms * 1e-3), "mean_leaf_depth": r["depth"],
                "timed_region_s": ms * 1e-3,
                "e2e": {"value": e2e_sims_all / (e2e_ms * 1e-3), "unit": "sims/s", "h2d_bytes_per_step": r["h2d"],
                        "d2h_bytes_per_step": r["d2h"]},
                "gpu_launches": int(launches_all), "search_graph_replays": r["graph_replays"],
                "roofline": r["roofline"],
                "frac_step": {"value": value / world * flops_recurrent(n) / 1e12 / peaks["bf16_sustained"],
                              "what": "whole step as tensor work: sims/s per GPU x F_rec / measured SUSTAINED bf16 peak "
                                      "(the step is seconds long)", "peak": peaks["bf16_sustained"], "unit": "TFLOP/s"},
                "roofline_tree": r["roofline_tree"], "roofline_rules": r["roofline_rules"],
                "cpu_baseline": cpu, "clocks": r["clocks"], "weight_broadcast_bytes": r["bytes_bcast"]}
        if second:
            wl9, r9, r9o = second
            (ms9, e2e9), (s9, m9, es9, _, _) = red9
            (ms9o, e2e9o), (s9o, m9o, es9o, _, _) = red9o
            line["secondary"] = {"config": config_dict(wl9, wl9["games"], wl9["sims"], args.precision),
                                 "value": s9 / (ms9 * 1e-3), "unit": "sims/s", "moves_per_sec": m9 / (ms9 * 1e-3),
                                 "steps": max(K, 250), "ms_per_step": ms9 / max(K, 250), "mean_leaf_depth": r9["depth"],
                                 "e2e": {"value": es9 / (e2e9 * 1e-3), "unit": "sims/s", "h2d_bytes_per_step": r9["h2d"],
                                         "d2h_bytes_per_step": r9["d2h"]},
                                 "roofline": r9["roofline"],
                                 "frac_step": s9 / (ms9 * 1e-3) / world * flops_recurrent(9) / 1e12 / peaks["bf16_sustained"],
                                 "from_empty_boards": {"positions": "every game starts from the empty board and plays on through whole games "
                                                                    "(self-play's own mix of phases)",
                                              "value": s9o / (ms9o * 1e-3), "unit": "sims/s", "ms_per_step": ms9o / max(K, 250),
                                              "mean_leaf_depth": r9o["depth"],
                                              "e2e": {"value": es9o / (e2e9o * 1e-3), "unit": "sims/s"}}}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_train(args):
    """Extra line for SURVEY 8(f) rank 1 (not the headline metric): the batched trainer on the reference's own
    training configuration (main.py:27-37: 6x6 board, C=128, batch of 128 trajectories, K=10 unroll), next to the
    reference's procedure (one forward / backward / Adam step per trajectory at batch size 1) on the host cores."""
    import torch
    import muzero_go_b200 as mz
    from muzero_go_b200.trainer import build_batch
    n, C, B, K = 6, LATENT, 128, 10
    A = n * n + 1
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    net = mz.MuZeroNet(C, A, precision="bf16").to(dev)
    T = int(n * n * 1.5)
    eng = mz.SelfPlayEngine(net, n, 512, 8, search=mz.SearchConfig(seed=3), traj_capacity=T, restart_finished=True)
    eng.run(T)
    torch.cuda.synchronize()
    episodes, _ = mz.episodes_from_rows(eng.trajectory_rows(0, T), n)
    rb = mz.MultiVersionReplayBuffer(5000, 1, True)
    for ep in episodes:
        rb.add(ep)
    tr = mz.BatchedTrainer(net, A)
    import random as _random
    np_seed = 0
    import numpy as np
    np.random.seed(np_seed)
    for _ in range(max(args.warmup, 3)):
        tr.train(rb, B)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        tr.train(rb, B)                                   # host batch assembly + upload + bootstrap + step
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    batch, _ = rb.sample(B)
    starts = [_random.randint(0, len(t["actions"]) - 1) for t in batch]
    bt = build_batch(net, batch, starts, tr.cfg, A, dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        tr.train_step(bt)
    e0.record()
    for _ in range(args.steps):
        tr.train_step(bt)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    # the same loop with the buffer, the sampling and the batch assembly on the device
    from muzero_go_b200.device_replay import DeviceReplay, train_on_device
    dr = DeviceReplay(5000, n, device=dev)
    dr.add_rows(eng.trajectory_rows(0, T))
    for _ in range(3):
        train_on_device(tr, dr, B)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        train_on_device(tr, dr, B)
    torch.cuda.synchronize()
    e2e_dev_s = time.perf_counter() - t0
    flops = 3 * B * (flops_initial(n) + K * flops_recurrent(n))          # forward + backward (dgrad, wgrad)
    peaks = measured_peaks()
    cpu = None
    if not args.no_cpu_baseline:
        from oracle import nets_ref, train_ref
        torch.set_num_threads(os.cpu_count() or 1)
        ref = nets_ref.MuZeroNetRef(C, A)
        opt = torch.optim.Adam(ref.parameters(), lr=1e-4)
        done, t0 = 0, time.perf_counter()
        while time.perf_counter() - t0 < 15.0:
            traj = batch[done % B]
            opt.zero_grad()
            loss, _ = train_ref.trajectory_loss(ref, traj, starts[done % B], A)
            loss.backward()
            torch.nn.utils.clip_grad_norm_(ref.parameters(), 1.0)
            opt.step()
            done += 1
        cpu = {"value": done / (time.perf_counter() - t0), "unit": "trajectories/s", "cores": os.cpu_count(), "kind": "port",
               "sample": "15 s of the reference's procedure (main.py:395-480: forward, backward, clip, Adam per trajectory "
                         "at batch size 1) with the oracle networks, torch CPU, all cores"}
    emit({"metric": "training_trajectories_per_sec", "value": B / (ms * 1e-3), "unit": "trajectories/s", "n_gpus": 1,
          "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
          "vs_baseline": None, "dtype": "f32", "data": "self-play episodes of a random-init network (device engine)",
          "config": {"workload": "main.py Config: 6x6, batch of 128 trajectories, K=10 unroll, C=128; one optimizer step per batch",
                     "episodes_in_buffer": len(rb)},
          "e2e": {"value": B * args.steps / e2e_dev_s, "unit": "trajectories/s",
                  "note": "train_on_device: sampling, batch assembly (gathers + rules-kernel planes), bootstrap forward and step "
                          "on the GPU, loss read back every step",
                  "host_assembly": B * args.steps / e2e_s},
          "roofline": {"bound": "tensor", "kernel": "csrc/train.cu: the 3x3 convolutions forward + data gradient + weight gradient in hand-written "
                                                    "float32 CUDA-core kernels (the trainer's parity contract is float32, 1e-5), against the bf16 "
                                                    "tensor peak for scale only",
                       "achieved": flops / (ms * 1e-3) / 1e12, "peak": peaks["bf16_sustained"], "unit": "TFLOP/s",
                       "frac": flops / (ms * 1e-3) / 1e12 / peaks["bf16_sustained"], "traffic": None},
          "cpu_baseline": cpu})


_JSON_OUT = None


def claim_stdout():
    """The contract is ONE JSON line on stdout, and libraries write there too (NCCL prints its version banner on
    the first communicator): file descriptor 1 is pointed at stderr for the run and the line goes to the saved one."""
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(line):
    out = _JSON_OUT if _JSON_OUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="19x19", choices=sorted(WORKLOADS),
                    help="19x19 = BASELINE configs[3], the north-star configuration (default); 9x9 = configs[1]")
    ap.add_argument("--no-secondary", action="store_true", help="skip the 9x9 figures printed beside the 19x19 line")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--streams", type=int, default=1,
                    help="CUDA streams (sets of search buffers) the waves of games are spread over; 1 = one wave at a time")
    ap.add_argument("--waves", type=int, default=0,
                    help="split the games into at least this many waves that take turns on one set of search buffers "
                         "(default: as few as the free HBM allows)")
    ap.add_argument("--games", type=int, default=0)
    ap.add_argument("--sims", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--train", action="store_true", help="extra line: the batched trainer (SURVEY 8(f) rank 1)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    wl = WORKLOADS[args.workload]
    if args.train:
        run_train(args)
    elif args.impl == "reference":
        run_reference(args, wl)
    else:
        run_ours(args, wl)


if __name__ == "__main__":
    main()
