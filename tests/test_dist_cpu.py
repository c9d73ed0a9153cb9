"""CPU, world_size 2 over gloo: the two exchanges of the multi-GPU self-play path (weight broadcast from the
learner rank, trajectory gather to the replay side) and the game sharding."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from muzero_go_b200 import dist as mzdist


def test_shard_range_partitions_all_games():
    for total in (0, 1, 7, 2048, 2049):
        for world in (1, 2, 3, 8):
            ranges = [mzdist.shard_range(total, r, world) for r in range(world)]
            assert ranges[0][0] == 0 and ranges[-1][1] == total
            assert all(ranges[i][1] == ranges[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in ranges]
            assert max(sizes) - min(sizes) <= 1


def fake_rows(T, G, A, seed):
    g = torch.Generator().manual_seed(seed)
    return {"boards": torch.randint(0, 2 ** 31 - 1, (T, G, 64), generator=g, dtype=torch.int32),
            "action": torch.randint(0, A, (T, G), generator=g, dtype=torch.int32),
            "policy": torch.rand((T, G, A), generator=g),
            "root_value": torch.rand((T, G), generator=g),
            "reward": torch.rand((T, G), generator=g),
            "done": torch.randint(0, 2, (T, G), generator=g, dtype=torch.uint8),
            "winner": torch.randint(-1, 2, (T, G), generator=g, dtype=torch.int8)}


def test_pack_unpack_round_trip():
    rows = fake_rows(3, 5, 82, 0)
    buf = mzdist.pack_rows(rows)
    assert buf.shape == (3, 5, 256 + 4 + 4 * 82 + 4 + 4 + 1 + 1)
    back = mzdist.unpack_rows(buf, 82)
    for k in rows:
        assert torch.equal(rows[k], back[k]), k


def test_row_ring_equals_pack_rows():
    A, depth, Gw = 82, 4, 3
    waves = [fake_rows(depth, Gw, A, seed=s) for s in (1, 2)]
    ring = mzdist.RowRing(2 * Gw, A, depth, "cpu")
    for t in range(depth):
        ring.write(t, [{k: v[t:t + 1] for k, v in w.items()} for w in waves])
    want = mzdist.pack_rows({k: torch.cat([w[k] for w in waves], dim=1) for k in waves[0]})
    assert ring.buf.shape[2] == mzdist.row_stride(A) and torch.equal(ring.buf, want)
    back = ring.rows()
    assert torch.equal(back["policy"][:, Gw:], waves[1]["policy"])


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import muzero_go_b200 as mz
        torch.manual_seed(100 + rank)                      # ranks start with DIFFERENT weights
        net = mz.MuZeroNet(16, 26)
        before = mzdist.flatten_params(net).clone()
        nbytes = mzdist.broadcast_weights(net, src=0)
        after = mzdist.flatten_params(net)
        torch.manual_seed(100)
        want = mzdist.flatten_params(mz.MuZeroNet(16, 26))
        ok_bcast = torch.equal(after, want) and nbytes == want.numel() * 4 and (rank == 0 or not torch.equal(before, after))
        rows = fake_rows(2, 3, 26, seed=rank)
        gathered, _ = mzdist.gather_trajectories(rows, 26)
        ok_gather = True
        for r in range(world):
            ref = fake_rows(2, 3, 26, seed=r)
            for k in ref:
                ok_gather &= torch.equal(gathered[k][:, 3 * r:3 * r + 3], ref[k])
        # the ring: `depth` moves of two waves packed with slice copies, gathered in one collective
        depth, Gw = 3, 2
        ring = mzdist.RowRing(2 * Gw, 26, depth, "cpu")
        for t in range(depth):
            ring.write(t, [{k: v[t:t + 1] for k, v in fake_rows(depth, Gw, 26, seed=10 * rank + w).items()} for w in range(2)])
        merged, nbytes_ring = mzdist.gather_ring(ring)
        ok_ring = nbytes_ring == world * depth * 2 * Gw * mzdist.row_stride(26)
        for r in range(world):
            for w in range(2):
                ref = fake_rows(depth, Gw, 26, seed=10 * r + w)
                lo = r * 2 * Gw + w * Gw
                for k in ref:
                    ok_ring &= torch.equal(merged[k][:, lo:lo + Gw], ref[k])
        out[rank] = bool(ok_bcast and ok_gather and ok_ring)
    finally:
        dist.destroy_process_group()


def test_broadcast_and_gather_world2_gloo():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    assert dict(out) == {0: True, 1: True}
