"""Multi-GPU self-play: one process per GPU, games sharded independently, no collective inside the search.

The reference is single-process (/root/reference/main.py:613); BASELINE.json's north_star adds two
exchanges outside the search, both implemented here over torch.distributed (NCCL on GPUs, gloo in the
CPU tests):
  * broadcast_weights: the learner rank's parameters to every self-play rank (one flat buffer);
  * gather_trajectories: fixed-stride trajectory rows of every rank to the learner rank's replay side.
"""
import torch
import torch.distributed as dist


def shard_range(total_games, rank, world_size):
    """Rank r owns games [lo, hi): contiguous, sizes differ by at most one."""
    base, rem = divmod(int(total_games), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def flatten_params(module):
    params = [p for p in module.parameters()]
    return torch.cat([p.detach().reshape(-1) for p in params]) if params else torch.empty(0)


@torch.no_grad()
def broadcast_weights(module, src=0, group=None):
    """One collective for the whole network (308,663 parameters at 19x19, C=128: 1.2 MB)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return 0
    flat = flatten_params(module)
    dist.broadcast(flat, src=src, group=group)
    off = 0
    for p in module.parameters():
        k = p.numel()
        p.copy_(flat[off:off + k].view_as(p))      # bumps p._version -> kernel-side layouts are re-derived
        off += k
    return flat.numel() * flat.element_size()


ROW_FIELDS = ("boards", "action", "policy", "root_value", "reward", "done", "winner")


def row_field_bytes(action_size):
    """Bytes of every field of one (move, game) trajectory record, in ROW_FIELDS order."""
    return [256, 4, 4 * int(action_size), 4, 4, 1, 1]


def row_stride(action_size):
    return sum(row_field_bytes(action_size))


class RowRing:
    """`depth` trajectory rows of G games as ONE preallocated uint8 [depth, G, stride] buffer in the wire layout of
    pack_rows / unpack_rows.  write() fills a row with slice copies straight from the engines' trajectory buffers
    (no torch.cat, no allocation per step); the buffer is what gets all-gathered, `depth` moves at a time
    ("batched per K moves so messages are >= 1 MB")."""

    def __init__(self, num_games, action_size, depth, device):
        self.G, self.A, self.depth = int(num_games), int(action_size), int(depth)
        self.sizes = row_field_bytes(action_size)
        self.buf = torch.zeros((self.depth, self.G, sum(self.sizes)), dtype=torch.uint8, device=device)

    def write(self, t, rows_per_engine):
        """rows_per_engine: one dict of [1, G_w, ...] tensors per engine (SelfPlayEngine.trajectory_rows(r, r + 1)),
        engines in game order."""
        g0 = 0
        for rows in rows_per_engine:
            gw = rows["action"].shape[1]
            off = 0
            for key, nbytes in zip(ROW_FIELDS, self.sizes):
                src = rows[key][0]
                self.buf[t, g0:g0 + gw, off:off + nbytes].copy_(src.contiguous().view(torch.uint8).reshape(gw, nbytes))
                off += nbytes
            g0 += gw
        assert g0 == self.G

    def write_from(self, t, group, row):
        """Row `row` of a SelfPlayEngine / WaveGroup into ring slot t with the engine's own packing kernel
        (mzgo_selfplay_pack_rows): one launch per wave instead of seven slice copies."""
        group.pack_rows(row, row + 1, self.buf[t:t + 1])

    def rows(self, lo=0, hi=None):
        """The ring's rows [lo, hi) as a dict of tensors (unpack_rows)."""
        return unpack_rows(self.buf[lo:self.depth if hi is None else hi], self.A)


def all_gather_rows(out, buf, group=None):
    """out [world, ...] <- every rank's buf [...] in one collective (the per-rank slices of `out` are contiguous views
    of one tensor, which NCCL gathers in place; works on gloo too)."""
    dist.all_gather(list(out.unbind(0)), buf, group=group)


def gather_ring(ring, group=None):
    """all_gather of a whole RowRing: -> dict of [depth, sum G_r, ...] tensors on every rank."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return ring.rows(), ring.buf.numel()
    world = dist.get_world_size(group)
    out = torch.empty((world,) + tuple(ring.buf.shape), dtype=torch.uint8, device=ring.buf.device)
    all_gather_rows(out, ring.buf, group)
    merged = out.permute(1, 0, 2, 3).reshape(ring.depth, world * ring.G, -1)
    return unpack_rows(merged.contiguous(), ring.A), out.numel()


def pack_rows(rows):
    """Trajectory rows (dict of [T,G,...] tensors, SelfPlayEngine.trajectory_rows) -> one uint8 [T,G,stride]."""
    T, G = rows["action"].shape[:2]
    parts = []
    for key in ("boards", "action", "policy", "root_value", "reward", "done", "winner"):
        t = rows[key].contiguous()
        parts.append(t.view(torch.uint8).reshape(T, G, -1))
    return torch.cat(parts, dim=2).contiguous()


def unpack_rows(buf, action_size):
    """Inverse of pack_rows."""
    T, G, _ = buf.shape
    sizes = [("boards", 256, torch.int32, (64,)), ("action", 4, torch.int32, ()), ("policy", 4 * action_size, torch.float32, (action_size,)),
             ("root_value", 4, torch.float32, ()), ("reward", 4, torch.float32, ()), ("done", 1, torch.uint8, ()),
             ("winner", 1, torch.int8, ())]
    out, off = {}, 0
    for key, nbytes, dtype, shape in sizes:
        chunk = buf[:, :, off:off + nbytes].contiguous()
        out[key] = chunk.view(dtype).reshape((T, G) + shape)
        off += nbytes
    assert off == buf.shape[2]
    return out


def gather_trajectories(rows, action_size, dst=0, group=None):
    """All ranks call; returns the concatenated rows ([T, sum G_r, ...]) on every rank (all_gather keeps
    the collective symmetric; the learner is whoever consumes it).  Ranks must hold equal G and T."""
    buf = pack_rows(rows)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return unpack_rows(buf, action_size), buf.numel()
    world = dist.get_world_size(group)
    out = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(out, buf, group=group)
    return unpack_rows(torch.cat(out, dim=1), action_size), buf.numel() * world
