"""One process per GPU: tiles (or whole images of a batch) are independent units, so they shard across ranks with
no data-path collective inside the forward pass.

restore_strips() is the north_star's path for one large image: every rank restores a contiguous row-major chunk of
the tile list, receives the seam rows of its neighbours' tiles, blends + crops + clamps + rounds the output rows it
owns locally (irb_stitch_rows) and sends its uint8 strip to rank 0. Per pixel the covering tiles are still summed
in the reference's (h_idx, w_idx) order from the raw fp32 tile values, so the N-rank image equals the 1-rank image
bit for bit. restore_sharded() is the older variant (all fp32 tiles gathered, rank 0 blends) kept for comparison.

The functions take the compute step as a callable so that the host logic is testable with the gloo backend on CPU.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Tuple

import torch
import torch.distributed as dist


def partition(n_units: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous, balanced ranges: the first n_units % world ranks get one extra unit."""
    base, extra = divmod(n_units, world)
    out, lo = [], 0
    for r in range(world):
        hi = lo + base + (1 if r < extra else 0)
        out.append((lo, hi))
        lo = hi
    return out


def gather_units(local: torch.Tensor, n_units: int, dst: int = 0, group=None) -> Optional[torch.Tensor]:
    """Gathers the per-rank slices (first dim = units, partitioned by partition()) to `dst`. Returns the
    concatenation [n_units, ...] on dst and None elsewhere. Ranks holding fewer units pad to the largest slice,
    so one collective (all ranks equal sizes) serves NCCL and gloo alike."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if world == 1:
        return local
    parts = partition(n_units, world)
    max_n = max(hi - lo for lo, hi in parts)
    unit_shape = tuple(local.shape[1:])
    buf = local
    if local.shape[0] != max_n:
        buf = torch.zeros((max_n,) + unit_shape, dtype=local.dtype, device=local.device)
        buf[:local.shape[0]] = local
    buf = buf.contiguous()
    if rank == dst:
        recv = [torch.empty_like(buf) for _ in range(world)]
        dist.gather(buf, recv, dst=dst, group=group)
        return torch.cat([recv[r][:hi - lo] for r, (lo, hi) in enumerate(parts)], dim=0)
    dist.gather(buf, None, dst=dst, group=group)
    return None


def sharded_forward(forward_fn: Callable[[torch.Tensor], torch.Tensor], units: torch.Tensor, dst: int = 0,
                    group=None) -> Optional[torch.Tensor]:
    """Every rank holds all `units` (tiles [nt, th, tw, c] or images); each processes its contiguous range with
    forward_fn and the results are gathered on dst."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    n = units.shape[0]
    lo, hi = partition(n, world)[rank]
    local = forward_fn(units[lo:hi])
    if world == 1:
        return local
    return gather_units(local, n, dst=dst, group=group)


def restore_sharded(engine, img: torch.Tensor, tile=None, tile_overlap: int = 32, bgr: bool = False, dst: int = 0,
                    group=None, want_f32: bool = False):
    """Multi-rank version of Engine.restore_device: every rank pads + cuts the tiles (cheap), restores its own tile
    range, rank `dst` gathers the fp32 tiles over NCCL/NVLink and runs the fused blend/crop/clamp/round kernel.
    Returns (u8, f32) on dst, (None, None) elsewhere."""
    if img.dtype == torch.uint8:
        h, w, _ = img.shape
    else:
        _, h, w = img.shape
    plan = engine.plan(h, w, tile, tile_overlap)
    tiles = engine.prepare_tiles(img, plan, tile, tile_overlap, bgr)
    out_tiles = sharded_forward(engine.forward_cl, tiles, dst=dst, group=group)
    if out_tiles is None:
        return None, None
    return engine.stitch(out_tiles, plan, bgr=bgr, want_u8=True, want_f32=want_f32)


def restore_batch_sharded(engine, imgs: List[torch.Tensor], tile=None, tile_overlap: int = 32, bgr: bool = False,
                          dst: int = 0, group=None):
    """Batch of same-sized images: contiguous image ranges per rank, uint8 results gathered on dst."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    lo, hi = partition(len(imgs), world)[rank]
    outs = [engine.restore_device(imgs[i], tile, tile_overlap, bgr)[0] for i in range(lo, hi)]
    s, c = engine.scale, engine.channels
    h, w = (imgs[0].shape[0], imgs[0].shape[1]) if imgs[0].dtype == torch.uint8 else imgs[0].shape[1:]
    local = torch.stack(outs) if outs else torch.empty((0, h * s, w * s, c), dtype=torch.uint8, device=imgs[0].device)
    if world == 1:
        return local
    return gather_units(local, len(imgs), dst=dst, group=group)


# ------------------------------------------------------------------------------------------------------------
# Strip path (north_star): local blend of owned rows, seam exchange, uint8 strip gather
# ------------------------------------------------------------------------------------------------------------
def strip_rows(crop_h: int, world: int) -> List[Tuple[int, int]]:
    """Output rows [r0, r1) owned by each rank: contiguous and balanced."""
    return partition(crop_h, world)


def seam_messages(plan: dict, world: int) -> List[Tuple[int, int, int, int, int]]:
    """(src_rank, dst_rank, tile, a, b): rows [a, b) of tile `tile`'s OUTPUT (tile-local, already scaled) that the
    owner `dst_rank` of some output rows needs from the rank that computed the tile. The list is in one global order
    (dst, tile), which both ends of every pair walk identically."""
    s, th = plan["scale"], plan["th"]
    ys, nx = plan["ys"], len(plan["xs"])
    nt = len(ys) * nx
    crop_h = plan["h"] * s
    owner_of_tile = [r for r, (lo, hi) in enumerate(partition(nt, world)) for _ in range(lo, hi)]
    out = []
    for q, (r0, r1) in enumerate(strip_rows(crop_h, world)):
        if r1 <= r0:
            continue
        for t in range(nt):
            top = ys[t // nx] * s
            a, b = max(r0, top), min(r1, top + th * s, crop_h)
            if b > a and owner_of_tile[t] != q:
                out.append((owner_of_tile[t], q, t, a - top, b - top))
    return out


def restore_strips(engine, img: torch.Tensor, tile=None, tile_overlap: int = 32, bgr: bool = False, dst: int = 0,
                   group=None, solo: bool = False):
    """One image over all ranks. `engine` needs plan / prepare_tiles / forward_cl(tiles, out=) / stitch_rows
    (Engine, or a CPU stand-in in the gloo tests). Returns the uint8 image [H*s, W*s, c] on dst, None elsewhere,
    plus a dict of the bytes that crossed the wire. solo=True runs the same code as a single rank even inside an
    initialised process group (bench.py's N = 1 leg of the strong-scaling pair)."""
    world = dist.get_world_size(group) if dist.is_initialized() and not solo else 1
    rank = dist.get_rank(group) if dist.is_initialized() and not solo else 0
    if img.dtype == torch.uint8:
        h, w, _ = img.shape
    else:
        _, h, w = img.shape
    plan = engine.plan(h, w, tile, tile_overlap)
    s, c = plan["scale"], plan["c"]
    nt = len(plan["ys"]) * len(plan["xs"])
    tiles = engine.prepare_tiles(img, plan, tile, tile_overlap, bgr)   # every rank cuts all tiles (a few MB, cheap)
    lo, hi = partition(nt, world)[rank]
    # one slot per GLOBAL tile number; only this rank's tiles and its neighbours' seam rows are ever filled
    slots = torch.empty((nt, plan["th"] * s, plan["tw"] * s, c), dtype=torch.float32, device=tiles.device)
    engine.forward_cl(tiles[lo:hi], out=slots[lo:hi])
    stats = {"seam_bytes": 0, "gather_bytes": 0}
    if world > 1:
        p2p = []
        for src, q, t, a, b in seam_messages(plan, world):
            if src == rank:
                p2p.append(dist.P2POp(dist.isend, slots[t, a:b], q, group))
            elif q == rank:
                p2p.append(dist.P2POp(dist.irecv, slots[t, a:b], src, group))
                stats["seam_bytes"] += slots[t, a:b].numel() * 4
        if p2p:
            for req in dist.batch_isend_irecv(p2p):
                req.wait()
    rows = strip_rows(h * s, world)
    r0, r1 = rows[rank]
    if rank == dst:
        out = torch.empty((h * s, w * s, c), dtype=torch.uint8, device=tiles.device)
        engine.stitch_rows(slots, plan, r0, r1 - r0, bgr=bgr, out_u8=out[r0:r1])
        if world > 1:
            p2p = [dist.P2POp(dist.irecv, out[a:b], q, group) for q, (a, b) in enumerate(rows) if q != dst and b > a]
            stats["gather_bytes"] = sum(out[a:b].numel() for q, (a, b) in enumerate(rows) if q != dst)
            if p2p:
                for req in dist.batch_isend_irecv(p2p):
                    req.wait()
        return out, stats
    strip, _ = engine.stitch_rows(slots, plan, r0, r1 - r0, bgr=bgr)
    if r1 > r0:
        for req in dist.batch_isend_irecv([dist.P2POp(dist.isend, strip, dst, group)]):
            req.wait()
    return None, stats


class AsyncGather:
    """Weak-scaling gather of per-rank uint8 results to rank 0 that overlaps the NEXT step's compute: the collective
    is launched with async_op=True (NCCL's own stream) and only waited for when its buffers are needed again, two
    steps later. flush() waits for everything (call it before the closing timestamp)."""

    def __init__(self, like: torch.Tensor, world: int, rank: int, dst: int = 0, depth: int = 2, group=None):
        self.world, self.rank, self.dst, self.group, self.depth = world, rank, dst, group, depth
        self.send = [torch.empty_like(like) for _ in range(depth)]
        self.recv = [[torch.empty_like(like) for _ in range(world)] if rank == dst else None for _ in range(depth)]
        self.work = [None] * depth
        self.k = 0

    def submit(self, u8: torch.Tensor):
        i = self.k % self.depth
        if self.work[i] is not None:
            self.work[i].wait()
        self.send[i].copy_(u8, non_blocking=True)
        self.work[i] = dist.gather(self.send[i], self.recv[i], dst=self.dst, group=self.group, async_op=True)
        self.k += 1
        return i

    def flush(self):
        for i, wk in enumerate(self.work):
            if wk is not None:
                wk.wait()
                self.work[i] = None

    def last(self):
        """The most recently gathered list of per-rank results on dst (after flush())."""
        return self.recv[(self.k - 1) % self.depth]
