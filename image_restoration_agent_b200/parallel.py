"""One process per GPU: tiles (or whole images of a batch) are independent units, so they shard across ranks with
no data-path collective; the only exchange is the gather of finished fp32 tiles (or uint8 images) to rank 0
(SURVEY.md §8e). Rank 0 then blends all tiles in the reference's accumulation order, so the N-rank result is
bit-identical to the 1-rank result.

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
