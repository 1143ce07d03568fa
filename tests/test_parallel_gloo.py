"""CPU, world_size 2 (gloo): the multi-rank host logic — tile partition, gather to rank 0 and blend — gives exactly
the single-process result. The compute step is the oracle model so the test needs no GPU."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_units, out_path):
    sys.path.insert(0, ROOT)
    from image_restoration_agent_b200 import parallel
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    units = torch.rand(n_units, 6, 5, 3)
    fwd = lambda t: t * 2.0 + t.flip(1)  # any per-unit function
    got = parallel.sharded_forward(fwd, units, dst=0)
    if rank == 0:
        torch.save({"got": got, "want": fwd(units)}, out_path)
    else:
        assert got is None
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n_units", [7, 2, 1])
def test_sharded_forward_equals_single_process(tmp_path, n_units):
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), n_units, out), nprocs=2, join=True)
    r = torch.load(out)
    assert torch.equal(r["got"], r["want"])


def test_partition_properties():
    from image_restoration_agent_b200.parallel import partition
    for n in (0, 1, 5, 24, 45, 576):
        for w in (1, 2, 4, 8):
            p = partition(n, w)
            assert len(p) == w and p[0][0] == 0 and p[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(p, p[1:]))
            sizes = [hi - lo for lo, hi in p]
            assert max(sizes) - min(sizes) <= 1
    assert partition(45, 8)[0] == (0, 6)  # 45 tiles / 8 ranks -> max 6 per rank


def _tiled_worker(rank, world, port, out_path):
    """Whole tool body sharded over 2 ranks with the ORACLE as the per-tile model; rank 0 blends."""
    sys.path.insert(0, ROOT)
    from image_restoration_agent_b200 import configs, parallel, synth
    from oracle import restormer_ref, tiling
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    cfg = configs.small_restormer_config()
    model = restormer_ref.Model(synth.restormer_state_dict(cfg, 1), cfg)
    img = synth.synthetic_image_u8(40, 56, 3, seed=2)
    x = torch.from_numpy(img).float().div(255.0).permute(2, 0, 1).unsqueeze(0)
    tile, ov = 24, 8
    ys, xs = tiling.tile_indices(40, tile, ov), tiling.tile_indices(56, tile, ov)
    tiles = torch.stack([x[0, :, y:y + tile, xx:xx + tile] for y in ys for xx in xs])
    outs = parallel.sharded_forward(lambda t: model(t) if t.shape[0] else t, tiles, dst=0)
    if rank == 0:
        E = torch.zeros(3, 40, 56)
        W = torch.zeros_like(E)
        k = 0
        for y in ys:
            for xx in xs:
                E[:, y:y + tile, xx:xx + tile] += outs[k]
                W[:, y:y + tile, xx:xx + tile] += 1
                k += 1
        want = tiling.restormer_restore_image(img, model, tile, ov, return_float=True)
        torch.save({"got": E / W, "want": want}, out_path)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_tiled_restore_matches_reference_loop(tmp_path):
    out = str(tmp_path / "r0.pt")
    mp.spawn(_tiled_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    r = torch.load(out)
    assert (r["got"] - r["want"]).abs().max() <= 1e-6


# ------------------------------------------------------------------------------------------------------------
# strip path: local blend of owned rows + seam exchange + uint8 strip gather (parallel.restore_strips)
# ------------------------------------------------------------------------------------------------------------
class _CpuEngine:
    """CPU stand-in with the Engine methods restore_strips() uses. The 'model' is a fixed per-tile function with an
    upscale factor, the blend is the reference's E/W accumulation restricted to a row range."""

    def __init__(self, scale=1, c=3):
        self.scale, self.channels = scale, c

    def plan(self, h, w, tile, overlap):
        from oracle import tiling
        t = min(tile, h, w)
        return dict(hp=h, wp=w, th=t, tw=t, ys=tiling.tile_indices(h, t, overlap), xs=tiling.tile_indices(w, t, overlap),
                    scale=self.scale, c=self.channels, h=h, w=w)

    def prepare_tiles(self, img, plan, tile, overlap, bgr=False):
        x = img.float() / 255.0
        t = plan["th"]
        return torch.stack([x[y:y + t, xx:xx + t] for y in plan["ys"] for xx in plan["xs"]]).contiguous()

    def forward_cl(self, tiles, out=None):
        s = self.scale
        y = (tiles * 0.75 + tiles.flip(1) * 0.5 - 0.1 * tiles.mean(dim=(1, 2, 3), keepdim=True))
        y = y.repeat_interleave(s, 1).repeat_interleave(s, 2)
        if out is None:
            return y
        out.copy_(y)
        return out

    def stitch_rows(self, slots, plan, row0, rows, bgr=False, want_f32=False, out_u8=None):
        s, t = plan["scale"], plan["th"] * plan["scale"]
        H, W, c = plan["h"] * s, plan["w"] * s, plan["c"]
        E = torch.zeros(rows, W, c)
        cnt = torch.zeros(rows, W, 1)
        k = 0
        for y in plan["ys"]:
            for xx in plan["xs"]:
                a, b = max(row0, y * s), min(row0 + rows, y * s + t)
                if b > a:
                    E[a - row0:b - row0, xx * s:xx * s + t] += slots[k, a - y * s:b - y * s]
                    cnt[a - row0:b - row0, xx * s:xx * s + t] += 1
                k += 1
        f = E / cnt
        u8 = torch.round(f.clamp(0, 1) * 255.0).to(torch.uint8)
        if bgr:
            u8 = u8.flip(2)
        if out_u8 is not None:
            out_u8.copy_(u8)
            u8 = out_u8
        return u8, (f if want_f32 else None)


def _strip_worker(rank, world, port, h, w, tile, ov, scale, out_path):
    sys.path.insert(0, ROOT)
    from image_restoration_agent_b200 import parallel, synth
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    eng = _CpuEngine(scale=scale)
    img = torch.from_numpy(synth.synthetic_image_u8(h, w, 3, seed=4))
    got, stats = parallel.restore_strips(eng, img, tile, ov, bgr=True)
    if rank == 0:
        # single-process result: every tile computed here, the whole image blended at once
        plan = eng.plan(h, w, tile, ov)
        slots = eng.forward_cl(eng.prepare_tiles(img, plan, tile, ov))
        want, _ = eng.stitch_rows(slots, plan, 0, h * scale, bgr=True)
        torch.save({"got": got, "want": want, "stats": stats}, out_path)
    else:
        assert got is None
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,h,w,tile,ov,scale", [(2, 40, 56, 24, 8, 1), (2, 33, 47, 16, 4, 2), (3, 64, 40, 24, 8, 1),
                                                     (2, 24, 24, 24, 8, 1)])
def test_strip_path_bit_identical_to_single_process(tmp_path, world, h, w, tile, ov, scale):
    out = str(tmp_path / "r0.pt")
    mp.spawn(_strip_worker, args=(world, _free_port(), h, w, tile, ov, scale, out), nprocs=world, join=True)
    r = torch.load(out)
    assert r["got"].shape == (h * scale, w * scale, 3) and torch.equal(r["got"], r["want"])
    if h > tile:
        assert r["stats"]["gather_bytes"] > 0


def test_seam_messages_cover_exactly_the_foreign_rows():
    """Every (output row, covering tile) pair is either local to the row's owner or delivered by exactly one message."""
    from image_restoration_agent_b200.parallel import partition, seam_messages, strip_rows
    plan = dict(scale=1, th=512, tw=512, ys=[0, 448, 896, 1344, 1648],
                xs=[0, 448, 896, 1344, 1792, 2240, 2688, 3136, 3328], h=2160, w=3840, c=3)   # cfg5
    for world in (2, 4, 8):
        nt = 45
        owner = [r for r, (lo, hi) in enumerate(partition(nt, world)) for _ in range(lo, hi)]
        have = {}
        for src, q, t, a, b in seam_messages(plan, world):
            assert owner[t] == src and src != q and 0 <= a < b <= 512
            for row in (a, b - 1):
                assert (q, t, row) not in have
                have[(q, t, row)] = 1
        for q, (r0, r1) in enumerate(strip_rows(2160, world)):
            for t in range(nt):
                top = plan["ys"][t // 9]
                a, b = max(r0, top), min(r1, top + 512)
                if b > a and owner[t] != q:
                    assert (q, t, a - top) in have and (q, t, b - 1 - top) in have
