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
