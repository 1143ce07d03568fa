"""CPU: host-side logic behind the C ABI — tile planner, pad sizes, error behaviour, symbol export, weights blob."""
import ctypes
import os
import re
import struct

import numpy as np
import pytest

from image_restoration_agent_b200 import _lib, configs, synth, weights
from oracle import tiling

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _plan(lib, n, tile, overlap):
    cnt = ctypes.c_int()
    buf = (ctypes.c_int * 4096)()
    _lib.check(lib.irb_plan_tiles(n, tile, overlap, buf, ctypes.byref(cnt)))
    return list(buf[:cnt.value])


def test_every_declared_symbol_is_exported(lib):
    header = open(os.path.join(ROOT, "include", "irb200.h")).read()
    declared = set(re.findall(r"\b(irb_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations found"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/irb200.h but not exported by libirb200.so"
    assert declared == set(_lib.PROTOTYPES), "ctypes prototypes and header disagree"


@pytest.mark.parametrize("n,tile,overlap", [
    (720, 256, 32), (1280, 256, 32), (2160, 512, 64), (3840, 512, 64), (520, 256, 32),  # BASELINE configs
    (256, 256, 32), (257, 256, 32), (480, 256, 32), (48, 16, 8), (64, 8, 0), (1000, 8, 7), (33, 32, 31)])
def test_plan_tiles_matches_reference_index_list(lib, n, tile, overlap):
    assert _plan(lib, n, tile, overlap) == tiling.tile_indices(n, tile, overlap)


def test_plan_tiles_randomised(lib):
    rng = np.random.RandomState(0)
    for _ in range(300):
        n = int(rng.randint(8, 5000))
        tile = int(rng.randint(1, n + 1))
        overlap = int(rng.randint(0, tile))
        idx = _plan(lib, n, tile, overlap)
        assert idx == tiling.tile_indices(n, tile, overlap)
        cover = np.zeros(n, dtype=int)
        for i in idx:
            cover[i:i + tile] += 1
        assert cover.min() >= 1                       # every pixel covered
        assert (cover == tiling.coverage_count(n, tile, overlap)).all()


def test_plan_tiles_rejects_bad_arguments(lib):
    cnt = ctypes.c_int()
    assert lib.irb_plan_tiles(100, 200, 32, None, ctypes.byref(cnt)) == 1     # tile > n: caller must clamp first
    assert lib.irb_plan_tiles(100, 32, 32, None, ctypes.byref(cnt)) == 1      # stride 0
    assert lib.irb_plan_tiles(0, 32, 8, None, ctypes.byref(cnt)) == 1
    assert b"irb_plan_tiles" in lib.irb_last_error()


@pytest.mark.parametrize("h,w", [(720, 1280), (45, 61), (256, 256), (7, 9), (2160, 3840)])
def test_padded_size_rules(lib, h, w):
    hp, wp = ctypes.c_int(), ctypes.c_int()
    _lib.check(lib.irb_padded_size(_lib.ARCH_RESTORMER, 0, h, w, ctypes.byref(hp), ctypes.byref(wp)))
    m = 8  # demo.py:122-126
    H, W = ((h + m) // m) * m, ((w + m) // m) * m
    assert (hp.value, wp.value) == (h + (H - h if h % m else 0), w + (W - w if w % m else 0))
    for ws in (7, 8):
        _lib.check(lib.irb_padded_size(_lib.ARCH_SWINIR, ws, h, w, ctypes.byref(hp), ctypes.byref(wp)))
        assert (hp.value, wp.value) == ((h // ws + 1) * ws, (w // ws + 1) * ws)  # inference.py:76-78


def test_compute_entry_fails_loudly_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    rc = lib.irb_layernorm(None, 0, 0, None, 0, 0, 48, None, None, 0, 1e-5, None)
    assert rc == 2 and b"no CPU fallback" in lib.irb_last_error()
    desc = _lib.ModelDesc()
    h = ctypes.c_void_p()
    rc = lib.irb_create(ctypes.byref(desc), b"IRBW", 4, 0, ctypes.byref(h))
    assert rc == 2 and b"no CPU fallback" in lib.irb_last_error()
    from image_restoration_agent_b200.engine import Engine
    cfg = configs.small_restormer_config()
    with pytest.raises(_lib.IrbError):
        Engine("restormer", cfg, synth.restormer_state_dict(cfg, 0))


def test_weights_blob_layout():
    cfg = configs.small_swinir_config("color_dn")
    sd = synth.swinir_state_dict(cfg, seed=3)
    sd["layers.0.residual_group.blocks.1.attn_mask"] = np.zeros((4, 64, 64), np.float32)  # buffers are dropped
    blob = weights.state_dict_to_blob(sd)
    magic, ver, n, _ = struct.unpack_from("<4sIII", blob, 0)
    assert magic == b"IRBW" and ver == 1 and n == len(sd) - 1
    ent = struct.Struct("<96sI4IQQ")
    seen = {}
    for i in range(n):
        name, ndim, d0, d1, d2, d3, off, numel = ent.unpack_from(blob, 16 + i * ent.size)
        name = name.rstrip(b"\0").decode()
        assert off % 16 == 0 and d0 * d1 * d2 * d3 == numel
        seen[name] = np.frombuffer(blob, dtype="<f4", count=numel, offset=off).reshape([d0, d1, d2, d3][:ndim])
    assert "layers.0.residual_group.blocks.1.attn_mask" not in seen
    for k, v in sd.items():
        if k.endswith("attn_mask"):
            continue
        assert np.array_equal(seen[k], v.numpy())


def test_state_dict_shapes_match_reference_counts():
    # Restormer state_dict has 494 entries (WithBias) ; SwinIR-M classical x4 has 550 incl. 108+18 buffers
    sd = synth.restormer_state_dict(configs.restormer_config("Motion_Deblurring"), 0)
    assert len(sd) == 494
    assert sum(v.numel() for v in sd.values()) == 26126644  # 26.13 M params
    sw = synth.swinir_state_dict(configs.swinir_config("classical_sr", scale=4), 0)
    assert sum(v.numel() for v in sw.values()) == 11900199  # 11.90 M params


def test_synthetic_image_is_deterministic():
    a = synth.synthetic_image_u8(37, 53, 3, seed=4)
    b = synth.synthetic_image_u8(37, 53, 3, seed=4)
    assert a.dtype == np.uint8 and a.shape == (37, 53, 3) and np.array_equal(a, b)
    assert a.min() == 0 and a.max() == 255
