"""CPU: the oracle restatements reproduce the reference outputs stored in tests/golden/ (generated from the
unmodified reference modules by tools/make_golden.py). This is what pins the oracle."""
import glob
import os

import numpy as np
import pytest
import torch

from image_restoration_agent_b200 import configs, synth
from oracle import restormer_ref, swinir_ref, tiling

from cases import R_CASES, S_CASES, tool_case  # noqa: E402  (tests/cases.py)


@pytest.mark.parametrize("name", sorted(R_CASES))
def test_restormer_oracle_matches_reference(name, golden_dir):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    cfg = R_CASES[name]()
    sd = synth.restormer_state_dict(cfg, seed=int(g["seed"]))
    y = restormer_ref.Model(sd, cfg)(torch.from_numpy(g["x"])).numpy()
    assert np.abs(y - g["y"]).max() <= 2e-5


@pytest.mark.parametrize("name", sorted(S_CASES))
def test_swinir_oracle_matches_reference(name, golden_dir):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    cfg = S_CASES[name]()
    sd = synth.swinir_state_dict(cfg, seed=int(g["seed"]))
    y = swinir_ref.Model(sd, cfg)(torch.from_numpy(g["x"])).numpy()
    assert np.abs(y - g["y"]).max() <= 5e-5 * max(1.0, cfg["img_range"] / 50)


def _tool_files(golden_dir):
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(golden_dir, "*_tool_*.npz")))


@pytest.mark.parametrize("name", _tool_files(os.path.join(os.path.dirname(__file__), "golden")))
def test_tool_oracle_matches_reference(name, golden_dir):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    arch, cfg, sd, tile, overlap, scale = tool_case(name, g)
    if arch == "restormer":
        u8 = tiling.restormer_restore_image(g["img"], restormer_ref.Model(sd, cfg), tile, overlap)
        f = tiling.restormer_restore_image(g["img"], restormer_ref.Model(sd, cfg), tile, overlap, return_float=True)
    else:
        m = swinir_ref.Model(sd, cfg)
        u8 = tiling.swinir_process_image(g["img"], m, scale, cfg["window_size"], tile, overlap)
        f = tiling.swinir_process_image(g["img"], m, scale, cfg["window_size"], tile, overlap, return_float=True)
    assert np.abs(f.numpy() - g["out_f32"]).max() <= 2e-5
    assert np.abs(u8.astype(int) - g["out_u8"].astype(int)).max() <= 1
    assert u8.shape == g["out_u8"].shape and u8.dtype == np.uint8


from cases import X_CASES  # noqa: E402
from oracle import xrestormer_ref  # noqa: E402


@pytest.mark.parametrize("name", sorted(X_CASES))
def test_xrestormer_oracle_matches_reference(name, golden_dir):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    cfg = X_CASES[name]()
    sd = synth.xrestormer_state_dict(cfg, seed=int(g["seed"]))
    y = xrestormer_ref.Model(sd, cfg)(torch.from_numpy(g["x"])).numpy()
    assert np.abs(y - g["y"]).max() <= 2e-5


def test_xrestormer_tool_oracle_matches_reference(golden_dir):
    g = np.load(os.path.join(golden_dir, "xrestormer_pipeline_small.npz"))
    cfg = configs.small_xrestormer_config("WithBias")
    m = xrestormer_ref.Model(synth.xrestormer_state_dict(cfg, seed=int(g["seed"])), cfg)
    assert np.abs(tiling.xrestormer_restore_image(g["img"], m, return_float=True).numpy() - g["out_f32"]).max() <= 2e-5
    assert np.abs(tiling.xrestormer_restore_image(g["img"], m).astype(int) - g["out_u8"].astype(int)).max() <= 1


from cases import H_CASES  # noqa: E402
from oracle import hat_ref  # noqa: E402


@pytest.mark.parametrize("name", sorted(H_CASES))
def test_hat_oracle_matches_reference(name, golden_dir):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    cfg = H_CASES[name]()
    y = hat_ref.Model(synth.hat_state_dict(cfg, seed=int(g["seed"])), cfg)(torch.from_numpy(g["x"])).numpy()
    assert np.abs(y - g["y"]).max() <= 2e-5


@pytest.mark.parametrize("name", ["hat_pipeline_small_t32_p16", "hat_pipeline_small_whole"])
def test_hat_tool_oracle_matches_reference(name, golden_dir):
    """HATModel.pre_process / tile_process / post_process (ragged last tile row and column, clipped tile_pad)."""
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    cfg = configs.small_hat_config()
    m = hat_ref.Model(synth.hat_state_dict(cfg, seed=int(g["seed"])), cfg)
    ts = None if int(g["tile"]) < 0 else int(g["tile"])
    f = tiling.hat_restore_image(g["img"], m, 4, 16, ts, int(g["overlap"]), return_float=True).numpy()
    assert f.shape == g["out_f32"].shape and np.abs(f - g["out_f32"]).max() <= 2e-5
    u8 = tiling.hat_restore_image(g["img"], m, 4, 16, ts, int(g["overlap"]))
    assert np.abs(u8.astype(int) - g["out_u8"].astype(int)).max() <= 1


def test_hat_oca_index_goes_negative_like_the_reference():
    """calculate_rpi_oca (hat_arch.py:896-919) shifts by ws - ows + 1, so part of the index is negative and addresses the
    table from its end; the oracle (and the kernel) must keep that, not 'fix' it."""
    idx = hat_ref.rpi_oca(16, 24)
    assert idx.shape == (256, 576) and int(idx.min()) == -880 and int(idx.max()) == 640


def test_hat_tile_plan_covers_the_image_once():
    plan = tiling.hat_tile_plan(80, 112, 32, 16)
    cover = np.zeros((80, 112), int)
    for (y0, y1, x0, x1, y0p, y1p, x0p, x1p) in plan:
        cover[y0:y1, x0:x1] += 1
        assert y0p <= y0 < y1 <= y1p and x0p <= x0 < x1 <= x1p and y0p >= 0 and y1p <= 80 and x1p <= 112
    assert (cover == 1).all() and len(plan) == 3 * 4
