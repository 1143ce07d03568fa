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
