"""GPU (-m gpu): the drop-in surface — Tool(input_dir, output_dir) -> output.png, process_image, restore_image —
checked against the oracle's restatement of the reference tool scripts on PNG files."""
import os

import numpy as np
import pytest
import torch

from image_restoration_agent_b200 import configs, synth

pytestmark = pytest.mark.gpu


def _dirs(tmp_path, img_bgr):
    import cv2
    i, o = tmp_path / "in", tmp_path / "out"
    i.mkdir()
    o.mkdir()
    cv2.imwrite(str(i / "input.png"), img_bgr)
    return i, o


def test_restormer_tool_call(tmp_path):
    import cv2
    from image_restoration_agent_b200.tool_api import RestormerB200, clear_engines
    from oracle import restormer_ref, tiling
    cfg = configs.small_restormer_config()
    sd = synth.restormer_state_dict(cfg, 1)
    img_bgr = synth.synthetic_image_u8(50, 70, 3, seed=9)
    i, o = _dirs(tmp_path, img_bgr)
    tool = RestormerB200("motion_deblurring", state_dict=sd, cfg=cfg, precision="fp32")
    assert tool.tool_name == "restormer"
    tool(i, o, silent=True)
    assert os.listdir(o) == ["output.png"]
    got = cv2.imread(str(o / "output.png"), cv2.IMREAD_COLOR)
    rgb = cv2.cvtColor(img_bgr, cv2.COLOR_BGR2RGB)                      # demo.py:40-41
    want = tiling.restormer_restore_image(rgb, restormer_ref.Model(sd, cfg))
    want = cv2.cvtColor(want, cv2.COLOR_RGB2BGR)                        # demo.py:43-44
    assert got.shape == want.shape and np.abs(got.astype(int) - want.astype(int)).max() <= 1
    with pytest.raises(AssertionError):  # tool.py:51-54: output dir must be empty
        tool(i, o, silent=True)
    clear_engines()


def test_swinir_tool_call_and_process_image(tmp_path):
    import cv2
    from image_restoration_agent_b200.tool_api import SwinIRB200, clear_engines, process_image
    from oracle import swinir_ref, tiling
    cfg = configs.small_swinir_config("color_dn")
    sd = synth.swinir_state_dict(cfg, 2)
    img_bgr = synth.synthetic_image_u8(41, 37, 3, seed=8)
    i, o = _dirs(tmp_path, img_bgr)
    tool = SwinIRB200("denoising", "15", state_dict=sd, cfg=cfg, precision="fp32")
    assert tool.tool_name == "swinir_15"
    tool(i, o, silent=True)
    got = cv2.imread(str(o / "output.png"), cv2.IMREAD_COLOR)
    want = tiling.swinir_process_image(img_bgr, swinir_ref.Model(sd, cfg), 1, 8)
    assert np.abs(got.astype(int) - want.astype(int)).max() <= 1
    eng = tool._engine()
    got2 = process_image(img_bgr.astype(np.float32) / 255.0, eng, 1, 8, tile=16, tile_overlap=8)
    want2 = tiling.swinir_process_image(img_bgr, swinir_ref.Model(sd, cfg), 1, 8, 16, 8)
    assert np.abs(got2.astype(int) - want2.astype(int)).max() <= 1
    clear_engines()


def test_restormer_service_restore_image_bytes():
    import cv2
    from image_restoration_agent_b200.tool_api import RestormerService, clear_engines
    from oracle import restormer_ref, tiling
    cfg = configs.small_restormer_config("BiasFree", 1)
    sd = synth.restormer_state_dict(cfg, 4)
    gray = synth.synthetic_image_u8(33, 47, 1, seed=6)[:, :, 0]
    ok, enc = cv2.imencode(".png", gray)
    svc = RestormerService("Gaussian_Gray_Denoising", state_dict=sd, cfg=cfg, precision="fp32")
    out_bytes = svc.restore_image(enc.tobytes(), "Gaussian_Gray_Denoising", tile=16, tile_overlap=8)
    got = cv2.imdecode(np.frombuffer(out_bytes, np.uint8), cv2.IMREAD_GRAYSCALE)
    want = tiling.restormer_restore_image(gray[:, :, None], restormer_ref.Model(sd, cfg), 16, 8)[:, :, 0]
    assert np.abs(got.astype(int) - want.astype(int)).max() <= 1
    clear_engines()


def test_single_rank_sharded_restore_equals_plain():
    from image_restoration_agent_b200 import parallel
    from image_restoration_agent_b200.engine import Engine
    cfg = configs.small_restormer_config()
    eng = Engine("restormer", cfg, synth.restormer_state_dict(cfg, 1), precision="fp32")
    img = torch.from_numpy(synth.synthetic_image_u8(72, 90, 3, seed=1)).cuda()
    u8a, fa = eng.restore_device(img, 32, 8, want_f32=True)
    u8b, fb = parallel.restore_sharded(eng, img, 32, 8, want_f32=True)
    assert torch.equal(u8a, u8b) and (fa - fb).abs().max().item() <= 1e-6
    eng.close()


def test_tool_chain_derain_denoise_sr(tmp_path):
    """BASELINE config 4 (mixed-degradation chain): Restormer derain -> SwinIR denoise -> SwinIR x4 SR, each tool reading
    the previous tool's output.png (AgenticIR/executor/tool.py:33-62). Every stage is checked against the oracle run
    on the SAME input file the stage saw, so a 1-LSB difference of one stage cannot compound."""
    import cv2
    from image_restoration_agent_b200.tool_api import RestormerB200, SwinIRB200, clear_engines
    from oracle import restormer_ref, swinir_ref, tiling
    rcfg = configs.small_restormer_config()
    rsd = synth.restormer_state_dict(rcfg, 3)
    dcfg = configs.small_swinir_config("color_dn")
    dsd = synth.swinir_state_dict(dcfg, 4)
    scfg = configs.small_swinir_config("real_sr", scale=4, large_model=True)
    ssd = synth.swinir_state_dict(scfg, 5)
    img = synth.synthetic_image_u8(40, 56, 3, seed=11)
    d0 = tmp_path / "s0"
    d0.mkdir()
    cv2.imwrite(str(d0 / "input.png"), img)
    stages = [
        (RestormerB200("deraining", state_dict=rsd, cfg=rcfg, precision="fp32"),
         lambda bgr: cv2.cvtColor(tiling.restormer_restore_image(cv2.cvtColor(bgr, cv2.COLOR_BGR2RGB),
                                                                 restormer_ref.Model(rsd, rcfg)), cv2.COLOR_RGB2BGR)),
        (SwinIRB200("denoising", "15", state_dict=dsd, cfg=dcfg, precision="fp32"),
         lambda bgr: tiling.swinir_process_image(bgr, swinir_ref.Model(dsd, dcfg), 1, 8)),
        (SwinIRB200("super_resolution", "psnr", state_dict=ssd, cfg=scfg, precision="fp32"),
         lambda bgr: tiling.swinir_process_image(bgr, swinir_ref.Model(ssd, scfg), 4, 8)),
    ]
    src = d0
    for k, (tool, oracle_fn) in enumerate(stages):
        dst = tmp_path / f"s{k + 1}"
        dst.mkdir()
        tool(src, dst, silent=True)
        seen = cv2.imread(str(sorted(src.glob("*"))[0]), cv2.IMREAD_COLOR)
        got = cv2.imread(str(dst / "output.png"), cv2.IMREAD_COLOR)
        want = oracle_fn(seen)
        assert got.shape == want.shape, (k, got.shape, want.shape)
        assert np.abs(got.astype(int) - want.astype(int)).max() <= 1, k
        src = dst
    assert got.shape == (160, 224, 3)
    clear_engines()
