"""CPU: the drop-in boundary around the C ABI — CLI shims parse exactly what the reference's Tool._get_cmd_opts
builds, the B200 tools subclass the reference's own Tool when it is importable, the engine cache keys, and the
shipped library carries no diagnosis switch."""
import os
import subprocess
import sys

import pytest

from image_restoration_agent_b200 import _lib
from image_restoration_agent_b200.shims import demo as demo_shim
from image_restoration_agent_b200.shims import inference as inference_shim

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_AGENT = "/root/reference/AgenticIR"


def test_inference_shim_parses_the_executor_command_lines():
    """multitask_tools.py:120-133: the option list SwinIR._get_cmd_opts() hands to inference.py."""
    base = ["--task", "real_sr", "--model_path", "SwinIR/model_zoo/swinir/003_realSR_BSRGAN_DFOWMFC_s64w8_SwinIR-L_x4_GAN.pth",
            "--folder_lq", "/w/in", "--save_dir", "/w/out", "--scale", "4", "--large_model"]
    a = inference_shim.build_parser().parse_args(base)
    assert (a.task, a.scale, a.large_model, a.folder_lq, a.save_dir, a.tile, a.tile_overlap) == \
        ("real_sr", 4, True, "/w/in", "/w/out", None, 32)
    assert inference_shim.output_name(a.model_path) == "output_GAN"            # inference.py:244
    a = inference_shim.build_parser().parse_args(
        ["--task", "color_dn", "--model_path", "SwinIR/model_zoo/swinir/005_colorDN_DFWB_s128w8_SwinIR-M_noise15.pth",
         "--folder_lq", "i", "--save_dir", "o", "--tile", "256", "--tile_overlap", "16"])
    assert (a.scale, a.large_model, a.tile, a.tile_overlap) == (1, False, 256, 16)
    assert inference_shim.output_name(a.model_path) == "output_noise15"
    with pytest.raises(AssertionError):                                        # inference.py:38
        inference_shim.main(["--task", "gray_dn", "--folder_lq", "i", "--save_dir", "o"])


def test_demo_shim_parses_the_executor_command_lines(tmp_path):
    """multitask_tools.py:157-163 / demo.py:26-38."""
    a = demo_shim.build_parser().parse_args(["--task", "Deraining", "--input_dir", "/w/in", "--result_dir", "/w/out"])
    assert (a.task, a.input_dir, a.result_dir, a.tile, a.tile_overlap) == ("Deraining", "/w/in", "/w/out", None, 32)
    with pytest.raises(SystemExit):                                            # argparse choices, demo.py:29-34
        demo_shim.build_parser().parse_args(["--task", "Dehazing"])
    for n in ("img10.png", "img9.png", "img1.JPG", "notes.txt"):
        (tmp_path / n).write_bytes(b"")
    got = [os.path.basename(f) for f in demo_shim.find_files(str(tmp_path))]
    assert got == ["img1.JPG", "img9.png", "img10.png"]                        # natsorted, image extensions only
    assert demo_shim.find_files("/x/one.png") == ["/x/one.png"]                # :78-79 single-file form
    with pytest.raises(Exception, match="No files found"):
        demo_shim.main(["--task", "Deraining", "--input_dir", str(tmp_path / "none"), "--result_dir", str(tmp_path)])


@pytest.mark.skipif(not os.path.isdir(REF_AGENT), reason="reference tree not present (GPU box)")
def test_b200_tools_subclass_the_reference_tool():
    """In the reference process (AgenticIR on sys.path) isinstance(tool, executor.tool.Tool) must hold."""
    code = (
        "import sys; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "from executor.tool import Tool as RefTool\n"
        "from image_restoration_agent_b200 import tool_api\n"
        "assert tool_api.Tool is RefTool\n"
        "t = tool_api.SwinIRB200('denoising', '15', state_dict={}, cfg=None)\n"
        "r = tool_api.RestormerB200('deraining', state_dict={})\n"
        "assert isinstance(t, RefTool) and isinstance(r, RefTool)\n"
        "assert (t.tool_name, t.subtask, r.tool_name) == ('swinir_15', 'denoising', 'restormer')\n"
        "print('ok')\n" % (REF_AGENT, ROOT))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd="/tmp")
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr[-2000:]


def test_local_tool_mirror_contract(tmp_path):
    """Without the reference on sys.path the local mirror keeps tool.py:51-62's pre/post-conditions."""
    from image_restoration_agent_b200 import tool_api

    class Copy(tool_api._LocalTool):
        def _invoke(self):
            (self.output_dir / "whatever.png").write_bytes((sorted(self.input_dir.glob("*"))[0]).read_bytes())

    i, o = tmp_path / "i", tmp_path / "o"
    i.mkdir(), o.mkdir()
    (i / "input.png").write_bytes(b"x")
    Copy("copy", "none")(i, o, silent=True)
    assert os.listdir(o) == ["output.png"]                                     # renamed (:59-62)
    with pytest.raises(AssertionError, match="should be empty"):
        Copy("copy", "none")(i, o, silent=True)


def test_engine_cache_key_separates_configs_and_sources(monkeypatch):
    from image_restoration_agent_b200 import tool_api
    built = []

    class FakeEngine:
        def __init__(self, arch, cfg, sd, precision="bf16", device=0):
            built.append((arch, cfg, id(sd), precision, device))

        def close(self):
            pass

    monkeypatch.setattr(tool_api, "Engine", FakeEngine)
    tool_api.clear_engines()
    sd1, sd2 = {"a": 1}, {"a": 2}
    e1 = tool_api.get_engine("swinir", sd1, {"d": 1}, lambda: sd1)
    assert tool_api.get_engine("swinir", sd1, {"d": 1}, lambda: sd1) is e1
    assert tool_api.get_engine("swinir", sd1, {"d": 2}, lambda: sd1) is not e1          # cfg is part of the key
    assert tool_api.get_engine("swinir", sd2, {"d": 1}, lambda: sd2) is not e1          # other weights object
    assert tool_api.get_engine("swinir", "/z/a.pth", {"d": 1}, lambda: sd1) is not \
        tool_api.get_engine("swinir", "/y/a.pth", {"d": 1}, lambda: sd1)                # checkpoint directory matters
    assert tool_api.get_engine("swinir", sd1, {"d": 1}, lambda: sd1, device=1) is not e1
    assert len(built) == 6
    tool_api.clear_engines()


def test_shipped_library_has_no_diagnosis_switches(lib):
    """VERDICT r1 hygiene #14: no environment variable may skip stores / MMAs / epilogues in the product library."""
    blob = open(_lib.LIB_PATH, "rb").read()
    for name in (b"IRB_GEMM_DBG", b"IRB_GEMM_SPEC", b"IRB_PDL", b"IRB_FUSED_FFN", b"IRB_TMA_PROMO", b"IRB_GELU_TANH"):
        assert name not in blob, name
    csrc = os.path.join(ROOT, "image_restoration_agent_b200", "csrc")
    for f in os.listdir(csrc):
        src = open(os.path.join(csrc, f)).read()
        if "getenv" in src:
            assert f == "api_ops.cu" and "#ifdef IRB_DIAG" in src, f"getenv outside the IRB_DIAG block in {f}"
