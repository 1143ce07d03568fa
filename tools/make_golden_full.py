"""Round-2 golden fixtures: the UNMODIFIED reference at the sizes and depths that bench.py measures.

Runs only in the build container (needs /root/reference); `python tools/make_golden_full.py [group ...]`.
Groups:
  restormer_tiles : Restormer Motion_Deblurring on the first 256x256 tile of the cfg2 image (Gram K = 65 536) and
                    Real_Denoising (BiasFree) on the first 512x512 tile of the cfg5 image (K = 262 144)
  swinir_full     : full-depth SwinIR-M color_dn, SwinIR-M classical_sr x4 (cfg3's model) and SwinIR-L real_sr x4
                    (9 RSTB, C = 240, 3conv) on 64x64 image-like inputs
  tool_cfg2       : tool-level crop of cfg2 (480x480, tile 256 / overlap 32 -> 2x2 tiles), full Restormer
  swinir_test     : the SwinIR tool fixtures with the tile loop run by the reference's OWN test()
                    (SwinIR/inference.py:294-322, imported with the timm shim), pad / crop / quantise of
                    inference.py:76-90 around it; oracle/tiling.py must reproduce them bit for bit
Every case asserts oracle == reference before the fixture is written.
"""
import os
import sys
import time
import types

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden as mg  # noqa: E402  (sets sys.path, REF paths, shim)

from image_restoration_agent_b200 import configs, synth  # noqa: E402
from oracle import restormer_ref, swinir_ref, tiling  # noqa: E402


def image_like(h, w, c, seed):
    """[1, c, h, w] fp32 in [0,1] from the bench's synthetic uint8 image (same generator, same seed)."""
    img = synth.synthetic_image_u8(h, w, c, seed=seed)
    return torch.from_numpy(img).float().div(255.0).permute(2, 0, 1).unsqueeze(0).contiguous()


def restormer_tiles():
    for name, task, full, t in (("restormer_cfg2_tile256", "Motion_Deblurring", (720, 1280), 256),
                                ("restormer_cfg5_tile512", "Real_Denoising", (2160, 3840), 512)):
        print(name)
        cfg = configs.restormer_config(task)
        sd = synth.restormer_state_dict(cfg, seed=0)
        ref = mg.ref_restormer(cfg, sd)
        x = image_like(full[0], full[1], 3, seed=0)[..., :t, :t].contiguous()
        t0 = time.time()
        with torch.no_grad():
            y_ref = ref(x)
        print(f"  reference forward {time.time() - t0:.1f} s")
        y_or = restormer_ref.Model(sd, cfg)(x)
        mg.check(name, y_or, y_ref)
        mg.save(name, x=x.numpy(), y=y_ref.numpy(), seed=np.int64(0))


def swinir_full():
    cases = [
        ("swinir_full_color_dn", configs.swinir_config("color_dn"), 0),
        ("swinir_full_classical_x4", configs.swinir_config("classical_sr", scale=4), 0),
        ("swinir_full_realsr_large_x4", configs.swinir_config("real_sr", scale=4, large_model=True), 0),
    ]
    for name, cfg, seed in cases:
        print(name)
        sd = synth.swinir_state_dict(cfg, seed=seed)
        ref = mg.ref_swinir(cfg, sd)
        x = image_like(512, 512, 3, seed=0)[..., 100:164, 200:264].contiguous()
        t0 = time.time()
        with torch.no_grad():
            y_ref = ref(x)
        print(f"  reference forward {time.time() - t0:.1f} s, out {tuple(y_ref.shape)}")
        y_or = swinir_ref.Model(sd, cfg)(x)
        mg.check(name, y_or, y_ref, tol=5e-5)
        mg.save(name, x=x.numpy(), y=y_ref.numpy(), seed=np.int64(seed))


def tool_cfg2():
    name = "restormer_tool_cfg2crop_t256_o32"
    print(name)
    cfg = configs.restormer_config("Motion_Deblurring")
    sd = synth.restormer_state_dict(cfg, seed=0)
    ref = mg.ref_restormer(cfg, sd)
    img = np.ascontiguousarray(synth.synthetic_image_u8(720, 1280, 3, seed=0)[:480, :480])
    u8_ref = tiling.restormer_restore_image(img, ref, 256, 32)
    f_ref = tiling.restormer_restore_image(img, ref, 256, 32, return_float=True)
    u8_or = tiling.restormer_restore_image(img, restormer_ref.Model(sd, cfg), 256, 32)
    d = np.abs(u8_or.astype(int) - u8_ref.astype(int)).max()
    print(f"  uint8 oracle vs reference: max LSB diff = {d}")
    assert d <= 1
    mg.save(name, img=img, out_u8=u8_ref, out_f32=f_ref.numpy(), seed=np.int64(0), tile=np.int64(256),
            overlap=np.int64(32))


def reference_swinir_tool(img_bgr_u8, model, scale, window_size, tile, overlap, ref_test):
    """inference.py:69-90 around the reference's own test(): returns (uint8 HWC BGR, float CHW RGB un-clamped)."""
    args = types.SimpleNamespace(tile=tile, tile_overlap=overlap, scale=scale)
    img_lq = img_bgr_u8.astype(np.float32) / 255.0                                            # :240
    img_lq = np.transpose(img_lq if img_lq.shape[2] == 1 else img_lq[:, :, [2, 1, 0]], (2, 0, 1))  # :70
    img_lq = torch.from_numpy(np.ascontiguousarray(img_lq)).float().unsqueeze(0)                # :71
    with torch.no_grad():
        _, _, h_old, w_old = img_lq.size()
        h_pad = (h_old // window_size + 1) * window_size - h_old
        w_pad = (w_old // window_size + 1) * window_size - w_old
        img_lq = torch.cat([img_lq, torch.flip(img_lq, [2])], 2)[:, :, :h_old + h_pad, :]
        img_lq = torch.cat([img_lq, torch.flip(img_lq, [3])], 3)[:, :, :, :w_old + w_pad]
        output = ref_test(img_lq, model, args, window_size)                                    # :81 the reference's
        output = output[..., :h_old * scale, :w_old * scale]
    f = output[0].clone()
    o = output.data.squeeze().float().cpu().clamp_(0, 1).numpy()
    if o.ndim == 3:
        o = np.transpose(o[[2, 1, 0], :, :], (1, 2, 0))
    return (o * 255.0).round().astype(np.uint8), f


def swinir_test():
    mg.install_timm_shim()
    sys.path.insert(0, mg.SW)
    import inference as ref_inference  # the reference script; main() is behind a __main__ guard

    for task, scale, tile, ov, full in (("color_dn", 1, None, 32, False), ("color_dn", 1, 16, 8, False),
                                        ("classical_sr", 4, 16, 8, False), ("color_dn", 1, 64, 32, True)):
        if full:  # full-depth SwinIR-M, 104x88 image -> 112x96 padded, tile 64 / overlap 32 -> 3x2 tiles
            cfg = configs.swinir_config(task, scale=scale)
            img = np.ascontiguousarray(synth.synthetic_image_u8(512, 512, 3, seed=0)[40:144, 60:148])
            name = f"swinir_tool_full_{task}_x{scale}_t{tile}_o{ov}"
        else:
            cfg = configs.small_swinir_config(task, scale=scale)
            img = synth.synthetic_image_u8(30, 27, 3, seed=5)
            name = f"swinir_tool_small_{task}_x{scale}_t{tile}_o{ov}"
        print(name)
        sd = synth.swinir_state_dict(cfg, seed=7)
        ref = mg.ref_swinir(cfg, sd)
        u8_ref, f_ref = reference_swinir_tool(img, ref, scale, cfg["window_size"], tile, ov, ref_inference.test)
        # the restated tile loop must be the same function
        u8_t = tiling.swinir_process_image(img, ref, scale, cfg["window_size"], tile, ov)
        f_t = tiling.swinir_process_image(img, ref, scale, cfg["window_size"], tile, ov, return_float=True)
        assert np.array_equal(u8_t, u8_ref) and torch.equal(f_t, f_ref), "oracle/tiling.py != inference.test()"
        u8_or = tiling.swinir_process_image(img, swinir_ref.Model(sd, cfg), scale, cfg["window_size"], tile, ov)
        d = np.abs(u8_or.astype(int) - u8_ref.astype(int)).max()
        print(f"  tiling.py == inference.test() bit-exact; uint8 oracle model vs reference: max LSB diff = {d}")
        assert d <= 1
        mg.save(name, img=img, out_u8=u8_ref, out_f32=f_ref.numpy(), seed=np.int64(7),
                tile=np.int64(-1 if tile is None else tile), overlap=np.int64(ov), scale=np.int64(scale),
                full=np.int64(int(full)))


GROUPS = {"restormer_tiles": restormer_tiles, "swinir_full": swinir_full, "tool_cfg2": tool_cfg2,
          "swinir_test": swinir_test}



# ---------------- X-Restormer ----------------
XR = f"{mg.REF}/AgenticIR/executor/dehazing/tools/X-Restormer"


def ref_xrestormer(cfg, sd):
    """The reference XRestormer class, imported from its file with stand-ins for the two imports that are absent here
    (basicsr's ARCH_REGISTRY decorator and torchstat, neither touches the arithmetic)."""
    import importlib.util
    basicsr = types.ModuleType("basicsr")
    utils = types.ModuleType("basicsr.utils")
    registry = types.ModuleType("basicsr.utils.registry")

    class _Reg:
        def register(self, *a, **k):
            return lambda cls: cls

    registry.ARCH_REGISTRY = _Reg()
    torchstat = types.ModuleType("torchstat")
    torchstat.stat = lambda *a, **k: None
    sys.modules.update({"basicsr": basicsr, "basicsr.utils": utils, "basicsr.utils.registry": registry,
                        "torchstat": torchstat})
    spec = importlib.util.spec_from_file_location("xrestormer_arch_ref", f"{XR}/xrestormer/archs/xrestormer_arch.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    m = mod.XRestormer(**cfg)
    m.load_state_dict(sd, strict=True)
    return m.eval()


def xrestormer():
    from oracle import xrestormer_ref
    cases = [("xrestormer_small_withbias", configs.small_xrestormer_config("WithBias"), (1, 3, 64, 128), 1),
             ("xrestormer_small_biasfree", configs.small_xrestormer_config("BiasFree"), (2, 3, 64, 64), 2),
             ("xrestormer_full", configs.xrestormer_config(), (1, 3, 128, 64), 0)]
    for name, cfg, shape, seed in cases:
        print(name)
        sd = synth.xrestormer_state_dict(cfg, seed=seed)
        ref = ref_xrestormer(cfg, sd)
        x = image_like(512, 512, 3, seed=seed)[..., 30:30 + shape[2], 70:70 + shape[3]].repeat(shape[0], 1, 1, 1).contiguous()
        if shape[0] > 1:
            x[1] = x[1].flip(-1)
        with torch.no_grad():
            y_ref = ref(x)
        y_or = xrestormer_ref.Model(sd, cfg)(x)
        mg.check(name, y_or, y_ref, tol=2e-5)
        mg.save(name, x=x.numpy(), y=y_ref.numpy(), seed=np.int64(seed))
    # tool level: reflect pad to a multiple of 64, whole image, crop, tensor2img (xrestormer_model.py:56-66, :139-142)
    name = "xrestormer_pipeline_small"
    print(name)
    cfg = configs.small_xrestormer_config("WithBias")
    sd = synth.xrestormer_state_dict(cfg, seed=3)
    ref = ref_xrestormer(cfg, sd)
    img = synth.synthetic_image_u8(70, 100, 3, seed=4)
    u8_ref = tiling.xrestormer_restore_image(img, ref)
    f_ref = tiling.xrestormer_restore_image(img, ref, return_float=True)
    u8_or = tiling.xrestormer_restore_image(img, xrestormer_ref.Model(sd, cfg))
    d = np.abs(u8_or.astype(int) - u8_ref.astype(int)).max()
    print(f"  uint8 oracle vs reference: max LSB diff = {d}")
    assert d <= 1
    mg.save(name, img=img, out_u8=u8_ref, out_f32=f_ref.numpy(), seed=np.int64(3), tile=np.int64(-1), overlap=np.int64(0))


GROUPS["xrestormer"] = xrestormer


def xrestormer_sr():
    """The scale-4 SR config (executor/super_resolution/configs/xrestormer.yml: scale 4, padding_size 16)."""
    from oracle import xrestormer_ref
    name = "xrestormer_small_sr_x4"
    print(name)
    cfg = configs.small_xrestormer_config("WithBias", scale=4)
    sd = synth.xrestormer_state_dict(cfg, seed=5)
    ref = ref_xrestormer(cfg, sd)
    x = image_like(512, 512, 3, seed=5)[..., 10:42, 20:68].contiguous()      # 32 x 48 -> 128 x 192
    with torch.no_grad():
        y_ref = ref(x)
    y_or = xrestormer_ref.Model(sd, cfg)(x)
    mg.check(name, y_or, y_ref, tol=2e-5)
    mg.save(name, x=x.numpy(), y=y_ref.numpy(), seed=np.int64(5))
    name = "xrestormer_pipeline_small_sr_x4"
    print(name)
    img = synth.synthetic_image_u8(30, 41, 3, seed=6)                          # pad 16 -> 32 x 48, out 120 x 164
    u8_ref = tiling.xrestormer_restore_image(img, ref, padding_size=16, scale=4)
    f_ref = tiling.xrestormer_restore_image(img, ref, return_float=True, padding_size=16, scale=4)
    u8_or = tiling.xrestormer_restore_image(img, xrestormer_ref.Model(sd, cfg), padding_size=16, scale=4)
    assert np.abs(u8_or.astype(int) - u8_ref.astype(int)).max() <= 1
    mg.save(name, img=img, out_u8=u8_ref, out_f32=f_ref.numpy(), seed=np.int64(5), tile=np.int64(-1), overlap=np.int64(0))


GROUPS["xrestormer_sr"] = xrestormer_sr


# ---------------- DiffBIR stage-1 SwinIR ----------------
DB = f"{mg.REF}/AgenticIR/executor/super_resolution/tools/DiffBIR"


def ref_diffbir_swinir(cfg, sd):
    """DiffBIR/model/swinir.py's SwinIR, imported from its file with stand-ins for what is absent here and touches no
    arithmetic: pytorch_lightning (LightningModule -> nn.Module), utils.metrics (LPIPS / PSNR, training-time logging
    only), the logging mixin, timm's three layer helpers."""
    import importlib.util
    mg.install_timm_shim()
    pl = types.ModuleType("pytorch_lightning")
    pl.LightningModule = torch.nn.Module
    plu = types.ModuleType("pytorch_lightning.utilities")
    plt = types.ModuleType("pytorch_lightning.utilities.types")
    plt.STEP_OUTPUT = object
    utils = types.ModuleType("utils")
    metrics = types.ModuleType("utils.metrics")
    metrics.calculate_psnr_pt = lambda *a, **k: None

    class LPIPS(torch.nn.Module):
        def __init__(self, *a, **k):
            super().__init__()

    metrics.LPIPS = LPIPS
    pkg = types.ModuleType("model")
    pkg.__path__ = [f"{DB}/model"]
    mixins = types.ModuleType("model.mixins")

    class ImageLoggerMixin:
        pass

    mixins.ImageLoggerMixin = ImageLoggerMixin
    sys.modules.update({"pytorch_lightning": pl, "pytorch_lightning.utilities": plu,
                        "pytorch_lightning.utilities.types": plt, "utils": utils, "utils.metrics": metrics,
                        "model": pkg, "model.mixins": mixins})
    spec = importlib.util.spec_from_file_location("model.swinir", f"{DB}/model/swinir.py")
    mod = importlib.util.module_from_spec(spec)
    sys.modules["model.swinir"] = mod
    spec.loader.exec_module(mod)
    kw = dict(cfg)
    kw["sf"] = kw.pop("upscale")
    kw["patch_size"] = 1
    m = mod.SwinIR(**kw)
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all(k.endswith("relative_position_index") or k.endswith("attn_mask") for k in missing), missing
    return m.eval()


def diffbir():
    for name, cfg, shape, seed in (("diffbir_swinir_small", configs.small_diffbir_swinir_config(), (1, 3, 64, 128), 1),
                                   ("diffbir_swinir_full", configs.diffbir_swinir_config(), (1, 3, 128, 128), 0)):
        print(name)
        sd = synth.swinir_state_dict(cfg, seed=seed)
        ref = ref_diffbir_swinir(cfg, sd)
        x = image_like(512, 512, 3, seed=seed)[..., 64:64 + shape[2], 128:128 + shape[3]].contiguous()
        with torch.no_grad():
            y_ref = ref(x)
        y_or = swinir_ref.Model(sd, cfg)(x)
        mg.check(name, y_or, y_ref, tol=5e-5)
        mg.save(name, x=x.numpy(), y=y_ref.numpy(), seed=np.int64(seed))


GROUPS["diffbir"] = diffbir



# ---------------- HAT ----------------
HATDIR = f"{mg.REF}/AgenticIR/executor/super_resolution/tools/HAT"


def _hat_shims():
    """Stand-ins for the basicsr imports of hat_arch.py / hat_model.py (registry decorators, to_2tuple, trunc_normal_,
    the SRModel base class, metric / image helpers): none of them touches the arithmetic of forward / tile_process."""
    def mod(name, **attrs):
        m = types.ModuleType(name)
        for k, v in attrs.items():
            setattr(m, k, v)
        sys.modules[name] = m
        return m

    class _Reg:
        def register(self, *a, **k):
            return lambda cls: cls

    mod("basicsr")
    mod("basicsr.utils", imwrite=None, tensor2img=None)
    mod("basicsr.utils.registry", ARCH_REGISTRY=_Reg(), MODEL_REGISTRY=_Reg())
    mod("basicsr.archs")
    mod("basicsr.archs.arch_util", to_2tuple=lambda x: (x, x) if not isinstance(x, tuple) else x,
        trunc_normal_=lambda t, std=1.0, **k: torch.nn.init.trunc_normal_(t, std=std))
    mod("basicsr.models")
    mod("basicsr.models.sr_model", SRModel=type("SRModel", (), {}))
    mod("basicsr.metrics", calculate_metric=None)


def ref_hat(cfg, sd):
    import importlib.util
    _hat_shims()
    spec = importlib.util.spec_from_file_location("hat_arch_ref", f"{HATDIR}/hat/archs/hat_arch.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    m = mod.HAT(**cfg)
    missing, unexpected = m.load_state_dict(sd, strict=False)  # the two rpi buffers are not parameters
    assert not unexpected and all("relative_position_index" in k for k in missing), (missing, unexpected)
    return m.eval()


def ref_hat_tool(img_bgr_u8, net, scale, window_size, tile_size, tile_pad, return_float=False):
    """The reference's OWN HATModel.pre_process / tile_process / process / post_process (hat_model.py:16-109), run as
    unbound functions on a stand-in object; img2tensor / tensor2img of basicsr around them."""
    import importlib.util
    _hat_shims()
    spec = importlib.util.spec_from_file_location("hat_model_ref", f"{HATDIR}/hat/models/hat_model.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    M = mod.HATModel
    o = types.SimpleNamespace()
    o.opt = {"network_g": {"window_size": window_size}, "scale": scale}
    if tile_size is not None:
        o.opt["tile"] = {"tile_size": tile_size, "tile_pad": tile_pad}
    o.net_g = net
    o.lq = torch.from_numpy(np.ascontiguousarray(img_bgr_u8[:, :, ::-1])).float().div(255.0).permute(2, 0, 1).unsqueeze(0)
    M.pre_process(o)
    if tile_size is not None:
        M.tile_process(o)
    else:
        M.process(o)
    M.post_process(o)
    y = o.output
    if return_float:
        return y[0].clone()
    a = y[0].float().clamp_(0, 1).numpy()
    a = np.transpose(a[[2, 1, 0], :, :], (1, 2, 0))
    return (a * 255.0).round().astype(np.uint8)


def hat():
    from oracle import hat_ref
    cases = [("hat_small", configs.small_hat_config(), (1, 3, 32, 48), 1),
             ("hat_small_w8_x2", configs.small_hat_config(window_size=8, upscale=2), (2, 3, 32, 24), 2),
             ("hat_full", configs.hat_config(), (1, 3, 32, 48), 0)]
    for name, cfg, shape, seed in cases:
        print(name)
        sd = synth.hat_state_dict(cfg, seed=seed)
        ref = ref_hat(cfg, sd)
        x = image_like(512, 512, 3, seed=seed)[..., 30:30 + shape[2], 70:70 + shape[3]].repeat(shape[0], 1, 1, 1).contiguous()
        if shape[0] > 1:
            x[1] = x[1].flip(-1)
        t0 = time.time()
        with torch.no_grad():
            y_ref = ref(x)
        print(f"  reference forward {time.time() - t0:.1f} s")
        y_or = hat_ref.Model(sd, cfg)(x)
        mg.check(name, y_or, y_ref, tol=2e-5)
        mg.save(name, x=x.numpy(), y=y_ref.numpy(), seed=np.int64(seed))
    # tool level: reflect pad to the window, crop-with-pad tiles (ragged last row / column), paste, crop, tensor2img
    cfg = configs.small_hat_config()
    sd = synth.hat_state_dict(cfg, seed=3)
    ref = ref_hat(cfg, sd)
    orc = hat_ref.Model(sd, cfg)
    for name, shape, ts, tp in (("hat_pipeline_small_t32_p16", (70, 100), 32, 16), ("hat_pipeline_small_whole", (40, 52), None, 0)):
        print(name)
        img = synth.synthetic_image_u8(shape[0], shape[1], 3, seed=4)
        u8_ref = ref_hat_tool(img, ref, 4, 16, ts, tp)
        f_ref = ref_hat_tool(img, ref, 4, 16, ts, tp, return_float=True)
        u8_same = tiling.hat_restore_image(img, ref, 4, 16, ts, tp)      # oracle tile loop around the reference network
        assert np.array_equal(u8_same, u8_ref), "oracle tile loop differs from HATModel.tile_process"
        u8_or = tiling.hat_restore_image(img, orc, 4, 16, ts, tp)
        d = np.abs(u8_or.astype(int) - u8_ref.astype(int)).max()
        print(f"  uint8 oracle vs reference: max LSB diff = {d}")
        assert d <= 1
        mg.save(name, img=img, out_u8=u8_ref, out_f32=f_ref.numpy(), seed=np.int64(3), tile=np.int64(ts or -1),
                overlap=np.int64(tp))


GROUPS["hat"] = hat


# ---------------- SwinIR lightweight SR: the 'pixelshuffledirect' upsampler (inference.py:144-150) ----------------
def lightweight():
    cases = [("swinir_small_lightweight_x2", configs.small_swinir_config("lightweight_sr", scale=2), (1, 3, 24, 40), 1),
             ("swinir_small_lightweight_x3", configs.small_swinir_config("lightweight_sr", scale=3), (2, 3, 21, 30), 2),
             ("swinir_full_lightweight_x4", configs.swinir_config("lightweight_sr", scale=4), (1, 3, 48, 64), 0)]
    for name, cfg, shape, seed in cases:
        print(name)
        sd = synth.swinir_state_dict(cfg, seed=seed)
        ref = mg.ref_swinir(cfg, sd)
        x = image_like(512, 512, 3, seed=seed)[..., 30:30 + shape[2], 70:70 + shape[3]].repeat(shape[0], 1, 1, 1).contiguous()
        if shape[0] > 1:
            x[1] = x[1].flip(-1)
        with torch.no_grad():
            y_ref = ref(x)
        y_or = swinir_ref.Model(sd, cfg)(x)
        mg.check(name, y_or, y_ref)
        mg.save(name, x=x.numpy(), y=y_ref.numpy(), seed=np.int64(seed))


GROUPS["lightweight"] = lightweight


if __name__ == "__main__":
    torch.manual_seed(0)
    torch.set_num_threads(os.cpu_count() or 8)
    for g in (sys.argv[1:] or list(GROUPS)):
        GROUPS[g]()
    print("done")
