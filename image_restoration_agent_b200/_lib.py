"""ctypes binding of libirb200.so (include/irb200.h). No torch types cross this boundary: only raw device
pointers (tensor.data_ptr()), sizes and a cudaStream_t handle.

The library is the ONLY compute path: if it is missing or fails, calls raise — there is no fallback.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import os

PKG = Path(__file__).resolve().parent
# IRB_LIB selects another build of the same ABI (the -DIRB_DIAG library of tools/gpu_ab.sh); default = the product
LIB_PATH = Path(os.environ["IRB_LIB"]).resolve() if os.environ.get("IRB_LIB") else PKG / "libirb200.so"

ARCH_RESTORMER, ARCH_SWINIR, ARCH_XRESTORMER, ARCH_HAT = 1, 2, 3, 4
F32, BF16 = 0, 1
UPSAMPLER_NONE, UPSAMPLER_PIXELSHUFFLE, UPSAMPLER_NEAREST_CONV = 0, 1, 2
ACT_NONE, ACT_GELU, ACT_LRELU = 0, 1, 2


class ModelDesc(C.Structure):
    _fields_ = [
        ("arch", C.c_int), ("precision", C.c_int), ("in_chans", C.c_int),
        ("dim", C.c_int), ("num_blocks", C.c_int * 4), ("num_refinement_blocks", C.c_int),
        ("heads", C.c_int * 4), ("ffn_expansion_factor", C.c_float), ("ln_bias_free", C.c_int),
        ("img_size", C.c_int), ("window_size", C.c_int), ("embed_dim", C.c_int), ("num_layers", C.c_int),
        ("depths", C.c_int * 16), ("num_heads", C.c_int * 16), ("mlp_ratio", C.c_int), ("upscale", C.c_int),
        ("img_range", C.c_float), ("upsampler", C.c_int), ("resi_3conv", C.c_int),
        ("spatial_heads", C.c_int * 4), ("spatial_dim_head", C.c_int), ("overlap_ratio", C.c_float),
        ("unshuffle_scale", C.c_int),
        ("compress_ratio", C.c_int), ("squeeze_factor", C.c_int), ("conv_scale", C.c_float),
    ]


_vp, _i, _ll, _f, _sz = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_size_t
_ip = C.POINTER(C.c_int)

# name -> (restype, argtypes); mirrors include/irb200.h one to one
PROTOTYPES = {
    "irb_last_error": (C.c_char_p, []),
    "irb_version": (_i, []),
    "irb_plan_tiles": (_i, [_i, _i, _i, _ip, _ip]),
    "irb_padded_size": (_i, [_i, _i, _i, _i, _ip, _ip]),
    "irb_create": (_i, [C.POINTER(ModelDesc), _vp, _sz, _i, C.POINTER(_vp)]),
    "irb_destroy": (None, [_vp]),
    "irb_set_option": (_i, [_vp, C.c_char_p, _ll]),
    "irb_scale": (_i, [_vp]),
    "irb_tile_batch": (_i, [_vp, _i, _i, _i, _ip]),
    "irb_workspace_bytes": (_i, [_vp, _i, _i, _i, C.POINTER(_sz)]),
    "irb_forward_tiles": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _sz, _vp]),
    "irb_restore_workspace_bytes": (_i, [_vp, _i, _i, _i, _i, C.POINTER(_sz)]),
    "irb_restore_image": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _sz, _vp]),
    "irb_restore_plan": (_i, [_vp, _i, _i, _i, _i, _ip, _ip, _ip, _ip, _ip, _ip, _ip, _ip]),
    "irb_tiles_prepare": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp]),
    "irb_forward_tiles_cl": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _sz, _vp]),
    "irb_layernorm": (_i, [_vp, _i, _ll, _vp, _ll, _ll, _i, _vp, _vp, _i, _f, _vp]),
    "irb_gemm1x1": (_i, [_vp, _i, _ll, _i, _i, _i, _vp, _i, _i, _i, _i, _vp, _vp, _ll, _i, _f, _vp, _ll, _vp]),
    "irb_conv3x3": (_i, [_vp, _i, _ll, _i, _i, _i, _i, _vp, _i, _i, _i, _vp, _vp, _ll, _i, _f, _i, _vp, _ll, _vp]),
    "irb_dwconv3x3": (_i, [_vp, _i, _ll, _vp, _ll, _i, _i, _i, _i, _vp, _vp, _i, _vp]),
    "irb_gdfn_dw_gate": (_i, [_vp, _i, _ll, _vp, _ll, _i, _i, _i, _i, _vp, _vp]),
    "irb_pwdw_fused": (_i, [_i, _vp, _ll, _vp, _vp, _i, _i, _vp, _vp, _vp, _i, _vp, _ll, _i, _i, _i, _i, _vp]),
    "irb_mlp_fused": (_i, [_vp, _ll, _ll, _i, _i, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _f, _vp]),
    "irb_hat_attention": (_i, [_vp, _i, _ll, _vp, _ll, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "irb_restore_crop_workspace_bytes": (_i, [_vp, _i, _i, _i, _i, C.POINTER(_sz)]),
    "irb_restore_image_crop": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _sz, _vp]),
    "irb_ocab": (_i, [_vp, _i, _ll, _vp, _ll, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp]),
    "irb_mdta_nchunks": (_i, [_i, _i, _i]),
    "irb_mdta_gram": (_i, [_vp, _i, _ll, _i, _i, _i, _i, _i, _vp, _vp, _vp]),
    "irb_mdta_attn_static_max_bytes": (_sz, []),
    "irb_mdta_attn_static": (_i, [_vp, _i, _ll, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "irb_mdta_attn": (_i, [_vp, _vp, _i, _vp, _vp, _i, _i, _i, _vp, _i, _vp, _vp]),
    "irb_stitch": (_i, [_vp, _i, _i, _i, _i, _ip, _ip, _i, _i, _i, _i, _i, _i, _vp, _vp, _i, _vp]),
    "irb_stitch_rows": (_i, [_vp, _i, _i, _i, _i, _ip, _ip, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _i, _vp]),
    "irb_profile_begin": (_i, [_vp]),
    "irb_profile_end": (_i, [_vp]),
    "irb_profile_count": (_i, []),
    "irb_profile_entry": (_i, [_i, C.c_char_p, _ip, C.POINTER(C.c_double), C.POINTER(C.c_double),
                               C.POINTER(C.c_double)]),
    "irb_window_attention": (_i, [_vp, _i, _ll, _vp, _ll, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _vp]),
}

_lib = None


class IrbError(RuntimeError):
    pass


def load(strict: bool = True) -> C.CDLL:
    """Loads the C-ABI library, failing loudly when it has not been built.

    strict=False (kernel bring-up scripts only) tolerates a library that does not export every symbol yet.
    """
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise IrbError(f"{LIB_PATH} is missing: run `python -m image_restoration_agent_b200.build` "
                           "(there is no CPU or PyTorch fallback)")
        lib = C.CDLL(str(LIB_PATH))
        for name, (res, args) in PROTOTYPES.items():
            if not strict and not hasattr(lib, name):
                continue
            fn = getattr(lib, name)  # AttributeError if the header and the library disagree
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        msg = load().irb_last_error().decode(errors="replace")
        if rc == 1:
            raise ValueError(f"irb200: {msg}")
        raise IrbError(f"irb200 (code {rc}): {msg}")


def int_array(values):
    arr = (C.c_int * len(values))(*values)
    return arr
