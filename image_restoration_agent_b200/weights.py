"""Reference state_dict -> IRBW v1 weights blob (the only weights format the C ABI accepts).

Accepts exactly what the reference loads: `torch.load(path)['params']` (Restormer, demo.py:101-102) or
`['params'|'params_ema']` (SwinIR, inference.py:197-198). Buffers (`relative_position_index`, `attn_mask`) are
dropped: the library derives them from token coordinates.

Layout (little endian):
  header : char magic[4]="IRBW"; u32 version=1; u32 n_tensors; u32 reserved
  entry  : char name[96]; u32 ndim; u32 dims[4]; u64 offset (from blob start, 16-byte aligned); u64 numel
  data   : fp32
"""
from __future__ import annotations

import struct

import numpy as np

_SKIP_SUFFIXES = ("relative_position_index", "attn_mask",
                  "relative_position_index_SA", "relative_position_index_OCA")  # HAT's two index buffers
_ENTRY = struct.Struct("<96sI4IQQ")


def state_dict_to_blob(state_dict) -> bytes:
    items = []
    for name, t in state_dict.items():
        if name.endswith(_SKIP_SUFFIXES):
            continue
        a = t.detach().cpu().float().numpy() if hasattr(t, "detach") else np.asarray(t, dtype=np.float32)
        a = np.ascontiguousarray(a, dtype="<f4")
        if a.ndim > 4:
            raise ValueError(f"{name}: more than 4 dims")
        if len(name.encode()) > 95:
            raise ValueError(f"{name}: name too long")
        items.append((name, a))
    table_bytes = 16 + _ENTRY.size * len(items)
    off = (table_bytes + 15) // 16 * 16
    head = [struct.pack("<4sIII", b"IRBW", 1, len(items), 0)]
    chunks = []
    for name, a in items:
        dims = list(a.shape) + [1] * (4 - a.ndim)
        head.append(_ENTRY.pack(name.encode(), a.ndim, *dims, off, a.size))
        raw = a.tobytes()
        pad = (-len(raw)) % 16
        chunks.append(raw + b"\0" * pad)
        off += len(raw) + pad
    table = b"".join(head)
    table += b"\0" * ((-len(table)) % 16)
    return table + b"".join(chunks)


def load_checkpoint(path: str, keys=("params", "params_ema")):
    """torch.load + the reference's key selection."""
    import torch

    ckpt = torch.load(path, map_location="cpu")
    for k in keys:
        if isinstance(ckpt, dict) and k in ckpt:
            return ckpt[k]
    return ckpt
