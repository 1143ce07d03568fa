"""Per-launch device timing of the irb200 kernels (irb_profile_* in include/irb200.h).

    with KernelProfile() as prof:
        engine.restore_device(img, 256, 32)
    prof.entries -> [{"name", "launches", "ms", "flops", "bytes"}, ...] sorted by device time

Every kernel the library launches on the current stream between enter and exit is bracketed by CUDA events; the
flops / bytes are the ALGORITHMIC work of each launch (operands read once, results written once), which is what
bench.py's roofline divides by the measured duration.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import check


class KernelProfile:
    def __init__(self):
        self.entries: list[dict] = []

    def __enter__(self):
        self._stream = torch.cuda.current_stream().cuda_stream
        check(_lib.load().irb_profile_begin(self._stream))
        return self

    def __exit__(self, *exc):
        lib = _lib.load()
        check(lib.irb_profile_end(self._stream))
        name = C.create_string_buffer(56)
        n, ms, fl, by = C.c_int(), C.c_double(), C.c_double(), C.c_double()
        out = []
        for i in range(lib.irb_profile_count()):
            check(lib.irb_profile_entry(i, name, C.byref(n), C.byref(ms), C.byref(fl), C.byref(by)))
            out.append(dict(name=name.value.decode(), launches=n.value, ms=ms.value, flops=fl.value, bytes=by.value))
        self.entries = sorted(out, key=lambda e: -e["ms"])
        return False

    @property
    def total_ms(self) -> float:
        return sum(e["ms"] for e in self.entries)

    @property
    def launches(self) -> int:
        return sum(e["launches"] for e in self.entries)

    def by_kernel(self) -> list[dict]:
        """Entries merged over shape classes: name[...] -> name."""
        agg: dict[str, dict] = {}
        for e in self.entries:
            k = e["name"].split("[")[0]
            a = agg.setdefault(k, dict(name=k, launches=0, ms=0.0, flops=0.0, bytes=0.0))
            for f in ("launches", "ms", "flops", "bytes"):
                a[f] += e[f]
        return sorted(agg.values(), key=lambda e: -e["ms"])
