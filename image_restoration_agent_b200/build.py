"""Builds libirb200.so (hand-written sm_100a CUDA behind the C ABI in include/irb200.h) in-tree with nvcc.

nvcc cross-compiles for sm_100a without a GPU; the .so travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
BUILD = PKG / "_build"
LIB = PKG / "libirb200.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def _digest(paths) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(str(p.name).encode())
        h.update(p.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False, diag: bool = False) -> Path:
    """diag=True builds libirb200_diag.so with -DIRB_DIAG: the A/B environment switches of DESIGN.md §5 exist only
    there (IRB_LIB=<path> makes _lib.load() pick it up); the shipped library has them compiled out."""
    if diag:
        return _build(force, verbose, PKG / "libirb200_diag.so", PKG / "_build_diag", ["-DIRB_DIAG"])
    return _build(force, verbose, LIB, BUILD, [])


def _build(force: bool, verbose: bool, LIB: Path, BUILD: Path, extra) -> Path:
    NVCC_FLAGS = globals()["NVCC_FLAGS"] + list(extra)
    sources = sorted(CSRC.glob("*.cu"))
    headers = sorted(CSRC.glob("*.cuh")) + sorted(CSRC.glob("*.h")) + [PKG.parent / "include" / "irb200.h"]
    BUILD.mkdir(exist_ok=True)
    stamp = BUILD / "stamp"
    dig = _digest(sources + headers) + " ".join(extra)
    if not force and LIB.exists() and stamp.exists() and stamp.read_text() == dig:
        return LIB
    nvcc = _nvcc()

    def compile_one(src: Path) -> Path:
        obj = BUILD / (src.stem + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(sources))) as ex:
        objs = list(ex.map(compile_one, sources))
    cmd = [nvcc, "-shared", "-o", str(LIB), *map(str, objs), "-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(dig)
    return LIB


if __name__ == "__main__":
    # --variant NAME -DFOO=1 ...  builds libirb200_NAME.so with extra defines (same-box A/B of two kernel variants:
    # tools/gpu_round.sh TAG ab:<lib_a>:<lib_b>:<workload>)
    if "--variant" in sys.argv:
        name = sys.argv[sys.argv.index("--variant") + 1]
        defs = [a for a in sys.argv if a.startswith("-D")]
        print(_build("--force" in sys.argv, "-v" in sys.argv, PKG / f"libirb200_{name}.so", PKG / f"_build_{name}", defs))
    else:
        print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, diag="--diag" in sys.argv))
