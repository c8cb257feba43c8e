"""In-tree build of libb200splat.so (hand-written CUDA for sm_100a, C ABI only — no torch headers).

The shared library lands next to this file so that it travels to the GPU box with the repo
snapshot and shows up as an in-tree .so in the loaded-library record.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
CSRC = PKG_DIR / "csrc"
INCLUDE = PKG_DIR.parent / "include"
OBJ_DIR = PKG_DIR / "_obj"
LIB_PATH = PKG_DIR / "libb200splat.so"

SOURCES = ["api.cu", "preprocess.cu", "binning.cu", "blend_fwd.cu", "blend_bwd.cu", "preprocess_bwd.cu", "idxmap.cu"]
HEADERS = [CSRC / "common.cuh", CSRC / "kernels.cuh", CSRC / "blend_common.cuh", INCLUDE / "b200splat.h"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "--expt-relaxed-constexpr", "--extended-lambda",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
] + (os.environ.get("B200S_NVCC_EXTRA", "").split())


def find_nvcc() -> str:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(cand).exists():
        raise RuntimeError("nvcc not found: cannot build libb200splat.so (no CPU fallback exists)")
    return cand


def _stale(target: Path, deps) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(d).stat().st_mtime > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every kernel translation unit for sm_100a and link the shared library."""
    nvcc = find_nvcc()
    OBJ_DIR.mkdir(exist_ok=True)
    logs = {}

    def compile_one(src: str):
        s = CSRC / src
        o = OBJ_DIR / (Path(src).stem + ".o")
        if force or _stale(o, [s, *HEADERS]):
            cmd = [nvcc, *NVCC_FLAGS, "-I", str(INCLUDE), "-c", str(s), "-o", str(o)]
            r = subprocess.run(cmd, capture_output=True, text=True)
            logs[src] = r.stderr
            if r.returncode != 0:
                raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        return o

    with ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    if force or _stale(LIB_PATH, objs):
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(LIB_PATH), *map(str, objs)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    if verbose:
        for k, v in logs.items():
            print(f"==== {k}\n{v}", file=sys.stderr)
    if logs:
        (OBJ_DIR / "ptxas.log").write_text("".join(f"==== {k}\n{v}\n" for k, v in logs.items()))
    return LIB_PATH


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose=True)
    print(p)
