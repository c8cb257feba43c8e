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

SOURCES = ["api.cu", "preprocess.cu", "binning.cu", "blend_fwd.cu", "blend_bwd.cu", "preprocess_bwd.cu", "idxmap.cu",
           "peer_reduce.cu", "sh_reconstruct.cu", "optimizer.cu", "loss.cu"]
HEADERS = [CSRC / "common.cuh", CSRC / "kernels.cuh", CSRC / "blend_common.cuh", CSRC / "sh_grad.cuh",
           INCLUDE / "b200splat.h"]

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


def build(force: bool = False, verbose: bool = False, extra_flags=(), tag: str = "") -> Path:
    """Compile every kernel translation unit for sm_100a and link the shared library. `tag` builds a tuning
    variant (extra -D flags) into _variants/libb200splat_<tag>.so without touching the product library."""
    nvcc = find_nvcc()
    global OBJ_DIR, LIB_PATH
    obj_dir, lib_path = OBJ_DIR, LIB_PATH
    if tag:
        obj_dir = PKG_DIR / "_obj" / f"variant_{tag}"
        (PKG_DIR / "_variants").mkdir(exist_ok=True)
        lib_path = PKG_DIR / "_variants" / f"libb200splat_{tag}.so"
        force = True
    obj_dir.mkdir(parents=True, exist_ok=True)
    logs = {}

    def compile_one(src: str):
        s = CSRC / src
        o = obj_dir / (Path(src).stem + ".o")
        if force or _stale(o, [s, *HEADERS]):
            cmd = [nvcc, *NVCC_FLAGS, *extra_flags, "-I", str(INCLUDE), "-c", str(s), "-o", str(o)]
            r = subprocess.run(cmd, capture_output=True, text=True)
            logs[src] = r.stderr
            if r.returncode != 0:
                raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        return o

    with ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    if force or _stale(lib_path, objs):
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(lib_path), *map(str, objs)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    if verbose:
        for k, v in logs.items():
            print(f"==== {k}\n{v}", file=sys.stderr)
    if logs:
        (obj_dir / "ptxas.log").write_text("".join(f"==== {k}\n{v}\n" for k, v in logs.items()))
    return lib_path


if __name__ == "__main__":
    # python build.py [--force] [--tag NAME -DFOO=1 ...]
    argv = sys.argv[1:]
    tag = argv[argv.index("--tag") + 1] if "--tag" in argv else ""
    p = build(force="--force" in argv, verbose=not tag, extra_flags=[a for a in argv if a.startswith("-D")], tag=tag)
    print(p)
