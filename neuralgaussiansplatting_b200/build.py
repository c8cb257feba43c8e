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
    if not tag and len(logs) == len(SOURCES) and not extra_flags and not os.environ.get("B200S_NVCC_EXTRA"):
        # a full build of the product library refreshes the committed resource record, so it cannot go stale
        try:
            write_ptxas_record(logs, PKG_DIR.parent / "profiles" / "ptxas_current.txt")
        except OSError:
            pass
    return lib_path


def write_ptxas_record(logs, path: Path) -> None:
    """One line per kernel from `-Xptxas -v`: registers, spill bytes, static shared memory (no compile times, so the
    file only changes when the code generation does). Kernels that spill are listed again at the end."""
    import re
    rows, spills = [], []
    for src in SOURCES:
        name = None
        stack = st = ld = 0
        for line in logs.get(src, "").splitlines():
            m = re.search(r"Function properties for (\S+)", line)
            if m:
                name = m.group(1)
                continue
            m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
            if m:
                stack, st, ld = map(int, m.groups())
                continue
            m = re.search(r"Used (\d+) registers", line)
            if m and name:
                sm = re.search(r"(\d+) bytes smem", line)
                regs, smem = int(m.group(1)), int(sm.group(1)) if sm else 0
                r = subprocess.run(["c++filt", "-p", name], capture_output=True, text=True)
                nice = r.stdout.strip() if r.returncode == 0 and r.stdout.strip() else name
                rows.append(f"{src:20s} {regs:4d} regs {smem:6d} B smem {stack:4d} B stack {st:4d} B spill st {ld:4d} B spill ld  {nice}")
                if st or ld:
                    spills.append(rows[-1])
                name = None
    path.parent.mkdir(exist_ok=True)
    path.write_text("# written by neuralgaussiansplatting_b200/build.py on every full build of libb200splat.so "
                    "(nvcc -Xptxas -v, sm_100a)\n" + "\n".join(rows) + "\n\n# kernels that spill: "
                    + (str(len(spills)) + "\n" + "\n".join(spills) if spills else "none") + "\n")


if __name__ == "__main__":
    # python build.py [--force] [--tag NAME -DFOO=1 ...]
    argv = sys.argv[1:]
    tag = argv[argv.index("--tag") + 1] if "--tag" in argv else ""
    p = build(force="--force" in argv, verbose=not tag, extra_flags=[a for a in argv if a.startswith("-D")], tag=tag)
    print(p)
