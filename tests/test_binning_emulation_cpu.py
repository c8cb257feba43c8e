"""Host-side check of the index arithmetic of the counting-rank tile partition (csrc/binning.cu: cr_sort_tile — per-thread
byte counters, 16-bit packed warp scan, nibble bases), restated thread by thread in tools/emulate_binning.py, against
numpy's stable sort. Small cases only; the GPU tests (tests/test_binning_modes_gpu.py) are the gate for the kernel."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tools"))
import emulate_binning as E  # noqa: E402


def test_counting_rank_partition_matches_stable_sort():
    rng = np.random.default_rng(3)
    for tile_n, shift, few in ((4096, 0, False), (777, 8, True), (1, 0, False)):
        keys = (rng.integers(0, 19, tile_n, dtype=np.uint64) << 8 | rng.integers(0, 3, tile_n, dtype=np.uint64)) if few \
            else rng.integers(0, 2 ** 32, tile_n, dtype=np.uint64)
        vals = np.arange(tile_n, dtype=np.uint64)
        ok, ov, ds = E.tile_sort(keys, vals, tile_n, shift)
        dig = (keys >> shift) & 255
        order = np.argsort(dig, kind="stable")
        assert (ok == keys[order]).all() and (ov == vals[order]).all()
        assert (np.diff(ds) == np.bincount(dig.astype(np.int64), minlength=256)).all()
