"""The opt-in binning paths (csrc/binning.cu: counting-rank sort passes, tile-id keys instead of (x, y) keys, expansion
fused into the first tile pass) must produce the SAME sorted lists, ranges, n_contrib and image bits as the default
path. The library reads its B200S_* switches once per process, so every mode runs in a child process; the default
mode's digests are the golden (the default itself is compared with the reference in test_parity_gpu.py)."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
MODES = {
    "tile_id_keys": {"B200S_XY_KEYS": "0"},
    "counting_rank_all_passes": {"B200S_RANK_COUNT": "0xff"},
    "counting_rank_tile_id_keys": {"B200S_RANK_COUNT": "0xff", "B200S_XY_KEYS": "0"},
    "fused_expansion": {"B200S_RANK_COUNT": "0x10", "B200S_FUSE_EXPAND": "1"},
    "ballot_rank": {"B200S_RANK_MATCH": "0"},
}


def run_child(extra_env):
    env = {k: v for k, v in os.environ.items() if not k.startswith("B200S_") or k in ("B200S_LIB",)}
    env.update(extra_env)
    r = subprocess.run([sys.executable, str(ROOT / "tests" / "_binning_mode_child.py")], env=env, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return json.loads(r.stdout.strip().splitlines()[-1])


@pytest.fixture(scope="module")
def golden():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return run_child({})


@pytest.mark.parametrize("mode", sorted(MODES))
def test_binning_mode_is_bit_identical_to_default(golden, mode):
    got = run_child(MODES[mode])
    assert got == golden, {k: (got[k], golden[k]) for k in golden if got.get(k) != golden[k]}
