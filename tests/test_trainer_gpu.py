"""The whole iteration through this repo's kernels (trainer.ViewParallelTrainer): a perturbed scene fitted to renders of
the original must lower the L1+SSIM loss; the pieces it chains are parity-tested individually elsewhere."""
import json
import subprocess
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
pytestmark = pytest.mark.gpu


def _run(nproc, *extra):
    cmd = [sys.executable, str(ROOT / "tools" / "train_synthetic.py"), *extra] if nproc == 1 else \
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr",
         "127.0.0.1", "--master-port", "29537", str(ROOT / "tools" / "train_synthetic.py"), *extra]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-3000:]
    return json.loads(r.stdout.strip().splitlines()[-1])


def test_training_iteration_lowers_the_loss():
    rep = _run(1, "--cfg", "T1", "--iters", "60")
    assert rep["loss_last5"] < 0.8 * rep["loss_first5"], rep


def test_two_rank_training_keeps_replicas_bit_identical():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    rep = _run(2, "--cfg", "T1", "--iters", "40")
    assert rep["replicas_in_sync"] and rep["collective"] in ("peer", "nccl"), rep


def test_densification_inside_the_training_loop_changes_P_and_training_continues():
    rep = _run(1, "--cfg", "T1", "--iters", "40", "--densify-at", "20")
    assert rep["densified"][0] == 20000 and rep["densified"][1] != 20000 and rep["P"] == rep["densified"][1], rep
