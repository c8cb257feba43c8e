"""Child process of tests/test_binning_modes_gpu.py: renders a few scenes with the binning mode selected by the
environment (the library reads its B200S_* switches once per process) and prints one JSON line of SHA-1 digests of
everything that must not depend on the mode: num_rendered, sorted list, ranges, n_contrib, image, radii."""
import hashlib
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import helpers as Hh  # noqa: E402


def digest(t: torch.Tensor) -> str:
    return hashlib.sha1(t.detach().contiguous().cpu().numpy().tobytes()).hexdigest()


if __name__ == "__main__":
    dev = torch.device("cuda:0")
    out_rows = {}
    for cfg_name, view in (("T1", 0), ("T1", 3), ("C1", 0), ("C2", 1)):
        cfg, sc, cam = Hh.scene_on(cfg_name, dev, view=view)
        bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
        args, out, state = Hh.ours_forward(cfg, sc, cam, bg, names=["point_list", "ranges", "n_contrib"])
        out_rows[f"{cfg_name}/{view}"] = {"num_rendered": int(out[0]), "image": digest(out[1]), "radii": digest(out[2]),
                                          **{k: digest(v) for k, v in state.items()}}
    print(json.dumps(out_rows))
