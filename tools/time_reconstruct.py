"""Time b200s_sh_grad_reconstruct alone (P Gaussians, `views` factor blocks at 43 % visibility each)."""
import ctypes as C
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from neuralgaussiansplatting_b200 import _lib  # noqa: E402

lib = _lib.load()
dev = torch.device("cuda:0")
P = 3_000_000
slot = -(-(3 * P + 4) // 64) * 64
means = torch.randn(P, 3, device=dev)
out = torch.empty(P, 16, 3, device=dev)
res = {}
for views in (1, 2, 4, 8):
    f = torch.zeros(views * slot, device=dev)
    for v in range(views):
        g = torch.randn(P, 3, device=dev)
        g[torch.rand(P, device=dev) > 0.43] = 0
        f[v * slot:v * slot + 3 * P] = g.reshape(-1)
        f[v * slot + 3 * P:v * slot + 3 * P + 3] = torch.randn(3, device=dev) * 4
    a = _lib.ShGradArgs()
    a.P, a.D, a.M, a.views = P, 3, 16, views
    a.means3D, a.factors, a.stride_floats, a.dL_dsh = means.data_ptr(), f.data_ptr(), slot, out.data_ptr()
    st = torch.cuda.current_stream().cuda_stream
    for fast in (0, 1):
        a.fast = fast
        for _ in range(3):
            lib.b200s_sh_grad_reconstruct(C.byref(a), st)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            lib.b200s_sh_grad_reconstruct(C.byref(a), st)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        nbytes = P * (12 * views + 12 + 192)
        res[f"views{views}_{'fast' if fast else 'exact'}"] = {"ms": round(ms, 4), "GBps": round(nbytes / ms / 1e6, 1)}
print(json.dumps(res))
