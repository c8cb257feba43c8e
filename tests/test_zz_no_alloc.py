"""b200s_idxmap_forward_static: the idxmap forward for hosts that have no workspace to pass.

Who needs it: a render loop that lives inside an application whose device memory already belongs, all of it, to
some other allocator (a fixed arena, a framework pool capped at a memory fraction), or one that must not stall on
cudaMalloc's device-wide synchronisation. Such a host cannot produce the 8-byte-per-pixel z-buffer the regular
b200s_idxmap_forward asks for, so this variant keeps that z-buffer inside the library (one array per GPU, views of
at most B200S_IDXMAP_STATIC_MAX_PIXELS pixels).

What is checked, straight through the C ABI (ctypes, raw device pointers): the variant writes the same maps as the
workspace entry point on two calls in a row. The second call renders a DIFFERENT, smaller view, so anything the
first call left in the shared array and the second failed to clear would show up as wrong winners. Index, depth
and the gathered channels are compared bit for bit, the encoded channels within the 1e-5 the other idxmap tests
use. A view beyond the limit has to be refused with B200S_E_WORKSPACE and leave the outputs alone.

The file name sorts last on purpose: `pytest -x` stops at the first failure, so this test must not be able to hide
the parity tests. It passes alone in a fresh process (the first call on the device is then the one that looks up
the array's address) and after the rest of the suite."""
import ctypes as C
import re
from pathlib import Path

import pytest
import torch

import helpers as Hh
from neuralgaussiansplatting_b200 import _lib

pytestmark = pytest.mark.gpu

HEADER = (Path(__file__).resolve().parent.parent / "include" / "b200splat.h").read_text()
MAX_PIXELS = int(re.search(r"#define B200S_IDXMAP_STATIC_MAX_PIXELS (\d+)", HEADER).group(1))


def _run(lib, entry, sc, cam, W, H, dev, with_ws):
    """One idxmap forward through `entry`; every output starts from a sentinel so unwritten elements are seen."""
    P = sc["means3D"].size(0)
    f32 = dict(dtype=torch.float32, device=dev)
    out = {
        "idx": torch.full((H, W), -7, dtype=torch.int32, device=dev),
        "color": torch.full((H, W, 3), -7.0, **f32),
        "depth": torch.full((H, W), -7.0, **f32),
        "feature": torch.full((H, W, 64), -7.0, **f32),
        "radii": torch.full((P,), -7, dtype=torch.int32, device=dev),
        "num_rendered": torch.full((1,), -7, dtype=torch.int32, device=dev),
    }
    keep = [t.contiguous() for t in (sc["means3D"], cam.world_view_transform, cam.full_proj_transform,
                                     cam.camera_center, sc["feature_vector"])]
    a = _lib.IdxmapArgs()
    a.P, a.W, a.H, a.debug = P, W, H, 0
    a.means3D, a.viewmatrix, a.projmatrix, a.campos, a.feature_vector = [t.data_ptr() for t in keep]
    a.out_idx, a.out_color, a.out_depth, a.out_feature = (out["idx"].data_ptr(), out["color"].data_ptr(),
                                                          out["depth"].data_ptr(), out["feature"].data_ptr())
    a.radii, a.num_rendered_dev = out["radii"].data_ptr(), out["num_rendered"].data_ptr()
    ws = None
    if with_ws:
        nbytes = lib.b200s_idxmap_workspace_size(P, W, H)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        a.ws, a.ws_bytes = ws.data_ptr(), nbytes
    else:
        a.ws, a.ws_bytes = None, 0          # nothing to give: that is the point of the variant
    rc = entry(C.byref(a), torch.cuda.current_stream(dev).cuda_stream)
    torch.cuda.synchronize(dev)
    del keep, ws
    return rc, out


def _same(static, regular):
    assert torch.equal(static["idx"], regular["idx"])
    assert torch.equal(static["depth"].view(torch.int32), regular["depth"].view(torch.int32))
    assert torch.equal(static["radii"], regular["radii"])
    assert torch.equal(static["num_rendered"], regular["num_rendered"])
    assert float((static["color"] - regular["color"]).abs().max()) < 1e-6
    assert float((static["feature"][..., :25] - regular["feature"][..., :25]).abs().max()) < 1e-5
    assert torch.equal(static["feature"][..., 25:], regular["feature"][..., 25:])


def test_static_scratch_idxmap_equals_the_workspace_entry_point_twice_in_a_row(dev):
    lib = _lib.load()
    with torch.cuda.device(dev):
        # first call: T1 (400x304); second call right after it: T0 (160x96, another scene and camera), which lands
        # on pixels of the shared z-buffer the first call has just filled with its winners
        for cfg_name in ("T1", "T0"):
            cfg, sc, cam = Hh.scene_on(cfg_name, dev)
            assert cfg.width * cfg.height <= MAX_PIXELS
            rc, static = _run(lib, lib.b200s_idxmap_forward_static, sc, cam, cfg.width, cfg.height, dev, False)
            assert rc == 0, rc
            rc, regular = _run(lib, lib.b200s_idxmap_forward, sc, cam, cfg.width, cfg.height, dev, True)
            assert rc == 0, rc
            assert int((regular["idx"] >= 0).sum()) > 0 and int((regular["idx"] < 0).sum()) > 0   # hits and holes
            for k, t in static.items():
                assert not bool((t == -7).any()), f"{k}: elements left unwritten"
            _same(static, regular)


def test_static_scratch_idxmap_refuses_views_beyond_its_limit(dev):
    lib = _lib.load()
    with torch.cuda.device(dev):
        cfg, sc, cam = Hh.scene_on("T0", dev)
        W, H = 2048, MAX_PIXELS // 2048 + 1           # one row more than the array holds
        P = sc["means3D"].size(0)
        keep = [t.contiguous() for t in (sc["means3D"], cam.world_view_transform, cam.full_proj_transform,
                                         cam.camera_center, sc["feature_vector"])]
        # outputs of 4 bytes each stand in for the full maps: a refused call must not touch them at all
        outs = [torch.full((1,), -7, dtype=torch.int32, device=dev) for _ in range(6)]
        a = _lib.IdxmapArgs()
        a.P, a.W, a.H, a.debug = P, W, H, 0
        a.means3D, a.viewmatrix, a.projmatrix, a.campos, a.feature_vector = [t.data_ptr() for t in keep]
        a.out_idx, a.out_color, a.out_depth, a.out_feature, a.radii, a.num_rendered_dev = [t.data_ptr() for t in outs]
        rc = lib.b200s_idxmap_forward_static(C.byref(a), torch.cuda.current_stream(dev).cuda_stream)
        torch.cuda.synchronize(dev)
        assert rc == -2                                # B200S_E_WORKSPACE
        assert all(int(t.item()) == -7 for t in outs)
        with pytest.raises(RuntimeError, match="B200S_E_WORKSPACE"):
            _lib.check(rc, "b200s_idxmap_forward_static")
