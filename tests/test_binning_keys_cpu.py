"""Host-side statement of the two facts the (x, y) tile keys of csrc/binning.cu rest on, checked on the CPU oracle's
rectangles of a real synthetic view (oracle/oracle_cpu.c, test infrastructure):
  1. row << 8 | column orders instances exactly like the tile id row * grid_x + column (stable sort => same lists);
  2. the digit histograms of those keys are range-adds per Gaussian: a w x h rectangle puts h instances on each of its
     columns and w on each of its rows, so the difference arrays (+h at x0, -h at x0 + w; +w at y0, -w at y0 + h),
     integrated modulo 2^32 as sort_scan_hist_kernel does, equal the per-instance histograms."""
import numpy as np
import torch

import oracle_cpu
from neuralgaussiansplatting_b200 import synthetic as syn


def test_xy_keys_order_and_range_add_histograms():
    cfg = syn.CONFIGS["T1"]
    sc = syn.make_scene(cfg)
    cam = syn.make_camera(cfg, 0)
    o = oracle_cpu.render_scene(cfg, sc, cam, torch.zeros(3), as_torch=False, stop_after="preprocess")
    gx, gy = (cfg.width + 15) // 16, (cfg.height + 15) // 16
    assert gx <= 256 and gy <= 256 and gx * gy > 256            # the 2-pass case the keys are used for
    vis = np.nonzero(o["radii"] > 0)[0]
    rects = o["rects"][vis].astype(np.int64)                    # x0, y0, x1, y1 in tiles
    x0, y0, w, h = rects[:, 0], rects[:, 1], rects[:, 2] - rects[:, 0], rects[:, 3] - rects[:, 1]
    keep = (w > 0) & (h > 0)
    x0, y0, w, h = x0[keep], y0[keep], w[keep], h[keep]
    # instances in emission order: Gaussian-major, rows then columns (expand_kernel)
    xs = np.concatenate([np.tile(np.arange(a, a + ww), hh) for a, ww, hh in zip(x0, w, h)])
    ys = np.concatenate([np.repeat(np.arange(b, b + hh), ww) for b, ww, hh in zip(y0, w, h)])
    tile_id, xy_key = ys * gx + xs, (ys << 8) | xs
    assert (np.argsort(tile_id, kind="stable") == np.argsort(xy_key, kind="stable")).all()
    # range-add difference arrays, modulo 2^32 like the device code
    dx, dy = np.zeros(257, np.uint32), np.zeros(257, np.uint32)
    np.add.at(dx, x0, h.astype(np.uint32)); np.add.at(dx, x0 + w, (0 - h).astype(np.uint32))
    np.add.at(dy, y0, w.astype(np.uint32)); np.add.at(dy, y0 + h, (0 - w).astype(np.uint32))
    assert (np.cumsum(dx[:256], dtype=np.uint32) == np.bincount(xs, minlength=256).astype(np.uint32)).all()
    assert (np.cumsum(dy[:256], dtype=np.uint32) == np.bincount(ys, minlength=256).astype(np.uint32)).all()
