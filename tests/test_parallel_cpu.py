"""world_size-2 gloo tests of the view-parallel plumbing (parallel.py): view sharding, the bucketed gradient
all-reduce, replica consistency, densification-statistics reduction."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from neuralgaussiansplatting_b200 import parallel as par


def test_shard_views_round_robin():
    assert par.shard_views(8, 0, 8) == [0] and par.shard_views(8, 7, 8) == [7]
    assert par.shard_views(64, 3, 8) == list(range(3, 64, 8))
    assert par.shard_views(3, 3, 4) == []
    got = sorted(v for r in range(4) for v in par.shard_views(10, r, 4))
    assert got == list(range(10))
    assert all(par.owner_of_view(v, 4) == r for r in range(4) for v in par.shard_views(10, r, 4))
    with pytest.raises(ValueError):
        par.shard_views(8, 8, 8)


def test_payload_matches_survey_figure():
    assert par.allreduce_payload_bytes(3_000_000) == 708_000_000      # 236 B/Gaussian
    assert par.GRAD_FLOATS_PER_GAUSSIAN * 4 == 236


def test_buckets_cover_every_element_once():
    P = 1000
    g = dict(means3D=torch.zeros(P, 3), shs=torch.zeros(P, 16, 3), opacities=torch.zeros(P, 1),
             scales=torch.zeros(P, 3), rotations=torch.zeros(P, 4))
    ar = par.GradientAllReduce(max_bucket_bytes=40_000)
    bks = ar.buckets(g)
    assert sum(b.numel() for b in bks) == P * 59
    assert max(b.numel() * 4 for b in bks) <= 40_000 + 16 * 3 * 4 * 8
    for b in bks:
        b += 1.0
    assert all(float(v.min()) == 1.0 and float(v.max()) == 1.0 for v in g.values())   # views, not copies


def test_gradient_layout_is_aligned_and_disjoint():
    for P in (1, 7, 257, 10_007, 3_000_000):
        lay, total = par.gradient_layout(P, 16)
        assert total % 64 == 0 and set(lay) == set(par.PARAM_ORDER)
        spans = []
        for name, (off, shape) in lay.items():
            n = 1
            for d in shape:
                n *= d
            assert off % 64 == 0                       # 256-byte aligned sub-tensors (16-byte vector stores)
            spans.append((off, off + n))
        spans.sort()
        assert all(a[1] <= b[0] for a, b in zip(spans, spans[1:])) and spans[-1][1] <= total
        assert sum(b - a for a, b in spans) == P * par.GRAD_FLOATS_PER_GAUSSIAN
    lay, total = par.gradient_layout(3_000_000)
    assert total * 4 == par.allreduce_payload_bytes(3_000_000)     # no padding at the bench size


def test_peer_collective_partitions_cover_every_element_exactly_once():
    from neuralgaussiansplatting_b200 import _lib
    lib = _lib.load()
    for world in (2, 3, 4, 8):
        for n4 in (0, 1, 7, 8, 1000, 44_250_000):
            sl = par.peer_owner_slices(n4, world)
            assert sl[0][0] == 0 and sl[-1][1] == n4
            assert all(a[1] == b[0] for a, b in zip(sl, sl[1:])) and all(b >= a for a, b in sl)
        for rows in (1, 1023, 1024, 1025, 10_007, 3_000_000):
            rows_per = lib.b200s_peer_allreduce_rows_per_rank(rows, world)
            assert rows_per % 1024 == 0 and rows_per * world >= rows and rows_per == -(-(-(-rows // world)) // 1024) * 1024
            assert lib.b200s_peer_allreduce_vis_bytes(rows, world) == rows_per * world
            for ctas in (1, 8, 16):
                spans = sorted(sp for lst in par.peer_sparse_chunks(rows, world, ctas).values() for sp in lst)
                assert spans[0][0] == 0 and spans[-1][1] == rows
                assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))          # no gap, no overlap
    lay, total = par.gradient_layout(3_000_000)
    assert lay["shs"][0] % 4 == 0 and lay["shs"][0] + 3_000_000 * 48 == total      # sparse tail = last tensor


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        P = 257
        torch.manual_seed(0)                      # identical replicas
        params = dict(means3D=torch.randn(P, 3), shs=torch.randn(P, 16, 3), opacities=torch.randn(P, 1),
                      scales=torch.randn(P, 3), rotations=torch.randn(P, 4))
        assert par.replicas_in_sync(params.values())
        # each rank "renders" its own views: deterministic per-view gradients, accumulated locally
        views = par.shard_views(5, rank, world)
        total = {}
        for v in views:
            gen = torch.Generator().manual_seed(100 + v)
            part = {k: torch.randn(p.shape, generator=gen) for k, p in params.items()}
            par.accumulate_(total, part)
        ar = par.GradientAllReduce(max_bucket_bytes=20_000)
        ar.start(total)
        ar.wait()
        # expected: sum over all 5 views
        exp = {}
        for v in range(5):
            gen = torch.Generator().manual_seed(100 + v)
            par.accumulate_(exp, {k: torch.randn(p.shape, generator=gen) for k, p in params.items()})
        ok = all(torch.allclose(total[k], exp[k], atol=1e-5) for k in params)
        for k in params:                          # identical SGD step on every rank
            params[k] -= 0.01 * total[k]
        ok = ok and par.replicas_in_sync(params.values())
        st = par.DensificationStats(P, "cpu")
        vis = torch.zeros(P, dtype=torch.bool); vis[rank::world] = True
        st.add_view(torch.ones(P, 3), vis)
        st.allreduce()
        ok = ok and float(st.denom.sum()) == P and abs(float(st.accum.sum()) - P * 2 ** 0.5) < 1e-3
        # single-call flat-buffer all-reduce: gradients written into views of the flat buffer (no copy), strays copied
        flat = par.FlatGradientAllReduce(P, 16, "cpu")
        for step in range(3):                     # buffers rotate
            views_ = flat.grad_out()
            for k, t in views_.items():
                t.fill_(float(rank + 1) * (step + 1))
            stray = torch.full((P, 3), float(rank + 1))
            flat.start({**views_, "scales": stray})
            flat.wait()
            want = sum(r + 1 for r in range(world)) * (step + 1)
            ok = ok and all(float(t.min()) == want and float(t.max()) == want
                            for k, t in views_.items() if k != "scales")
            ok = ok and float(stray.min()) == sum(r + 1 for r in range(world))
        params["means3D"][0, 0] += float(rank)    # deliberately diverge -> must be detected
        ok = ok and not par.replicas_in_sync(params.values())
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


def test_gradient_allreduce_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(res) == [(0, True), (1, True)]


def test_sh_gradient_of_a_view_is_rank_one():
    """The identity the factored collective rests on (csrc/sh_reconstruct.cu): for one view,
    d(colour . w)/d sh[k][c] = basis_k(dir) * w[c], so 3 floats per Gaussian and view determine the 48-float row and
    the sum over views can be rebuilt from the gathered factors. Checked with autograd through the reference's own
    eval_sh polynomial (utils/sh_utils.py:57-112 as restated in sh_utils.py)."""
    from neuralgaussiansplatting_b200.sh_utils import eval_sh
    g = torch.Generator().manual_seed(5)
    P, V = 64, 3
    means = torch.randn(P, 3, generator=g, dtype=torch.float64)
    total_dense = torch.zeros(P, 16, 3, dtype=torch.float64)
    total_fact = torch.zeros(P, 16, 3, dtype=torch.float64)
    for v in range(V):
        campos = torch.randn(3, generator=g, dtype=torch.float64) * 4
        w = torch.randn(P, 3, generator=g, dtype=torch.float64)
        d = means - campos
        d = d / d.norm(dim=1, keepdim=True)
        sh = torch.randn(P, 16, 3, generator=g, dtype=torch.float64, requires_grad=True)
        col = eval_sh(3, sh.transpose(1, 2), d)
        (col * w).sum().backward()
        total_dense += sh.grad
        # basis_k(dir) = d colour_c / d sh[k][c] for any c: read it off with a one-hot coefficient tensor
        eye = torch.eye(16, dtype=torch.float64)[None, :, :, None].expand(P, 16, 16, 1)    # [P, k, coeff, 1]
        basis = torch.stack([eval_sh(3, eye[:, k].expand(P, 16, 3).transpose(1, 2), d)[:, 0] for k in range(16)], dim=1)
        total_fact += basis[:, :, None] * w[:, None, :]
    assert torch.allclose(total_dense, total_fact, atol=1e-12)


def test_factored_layout_of_the_peer_buffer():
    """Gather slots sit behind the five tensors, 256-byte aligned, one per rank, each holding 3P factors + the camera
    centre; the dense prefix the collective sums is everything before dL/dsh."""
    for P in (1, 257, 10_007, 3_000_000):
        lay, total = par.gradient_layout(P, 16)
        slot = -(-(3 * P + 4) // 64) * 64
        assert slot % 64 == 0 and slot >= 3 * P + 4
        assert lay["shs"][0] % 64 == 0 and lay["shs"][0] >= 11 * P
        assert lay["shs"][0] + 48 * P <= total
    # NVLink bytes per GPU and direction at the bench size, world 8: dense two-shot vs factored
    P, W = 3_000_000, 8
    dense = 2 * (W - 1) / W * 236 * P
    fact = 2 * (W - 1) / W * 44 * P + (W - 1) * 12 * P
    assert dense / fact > 2.5
