"""torchrun --nproc-per-node N tools/peer_check.py : correctness + timing of b200s_peer_allreduce_sum_f32 against NCCL.

Exact check: every element of the peer-kernel result equals the fixed-order (rank 0, 1, ..) fp32 sum computed
locally from an all_gather of the inputs; replicas are compared bit for bit."""
import argparse
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from neuralgaussiansplatting_b200 import parallel  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--P", type=int, default=3_000_000)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--backend", default="auto")
    ap.add_argument("--quick", action="store_true", help="exactness checks only")
    ap.add_argument("--only-factored", action="store_true", help="time the factored variants only")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    report = {"world": world}

    # ---- exactness on a small problem (all_gather of the inputs is affordable)
    small = parallel.PeerGradientAllReduce(10_007, 16, dev, ctas=4, backend=a.backend, factored=False)
    report["backend"] = small.mem.backend
    report["multicast"] = bool(small.mem.multicast_ptr)
    for trial in range(3):
        views = small.grad_out()
        g = torch.Generator(device=dev).manual_seed(100 * trial + rank)
        for v in views.values():
            v.copy_(torch.randn(v.shape, device=dev, generator=g))
        flat = small._flat[small._cur]
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat.clone())
        expect = gathered[0].clone()
        for r in range(1, world):
            expect += gathered[r]
        small.start(views)
        small.wait()
        torch.cuda.synchronize()
        exact = bool(torch.equal(flat, expect))
        allf = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(allf, flat)
        same = all(torch.equal(allf[0], x) for x in allf[1:])
        report[f"small_trial{trial}"] = {"exact_fixed_order_sum": exact, "replicas_identical": same}
        assert exact and same, report
    # visibility-sparse tail: rows with radii <= 0 are zero on that rank; result must equal the dense fixed-order sum
    for trial in range(3):
        views = small.grad_out()
        g = torch.Generator(device=dev).manual_seed(1000 + 100 * trial + rank)
        radii = (torch.rand(10_007, device=dev, generator=g) < (0.43 if trial < 2 else 0.02)).to(torch.int32) * 7
        for v in views.values():
            v.copy_(torch.randn(v.shape, device=dev, generator=g))
            v[radii <= 0] = 0.0
        flat = small._flat[small._cur]
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat.clone())
        expect = gathered[0].clone()
        for r in range(1, world):
            expect += gathered[r]
        small.start(views, radii=radii)
        small.wait()
        torch.cuda.synchronize()
        exact = bool(torch.equal(flat, expect))
        allf = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(allf, flat)
        same = all(torch.equal(allf[0], x) for x in allf[1:])
        report[f"sparse_trial{trial}"] = {"equals_dense_fixed_order_sum": exact, "replicas_identical": same}
        assert exact and same, report
    # stray (non-aliasing) gradients are copied in and back
    stray = {k: torch.full(v.shape, float(rank + 1), device=dev) for k, v in small._views(0).items()}
    small.grad_out()
    small.start(stray)
    small.wait()
    torch.cuda.synchronize()
    tot = float(sum(range(1, world + 1)))
    assert all(bool((v == tot).all()) for v in stray.values())
    report["stray_copy_path"] = True
    del small
    # ---- factored SH gradient: 11 dense floats summed, 3-float factors gathered, dL/dsh rebuilt locally. Expected
    # rows: each rank's own dense row block (b200s_sh_grad_reconstruct on ITS slot alone == what b200s_backward writes,
    # tests/test_factored_sh_gpu.py), all-gathered and summed in rank order with torch.
    import ctypes as C
    from neuralgaussiansplatting_b200 import _lib
    lib = _lib.load()
    Pf = 10_007
    fact = parallel.PeerGradientAllReduce(Pf, 16, dev, ctas=4, backend=a.backend, factored=True)
    gm = torch.Generator(device=dev).manual_seed(4242)                 # replicated positions: same seed everywhere
    means = torch.randn(Pf, 3, device=dev, generator=gm) * 2.0
    for trial in range(3):
        views = fact.grad_out()
        g = torch.Generator(device=dev).manual_seed(5000 + 100 * trial + rank)
        for k, v in views.items():
            if k not in ("shs", "sh_factors"):
                v.copy_(torch.randn(v.shape, device=dev, generator=g))
        views["shs"].fill_(float("nan"))                                   # must be fully overwritten
        slot = views["sh_factors"]
        rgb = torch.randn(Pf, 3, device=dev, generator=g)
        rgb[torch.rand(Pf, device=dev, generator=g) < 0.57] = 0.0          # culled in this rank's view
        slot[:3 * Pf].copy_(rgb.reshape(-1))
        slot[3 * Pf:3 * Pf + 3].copy_(torch.randn(3, device=dev, generator=g) * 4.0)
        own_rows = torch.empty(Pf, 16, 3, device=dev)
        sg = _lib.ShGradArgs()
        sg.P, sg.D, sg.M, sg.views = Pf, 3, 16, 1
        sg.means3D, sg.factors, sg.stride_floats, sg.dL_dsh = means.data_ptr(), slot.data_ptr(), fact.slot_floats, own_rows.data_ptr()
        _lib.check(lib.b200s_sh_grad_reconstruct(C.byref(sg), torch.cuda.current_stream().cuda_stream), "reconstruct")
        expect = {}
        for k, v in list(views.items()) + [("shs_rows", own_rows)]:
            if k in ("shs", "sh_factors"):
                continue
            parts = [torch.empty_like(v) for _ in range(world)]
            dist.all_gather(parts, v.clone())
            e = parts[0].clone()
            for q in parts[1:]:
                e += q
            expect[k] = e
        fact.start(views, means3D=means, sh_degree=3)
        fact.wait()
        torch.cuda.synchronize()
        exact = all(bool(torch.equal(views[k], expect[k])) for k in expect if k != "shs_rows")
        exact = exact and bool(torch.equal(views["shs"], expect["shs_rows"]))
        chk = torch.stack([views[k].double().sum() for k in ("means3D", "opacities", "scales", "rotations", "shs")])
        allc = [torch.empty_like(chk) for _ in range(world)]
        dist.all_gather(allc, chk)
        same = all(torch.equal(allc[0], x) for x in allc[1:])
        report[f"factored_trial{trial}"] = {"exact_fixed_order_sum": exact, "replicas_identical": same}
        assert exact and same, report
    del fact
    if a.quick:
        if rank == 0:
            print(json.dumps(report))
        dist.destroy_process_group()
        return

    # ---- timing at the bench payload
    def time_it(fn, iters):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record(); torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    layout, n_floats = parallel.gradient_layout(a.P)
    nbytes = n_floats * 4
    times = {}
    g = torch.Generator(device=dev).manual_seed(7 + rank)
    radii43 = (torch.rand(a.P, device=dev, generator=g) < 0.43).to(torch.int32)
    means_big = torch.randn(a.P, 3, device=dev, generator=torch.Generator(device=dev).manual_seed(1))
    variants = [(False, True, True), (False, True, False), (True, True, True)]
    if not a.only_factored:
        variants += [(False, False, False), ("sparse", False, False), (True, False, False)]
    for mm, fct, push in variants:
        for ctas in (4, 8, 16, 32):
            red = parallel.PeerGradientAllReduce(a.P, 16, dev, ctas=ctas, multimem=(mm is True), n_buffers=1,
                                                 backend=a.backend, factored=fct, gather_push=push)
            if mm is True and not red.multimem:
                del red
                break
            views = red.grad_out()
            rad = None
            if mm == "sparse":
                rad = radii43
                for v in views.values():
                    v[rad <= 0] = 0.0
            if fct:
                views["sh_factors"].zero_()
                views["sh_factors"][:3 * a.P].copy_(torch.randn(3 * a.P, device=dev))
                views["sh_factors"][:3 * a.P].view(a.P, 3)[radii43 <= 0] = 0.0     # 43 % of the rows visible per view

            def fn():
                if fct:
                    red._factors_fresh = True
                    red.start(None, means3D=means_big, sh_degree=3)
                else:
                    red.start(None, radii=rad)
                red.wait()
            ms = time_it(fn, a.iters)
            name = ("factored_" + ("push_" if red.gather_push else "pull_") if fct else "") + (
                "multimem" if mm is True else "sparse43" if mm else "p2p")
            times[f"{name}_ctas{ctas}"] = {"ms": round(ms, 4), "algbw_GBps": round(nbytes / ms / 1e6, 1)}
            del red
            torch.cuda.empty_cache()
    x = torch.zeros(n_floats, device=dev)
    ms = time_it(lambda: dist.all_reduce(x), a.iters)
    times["nccl_single_call"] = {"ms": round(ms, 4), "algbw_GBps": round(nbytes / ms / 1e6, 1)}
    report["payload_MB"] = round(nbytes / 1e6, 1)
    report["times"] = times
    if rank == 0:
        print(json.dumps(report))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
