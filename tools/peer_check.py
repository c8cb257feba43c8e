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
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    report = {"world": world}

    # ---- exactness on a small problem (all_gather of the inputs is affordable)
    small = parallel.PeerGradientAllReduce(10_007, 16, dev, ctas=4, backend=a.backend)
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
    for mm in (False, "sparse", True):
        for ctas in (8, 16, 32):
            red = parallel.PeerGradientAllReduce(a.P, 16, dev, ctas=ctas, multimem=(mm is True), n_buffers=1,
                                                 backend=a.backend)
            if mm is True and not red.multimem:
                del red
                break
            views = red.grad_out()
            rad = None
            if mm == "sparse":
                rad = radii43
                for v in views.values():
                    v[rad <= 0] = 0.0

            def fn():
                red.start(None, radii=rad); red.wait()
            ms = time_it(fn, a.iters)
            times[f"{'multimem' if mm is True else 'sparse43' if mm else 'p2p'}_ctas{ctas}"] = {"ms": round(ms, 4), "algbw_GBps": round(nbytes / ms / 1e6, 1)}
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
