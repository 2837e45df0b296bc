#!/usr/bin/env python3
"""Phase timing of the multi-view gradient exchange under torchrun (diagnostic, not a bench):
barrier / NCCL all-reduce of the small block / fused SH kernel / each alone and combined, plus a plain all-reduce
of the whole live arena for comparison.  Prints one JSON line from rank 0 (max over ranks per phase)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting_cc-comments_b200")):
    sys.path.insert(0, p)
import numpy as np
import torch
import torch.distributed as dist


def main():
    os.environ.setdefault("NCCL_DEBUG", "WARN")
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import diff_gaussian_rasterization as dgr
    import gsr_multi_gpu as MG
    P, M = 3_000_000, 16
    g = torch.Generator(device="cpu").manual_seed(rank)
    means = torch.randn(P, 3, generator=g).to(dev) + torch.tensor([0, 0, 5.0], device=dev)
    dsh = torch.empty(P, M, 3, device=dev)
    small = torch.randn(14 * P, device=dev)
    arena = torch.randn(62 * P, device=dev)
    filler = torch.randn(120_000_000, device=dev)      # stands in for the per-Gaussian backward on the main stream (HBM bound)

    def timeit(fn, n=15):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / n], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return round(float(t.item()), 4)

    C = dgr._C
    res = {"world": world}
    res["allreduce_small_168MB"] = timeit(lambda: dist.all_reduce(small))
    res["allreduce_arena_744MB"] = timeit(lambda: dist.all_reduce(arena))
    res["filler_mul"] = timeit(lambda: filler.mul_(1.0001))
    for mode in ("peer", "allgather"):
        ex = MG.MultiViewExchange(P, dev, mode=mode)
        ex.block[4:4 + 3 * P] = torch.randn(3 * P, generator=g).to(dev)
        ex.block[:3] = torch.tensor([0.1 * rank, 0.0, -0.5], device=dev)
        n = 4 + 3 * P
        if mode == "peer":
            res["sh_kernel_local_only"] = timeit(lambda: C.multiview_sh_grad([ex.block.data_ptr()] * world, means, dsh, 3))
            rot = [ex.peer_ptrs[(rank + 1 + i) % world] for i in range(world)]
            res["sh_kernel_peers_same_order"] = timeit(lambda: C.multiview_sh_grad(ex.peer_ptrs, means, dsh, 3))
            res["sh_kernel_peers_staggered"] = timeit(lambda: C.multiview_sh_grad(rot, means, dsh, 3))
        else:
            res["allgather_36MB_per_rank"] = timeit(lambda: dist.all_gather_into_tensor(ex.stage[: world * n], ex.block[:n]))

        def combined():
            ex.ev_drgb.record()
            ex.exchange(C, means, 3, dsh, [small])

        def combined_with_work():
            ex.ev_drgb.record()
            filler.mul_(1.0001)
            ex.exchange(C, means, 3, dsh, [small])
        res["exchange_%s" % mode] = timeit(combined)
        res["exchange_%s_beside_filler" % mode] = timeit(combined_with_work)
        del ex
    if rank == 0:
        print(json.dumps(res), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
