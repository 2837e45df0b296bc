"""Multi-view training exchange on TWO real GPUs (NCCL + symmetric memory): gradients returned by backward() with the
fused exchange enabled must equal the sum of the two ranks' single-view gradients.  Skipped with fewer than 2 GPUs."""
import os
import socket

import numpy as np
import pytest

import helpers as Hh
from helpers import S

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), NCCL_DEBUG="WARN")
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import bench as B
        import gsr_multi_gpu as MG
        mod = Hh.my_module()
        P, W, H = 40000, 320, 200
        scene = S.make_scene(P, S.make_camera(W, H), sh_degree=3, M=16, seed=4)
        cams = [B.ring_camera(S, W, H, r) for r in range(world)]
        dLs = [S.upstream_grad(c, seed=20 + i) for i, c in enumerate(cams)]
        dev = "cuda:%d" % rank
        # every rank computes every view locally (no exchange) -> the expected sum
        expect = None
        for c, dL in zip(cams, dLs):
            g = Hh.run_torch(mod, c, scene, dL, bg=(0.1, 0.2, 0.3), device=dev)["grads"]
            expect = g if expect is None else {k: (expect[k] + g[k] if g[k] is not None else None) for k in g}
        errs = {}
        for mode in ("allgather", "peer"):
            MG.enable_multiview_training(P, dev, mode=mode)
            got = Hh.run_torch(mod, cams[rank], scene, dLs[rank], bg=(0.1, 0.2, 0.3), device=dev)["grads"]
            MG.disable_multiview_training()
            errs.update({mode + "/" + k: float(Hh.rel_l2(got[k], expect[k]))
                         for k in ("means3D", "means2D", "shs", "opacities", "scales", "rotations")})
        q.put((rank, errs))
    finally:
        dist.destroy_process_group()


def test_exchange_equals_sum_of_views():
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    for _ in procs:
        rank, errs = q.get(timeout=10)
        for k, e in errs.items():
            assert e <= 1e-5, (rank, k, e)
