"""Multi-view training exchange on TWO real GPUs (NCCL + symmetric memory): the gradients after the step's last backward()
and end_step() with the fused exchange enabled must equal the sum of all views' single-view gradients, for one and two
views per rank.  Skipped with fewer than 2 GPUs (bench.py runs the same check before its timed region at N > 1)."""
import os
import sys
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
        mod = Hh.my_module()
        dev = torch.device("cuda", rank)
        errs = {}
        # bench.exchange_check: every rank computes every view locally (no exchange) -> the expected sum, then runs its own
        # views through the fused exchange twice.  One and two views per rank, both transports, odd P (block alignment).
        for mode, chunks in (("allgather", 1), ("allgather", 4), ("peer", 1), ("multicast", 1)):     # 4: the pipelined ranges (gsr_multiview_sh_grad_range)
            os.environ["GSR_MULTIVIEW_MODE"] = mode
            os.environ["GSR_MV_AG_CHUNKS"] = str(chunks)
            for vpr in (1, 2):
                r = B.exchange_check(mod, dev, rank, world, vpr, P=40001)
                errs["%s%d/vpr%d" % (mode, chunks, vpr)] = r["max_rel_l2"]
                if rank == 0 and r["transport"] != mode:
                    print("test_multiview_2gpu: mode %s ran as %s on this box" % (mode, r["transport"]), file=sys.stderr)
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
