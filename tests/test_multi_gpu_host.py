"""Host-side logic of the one-process-per-GPU partitioning (gsr_multi_gpu.py), on CPU with gloo, world size 2."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers as Hh  # noqa: F401  (sys.path)
import gsr_multi_gpu as MG


def test_split_even():
    assert MG.split_tile_rows(68, 8) == [(0, 9), (9, 18), (18, 27), (27, 36), (36, 44), (44, 52), (52, 60), (60, 68)]
    assert MG.split_tile_rows(135, 1) == [(0, 135)]
    assert MG.split_tile_rows(3, 8)[3:] == [(3, 3)] * 5
    assert MG.split_tile_rows(0, 2) == [(0, 0), (0, 0)]


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_split_covers_exactly_once(world):
    rng = np.random.default_rng(world)
    for n in (1, 7, 68, 135):
        for w in (None, rng.uniform(0, 1, n), np.r_[np.zeros(n // 2), np.ones(n - n // 2)]):
            bands = MG.split_tile_rows(n, world, w)
            assert len(bands) == world and bands[0][0] == 0 and bands[-1][1] == n
            assert all(b[1] >= b[0] for b in bands)
            assert all(bands[i][1] == bands[i + 1][0] for i in range(world - 1))


def test_weighted_split_balances_instances():
    w = np.r_[np.full(34, 1.0), np.full(34, 3.0)]
    bands = MG.split_tile_rows(68, 4, w)
    loads = [w[b:e].sum() for b, e in bands]
    assert max(loads) / (w.sum() / 4) < 1.1


def test_band_pixel_rows_clip_to_image():
    assert MG.band_pixel_rows((60, 68), 1080) == (960, 1080)
    assert MG.band_pixel_rows((0, 9), 1080) == (0, 144)
    assert MG.tile_rows(1080) == 68 and MG.tile_rows(2160) == 135 and MG.tile_rows(256) == 16


def test_cameras_for_rank():
    for P in (0, 1, 127, 128, 129, 4096, 40001, 3000000):
        for n in (1, 3, 4):
            ch = MG.exchange_chunks(P, n)
            assert ch[0][0] == 0 and ch[-1][1] == P and all(b % 128 == 0 for b, _ in ch) and len(ch) <= max(n, 1)
            assert all(ch[i][1] == ch[i + 1][0] for i in range(len(ch) - 1))
    assert MG.cameras_for_rank(8, 1, 4) == [1, 5]
    assert sorted(sum((MG.cameras_for_rank(8, r, 8) for r in range(8)), [])) == list(range(8))


def test_flat_span_detects_one_arena():
    flat = torch.arange(40, dtype=torch.float32)
    a, b = flat[0:12].view(4, 3), flat[16:32].view(4, 4)
    span = MG.flat_span([a, b])
    assert span is not None and span.numel() == 32 and span.data_ptr() == flat.data_ptr()
    assert MG.flat_span([a, torch.zeros(3)]) is None
    assert MG.flat_span([a.t()]) is None
    assert MG.flat_span([]) is None


def test_banded_settings_keeps_the_reference_fields():
    from diff_gaussian_rasterization import GaussianRasterizationSettings as GS
    s = GS(32, 48, 1.0, 1.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 3, torch.zeros(3), False, False)
    b = MG.banded_settings(s, (1, 2))
    assert tuple(b)[:12] == tuple(s) or all(x is y or x == y for x, y in zip(tuple(b)[:12], tuple(s)) if not torch.is_tensor(x))
    assert b.tile_rows == (1, 2) and b.image_width == 48


def test_exchange_blocks_stay_16_byte_aligned_for_any_P():
    """One view's block is campos (4 floats) + dL/dRGB (3P floats), padded: consecutive blocks of an all-gather must start on
    16-byte boundaries whatever P is (the SH kernel reads them with 16-byte loads; P changes at every densification)."""
    for P in list(range(0, 70)) + [40000, 40001, 40002, 40003, 2999999, 3000000, 3000001]:
        n = MG.exchange_block_floats(P)
        assert n % 4 == 0 and 4 + 3 * P <= n < 4 + 3 * P + 4
    for P in (0, 1, 127, 128, 129, 4096, 40001, 3000000):
        for n in (1, 3, 4):
            ch = MG.exchange_chunks(P, n)
            assert ch[0][0] == 0 and ch[-1][1] == P and all(b % 128 == 0 for b, _ in ch) and len(ch) <= max(n, 1)
            assert all(ch[i][1] == ch[i + 1][0] for i in range(len(ch) - 1))
    assert MG.cameras_for_rank(8, 1, 4) == [1, 5] and MG.cameras_for_rank(8, 7, 8) == [7] and MG.cameras_for_rank(8, 0, 1) == list(range(8))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # (1) multi-view gradient sum over ONE flat arena
        flat = torch.full((40,), float(rank + 1))
        grads = [flat[0:12].view(4, 3), flat[12:16].view(4, 1), flat[16:40].view(4, 6)]
        n = MG.allreduce_gradients(grads)
        ok1 = n == 1 and bool((flat == 3.0).all())
        # separate tensors: one collective each
        sep = [torch.full((4, 3), float(rank)), torch.full((2,), 1.0)]
        n2 = MG.allreduce_gradients(sep, average=True)
        ok2 = n2 == 2 and bool((sep[0] == 0.5).all()) and bool((sep[1] == 1.0).all())
        # (2) radii max
        r = MG.allreduce_radii(torch.tensor([rank, 5 - rank, 0], dtype=torch.int32))
        ok3 = r.tolist() == [1, 5, 0]
        # (3) band exchange with ragged band heights: H = 40 -> 3 tile rows -> bands (0,2),(2,3) = 32 + 8 pixel rows
        H, W = 40, 6
        bands = MG.split_tile_rows(MG.tile_rows(H), world)
        full = torch.arange(3 * H * W, dtype=torch.float32).view(3, H, W)
        p0, p1 = MG.band_pixel_rows(bands[rank], H)
        img = MG.gather_bands(full[:, p0:p1].contiguous(), bands, H)
        ok4 = bool((img == full).all())
        # (4) densification statistics of multi-view training: per-rank accumulators of per-view norms / counts are summed,
        #     max_radii2D max-reduced, so every rank takes the same clone / split / prune decisions
        import types
        m = types.SimpleNamespace(xyz_gradient_accum=torch.tensor([[1.0 + rank], [0.0], [2.0]]), denom=torch.tensor([[1.0], [0.0], [float(rank)]]),
                                  max_radii2D=torch.tensor([3.0 * rank, 7.0 - rank, 0.0]))
        MG.MultiViewExchange.reduce_densification_stats(types.SimpleNamespace(group=None), m)
        ok5 = m.xyz_gradient_accum.flatten().tolist() == [3.0, 0.0, 4.0] and m.denom.flatten().tolist() == [2.0, 0.0, 1.0] \
            and m.max_radii2D.tolist() == [3.0, 7.0, 0.0]
        q.put((rank, ok1, ok2, ok3, ok4, ok5))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    res = [q.get(timeout=10) for _ in procs]
    for r in res:
        assert all(r[1:]), r
