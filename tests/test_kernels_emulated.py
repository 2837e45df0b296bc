"""The Adam and k-NN CUDA kernels (csrc/gsr_adam.cu, csrc/gsr_knn.cu) compiled for the HOST through tests/cuda_emu/emu.h
(one OS thread per CUDA thread, real barriers, warp shuffles as rendezvous) and compared with the oracle bit for bit.

This is what stands in for a GPU in the build container: it exercises the kernels' indexing, barriers, reductions, box
pruning and ragged tails, not their speed or memory ordering.  The `-m gpu` tests remain the gate for the real thing.
CPU only; a few seconds."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import gs_next_oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
EMU_DIR = os.path.join(HERE, "cuda_emu")
SO = os.path.join(EMU_DIR, "libemu_next.so")
CSRC = os.path.join(os.path.dirname(HERE), "gaussian-splatting_cc-comments_b200", "csrc")


@pytest.fixture(scope="module")
def emu():
    deps = [os.path.join(EMU_DIR, "emu_next.cpp"), os.path.join(EMU_DIR, "emu.h"), os.path.join(CSRC, "gsr_adam.cu"),
            os.path.join(CSRC, "gsr_knn.cu"), os.path.join(CSRC, "gsr_internal.cuh")]
    if not os.path.exists(SO) or any(os.path.getmtime(SO) < os.path.getmtime(d) for d in deps):
        r = subprocess.run(["g++", "-std=c++20", "-O1", "-pthread", "-fPIC", "-shared", "-ffp-contract=off", "-mfma", "-w",
                            "-I/usr/local/cuda/include", deps[0], "-o", SO], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[-3000:]
    lib = C.CDLL(SO)
    lib.emu_knn_mean_dist2.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
    lib.emu_adam_step.argtypes = [C.c_int] + [C.c_void_p] * 8 + [C.c_double] * 3 + [C.c_void_p, C.c_int]
    return lib


def _knn(lib, pts):
    p = np.ascontiguousarray(pts, np.float32)
    out = np.full(p.shape[0], -1.0, np.float32)
    lib.emu_knn_mean_dist2(p.shape[0], p.ctypes.data, out.ctypes.data)
    return out


def _knn_cases():
    rng = np.random.default_rng(1)
    yield "one point", rng.normal(size=(1, 3))
    yield "two points", rng.normal(size=(2, 3))
    yield "three points", rng.normal(size=(3, 3))
    yield "four points", rng.normal(size=(4, 3))
    yield "one box minus one", rng.normal(size=(255, 3))
    yield "exactly one box", rng.normal(size=(256, 3))
    yield "one box plus a single-point tail box", rng.normal(size=(257, 3))
    yield "all coordinates negative", rng.uniform(-3, -1, size=(900, 3))
    yield "two clusters, far apart, plus 40 coincident points", np.concatenate(
        [rng.normal(size=(700, 3)) * 0.01, rng.normal(size=(700, 3)) * 10 + 50, np.zeros((40, 3))])
    yield "collinear (two axes have zero extent)", np.stack([np.linspace(0, 1, 1100), np.zeros(1100), np.zeros(1100)], 1)
    yield "uniform cube, 8 boxes", rng.uniform(size=(2000, 3))


@pytest.mark.parametrize("name,pts", list(_knn_cases()), ids=[n for n, _ in _knn_cases()])
def test_knn_kernels_match_the_oracle_bit_for_bit(emu, name, pts):
    got = _knn(emu, pts)
    want = O.knn_mean_dist2(pts)
    assert np.array_equal(got, want), "%s: %d of %d values differ" % (name, int((got != want).sum()), got.size)


def test_knn_non_finite_points_are_never_neighbours(emu):
    rng = np.random.default_rng(3)
    pts = rng.normal(size=(600, 3)).astype(np.float32)
    bad = [5, 77, 300]
    clean = np.delete(pts, bad, axis=0)
    pts[5, 0] = np.nan
    pts[77] = np.inf
    pts[300, 2] = -np.inf
    got = _knn(emu, pts)
    assert np.array_equal(np.delete(got, bad), O.knn_mean_dist2(clean))     # the others do not see them
    assert not np.isfinite(got[bad]).any()                                   # they have no neighbours themselves


def _adam_group(rng, P, w, lr, step, offset=0):
    def arr(fill):
        raw = np.zeros(P * w + 8, np.float32)
        a = raw[offset:offset + P * w]
        a[:] = fill
        return a
    g = rng.normal(size=P * w) * 10.0 ** rng.integers(-6, 1, size=P * w)
    g[rng.random(P * w) < 0.2] = 0.0
    return dict(p=arr(rng.normal(size=P * w)), g=arr(g), m=arr(rng.normal(size=P * w) * 0.1), v=arr(rng.random(P * w) * 0.01),
                lr=lr, step=step, w=w)


def _emu_adam(lib, groups, visible=None, force_scalar=0, betas=(0.9, 0.999), eps=1e-15):
    n = len(groups)
    ptrs = lambda k: (C.c_void_p * n)(*[g[k].ctypes.data for g in groups])
    N = (C.c_size_t * n)(*[g["p"].size for g in groups])
    LR = (C.c_double * n)(*[g["lr"] for g in groups])
    ST = (C.c_longlong * n)(*[g["step"] for g in groups])
    RW = (C.c_int * n)(*[g["w"] for g in groups])
    vis = None if visible is None else visible.ctypes.data_as(C.c_void_p)
    return lib.emu_adam_step(n, ptrs("p"), ptrs("g"), ptrs("m"), ptrs("v"), N, LR, ST, RW, betas[0], betas[1], eps, vis, force_scalar)


WIDTHS = [3, 3, 45, 1, 3, 4]                         # xyz, f_dc, f_rest, opacity, scaling, rotation
LRS = [1.6e-4 * 4.7, 2.5e-3, 2.5e-3 / 20, 0.05, 0.005, 0.001]


@pytest.mark.parametrize("P,offset,force_scalar", [(1, 0, 0), (3, 0, 0), (683, 0, 0), (2049, 0, 0), (777, 1, 0), (777, 0, 1), (0, 0, 0)])
def test_adam_kernel_matches_the_oracle_bit_for_bit(emu, P, offset, force_scalar):
    """Six groups in one launch; sizes that end inside a vector, inside a block, and exactly on a block; an unaligned
    view (offset 1 float) must take the element-wise kernel (mode 2) and still agree."""
    rng = np.random.default_rng(2)
    groups = [_adam_group(rng, P, w, lr, 7, offset) for w, lr in zip(WIDTHS, LRS)]
    ref = [{k: (x.copy() if isinstance(x, np.ndarray) else x) for k, x in g.items()} for g in groups]
    mode = _emu_adam(emu, groups, force_scalar=force_scalar)
    assert mode == (0 if P == 0 else (2 if (offset or force_scalar) else 1))
    for g, r in zip(groups, ref):
        O.adam_step(r["p"], r["g"], r["m"], r["v"], r["lr"], r["step"])
        for k in "pmv":
            assert np.array_equal(g[k], r[k]), k


def test_adam_kernel_visibility_mask_leaves_rows_untouched(emu):
    rng = np.random.default_rng(4)
    P = 1203
    vis = (rng.random(P) < 0.6).astype(np.uint8)
    groups = [_adam_group(rng, P, w, 1e-3, 3) for w in WIDTHS]
    ref = [{k: (x.copy() if isinstance(x, np.ndarray) else x) for k, x in g.items()} for g in groups]
    _emu_adam(emu, groups, visible=vis)
    for g, r in zip(groups, ref):
        before = {k: r[k].copy() for k in "pmv"}
        O.adam_step(r["p"], r["g"], r["m"], r["v"], r["lr"], r["step"])
        rows = np.repeat(vis.astype(bool), r["w"])
        for k in "pmv":
            assert np.array_equal(g[k], np.where(rows, r[k], before[k])), k
