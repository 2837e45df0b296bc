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
SO_DIR = os.environ.get("GSR_EMU_SO_DIR", EMU_DIR)      # tools/emu_sanitize.sh points this at sanitizer builds
SO = os.path.join(SO_DIR, "libemu_next.so")
SO_FUSED = os.path.join(SO_DIR, "libemu_fused.so")
SO_PRE = os.path.join(SO_DIR, "libemu_preprocess.so")
GOLDEN = os.path.join(HERE, "golden")
CSRC = os.path.join(os.path.dirname(HERE), "gaussian-splatting_cc-comments_b200", "csrc")


def _build(name, so):
    """Generated sources + g++ (tests/cuda_emu/build_emu.py); prebuilt sanitizer libraries are used as they are."""
    import sys
    sys.path.insert(0, EMU_DIR)
    import build_emu
    return build_emu.build(name) if so.startswith(EMU_DIR) else so


@pytest.fixture(scope="module")
def emu():
    lib = C.CDLL(_build("emu_next", SO))
    lib.emu_knn_mean_dist2.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
    lib.emu_adam_step.argtypes = [C.c_int] + [C.c_void_p] * 8 + [C.c_double] * 3 + [C.c_void_p, C.c_int]
    return lib


@pytest.fixture(scope="module")
def emu_fused():
    return C.CDLL(_build("emu_fused", SO_FUSED))


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


@pytest.mark.skipif(os.environ.get("GSR_SLOW") != "1", reason="2 minutes of host emulation: set GSR_SLOW=1 (ran green when it was written)")
def test_knn_more_than_256_boxes_walks_the_candidate_list_in_rounds(emu):
    """66 000 points = 258 boxes: the box-to-box pruning loop of knn_search_kernel takes a second round of 256 candidates."""
    rng = np.random.default_rng(5)
    P = 66000
    pts = np.concatenate([rng.normal(size=(P // 2, 3)) * 0.05, rng.uniform(-20, 20, size=(P - P // 2, 3))]).astype(np.float32)
    assert np.array_equal(_knn(emu, pts), O.knn_mean_dist2(pts))


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


# ------------------------------------------------------------------ raw-parameter helpers (csrc/gsr_fused_params.cuh)
def _aligned(n, dtype=np.float32, offset_floats=0):
    raw = np.zeros(n + 8 + offset_floats, dtype)
    start = (-raw.ctypes.data // 4) % 4                      # floats to the next 16-byte boundary
    return raw[start + offset_floats:start + offset_floats + n]


@pytest.mark.parametrize("M,D", [(16, 3), (16, 2), (16, 1), (16, 0), (9, 2), (4, 1), (2, 0), (1, 0)])
def test_split_sh_colour_equals_the_concatenated_layout(emu_fused, M, D):
    """Every row phase (rest rows are 12*(M-1) bytes apart: four alignments), vector and scalar path bit-identical, and
    equal to the dot product over the concatenated (P,M,3) tensor."""
    rng = np.random.default_rng(M * 10 + D)
    P = 41
    dc = _aligned(3 * P); dc[:] = rng.normal(size=3 * P)
    rest = _aligned(max(1, 3 * (M - 1) * P)); rest[:] = rng.normal(size=rest.size)
    k = np.ascontiguousarray(rng.normal(size=(P, 16)).astype(np.float32))
    out_v = np.zeros((P, 3), np.float32); out_s = np.zeros((P, 3), np.float32)
    args = lambda o, vec: (P, M, D, dc.ctypes.data_as(C.c_void_p), rest.ctypes.data_as(C.c_void_p), k.ctypes.data_as(C.c_void_p),
                           o.ctypes.data_as(C.c_void_p), vec)
    emu_fused.emu_sh_color_split(*args(out_v, 1))
    emu_fused.emu_sh_color_split(*args(out_s, 0))
    assert np.array_equal(out_v, out_s)
    K = (D + 1) ** 2
    shs = np.concatenate([dc.reshape(P, 1, 3), rest[:3 * (M - 1) * P].reshape(P, M - 1, 3)], axis=1).astype(np.float64)
    want = (shs[:, :K, :] * k[:, :K, None].astype(np.float64)).sum(axis=1)
    assert np.abs(out_v - want).max() <= 1e-5


@pytest.mark.parametrize("M,D,nrows", [(16, 3, 128), (16, 3, 77), (16, 1, 128), (9, 2, 128), (4, 1, 5), (2, 0, 128), (16, 0, 31)])
def test_backward_staging_and_split_write_out_index_maps(emu_fused, M, D, nrows):
    rng = np.random.default_rng(M + D + nrows)
    Lr, K = 3 * (M - 1), (D + 1) ** 2
    L = 3 * M
    S4 = (L + 3) // 4
    S = 4 * (S4 if S4 % 2 == 1 else S4 + 1)
    vis = (rng.random(128) < 0.7).astype(np.uint8)
    rest_block = _aligned(max(1, nrows * Lr)); rest_block[:] = rng.normal(size=rest_block.size)
    rows = np.full((128, S), np.nan, np.float32)
    emu_fused.emu_stage_rest_rows(nrows, M, 3 * K - 3, S, vis.ctypes.data_as(C.c_void_p), rest_block.ctypes.data_as(C.c_void_p),
                                  rows.ctypes.data_as(C.c_void_p))
    want = np.full((128, S), np.nan, np.float32)
    src = rest_block[:nrows * Lr].reshape(nrows, Lr)
    for r in range(nrows):
        if vis[r]:
            want[r, 3:3 + 3 * K - 3] = src[r, :3 * K - 3]
    assert np.array_equal(np.isnan(rows), np.isnan(want)) and np.array_equal(np.nan_to_num(rows), np.nan_to_num(want))
    # write-out: rows now hold dL_dsh[0..3K) (anything beyond is NOT to be trusted: poison it)
    rows = rng.normal(size=(128, S)).astype(np.float32)
    rows[:, 3 * K:] = np.nan
    for misalign in (0, 1):
        dc_out = _aligned(3 * nrows); dc_out[:] = -7
        rest_out = _aligned(max(1, nrows * Lr), offset_floats=misalign); rest_out[:] = -7
        emu_fused.emu_store_split_rows(nrows, M, 3 * K, S, vis.ctypes.data_as(C.c_void_p), rows.ctypes.data_as(C.c_void_p),
                                       dc_out.ctypes.data_as(C.c_void_p), rest_out.ctypes.data_as(C.c_void_p))
        v = vis[:nrows].astype(bool)
        assert np.array_equal(dc_out.reshape(nrows, 3), np.where(v[:, None], rows[:nrows, :3], 0.0))
        want_rest = np.zeros((nrows, Lr), np.float32)
        want_rest[:, :3 * K - 3] = np.where(v[:, None], rows[:nrows, 3:3 * K], 0.0)
        assert np.array_equal(rest_out[:nrows * Lr].reshape(nrows, Lr), want_rest), (misalign,)


def test_activations_and_their_derivatives_match_torch(emu_fused):
    import torch
    g = torch.Generator().manual_seed(0)
    n = 500
    raw_op = (torch.randn(n, generator=g) * 3).requires_grad_(True)
    raw_s = (torch.randn(n, 3, generator=g) * 2 - 3).requires_grad_(True)
    raw_q = torch.randn(n, 4, generator=g) * 10.0 ** torch.randint(-3, 3, (n, 1), generator=g).float()
    raw_q[0] = 0.0                                          # below eps: F.normalize clamps the denominator
    raw_q.requires_grad_(True)
    g_op, g_s, g_q = torch.randn(n, generator=g), torch.randn(n, 3, generator=g), torch.randn(n, 4, generator=g)
    op, s, q = torch.sigmoid(raw_op), torch.exp(raw_s), torch.nn.functional.normalize(raw_q)
    ((op * g_op).sum() + (s * g_s).sum() + (q * g_q).sum()).backward()
    outs = [np.zeros(x, np.float32) for x in (n, 3 * n, 4 * n, n, 3 * n, 4 * n)]
    p = lambda a: np.ascontiguousarray(a.detach().numpy()).ctypes.data_as(C.c_void_p)
    keep = [np.ascontiguousarray(t.detach().numpy()) for t in (raw_op, raw_s, raw_q, g_op, g_s, g_q)]
    emu_fused.emu_activations(n, *[k.ctypes.data_as(C.c_void_p) for k in keep], *[o.ctypes.data_as(C.c_void_p) for o in outs])
    for got, want in zip(outs[:5], (op, s, q, raw_op.grad, raw_s.grad)):
        w = want.detach().numpy().reshape(-1)
        assert np.all(np.abs(got - w) <= 2e-6 * np.maximum(np.abs(w), 1.0)), np.abs(got - w).max()
    # the normalize derivative cancels (g - q (q.g)): the error scales with |g| / |r| of the row, not with the result
    w = raw_q.grad.numpy()
    scale = (g_q.norm(dim=1) / raw_q.detach().norm(dim=1).clamp_min(1e-12)).numpy()[:, None]
    assert np.all(np.abs(outs[5].reshape(n, 4) - w) <= 4e-6 * scale), (np.abs(outs[5].reshape(n, 4) - w) / scale).max()


# ------------------------------------------------------------------ the per-Gaussian kernels themselves, on the host
@pytest.fixture(scope="module")
def emu_pre():
    return C.CDLL(_build("emu_preprocess", SO_PRE))


def _vp(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _al(a):
    """16-byte aligned float32 copy (the staged kernels are chosen on alignment, as on the device)."""
    a = np.asarray(a, np.float32)
    v = _aligned(max(1, a.size))[:a.size]
    v[:] = a.reshape(-1)
    return v.reshape(a.shape)


def _fwd(lib, z, D, smod=1.0, raw=0, split=False, scales=None, rots=None, opac=None, shs=None):
    P, W, H = z["in_means3D"].shape[0], int(z["cam_W"]), int(z["cam_H"])
    m = np.ascontiguousarray(z["in_means3D"])
    sc = np.ascontiguousarray(z["in_scales"] if scales is None else scales, np.float32)
    rt = _al(z["in_rotations"] if rots is None else rots)
    op = np.ascontiguousarray(z["in_opacities"] if opac is None else opac, np.float32)
    sh = _al(z["in_shs"] if shs is None else shs)
    M = sh.shape[1]
    dc, rest = _al(sh[:, :1, :]), _al(sh[:, 1:, :])
    o = dict(radii=np.zeros(P, np.int32), packed=np.zeros((P, 12), np.float32), tt=np.zeros(P, np.uint32), cl=np.zeros(P, np.uint8),
             dk=np.zeros(P, np.uint32))
    tot = np.zeros(1, np.uint32)
    view, proj, cp = (np.ascontiguousarray(z[k]) for k in ("cam_viewmatrix", "cam_projmatrix", "cam_campos"))
    lib.emu_preprocess_fwd(P, D, M, _vp(m), _vp(sc), C.c_float(smod), _vp(rt), _vp(op), None if split else _vp(sh), _vp(dc) if split else None,
                           _vp(rest) if split else None, _vp(view), _vp(proj), _vp(cp), W, H, C.c_float(float(z["cam_tanfovx"])),
                           C.c_float(float(z["cam_tanfovy"])), raw, _vp(o["radii"]), _vp(o["packed"]), _vp(o["tt"]), _vp(o["cl"]), _vp(o["dk"]), _vp(tot))
    o["total"] = int(tot[0])
    return o


@pytest.mark.parametrize("name,D,smod", [("ref_cuda_sh3_160x96", 3, 1.0), ("ref_cuda_sh1_mod_64x64", 1, 0.6)])
def test_forward_kernel_source_reproduces_the_reference_builds_b200_fixture_on_the_host(emu_pre, name, D, smod):
    """preprocess_fwd_kernel<false,false> -- the kernel source that runs on the GPU, compiled for the host -- against what
    the reference's own CUDA build wrote on a B200: radii, tile counts, num_rendered, means2D, conic, opacity, colours,
    clamp bits and depth keys, bit for bit (every rounding of this stage is pinned; it has no MUFU-dependent output)."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    o = _fwd(emu_pre, z, D, smod)
    vis = z["out_radii"] > 0
    assert np.array_equal(o["radii"], z["out_radii"]) and np.array_equal(o["tt"], z["out_tiles_touched"])
    assert o["total"] == int(z["out_num_rendered"])
    pk = o["packed"]
    assert np.array_equal(pk[vis, 0:2], z["out_means2D"][vis])
    assert np.array_equal(pk[vis][:, [2, 3, 4]], z["out_conic_opacity"][vis][:, :3]) and np.array_equal(pk[vis, 6], z["out_conic_opacity"][vis, 3])
    assert np.array_equal(pk[vis, 7:10], z["out_rgb"][vis])
    assert np.array_equal(o["dk"][vis].view(np.float32), z["out_depths"][vis])
    bits = z["out_clamped"][:, 0].astype(np.uint8) | (z["out_clamped"][:, 1].astype(np.uint8) << 1) | (z["out_clamped"][:, 2].astype(np.uint8) << 2)
    assert np.array_equal(o["cl"][vis], bits[vis])


def _raw_inputs(emu_fused, z, seed=0):
    rng = np.random.default_rng(seed)
    P = z["in_means3D"].shape[0]
    raw_s = np.ascontiguousarray(np.log(z["in_scales"]), np.float32)
    raw_q = _al(z["in_rotations"] * rng.uniform(0.2, 3, size=(P, 1)))
    raw_o = np.ascontiguousarray(np.log(z["in_opacities"] / (1 - z["in_opacities"])), np.float32).reshape(P)
    return raw_o, raw_s, raw_q


def _activate(emu_fused, raw_o, raw_s, raw_q, g_o=None, g_s=None, g_q=None):
    P = raw_o.shape[0]
    outs = [np.zeros(n, np.float32) for n in (P, 3 * P, 4 * P, P, 3 * P, 4 * P)]
    g = [np.ascontiguousarray(np.zeros(n, np.float32) if x is None else x, np.float32).reshape(-1) for x, n in ((g_o, P), (g_s, 3 * P), (g_q, 4 * P))]
    emu_fused.emu_activations(P, _vp(raw_o), _vp(raw_s), _vp(np.ascontiguousarray(raw_q)), _vp(g[0]), _vp(g[1]), _vp(g[2]), *[_vp(o) for o in outs])
    return outs


@pytest.mark.parametrize("M,D", [(16, 3), (16, 2), (16, 0), (9, 2), (4, 1), (1, 0)])
def test_raw_parameter_forward_kernels_equal_the_reference_shaped_kernel(emu_pre, emu_fused, M, D):
    """preprocess_fwd_kernel<RAW,SPLIT> on the raw tensors == <false,false> on the activated / concatenated ones, every
    output bit for bit (activations evaluated by the same host functions on both sides)."""
    z = np.load(os.path.join(GOLDEN, "ref_cuda_sh3_160x96.npz"))
    P = z["in_means3D"].shape[0]
    raw_o, raw_s, raw_q = _raw_inputs(emu_fused, z)
    a = _activate(emu_fused, raw_o, raw_s, raw_q)
    act_o, act_s, act_q = a[0].reshape(P, 1), a[1].reshape(P, 3), a[2].reshape(P, 4)
    sh = z["in_shs"][:, :M, :]
    base = _fwd(emu_pre, z, D, scales=act_s, rots=act_q, opac=act_o, shs=sh)
    assert (base["radii"] > 0).sum() > 0.5 * P
    for raw, split in ((1, False), (0, True), (1, True)):
        o = _fwd(emu_pre, z, D, raw=raw, split=split, scales=raw_s if raw else act_s, rots=raw_q if raw else act_q,
                 opac=raw_o.reshape(P, 1) if raw else act_o, shs=sh)
        for k in ("radii", "packed", "tt", "cl", "dk"):
            assert np.array_equal(o[k], base[k]), (raw, split, k)
        assert o["total"] == base["total"]


def _bwd(lib, z, fw, grad_acc, D, raw=0, split=False, scales=None, rots=None, shs=None, generic=0):
    P, W, H = z["in_means3D"].shape[0], int(z["cam_W"]), int(z["cam_H"])
    m = np.ascontiguousarray(z["in_means3D"])
    sc, rt, sh = np.ascontiguousarray(scales, np.float32), _al(rots), _al(shs)
    M = sh.shape[1]
    dc, rest = _al(sh[:, :1, :]), _al(sh[:, 1:, :])
    view, proj, cp = (np.ascontiguousarray(z[k]) for k in ("cam_viewmatrix", "cam_projmatrix", "cam_campos"))
    widths = dict(mean2D=3, conic=4, opacity=1, color=3, mean3D=3, cov3D=6, sh=3 * M, scale=3, rot=4, dc=3, rest=3 * (M - 1))
    out = {}
    for k, w in widths.items():
        out[k] = _aligned(max(1, P * w))
        out[k][:] = np.nan                                   # every element must be written by the kernel
    mode = lib.emu_preprocess_bwd(
        P, D, M, _vp(m), _vp(fw["radii"]), None if split else _vp(sh), _vp(fw["cl"]), _vp(sc), _vp(rt), C.c_float(1.0), _vp(view), _vp(proj), _vp(cp),
        W, H, C.c_float(float(z["cam_tanfovx"])), C.c_float(float(z["cam_tanfovy"])), _vp(grad_acc), _vp(fw["packed"]), raw,
        _vp(dc) if split else None, _vp(rest) if split else None, generic, _vp(out["mean2D"]), _vp(out["conic"]), _vp(out["opacity"]),
        _vp(out["color"]), _vp(out["mean3D"]), _vp(out["cov3D"]), None if split else _vp(out["sh"]), _vp(out["scale"]), _vp(out["rot"]),
        _vp(out["dc"]) if split else None, _vp(out["rest"]) if split else None)
    return mode, {k: v[:P * widths[k]] for k, v in out.items()}


@pytest.mark.parametrize("M,D", [(16, 3), (16, 1), (16, 0), (9, 2), (4, 1), (1, 0)])
def test_raw_parameter_backward_kernels_equal_the_reference_shaped_kernel_plus_the_chain_rule(emu_pre, emu_fused, M, D):
    """preprocess_bwd_kernel_staged<8,false,RAW,SPLIT> against the reference-shaped kernel on the same accumulators:
    geometry gradients identical, dL_dfeatures_dc / _rest == the two slices of dL_dsh, and the raw-parameter gradients ==
    the activated ones pushed through the activation derivatives.  Bit for bit where both sides run the same staged
    arithmetic; 1e-5 where the reference-shaped side is the generic kernel (M = 9, 1: another formulation of the SH term)."""
    z = np.load(os.path.join(GOLDEN, "ref_cuda_sh3_160x96.npz"))
    P = z["in_means3D"].shape[0]
    rng = np.random.default_rng(M + D)
    raw_o, raw_s, raw_q = _raw_inputs(emu_fused, z)
    a = _activate(emu_fused, raw_o, raw_s, raw_q)
    act_o, act_s, act_q = a[0].reshape(P, 1), a[1].reshape(P, 3), a[2].reshape(P, 4)
    sh = z["in_shs"][:, :M, :]
    fw = _fwd(emu_pre, z, D, scales=act_s, rots=act_q, opac=act_o, shs=sh)
    ga = _al(rng.normal(size=(P, 12)) * (fw["radii"] > 0)[:, None])
    mode_b, base = _bwd(emu_pre, z, fw, ga, D, scales=act_s, rots=act_q, shs=sh)
    mode_f, fused = _bwd(emu_pre, z, fw, ga, D, raw=1, split=True, scales=raw_s, rots=raw_q, shs=sh)
    assert mode_f == 3 and mode_b == (1 if (3 * M) % 4 == 0 else 2)
    same_math = mode_b == 1

    def eq(x, y, what):
        assert not np.isnan(x).any(), what + ": not every element was written"
        if same_math:
            assert np.array_equal(x, y), what
        else:
            assert np.abs(x - y).max() <= 1e-5 * max(1.0, np.abs(y).max()), what
    for k in ("mean2D", "conic", "color", "mean3D", "cov3D"):
        eq(fused[k], base[k], k)
    dsh = base["sh"].reshape(P, M, 3)
    eq(fused["dc"], dsh[:, :1, :].reshape(-1), "dL_dfeatures_dc")
    if M > 1:
        eq(fused["rest"], dsh[:, 1:, :].reshape(-1), "dL_dfeatures_rest")
    chain = _activate(emu_fused, raw_o, raw_s, raw_q, base["opacity"], base["scale"], base["rot"])
    eq(fused["opacity"], chain[3], "dL_d(raw opacity)")
    eq(fused["scale"], chain[4], "dL_d(raw scaling)")
    eq(fused["rot"], chain[5], "dL_d(raw rotation)")
    # culled Gaussians get exact zeros everywhere
    culled = fw["radii"] <= 0
    assert culled.any() and not fused["dc"].reshape(P, 3)[culled].any() and not fused["scale"].reshape(P, 3)[culled].any()
    # the two options are independent
    _, raw_only = _bwd(emu_pre, z, fw, ga, D, raw=1, split=False, scales=raw_s, rots=raw_q, shs=sh)
    assert np.array_equal(raw_only["scale"], fused["scale"]) and np.array_equal(raw_only["rot"], fused["rot"])
    eq(raw_only["sh"], base["sh"], "dL_dsh with raw activations only")
    _, split_only = _bwd(emu_pre, z, fw, ga, D, raw=0, split=True, scales=act_s, rots=act_q, shs=sh)
    assert np.array_equal(split_only["dc"], fused["dc"]) and np.array_equal(split_only["scale"], base["scale"])
