"""Densification (second half) on the CPU: the oracle against the fixture written by the reference's own
GaussianModel (tests/golden/make_golden_densify.py), and -- through the host-emulated C ABI (tests/cuda_emu) -- the
product's classify / apply / stats kernels against the oracle, bit for bit."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import gs_next_oracle as O   # noqa: E402

GOLD = np.load(os.path.join(ROOT, "tests", "golden", "ref_densify_cpu.npz"))
NAMES = O.GROUP_ORDER


def _case(tag):
    g = lambda k: GOLD[f"{tag}_{k}"]
    has = bool(g("with_moments"))
    before = {n: g("before_p_" + n) for n in NAMES}
    m = {n: g("before_m_" + n) for n in NAMES} if has else None
    v = {n: g("before_v_" + n) for n in NAMES} if has else None
    after = {n: g("after_p_" + n) for n in NAMES}
    am = {n: g("after_m_" + n) for n in NAMES} if has else None
    av = {n: g("after_v_" + n) for n in NAMES} if has else None
    args = dict(max_grad=float(g("max_grad")), min_opacity=float(g("min_opacity")), extent=float(g("extent")),
                percent_dense=float(g("percent_dense")), max_screen_size=float(g("max_screen_size")))
    return before, m, v, g("before_accum"), g("before_denom"), g("samples"), args, after, am, av, g


@pytest.mark.parametrize("tag", list(GOLD["cases"]))
def test_oracle_reproduces_the_references_densify_and_prune(tag):
    before, m, v, accum, denom, samples, args, after, am, av, g = _case(tag)
    p1, m1, v1 = O.densify_and_prune(before, m, v, accum, denom, samples=samples, **args)
    n = after["xyz"].shape[0]
    assert n > 0 and p1["xyz"].shape[0] == n
    for name in NAMES:
        assert p1[name].shape == after[name].shape, name
        if name in ("xyz", "scaling"):          # children are computed: bmm order / division vs reciprocal, <= a few ulp
            np.testing.assert_allclose(p1[name], after[name], rtol=2e-6, atol=2e-6, err_msg=name)
        else:                                    # copied rows: exact, in the reference's order
            assert np.array_equal(p1[name], after[name]), name
        if m is not None:
            assert np.array_equal(m1[name], am[name]) and np.array_equal(v1[name], av[name]), name
    # rows that are not split children are copies of xyz / scaling as well: exact
    n_child = samples.shape[0]
    if n_child < n:
        head = n - n_child          # children, if any survive, are at the end; at least the first n - n_child rows are copies
        assert np.array_equal(p1["xyz"][:head], after["xyz"][:head]) and np.array_equal(p1["scaling"][:head], after["scaling"][:head])
    # the statistics are all zeros afterwards (densification_postfix), of the new length
    for k in ("accum", "denom", "max_radii2D"):
        assert g("after_" + k).shape[0] == n and not g("after_" + k).any()


def test_oracle_reproduces_the_references_statistics_update():
    s = lambda k: GOLD["stats_" + k]
    accum, denom, mr = s("accum0").reshape(-1).copy(), s("denom0").reshape(-1).copy(), s("max_radii0").copy()
    O.densify_stats(s("vs_grad"), s("radii"), accum, denom, mr)
    assert np.array_equal(denom, s("denom1").reshape(-1)) and np.array_equal(mr, s("max_radii1"))
    np.testing.assert_allclose(accum, s("accum1").reshape(-1), rtol=2e-7, atol=0)      # 2-element norm: fma or not, 1 ulp
    untouched = s("radii") <= 0
    assert untouched.any() and np.array_equal(accum[untouched], s("accum0").reshape(-1)[untouched])


# ---- the product's kernels (host-emulated C ABI) against the oracle ---------------------------------------------------------
import ctypes as C   # noqa: E402

from test_abi import DensifyGroup, DensifyParams, densify_argtypes   # noqa: E402

WIDTHS = {"xyz": 3, "f_dc": 3, "f_rest": 45, "opacity": 1, "scaling": 3, "rotation": 4}


@pytest.fixture(scope="module")
def abi():
    sys.path.insert(0, os.path.join(ROOT, "tests", "cuda_emu"))
    import build_emu
    lib = C.CDLL(os.environ.get("GSR_EMU_ABI_SO") or build_emu.build("emu_abi"))
    lib.gsr_last_error.restype = C.c_char_p
    densify_argtypes(lib)
    return lib


def _ptr(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def _aligned_bytes(n):
    raw = np.zeros(n + 512, np.uint8)
    off = (-raw.ctypes.data) % 256
    return raw, raw.ctypes.data + off


def densify_through_abi(lib, params, m, v, accum, denom, args, draw_samples, N=2, names=NAMES):
    """classify -> read totals -> draw samples -> apply, as gsr_densify.densify_and_prune drives the library.
    draw_samples(split_mask) -> (N * n_split, 3).  Returns (params', m', v', flags, totals)."""
    P = params["xyz"].shape[0]
    flat = lambda a: np.ascontiguousarray(a, np.float32).reshape(P, -1)
    p = {n: flat(params[n]) for n in names}
    has = m is not None
    mm = {n: flat(m[n]) for n in names} if has else None
    vv = {n: flat(v[n]) for n in names} if has else None
    prm = DensifyParams(args["max_grad"], args["min_opacity"], args["extent"], args["percent_dense"], args["max_screen_size"] or 0.0, N)
    flags = np.zeros(max(P, 1), np.uint8)
    totals = np.full(5, -1, np.int32)
    keep, scratch = _aligned_bytes(lib.gsr_densify_scratch_bytes(P))
    a, d = (np.ascontiguousarray(x, np.float32).reshape(-1) for x in (accum, denom))
    rc = lib.gsr_densify_classify(P, _ptr(a), _ptr(d), _ptr(p["scaling"]), _ptr(p["opacity"]), C.byref(prm), _ptr(flags), _ptr(totals),
                                  C.c_void_p(scratch), None)
    assert rc == 0, lib.gsr_last_error()
    P_out, n_split = int(totals[0]), int(totals[1])
    samples = np.ascontiguousarray(draw_samples((flags[:P] & 16) != 0), np.float32).reshape(-1, 3)
    assert samples.shape[0] == N * n_split
    groups = (DensifyGroup * len(names))()
    outs = {}
    for g, n in zip(groups, names):
        w = p[n].shape[1]
        po = np.full((P_out + 1, w), np.nan, np.float32)            # one guard row: nothing may be written beyond P_out
        mo = np.full((P_out + 1, w), np.nan, np.float32) if has else None
        vo = np.full((P_out + 1, w), np.nan, np.float32) if has else None
        outs[n] = (po, mo, vo)
        g.param, g.param_out, g.row_width = p[n].ctypes.data, po.ctypes.data, w
        if has:
            g.exp_avg, g.exp_avg_sq, g.exp_avg_out, g.exp_avg_sq_out = mm[n].ctypes.data, vv[n].ctypes.data, mo.ctypes.data, vo.ctypes.data
    rc = lib.gsr_densify_apply(P, _ptr(flags), C.c_void_p(scratch), _ptr(totals), C.byref(prm), _ptr(samples), groups, len(names),
                               names.index("xyz"), names.index("scaling"), names.index("rotation"), None)
    assert rc == 0, lib.gsr_last_error()
    for n in names:
        for arr in outs[n]:
            if arr is not None:
                assert np.isnan(arr[P_out]).all(), "guard row of %s written" % n
                assert np.isfinite(arr[:P_out]).all(), "unwritten element in %s" % n
    shape = lambda n, x: x[:P_out].reshape((P_out,) + tuple(params[n].shape[1:]))
    pick = lambda k: {n: shape(n, outs[n][k]) for n in names}
    return pick(0), (pick(1) if has else None), (pick(2) if has else None), flags[:P], totals


@pytest.mark.parametrize("tag", list(GOLD["cases"]))
def test_emulated_kernels_equal_the_oracle_on_the_reference_fixture(abi, tag):
    before, m, v, accum, denom, samples, args, after, am, av, g = _case(tag)
    p1, m1, v1, flags, totals = densify_through_abi(abi, before, m, v, accum, denom, args, lambda split: samples)
    p0, m0, v0 = O.densify_and_prune(before, m, v, accum, denom, samples=samples, **args)
    cm, sm = O.densify_masks(accum.reshape(-1), denom.reshape(-1), before["scaling"], args["max_grad"], args["extent"], args["percent_dense"])
    assert np.array_equal((flags & 8) != 0, cm) and np.array_equal((flags & 16) != 0, sm)
    assert totals[0] == after["xyz"].shape[0] and totals[1] == sm.sum()
    for name in NAMES:                                   # same host libm on both sides: bit for bit, every tensor
        assert np.array_equal(p1[name], p0[name]), name
        if m is not None:
            assert np.array_equal(m1[name], m0[name]) and np.array_equal(v1[name], v0[name]), name
        # and against the reference's own output, as the oracle is pinned
        if name in ("xyz", "scaling"):
            np.testing.assert_allclose(p1[name], after[name], rtol=2e-6, atol=2e-6)
        else:
            assert np.array_equal(p1[name], after[name]), name


def _random_model(rng, P, with_moments=True):
    p = {n: rng.normal(size=(P, WIDTHS[n])).astype(np.float32) for n in NAMES}
    p["opacity"] = (p["opacity"] * 3 - 1).astype(np.float32)
    p["scaling"] = (p["scaling"] * 1.2 - 3.5).astype(np.float32)
    m = {n: (rng.normal(size=p[n].shape) * 1e-3).astype(np.float32) for n in NAMES} if with_moments else None
    v = {n: (rng.random(size=p[n].shape) * 1e-6).astype(np.float32) for n in NAMES} if with_moments else None
    denom = (rng.integers(1, 40, P) * (rng.random(P) > 0.15)).astype(np.float32)
    accum = (rng.random(P) * 6e-4 * denom).astype(np.float32)
    return p, m, v, accum, denom


@pytest.mark.parametrize("P,N,screen,with_moments", [(1, 2, 20.0, True), (255, 2, 20.0, True), (256, 2, 0.0, False), (257, 3, 20.0, True),
                                                     (1531, 2, 20.0, True), (2048, 1, 20.0, True)])
def test_emulated_kernels_equal_the_oracle_on_random_models(abi, P, N, screen, with_moments):
    """Block boundaries (256 rows per block), several blocks of offsets, N = 1..3 children, with and without Adam state."""
    rng = np.random.default_rng(P * 7 + N)
    p, m, v, accum, denom = _random_model(rng, P, with_moments)
    args = dict(max_grad=0.0002, min_opacity=0.005, extent=3.0, percent_dense=0.01, max_screen_size=screen)
    drawn = {}
    def draw(split):
        drawn["s"] = (rng.normal(size=(N * int(split.sum()), 3)) * np.tile(np.exp(p["scaling"][split]), (N, 1))).astype(np.float32)
        return drawn["s"]
    p1, m1, v1, flags, totals = densify_through_abi(abi, p, m, v, accum, denom, args, draw, N=N)
    p0, m0, v0 = O.densify_and_prune(p, m, v, accum, denom, samples=drawn["s"], N=N, **args)
    assert totals[0] == p0["xyz"].shape[0]
    for name in NAMES:
        assert np.array_equal(p1[name], p0[name]), name
        if with_moments:
            assert np.array_equal(m1[name], m0[name]) and np.array_equal(v1[name], v0[name]), name


def test_emulated_extremes_nothing_selected_everything_pruned_everything_split(abi):
    rng = np.random.default_rng(5)
    P = 300
    p, m, v, accum, denom = _random_model(rng, P)
    base = dict(min_opacity=0.005, extent=3.0, percent_dense=0.01, max_screen_size=20.0)
    cases = {"nothing selected": dict(base, max_grad=1e9), "all pruned": dict(base, max_grad=1e9, min_opacity=2.0),
             "all hot": dict(base, max_grad=-1.0, percent_dense=1e-9, max_screen_size=0.0, min_opacity=0.0)}
    for what, args in cases.items():
        drawn = {}
        def draw(split):
            drawn["s"] = rng.normal(size=(2 * int(split.sum()), 3)).astype(np.float32)
            return drawn["s"]
        p1, m1, v1, flags, totals = densify_through_abi(abi, p, m, v, accum, denom, args, draw)
        p0, m0, v0 = O.densify_and_prune(p, m, v, accum, denom, samples=drawn["s"], **args)
        assert totals[0] == p0["xyz"].shape[0], what
        if what == "all pruned":
            assert totals[0] == 0
        if what == "all hot":
            assert totals[1] == P and totals[0] == 2 * P        # every Gaussian is large and hot: all split, parents removed
        for name in NAMES:
            assert np.array_equal(p1[name], p0[name]), (what, name)
            assert np.array_equal(m1[name], m0[name]) and np.array_equal(v1[name], v0[name]), (what, name)


def test_emulated_statistics_kernel_equals_the_oracle_and_the_reference(abi):
    s = lambda k: GOLD["stats_" + k]
    accum, denom, mr = s("accum0").reshape(-1).copy(), s("denom0").reshape(-1).copy(), s("max_radii0").copy()
    vs, radii = np.ascontiguousarray(s("vs_grad"), np.float32), np.ascontiguousarray(s("radii"), np.int32)
    assert abi.gsr_densify_stats(radii.shape[0], _ptr(vs), _ptr(radii), _ptr(accum), _ptr(denom), _ptr(mr), None) == 0, abi.gsr_last_error()
    a0, d0, r0 = s("accum0").reshape(-1).copy(), s("denom0").reshape(-1).copy(), s("max_radii0").copy()
    O.densify_stats(vs, radii, a0, d0, r0)
    assert np.array_equal(accum, a0) and np.array_equal(denom, d0) and np.array_equal(mr, r0)
    assert np.array_equal(denom, s("denom1").reshape(-1)) and np.array_equal(mr, s("max_radii1"))
    np.testing.assert_allclose(accum, s("accum1").reshape(-1), rtol=2e-7, atol=0)
    # max_radii2D is optional
    a1, d1 = s("accum0").reshape(-1).copy(), s("denom0").reshape(-1).copy()
    assert abi.gsr_densify_stats(radii.shape[0], _ptr(vs), _ptr(radii), _ptr(a1), _ptr(d1), None, None) == 0
    assert np.array_equal(a1, a0) and np.array_equal(d1, d0)


# ---- host logic of gsr_densify (the drop-in for GaussianModel.densify_and_prune) -------------------------------------------
class _EmulatedExtension:
    """Stands in for the torch extension's three densify entry points with CPU tensors over the host-emulated C ABI, so
    that the Python glue (optimizer surgery, the sampling call, attribute hand-over) runs here.  Test double only."""

    def __init__(self, lib):
        self.lib = lib

    def densify_stats(self, grad, radii, accum, denom, max_radii2D=None):
        tp = lambda t: None if t is None else C.c_void_p(t.data_ptr())
        assert self.lib.gsr_densify_stats(radii.numel(), tp(grad), tp(radii), tp(accum), tp(denom), tp(max_radii2D), None) == 0

    def densify_classify(self, accum, denom, scaling, opacity, max_grad, min_opacity, extent, percent_dense, screen, copies):
        import torch
        P = scaling.shape[0]
        self.prm = DensifyParams(max_grad, min_opacity, extent, percent_dense, screen, copies)
        flags = torch.zeros(P, dtype=torch.uint8)
        totals = torch.zeros(5, dtype=torch.int32)
        n = self.lib.gsr_densify_scratch_bytes(P)
        raw = torch.zeros(n + 512, dtype=torch.uint8)
        off = (-raw.data_ptr()) % 256
        scratch = raw[off:off + n]
        tp = lambda t: C.c_void_p(t.data_ptr())
        rc = self.lib.gsr_densify_classify(P, tp(accum), tp(denom), tp(scaling), tp(opacity), C.byref(self.prm), tp(flags), tp(totals), tp(scratch), None)
        assert rc == 0, self.lib.gsr_last_error()
        return flags, totals, scratch

    def densify_apply(self, flags, scratch, totals, samples, params, ms, vs, xyz_g, scaling_g, rotation_g, *hyper):
        import torch
        P_out = totals[0]
        groups = (DensifyGroup * len(params))()
        out_p, out_m, out_v = [], [], []
        for g, p, m, v in zip(groups, params, ms, vs):
            shape = (P_out,) + tuple(p.shape[1:])
            out_p.append(torch.full(shape, float("nan")))
            g.param, g.param_out, g.row_width = p.data_ptr(), out_p[-1].data_ptr(), p.numel() // max(p.shape[0], 1)
            if m is not None:
                out_m.append(torch.full(shape, float("nan"))); out_v.append(torch.full(shape, float("nan")))
                g.exp_avg, g.exp_avg_sq, g.exp_avg_out, g.exp_avg_sq_out = m.data_ptr(), v.data_ptr(), out_m[-1].data_ptr(), out_v[-1].data_ptr()
            else:
                out_m.append(None); out_v.append(None)
        th = (C.c_int * 5)(*totals)
        tp = lambda t: C.c_void_p(t.data_ptr())
        rc = self.lib.gsr_densify_apply(flags.numel(), tp(flags), tp(scratch), th, C.byref(DensifyParams(*hyper)), tp(samples), groups,
                                        len(params), xyz_g, scaling_g, rotation_g, None)
        assert rc == 0, self.lib.gsr_last_error()
        return out_p, out_m, out_v


class _Model:
    """The attributes of the reference's GaussianModel that densification touches (scene/gaussian_model.py:44-62)."""
    percent_dense = 0.01


def _model_from_fixture(tag, optimizer_cls):
    import torch
    before, m, v, accum, denom, samples, args, after, am, av, g = _case(tag)
    attr = {"xyz": "_xyz", "f_dc": "_features_dc", "f_rest": "_features_rest", "opacity": "_opacity", "scaling": "_scaling", "rotation": "_rotation"}
    mdl = _Model()
    for n in NAMES:
        setattr(mdl, attr[n], torch.nn.Parameter(torch.from_numpy(before[n].copy())))
    mdl.percent_dense = args["percent_dense"]
    mdl.optimizer = optimizer_cls([{"params": [getattr(mdl, attr[n])], "lr": 1e-3, "name": n} for n in NAMES], lr=0.0, eps=1e-15)
    if m is not None:
        for n in NAMES:
            mdl.optimizer.state[getattr(mdl, attr[n])] = {"step": torch.tensor(2.0), "exp_avg": torch.from_numpy(m[n].copy()),
                                                          "exp_avg_sq": torch.from_numpy(v[n].copy())}
    mdl.xyz_gradient_accum = torch.from_numpy(accum.copy())
    mdl.denom = torch.from_numpy(denom.copy())
    mdl.max_radii2D = torch.from_numpy(g("before_max_radii2D").copy())
    return mdl, attr, args, after, am, av, int(g("sampling_seed"))


@pytest.mark.parametrize("tag", list(GOLD["cases"]))
def test_drop_in_densify_and_prune_reproduces_the_reference_model_including_its_random_draw(abi, tag, monkeypatch):
    """gsr_densify.densify_and_prune on the fixture's model, seeded like the reference run: the samples it draws are the
    ones the reference drew (same torch.normal call), and model + optimizer end up as the reference's."""
    import torch
    import gsr_densify
    monkeypatch.setattr(gsr_densify, "_C", _EmulatedExtension(abi))
    mdl, attr, args, after, am, av, seed = _model_from_fixture(tag, torch.optim.Adam)
    torch.manual_seed(seed)
    P0, P1, n_clone, n_split = gsr_densify.densify_and_prune(mdl, args["max_grad"], args["min_opacity"], args["extent"],
                                                             args["max_screen_size"] or None)
    assert P1 == after["xyz"].shape[0] and n_split * 2 == GOLD[f"{tag}_samples"].shape[0]
    assert [g["name"] for g in mdl.optimizer.param_groups] == list(NAMES)
    for n in NAMES:
        p = getattr(mdl, attr[n])
        assert isinstance(p, torch.nn.Parameter) and p.requires_grad and mdl.optimizer.param_groups[NAMES.index(n)]["params"][0] is p
        got = p.detach().numpy()
        assert got.shape == after[n].shape
        if n in ("xyz", "scaling"):
            np.testing.assert_allclose(got, after[n], rtol=2e-6, atol=2e-6)
        else:
            assert np.array_equal(got, after[n]), n
        if am is not None:
            st = mdl.optimizer.state[p]
            assert float(st["step"]) == 2.0 and len(mdl.optimizer.state) == len(NAMES)
            assert np.array_equal(st["exp_avg"].numpy(), am[n]) and np.array_equal(st["exp_avg_sq"].numpy(), av[n])
        else:
            assert len(mdl.optimizer.state) == 0
    assert mdl.xyz_gradient_accum.shape == (P1, 1) and mdl.denom.shape == (P1, 1) and mdl.max_radii2D.shape == (P1,)
    assert not mdl.xyz_gradient_accum.any() and not mdl.denom.any() and not mdl.max_radii2D.any()
    # the optimizer still steps on the new parameters
    for n in NAMES:
        getattr(mdl, attr[n]).grad = torch.ones_like(getattr(mdl, attr[n]))
    mdl.optimizer.step()


def test_drop_in_statistics_update_matches_the_reference(abi, monkeypatch):
    import torch
    import gsr_densify
    monkeypatch.setattr(gsr_densify, "_C", _EmulatedExtension(abi))
    s = lambda k: GOLD["stats_" + k]
    mdl = _Model()
    mdl.xyz_gradient_accum, mdl.denom = torch.from_numpy(s("accum0").copy()), torch.from_numpy(s("denom0").copy())
    mdl.max_radii2D = torch.from_numpy(s("max_radii0").copy())
    vsp = torch.zeros(s("vs_grad").shape, requires_grad=True)
    vsp.grad = torch.from_numpy(s("vs_grad").copy())
    gsr_densify.add_densification_stats(mdl, vsp, torch.from_numpy(s("radii").copy()))
    assert np.array_equal(mdl.denom.numpy(), s("denom1")) and np.array_equal(mdl.max_radii2D.numpy(), s("max_radii1"))
    np.testing.assert_allclose(mdl.xyz_gradient_accum.numpy(), s("accum1"), rtol=2e-7, atol=0)


@pytest.mark.skipif(bool(os.environ.get("GSR_EMU_ABI_SO")), reason="sanitizer run (tools/emu_sanitize.sh): the real torch extension is not "
                    "instrumented and throws through the preloaded runtime")
def test_drop_in_refuses_host_tensors_without_the_test_double():
    """The real extension has no CPU path: host tensors are rejected, nothing is computed."""
    import torch
    import gsr_densify
    mdl, attr, args, *_ = _model_from_fixture("a", torch.optim.Adam)
    with pytest.raises(RuntimeError, match="CUDA tensors only"):
        gsr_densify.densify_and_prune(mdl, args["max_grad"], args["min_opacity"], args["extent"], 20)
    with pytest.raises(RuntimeError, match="CUDA tensors only"):
        gsr_densify.add_densification_stats(mdl, torch.zeros(mdl._xyz.shape[0], 3), torch.ones(mdl._xyz.shape[0], dtype=torch.int32))


def test_training_loop_harness_runs_on_the_fused_densification(abi, monkeypatch):
    """tools/c3_trainer.py with fused_densify = gsr_densify (the b200 arm of tools/train_c3.py), on the CPU through the test
    double: the loop's densify events and statistics go through the library calls and leave a consistent model that
    grows like the torch arm's (not bit-equal across arms: the 2-element norm differs in the last bit, which moves
    the quantile threshold the harness derives from the statistics)."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import c3_trainer as T
    import gsr_densify
    import gsr_optim
    from test_c3_trainer_cpu import FakeRasterizerModule, _loss, _setup
    monkeypatch.setattr(gsr_densify, "_C", _EmulatedExtension(abi))
    sizes = {}
    for arm, fused in (("torch", None), ("fused", gsr_densify)):
        cam, g = _setup(300)
        g.fused_densify = fused
        extent = float(torch.exp(g.scaling.detach()).max(dim=1).values.median()) / T.OptArgs.percent_dense
        events = []
        T.train(g, [cam], [torch.zeros(3 * 8 * 8)], FakeRasterizerModule, _loss, iterations=31, densify_from=5, densify_every=10,
                target_P=700, extent=extent, bg=torch.zeros(3), timed=False, on_densify=lambda it, p0, p1, thr, st: events.append((it, p0, p1, st)))
        assert [e[0] for e in events] == [10, 20, 30]
        P = g.P
        sizes[arm] = P
        for n, w in zip(T.GROUPS, [(3,), (1, 3), (15, 3), (1,), (3,), (4,)]):
            p = getattr(g, n)
            assert p.shape == (P,) + w and p.requires_grad and isinstance(p, torch.nn.Parameter) and torch.isfinite(p).all()
            grp = [x for x in g.optimizer.param_groups if x["name"] == n][0]
            assert grp["params"][0] is p
            st = g.optimizer.state[p]
            assert st["exp_avg"].shape == p.shape and st["exp_avg_sq"].shape == p.shape
        assert len(g.optimizer.state) == 6
        assert g.xyz_gradient_accum.shape == (P, 1) and g.denom.shape == (P, 1) and g.max_radii2D.shape == (P,)
        assert sum(st[0] for *_, st in events) > 0 and sum(st[1] for *_, st in events) > 0
    assert abs(sizes["fused"] - sizes["torch"]) <= 0.02 * sizes["torch"], sizes
