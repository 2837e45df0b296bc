"""GPU parity of the SURVEY 8(f) rows built after the rasterizer: the fused Adam step (row 2) and distCUDA2 (row 4).

Written after round 1's GPU minutes were spent; first ran on a B200 in the driver's round-1 GPU pass (all but the
distCUDA2 cases passed; those failed on a module-name mismatch of the reference simple-knn build, fixed in
oracle/build_ref.py).  Plain gpu tests since round 2: a failure here makes `pytest -m gpu -x` non-zero.
They go through the public Python API -> torch binding -> C ABI.

Bars.  Adam: parameters and moments within 1e-6 relative to each tensor's largest magnitude of torch.optim.Adam's (the
reference's optimizer; the CPU fixture and torch's own CUDA kernels) and bit-exact against the oracle.  distCUDA2:
bit-exact against the oracle and, when oracle/_ref holds it, against the reference's own simple-knn build."""
import importlib.util
import os
import sys

import numpy as np
import pytest
import torch

from helpers import ROOT

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600, method="thread")]
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_adam_cpu.npz")
NAMES = ["xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation"]
SHAPES = {"xyz": (3,), "f_dc": (1, 3), "f_rest": (15, 3), "opacity": (1,), "scaling": (3,), "rotation": (4,)}


def _close(a, b, tol=1e-6):
    a = a.detach().cpu().numpy() if hasattr(a, "detach") else np.asarray(a)
    b = b.detach().cpu().numpy() if hasattr(b, "detach") else np.asarray(b)
    assert a.shape == b.shape
    return float(np.abs(a - b).max()) <= tol * max(float(np.abs(b).max()), 1e-30)


def _ref_knn():
    so = os.path.join(ROOT, "oracle", "_ref", "simple_knn", "ref_simple_knn_C.so")
    if not os.path.exists(so):
        return None
    spec = importlib.util.spec_from_file_location("ref_simple_knn_C", so)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


# ------------------------------------------------------------------------------------------------ Adam
def test_fused_adam_follows_the_golden_trajectory():
    import gsr_optim
    z = np.load(GOLD)
    steps = int(z["steps"])
    params = {n: torch.nn.Parameter(torch.tensor(z["init_" + n], device="cuda")) for n in NAMES}
    groups = [{"params": [params[n]], "lr": float(z["lr_" + n]), "name": n} for n in NAMES]
    opt = gsr_optim.FusedAdam(groups, lr=0.0, eps=1e-15)
    launches0 = _launches()
    for it in range(steps):
        if it == 3:      # cat_tensors_to_optimizer (scene/gaussian_model.py:447-471), as the fixture's generator did
            for group in opt.param_groups:
                ext = torch.tensor(z["ext_" + group["name"]], device="cuda")
                st = opt.state.get(group["params"][0], None)
                st["exp_avg"] = torch.cat((st["exp_avg"], torch.zeros_like(ext)), dim=0)
                st["exp_avg_sq"] = torch.cat((st["exp_avg_sq"], torch.zeros_like(ext)), dim=0)
                del opt.state[group["params"][0]]
                group["params"][0] = torch.nn.Parameter(torch.cat((group["params"][0].detach(), ext), dim=0).requires_grad_(True))
                opt.state[group["params"][0]] = st
        for group in opt.param_groups:
            if group["name"] == "xyz":
                group["lr"] = float(z["xyz_lr_step%d" % it])
            group["params"][0].grad = torch.tensor(z["grad_%s_step%d" % (group["name"], it)], device="cuda")
        opt.step()
        opt.zero_grad(set_to_none=True)
        for group in opt.param_groups:
            n = group["name"]
            st = opt.state[group["params"][0]]
            assert float(st["step"]) == it + 1
            assert _close(group["params"][0], z["param_%s_step%d" % (n, it)]), (n, it, "param")
            assert _close(st["exp_avg"], z["exp_avg_%s_step%d" % (n, it)]), (n, it, "exp_avg")
            assert _close(st["exp_avg_sq"], z["exp_avg_sq_%s_step%d" % (n, it)]), (n, it, "exp_avg_sq")
    assert _launches() - launches0 == steps          # ONE kernel per optimizer step, all six groups


def _launches():
    import ctypes as C
    lib = C.CDLL(os.path.join(ROOT, "gaussian-splatting_cc-comments_b200", "libgsr_b200.so"))
    lib.gsr_launch_count.restype = C.c_ulonglong
    return int(lib.gsr_launch_count())


@pytest.mark.parametrize("P", [1, 1000, 200_003])
def test_fused_adam_equals_torch_adam_on_the_gpu_and_the_oracle(P):
    import gsr_optim
    from oracle import gs_next_oracle as O
    g = torch.Generator(device="cpu").manual_seed(P)
    init = {n: torch.randn(P, *SHAPES[n], generator=g) for n in NAMES}
    lrs = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=1.25e-4, opacity=0.05, scaling=0.005, rotation=0.001)

    def make(cls):
        ps = {n: torch.nn.Parameter(init[n].clone().cuda()) for n in NAMES}
        return ps, cls([{"params": [ps[n]], "lr": lrs[n], "name": n} for n in NAMES], lr=0.0, eps=1e-15)
    mine_p, mine = make(gsr_optim.FusedAdam)
    ref_p, ref = make(torch.optim.Adam)
    cpu = {n: [init[n].numpy().copy().reshape(-1), np.zeros(init[n].numel(), np.float32), np.zeros(init[n].numel(), np.float32)] for n in NAMES}
    for it in range(4):
        for n in NAMES:
            gr = torch.randn(P, *SHAPES[n], generator=g) * (10.0 ** torch.randint(-6, 1, (P,) + (1,) * len(SHAPES[n]), generator=g).float())
            gr[torch.rand(P, generator=g) < 0.3] = 0.0
            mine_p[n].grad = gr.cuda()
            ref_p[n].grad = gr.cuda()
            O.adam_step(cpu[n][0], gr.numpy().reshape(-1), cpu[n][1], cpu[n][2], lrs[n], it + 1)
        mine.step()
        ref.step()
        for n in NAMES:
            assert _close(mine_p[n], ref_p[n]), (n, it)
            assert _close(mine.state[mine_p[n]]["exp_avg"], ref.state[ref_p[n]]["exp_avg"]), (n, it)
            assert _close(mine.state[mine_p[n]]["exp_avg_sq"], ref.state[ref_p[n]]["exp_avg_sq"]), (n, it)
            assert np.array_equal(mine_p[n].detach().cpu().numpy().reshape(-1), cpu[n][0]), (n, it, "oracle")
            assert np.array_equal(mine.state[mine_p[n]]["exp_avg_sq"].cpu().numpy().reshape(-1), cpu[n][2]), (n, it, "oracle")


def test_fused_adam_unaligned_views_and_visibility_mask():
    from diff_gaussian_rasterization import _C
    from oracle import gs_next_oracle as O
    g = torch.Generator().manual_seed(5)
    P, w = 1203, 3
    raw = [torch.randn(P * w + 1, generator=g).cuda() for _ in range(2)] + [torch.rand(P * w + 1, generator=g).cuda() * 0.01 for _ in range(2)]
    p, gr, m, v = [t[1:].view(P, w) for t in raw]                    # 4-byte aligned only: the element-wise kernel
    v.copy_(torch.rand(P, w, generator=g) * 0.01)
    host = [t.cpu().numpy().copy().reshape(-1) for t in (p, gr, m, v)]
    _C.adam_step([p], [gr], [m], [v], [0.01], [3], 0.9, 0.999, 1e-15)
    O.adam_step(host[0], host[1], host[2], host[3], 0.01, 3)
    assert np.array_equal(p.cpu().numpy().reshape(-1), host[0]) and np.array_equal(m.cpu().numpy().reshape(-1), host[2])
    # visibility mask: rows with 0 keep parameters and both moments
    vis = torch.rand(P, generator=g) < 0.5
    before = [t.clone() for t in (p, m, v)]
    _C.adam_step([p], [gr], [m], [v], [0.01], [4], 0.9, 0.999, 1e-15, vis.cuda())
    O.adam_step(host[0], host[1], host[2], host[3], 0.01, 4)
    rows = vis.numpy().repeat(w)
    for t, b, h in zip((p, m, v), before, (host[0], host[2], host[3])):
        assert np.array_equal(t.cpu().numpy().reshape(-1), np.where(rows, h, b.cpu().numpy().reshape(-1)))
    with pytest.raises(RuntimeError):
        _C.adam_step([p], [gr], [m], [v], [0.01], [0], 0.9, 0.999, 1e-15)          # step must be >= 1


# ------------------------------------------------------------------------------------------------ distCUDA2
def _points(P, seed):
    rng = np.random.default_rng(seed)
    kind = seed % 3
    if kind == 0:
        p = rng.normal(size=(P, 3))
    elif kind == 1:
        p = np.concatenate([rng.normal(size=(P // 2, 3)) * 0.02 - 4.0, rng.uniform(-30, 30, size=(P - P // 2, 3))])
    else:
        p = rng.uniform(size=(P, 3)) * np.array([100.0, 1.0, 0.01])
    p = p.astype(np.float32)
    if P >= 100:
        p[:20] = p[20:40]
    return p


@pytest.mark.parametrize("P,seed", [(1, 0), (2, 1), (3, 2), (4, 3), (255, 4), (256, 5), (257, 6), (5000, 7), (100_000, 8), (300_001, 9)])
def test_dist_cuda2_is_bit_exact_against_the_oracle_and_the_reference_build(P, seed):
    from simple_knn._C import distCUDA2
    from oracle import gs_next_oracle as O
    pts = _points(P, seed)
    dev = torch.tensor(pts, device="cuda")
    got = distCUDA2(dev)
    assert got.shape == (P,) and got.dtype == torch.float32 and got.is_cuda
    want = O.knn_mean_dist2(pts)
    assert np.array_equal(got.cpu().numpy(), want)
    ref = _ref_knn()
    if ref is not None:
        assert np.array_equal(ref.distCUDA2(dev).cpu().numpy(), got.cpu().numpy())


def test_dist_cuda2_reproduces_the_reference_builds_fixture_bit_for_bit():
    """tests/golden/ref_knn_cuda.npz: written on a B200 by the reference's own simple-knn (tests/golden/make_golden_knn_gpu.py)."""
    from simple_knn._C import distCUDA2
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_knn_cuda.npz"))
    names = [k[4:] for k in z.files if k.startswith("pts_")]
    assert len(names) == 6
    for n in names:
        got = distCUDA2(torch.tensor(z["pts_" + n], device="cuda")).cpu().numpy()
        assert np.array_equal(got, z["d2_" + n]), n


def test_dist_cuda2_empty_non_contiguous_and_wrong_inputs():
    from simple_knn._C import distCUDA2
    assert distCUDA2(torch.zeros(0, 3, device="cuda")).shape == (0,)
    base = torch.randn(500, 6, device="cuda")
    view = base[:, ::2]
    assert torch.equal(distCUDA2(view), distCUDA2(view.contiguous()))
    with pytest.raises(RuntimeError):
        distCUDA2(torch.zeros(5, 3))
    with pytest.raises(RuntimeError):
        distCUDA2(torch.zeros(5, 2, device="cuda"))
    # the expression of scene/gaussian_model.py:207-208 stays finite for coincident points
    d2 = torch.clamp_min(distCUDA2(torch.zeros(300, 3, device="cuda")), 0.0000001)
    assert torch.isfinite(torch.log(torch.sqrt(d2))).all()


def test_dist_cuda2_one_million_points_against_a_sampled_exact_search():
    from scipy.spatial import cKDTree
    from simple_knn._C import distCUDA2
    pts = _points(1_000_000, 10)
    got = distCUDA2(torch.tensor(pts, device="cuda")).cpu().numpy()
    idx = np.random.default_rng(0).choice(len(pts), 20000, replace=False)
    p64 = pts.astype(np.float64)
    d, _ = cKDTree(p64).query(p64[idx], k=4)
    want = (d[:, 1:] ** 2).mean(axis=1)
    assert np.all(np.abs(got[idx] - want) <= 1e-5 * np.maximum(want, 1e-12) + 1e-12)


# ------------------------------------------------------------------------------------------------ raw-parameter path (row 1)
def _raw_model(P, seed, W=160, H=96, sh_degree=3, M=16):
    from helpers import S
    cam = S.look_at_camera(W, H, eye=(0.2, -0.1, -0.4), target=(0.0, 0.0, 5.0))
    scene = S.make_scene(P, cam, sh_degree=sh_degree, M=M, seed=seed, s0_pix=3.0)
    t = lambda a: torch.tensor(a, device="cuda")
    g = torch.Generator().manual_seed(seed)
    raw = dict(xyz=t(scene["means3D"]), f_dc=t(scene["shs"][:, :1, :].copy()), f_rest=t(scene["shs"][:, 1:, :].copy()),
               opacity=torch.logit(t(scene["opacities"]).clamp(1e-4, 1 - 1e-4)), scaling=torch.log(t(scene["scales"])),
               rotation=t(scene["rotations"]) * (0.2 + 3.0 * torch.rand(P, 1, generator=g).cuda()))      # NOT unit length
    return cam, {k: v.contiguous().requires_grad_(True) for k, v in raw.items()}


def _settings(mod, cam, sh_degree):
    t = lambda a: torch.tensor(np.asarray(a), dtype=torch.float32, device="cuda")
    return mod.GaussianRasterizationSettings(image_height=cam["H"], image_width=cam["W"], tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
                                             bg=t([0.1, 0.2, 0.3]), scale_modifier=1.0, viewmatrix=t(cam["viewmatrix"]),
                                             projmatrix=t(cam["projmatrix"]), sh_degree=sh_degree, campos=t(cam["campos"]),
                                             prefiltered=False, debug=False)


@pytest.mark.parametrize("P,deg,M", [(5000, 3, 16), (1, 3, 16), (777, 1, 16), (300, 0, 16), (2000, 2, 9), (900, 0, 1), (129, 1, 4)])
def test_raw_parameter_path_equals_torch_activations_plus_the_rasterizer(P, deg, M):
    """forward_raw(raw tensors) against forward(sigmoid / exp / normalize / cat of the same tensors) through torch autograd.
    Bars: radii equal (at most 2 Gaussians may sit on a rounding boundary of the quaternion norm), image <= 1e-5 max-abs,
    every raw-parameter gradient <= 1e-4 relative L2 (north_star's bars for the rasterizer itself)."""
    import diff_gaussian_rasterization as mod
    from helpers import rel_l2
    cam, raw = _raw_model(P, seed=P + deg, sh_degree=deg, M=M)
    rast = mod.GaussianRasterizer(_settings(mod, cam, deg))
    w = torch.rand(3, cam["H"], cam["W"], device="cuda") * 2 - 1

    def run(fused):
        for v in raw.values():
            v.grad = None
        m2d = torch.zeros(P, 3, device="cuda", requires_grad=True)
        if fused:
            img, radii = rast.forward_raw(raw["xyz"], m2d, raw["f_dc"], raw["f_rest"], raw["opacity"], raw["scaling"], raw["rotation"])
        else:
            img, radii = rast(means3D=raw["xyz"], means2D=m2d, shs=torch.cat((raw["f_dc"], raw["f_rest"]), dim=1),
                              opacities=torch.sigmoid(raw["opacity"]), scales=torch.exp(raw["scaling"]),
                              rotations=torch.nn.functional.normalize(raw["rotation"]))
        (img * w).sum().backward()
        grads = {k: (v.grad.detach().clone() if v.grad is not None else torch.zeros_like(v)) for k, v in raw.items()}
        grads["means2D"] = m2d.grad.detach().clone()
        return img.detach(), radii, grads
    img_f, rad_f, g_f = run(True)
    img_u, rad_u, g_u = run(False)
    assert int((rad_f != rad_u).sum()) <= 2
    assert float((img_f - img_u).abs().max()) <= 1e-5
    for k in g_u:
        assert g_f[k].shape == g_u[k].shape
        if g_u[k].numel() == 0:
            continue
        if float(g_u[k].abs().max()) == 0.0:
            assert float(g_f[k].abs().max()) == 0.0, k
        else:
            assert rel_l2(g_f[k].cpu().numpy(), g_u[k].cpu().numpy()) <= 1e-4, k


def test_fused_render_is_a_drop_in_for_the_reference_render():
    import gsr_render
    from helpers import S
    P = 3000
    cam, raw = _raw_model(P, seed=4)

    class PC:          # the attributes of GaussianModel that render() touches (scene/gaussian_model.py:44-138)
        active_sh_degree = 3
        max_sh_degree = 3
        _xyz, _features_dc, _features_rest = raw["xyz"], raw["f_dc"], raw["f_rest"]
        _opacity, _scaling, _rotation = raw["opacity"], raw["scaling"], raw["rotation"]

        def __init__(self):      # instance attributes, as GaussianModel.setup_functions sets them (a class attribute would bind)
            self.scaling_activation, self.opacity_activation, self.rotation_activation = torch.exp, torch.sigmoid, torch.nn.functional.normalize

        get_xyz = property(lambda s: s._xyz)
        get_scaling = property(lambda s: s.scaling_activation(s._scaling))
        get_rotation = property(lambda s: s.rotation_activation(s._rotation))
        get_opacity = property(lambda s: s.opacity_activation(s._opacity))
        get_features = property(lambda s: torch.cat((s._features_dc, s._features_rest), dim=1))

    class Pipe:
        convert_SHs_python = False
        compute_cov3D_python = False
        debug = False

    class View:
        FoVx = 2 * np.arctan(cam["tanfovx"]); FoVy = 2 * np.arctan(cam["tanfovy"])
        image_height, image_width = cam["H"], cam["W"]
        world_view_transform = torch.tensor(cam["viewmatrix"], device="cuda")
        full_proj_transform = torch.tensor(cam["projmatrix"], device="cuda")
        camera_center = torch.tensor(cam["campos"], device="cuda")

    bg = torch.tensor([0.0, 0.0, 0.0], device="cuda")
    pc = PC()
    assert gsr_render._default_activations(pc)                # so this call takes the fused route (raw tensors into the kernels)
    fused = gsr_render.render(View, pc, Pipe, bg)
    assert set(fused) == {"render", "viewspace_points", "visibility_filter", "radii"}
    fused["render"].sum().backward()
    vs_grad = fused["viewspace_points"].grad.clone()
    g_fused = {k: v.grad.clone() for k, v in raw.items()}
    for v in raw.values():
        v.grad = None
    pc.scaling_activation = lambda x: torch.exp(x)            # not the default object: render() takes the reference's route
    assert not gsr_render._default_activations(pc)
    plain = gsr_render.render(View, pc, Pipe, bg)
    plain["render"].sum().backward()
    assert float((fused["render"] - plain["render"]).abs().max()) <= 1e-5
    assert int((fused["radii"] != plain["radii"]).sum()) <= 2
    assert torch.equal(fused["visibility_filter"], fused["radii"] > 0)
    assert float((vs_grad - plain["viewspace_points"].grad).norm() / plain["viewspace_points"].grad.norm()) <= 1e-4
    for k, v in raw.items():
        assert float((g_fused[k] - v.grad).norm() / v.grad.norm().clamp_min(1e-30)) <= 1e-4, k
    Pipe.convert_SHs_python = True                             # the python SH route (gsr_sh.eval_sh) gives the same picture
    py_sh = gsr_render.render(View, pc, Pipe, bg)
    assert float((py_sh["render"] - plain["render"]).abs().max()) <= 2e-5


# ------------------------------------------------------------------------------------------------ densification (row 2, second half)
DGOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_densify_cpu.npz")
ATTR = {"xyz": "_xyz", "f_dc": "_features_dc", "f_rest": "_features_rest", "opacity": "_opacity", "scaling": "_scaling", "rotation": "_rotation"}


class _DModel:
    percent_dense = 0.01


def _dmodel(tensors, moments, accum, denom, max_radii, percent_dense, opt_cls):
    m = _DModel()
    for n in NAMES:
        setattr(m, ATTR[n], torch.nn.Parameter(torch.as_tensor(tensors[n], device="cuda").clone().contiguous()))
    m.percent_dense = percent_dense
    m.optimizer = opt_cls([{"params": [getattr(m, ATTR[n])], "lr": 1e-3, "name": n} for n in NAMES], lr=0.0, eps=1e-15)
    if moments is not None:
        for n in NAMES:
            m.optimizer.state[getattr(m, ATTR[n])] = {"step": torch.tensor(2.0), "exp_avg": torch.as_tensor(moments[0][n], device="cuda").clone(),
                                                      "exp_avg_sq": torch.as_tensor(moments[1][n], device="cuda").clone()}
    m.xyz_gradient_accum = torch.as_tensor(accum, device="cuda").reshape(-1, 1).clone()
    m.denom = torch.as_tensor(denom, device="cuda").reshape(-1, 1).clone()
    m.max_radii2D = torch.as_tensor(max_radii, device="cuda").reshape(-1).clone()
    return m


@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_densify_and_prune_reproduces_the_reference_models_fixture(tag, monkeypatch):
    """The reference's GaussianModel.densify_and_prune, run by tests/golden/make_golden_densify.py, against the kernels:
    the fixture's recorded samples are injected in place of the random draw; selections, order, copied rows and moments
    exact, computed children (xyz, scaling) to 2e-6."""
    import gsr_densify
    z = np.load(DGOLD)
    g = lambda k: z[f"{tag}_{k}"]
    has = bool(g("with_moments"))
    before = {n: g("before_p_" + n) for n in NAMES}
    mom = ({n: g("before_m_" + n) for n in NAMES}, {n: g("before_v_" + n) for n in NAMES}) if has else None
    m = _dmodel(before, mom, g("before_accum"), g("before_denom"), g("before_max_radii2D"), float(g("percent_dense")), torch.optim.Adam)
    samples = torch.tensor(g("samples"), device="cuda")
    monkeypatch.setattr(gsr_densify.torch, "normal", lambda mean, std, generator=None: samples)
    launches0 = _launches()
    P0, P1, n_clone, n_split = gsr_densify.densify_and_prune(m, float(g("max_grad")), float(g("min_opacity")), float(g("extent")),
                                                             float(g("max_screen_size")) or None)
    assert _launches() - launches0 == 3                        # classify, scan, apply
    assert P1 == g("after_p_xyz").shape[0] and 2 * n_split == samples.shape[0]
    for n in NAMES:
        p = getattr(m, ATTR[n])
        got, want = p.detach().cpu().numpy(), g("after_p_" + n)
        assert got.shape == want.shape
        if n in ("xyz", "scaling"):
            np.testing.assert_allclose(got, want, rtol=2e-6, atol=2e-6)
        else:
            assert np.array_equal(got, want), n
        if has:
            st = m.optimizer.state[p]
            assert np.array_equal(st["exp_avg"].cpu().numpy(), g("after_m_" + n)) and np.array_equal(st["exp_avg_sq"].cpu().numpy(), g("after_v_" + n))
    assert m.xyz_gradient_accum.shape == (P1, 1) and not m.xyz_gradient_accum.any() and not m.denom.any() and not m.max_radii2D.any()


def _torch_reference_arm():
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import c3_trainer as T
    return T


@pytest.mark.parametrize("P,screen", [(100_000, 20), (333_333, None)])
def test_densify_and_prune_equals_the_torch_sequence_on_the_gpu_with_the_same_seed(P, screen):
    """Same seed, same torch.normal call: the three-round torch sequence (tools/c3_trainer.py, the restatement of
    scene/gaussian_model.py:366-597 used as the reference arm of the C3 harness) and the two library calls must give the
    same model -- every tensor and both moments of every group, same row order."""
    import gsr_densify
    import gsr_optim
    from helpers import S
    T = _torch_reference_arm()
    cam = S.make_camera(64, 64)
    scene = S.make_scene(P, cam, sh_degree=3, seed=3)
    arms = {}
    for arm, fused in (("torch", None), ("fused", gsr_densify)):
        g = T.Gaussians(scene, "cuda", torch.optim.Adam, gsr_optim.get_expon_lr_func, fused_densify=fused)
        gen = torch.Generator(device="cuda").manual_seed(5)
        for n in T.GROUPS:                                     # populate the moments
            getattr(g, n).grad = torch.randn(getattr(g, n).shape, device="cuda", generator=gen) * 1e-3
        g.optimizer.step()
        g.denom = (torch.randint(1, 40, (P, 1), device="cuda", generator=gen).float() * (torch.rand(P, 1, device="cuda", generator=gen) > 0.15))
        g.xyz_gradient_accum = torch.rand(P, 1, device="cuda", generator=gen) * 6e-4 * g.denom
        extent = float(torch.exp(g.scaling.detach()).max(dim=1).values.median()) / T.OptArgs.percent_dense
        torch.manual_seed(77)
        g.densify_and_prune(0.0002, 0.005, extent, screen)
        arms[arm] = g
    a, b = arms["torch"], arms["fused"]
    assert a.P == b.P and a.P != P
    for n in T.GROUPS:
        pa, pb = getattr(a, n).detach(), getattr(b, n).detach()
        if n in ("xyz", "scaling"):
            assert torch.allclose(pa, pb, rtol=2e-6, atol=2e-6), n
            assert float((pa != pb).float().mean()) < 0.6          # the copied rows are identical
        else:
            assert torch.equal(pa, pb), n
        sa, sb = a.optimizer.state[getattr(a, n)], b.optimizer.state[getattr(b, n)]
        assert torch.equal(sa["exp_avg"], sb["exp_avg"]) and torch.equal(sa["exp_avg_sq"], sb["exp_avg_sq"]), n
    assert torch.equal(a.denom, b.denom) and torch.equal(a.max_radii2D, b.max_radii2D)


def test_densification_statistics_kernel_against_torch_and_the_fixture():
    import gsr_densify
    z = np.load(DGOLD)
    s = lambda k: z["stats_" + k]
    m = _DModel()
    m.xyz_gradient_accum, m.denom = torch.tensor(s("accum0"), device="cuda"), torch.tensor(s("denom0"), device="cuda")
    m.max_radii2D = torch.tensor(s("max_radii0"), device="cuda")
    vsp = torch.zeros(s("vs_grad").shape, device="cuda", requires_grad=True)
    vsp.grad = torch.tensor(s("vs_grad"), device="cuda")
    launches0 = _launches()
    gsr_densify.add_densification_stats(m, vsp, torch.tensor(s("radii"), device="cuda"))
    assert _launches() - launches0 == 1
    assert np.array_equal(m.denom.cpu().numpy(), s("denom1")) and np.array_equal(m.max_radii2D.cpu().numpy(), s("max_radii1"))
    np.testing.assert_allclose(m.xyz_gradient_accum.cpu().numpy(), s("accum1"), rtol=2e-7, atol=0)
    # a large model against the torch statements of train.py:153-158 on the GPU
    P = 1_000_003
    gen = torch.Generator(device="cuda").manual_seed(1)
    accum, denom = torch.rand(P, 1, device="cuda", generator=gen), torch.randint(0, 9, (P, 1), device="cuda", generator=gen).float()
    mr = torch.rand(P, device="cuda", generator=gen) * 30
    radii = (torch.randint(-2, 60, (P,), device="cuda", generator=gen) * (torch.rand(P, device="cuda", generator=gen) > 0.3)).int()
    grad = torch.randn(P, 3, device="cuda", generator=gen) * 1e-3
    m.xyz_gradient_accum, m.denom, m.max_radii2D = accum.clone(), denom.clone(), mr.clone()
    gsr_densify.add_densification_stats(m, grad, radii)
    vis = radii > 0
    mr[vis] = torch.max(mr[vis], radii[vis].float())
    accum[vis] += torch.norm(grad[vis, :2], dim=-1, keepdim=True)
    denom[vis] += 1
    assert torch.equal(m.denom, denom) and torch.equal(m.max_radii2D, mr)
    assert torch.allclose(m.xyz_gradient_accum, accum, rtol=2e-7, atol=0)


def test_densify_rejects_host_tensors():
    import gsr_densify
    m = _DModel()
    with pytest.raises(RuntimeError, match="CUDA tensors only"):
        from diff_gaussian_rasterization import _C
        _C.densify_stats(torch.zeros(4, 3), torch.ones(4, dtype=torch.int32), torch.zeros(4), torch.zeros(4), None)
