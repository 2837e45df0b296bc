"""Pins the CPU oracle (oracle/gs_oracle.c) before anything is compared against it:
  * against golden vectors produced by the reference's own torch code (tests/golden/ref_torch_cpu.npz),
  * against golden vectors produced by the reference's CUDA rasterizer on a B200 (tests/golden/ref_cuda_*.npz),
  * against an independent float64 autograd model (tests/torch_ref.py),
  * and on the edge cases the domain has (nothing visible, near-plane culls, off-screen rects, ragged image sizes).
All CPU; runs in well under a minute."""
import glob
import os

import numpy as np
import pytest

import helpers as Hh
from helpers import S, O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


# ---- the reference's torch CPU paths ---------------------------------------------------------------
@pytest.fixture(scope="module")
def gold_cpu():
    return np.load(os.path.join(GOLD, "ref_torch_cpu.npz"))


@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_sh_to_rgb_matches_reference_eval_sh(gold_cpu, deg):
    rgb, clamped = O.sh_to_rgb(gold_cpu["means3D"], gold_cpu["campos"], gold_cpu["shs"], deg)
    ref = gold_cpu["rgb_deg%d" % deg]
    assert np.abs(rgb - ref).max() <= 2e-6          # fp32, different association order
    raw = gold_cpu["rgb_raw_deg%d" % deg]
    sure = np.abs(raw) > 1e-5                        # clamp decisions away from the rounding band
    assert ((raw < 0) == clamped)[sure].all()


@pytest.mark.parametrize("mod", [1.0, 0.7])
def test_cov3d_matches_reference_build_covariance(gold_cpu, mod):
    cov = O.cov3d(gold_cpu["scales"], gold_cpu["rotations"], mod)
    ref = gold_cpu["cov3D_mod%s" % mod]
    scale = np.abs(ref).max(axis=1, keepdims=True)
    assert (np.abs(cov - ref) / scale).max() <= 5e-6


# ---- the reference's CUDA rasterizer (fixtures made on a B200 by tests/golden/make_golden_gpu.py) ----
def _gold_cuda_files():
    return sorted(glob.glob(os.path.join(GOLD, "ref_cuda_*.npz")))


def load_gold_case(path):
    z = np.load(path)
    cam = {k[4:]: (z[k] if z[k].ndim else z[k].item()) for k in z.files if k.startswith("cam_")}
    cam["W"], cam["H"] = int(cam["W"]), int(cam["H"])
    scene = {k[3:]: z[k] for k in z.files if k.startswith("in_")}
    name = os.path.basename(path)[len("ref_cuda_"):-4]
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden_gpu", os.path.join(GOLD, "make_golden_gpu.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    P, W, H, deg, M, use, cov, bg, smod, _ = mg.CASES[name]
    scene["sh_degree"] = deg
    return z, cam, scene, dict(use=use, cov=cov, bg=tuple(float(x) for x in z["bg"]), scale_modifier=smod)


def check_against_gold(z, got, grads_tol=1e-4, image_tol=1e-5, exact_blend=False):
    """The north_star's bar: radii / sorted list / ranges bit-exact, image 1e-5 max-abs, gradients 1e-4 rel L2."""
    assert (got["radii"] == z["out_radii"]).all()
    assert int(got["num_rendered"]) == int(z["out_num_rendered"])
    assert (np.asarray(got["point_list"], np.uint32) == z["out_point_list"]).all()
    assert (np.asarray(got["ranges"], np.uint32) == z["out_ranges"]).all()
    vis = z["out_radii"] > 0
    assert (got["means2D"][vis].view(np.uint32) == z["out_means2D"][vis].view(np.uint32)).all()
    assert (got["conic_opacity"][vis].view(np.uint32) == z["out_conic_opacity"][vis].view(np.uint32)).all()
    assert np.abs(got["color"] - z["out_color"]).max() <= image_tol
    assert np.abs(got["final_T"] - z["out_final_T"]).max() <= 1e-6
    nc_diff = (np.asarray(got["n_contrib"], np.uint32) != z["out_n_contrib"]).mean()
    assert nc_diff == 0 if exact_blend else nc_diff <= 1e-3     # exp() differs in the last ulp host vs device
    if "grads" in got:
        for k in Hh.GRAD_KEYS:
            if "grad_" + k in z.files and got["grads"].get(k) is not None:
                assert Hh.rel_l2(got["grads"][k], z["grad_" + k]) <= grads_tol, k


@pytest.mark.parametrize("path", _gold_cuda_files() or [None])
def test_oracle_matches_reference_cuda_fixtures(path):
    if path is None:
        pytest.skip("no tests/golden/ref_cuda_*.npz yet (made on the GPU box by make_golden_gpu.py)")
    z, cam, scene, kw = load_gold_case(path)
    got = Hh.run_oracle(cam, scene, z["dL"], **kw)
    check_against_gold(z, got)


# ---- independent float64 autograd model ----------------------------------------------------------------
@pytest.mark.parametrize("use,cov,deg", [("sh", "sr", 3), ("sh", "sr", 1), ("colors", "cov3d", 0)])
def test_oracle_gradients_match_float64_autograd(use, cov, deg):
    import torch_ref as TR
    cam = S.look_at_camera(48, 40, eye=(0.2, 0.1, -0.3), target=(0, 0, 5.0))
    M = (deg + 1) ** 2
    scene = S.make_scene(120, cam, sh_degree=deg, M=M, seed=7, s0_pix=2.5)
    rng = np.random.default_rng(3)
    if use == "colors":
        scene["colors_precomp"] = rng.uniform(0, 1, (120, 3)).astype(np.float32)
    if cov == "cov3d":
        scene["cov3D_precomp"] = O.cov3d(scene["scales"], scene["rotations"], 1.0)
    dL = S.upstream_grad(cam, seed=4)
    bg = (0.1, 0.7, 0.4)
    got = Hh.run_oracle(cam, scene, dL, bg=bg, use=use, cov=cov)
    ref = TR.render(cam, scene, bg, use=use, cov=cov, dL=dL)
    assert (got["radii"].astype(np.float64) == ref["radii_f"]).mean() > 0.99   # float32 vs float64 ceil() edge
    assert np.abs(got["color"] - ref["color"]).max() <= 2e-5
    for k, tol in (("means2D", 2e-3), ("opacities", 2e-3), ("means3D", 5e-3), ("shs", 2e-3), ("colors_precomp", 2e-3),
                   ("scales", 5e-3), ("rotations", 5e-3), ("cov3D_precomp", 5e-3)):
        if ref["grads"].get(k) is None or got["grads"].get(k) is None:
            continue
        g = got["grads"][k]
        r = ref["grads"][k]
        if k == "shs":
            r = r[:, :M]
            g = g[:, :M]
        assert Hh.rel_l2(g, r) <= tol, (k, Hh.rel_l2(g, r))


# ---- structure and edge cases ------------------------------------------------------------------------------
def test_binning_invariants():
    cam = S.make_camera(250, 130)
    scene = S.make_scene(5000, cam, sh_degree=2, M=16, seed=1)
    f = Hh.run_oracle(cam, scene)
    R = f["num_rendered"]
    assert R == int(f["tiles_touched"].sum())
    keys = f["keys"]
    assert (np.diff(keys.astype(np.uint64)) >= 0).all()                 # sorted by (tile, depth bits)
    tiles = (keys >> np.uint64(32)).astype(np.int64)
    gx, gy = (250 + 15) // 16, (130 + 15) // 16
    assert tiles.max() < gx * gy
    rg = f["ranges"].astype(np.int64)
    for t in range(gx * gy):
        b, e = rg[t]
        if e > b:
            assert (tiles[b:e] == t).all()
            assert (b == 0 or tiles[b - 1] != t) and (e == R or tiles[e] != t)
        else:
            assert (tiles != t).all() and b == 0 and e == 0                # untouched tiles keep {0,0}
    # equal depths keep ascending Gaussian index (stable sort, emission order)
    d = f["depths"][f["point_list"]]
    same = (tiles[1:] == tiles[:-1]) & (d[1:] == d[:-1])
    assert (f["point_list"][1:][same] > f["point_list"][:-1][same]).all()
    # culled Gaussians emit nothing
    assert (f["radii"][scene["means3D"][:, 2] <= 0.2] == 0).all()


def test_nothing_visible_gives_background():
    cam = S.make_camera(40, 24)
    scene = S.make_scene(50, cam, sh_degree=0, M=1, seed=2)
    scene["means3D"][:, 2] = -1.0
    bg = (0.2, 0.3, 0.4)
    f = Hh.run_oracle(cam, scene, S.upstream_grad(cam), bg=bg)
    assert f["num_rendered"] == 0 and (f["radii"] == 0).all() and (f["ranges"] == 0).all()
    for ch in range(3):
        assert np.allclose(f["color"][ch], bg[ch])
    assert all(np.abs(g).max() == 0 for g in f["grads"].values() if g is not None and g.size)


def test_mark_visible_is_the_near_plane_test():
    cam = S.make_camera(64, 64)
    m = np.array([[0, 0, 0.2], [0, 0, 0.2000001], [0, 0, 5], [100, 100, 1], [0, 0, -3]], np.float32)
    vis = O.mark_visible(m, cam["viewmatrix"])
    assert vis.tolist() == [False, bool(np.float32(0.2000001) > np.float32(0.2)), True, True, False]


def test_higher_msb():
    L = O.lib()
    # rasterizer_impl.cu:37-52: number of key bits needed for the tile count
    for n in (1, 2, 255, 256, 257, 8160, 32400, 65535, 65536):
        bits = int(L.gso_higher_msb(n))
        assert (1 << bits) > n - 1 or n == 1
        assert bits <= 32
