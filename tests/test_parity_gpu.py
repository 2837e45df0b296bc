"""Parity of the sm_100a CUDA path (through the drop-in Python API -> torch binding -> C ABI) against
  (1) the UNMODIFIED reference CUDA rasterizer built for sm_100 (oracle/_ref) on identical inputs,
  (2) the committed golden fixtures the reference produced (tests/golden/ref_cuda_*.npz),
  (3) the CPU oracle.
Bars (north_star): radii, num_rendered, sorted point list, tile ranges, n_contrib bit-exact;
image <= 1e-5 max-abs; every gradient <= 1e-4 relative L2.
"""
import glob
import os

import numpy as np
import pytest

import helpers as Hh
from helpers import S, O

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
IMAGE_TOL = 1e-5          # max-abs, fp32
GRAD_TOL = 1e-4           # relative L2


@pytest.fixture(scope="module")
def mod():
    import torch
    assert torch.cuda.is_available()
    m = Hh.my_module()
    # the CUDA extension must be the thing that runs: no fallback exists, and the .so must be in-tree
    assert os.path.dirname(m._C.__file__).endswith("diff_gaussian_rasterization")
    return m


def assert_structural_parity(mine, ref):
    assert (mine["radii"] == ref["radii"]).all()
    assert int(mine["num_rendered"]) == int(ref["num_rendered"])
    assert (np.asarray(mine["point_list"], np.uint32) == np.asarray(ref["point_list"], np.uint32)).all()
    assert (np.asarray(mine["ranges"], np.uint32) == np.asarray(ref["ranges"], np.uint32)).all()


def assert_full_parity(mine, ref, exact_blend, grad_tol=GRAD_TOL):
    assert_structural_parity(mine, ref)
    vis = ref["radii"] > 0
    assert (mine["means2D"][vis].view(np.uint32) == ref["means2D"][vis].view(np.uint32)).all()
    assert (mine["conic_opacity"][vis].view(np.uint32) == ref["conic_opacity"][vis].view(np.uint32)).all()
    if exact_blend:                                    # GPU vs GPU: same MUFU.EX2, same contraction
        assert (np.asarray(mine["n_contrib"], np.uint32) == np.asarray(ref["n_contrib"], np.uint32)).all()
    else:
        assert (np.asarray(mine["n_contrib"], np.uint32) != np.asarray(ref["n_contrib"], np.uint32)).mean() <= 1e-3
    assert np.abs(mine["color"] - ref["color"]).max() <= IMAGE_TOL
    assert np.abs(mine["final_T"] - ref["final_T"]).max() <= 1e-6
    if "grads" in mine and "grads" in ref:
        for k in Hh.GRAD_KEYS:
            a, b = mine["grads"].get(k), ref["grads"].get(k)
            if a is not None and b is not None:
                assert Hh.rel_l2(a, b) <= grad_tol, (k, Hh.rel_l2(a, b))


def variant(scene, use, cov, seed=5):
    P = scene["means3D"].shape[0]
    if use == "colors":
        scene["colors_precomp"] = np.random.default_rng(seed).uniform(0, 1, (P, 3)).astype(np.float32)
    if cov == "cov3d":
        scene["cov3D_precomp"] = O.cov3d(scene["scales"], scene["rotations"], 1.0)
    return scene


CASES = [
    # name, P, W, H, degree, M, use, cov, bg, scale_modifier
    ("c1_10k_256_deg0", 10000, 256, 256, 0, 1, "sh", "sr", (0, 0, 0), 1.0),            # BASELINE configs[0]
    ("c1_random_bg", 10000, 256, 256, 0, 1, "sh", "sr", (0.3, 0.5, 0.7), 1.0),
    ("sh3_100k_640x360", 100000, 640, 360, 3, 16, "sh", "sr", (0.3, 0.5, 0.7), 1.0),
    ("ragged_250x130_deg2", 20000, 250, 130, 2, 16, "sh", "sr", (1, 1, 1), 1.0),        # W,H not multiples of 16; M > K
    ("deg1_M4_mod", 5000, 96, 64, 1, 4, "sh", "sr", (0, 0, 0), 0.5),
    ("precomputed_colour_and_cov", 8000, 200, 120, 0, 1, "colors", "cov3d", (0.2, 0.2, 0.2), 1.0),
    ("precomputed_colour_sr", 8000, 200, 120, 0, 1, "colors", "sr", (0, 0, 0), 1.0),
    ("sh_with_cov3d", 8000, 200, 120, 3, 16, "sh", "cov3d", (0, 0, 0), 1.0),
    ("tiny_1_gaussian", 1, 32, 32, 0, 1, "sh", "sr", (0, 0, 0), 1.0),
    ("dense_big_splats", 3000, 128, 128, 3, 16, "sh", "sr", (0, 0, 0), 1.0),
]


def build_case(c):
    name, P, W, H, deg, M, use, cov, bg, smod = c
    cam = S.make_camera(W, H) if "640" not in name else S.look_at_camera(W, H, eye=(0.3, -0.2, -0.5), target=(0, 0, 5.0))
    s0 = 12.0 if name == "dense_big_splats" else None
    scene = variant(S.make_scene(P, cam, sh_degree=deg, M=M, seed=0, s0_pix=s0, near_frac=0.0 if P == 1 else 0.02), use, cov)
    return cam, scene, S.upstream_grad(cam), dict(use=use, cov=cov, bg=bg, scale_modifier=smod)


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_matches_reference_cuda_build(mod, case):
    if not Hh.have_ref():
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    cam, scene, dL, kw = build_case(case)
    mine = Hh.run_torch(mod, cam, scene, dL, **kw)
    ref = Hh.run_torch(Hh.ref_module(), cam, scene, dL, is_ref=True, **kw)
    assert_full_parity(mine, ref, exact_blend=True)
    # means2D gradient: (P,3) with z == 0 (densification reads [:, :2], scene/gaussian_model.py:601)
    assert mine["grads"]["means2D"].shape == (case[1], 3) and (mine["grads"]["means2D"][:, 2] == 0).all()


@pytest.mark.parametrize("case", [c for c in CASES if c[1] <= 20000], ids=[c[0] for c in CASES if c[1] <= 20000])
def test_matches_cpu_oracle(mod, case):
    cam, scene, dL, kw = build_case(case)
    mine = Hh.run_torch(mod, cam, scene, dL, **kw)
    orc = Hh.run_oracle(cam, scene, dL, **kw)
    assert_full_parity(mine, orc, exact_blend=False)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "ref_cuda_*.npz"))) or [None])
def test_matches_committed_reference_fixtures(mod, path):
    if path is None:
        pytest.skip("no tests/golden/ref_cuda_*.npz yet")
    from test_oracle import load_gold_case, check_against_gold
    z, cam, scene, kw = load_gold_case(path)
    got = Hh.run_torch(mod, cam, scene, z["dL"], **kw)
    check_against_gold(z, got, exact_blend=True)


def test_debug_mode_runs_and_matches(mod):
    cam, scene, dL, kw = build_case(CASES[4])
    a = Hh.run_torch(mod, cam, scene, dL, **kw)
    b = Hh.run_torch(mod, cam, scene, dL, debug=True, **kw)
    assert (a["color"] == b["color"]).all() and (a["radii"] == b["radii"]).all()


def test_forward_is_deterministic_and_backward_close(mod):
    cam, scene, dL, kw = build_case(CASES[3])
    a = Hh.run_torch(mod, cam, scene, dL, **kw)
    b = Hh.run_torch(mod, cam, scene, dL, **kw)
    assert (a["color"].view(np.uint32) == b["color"].view(np.uint32)).all()
    assert (a["point_list"] == b["point_list"]).all()
    for k in ("means3D", "shs", "opacities"):
        assert Hh.rel_l2(a["grads"][k], b["grads"][k]) <= 1e-5


@pytest.mark.parametrize("W,H,P,s0", [(32, 32, 20000, 6.0), (48, 16, 3000, 40.0), (250, 130, 20000, 3.0), (1920, 1080, 300000, 2.0),
                                      (1920, 1088, 50000, 30.0), (2560, 1440, 100000, 3.0), (3840, 2160, 200000, 4.0)])
def test_counting_tile_sort_equals_the_radix_tile_sort(mod, W, H, P, s0):
    """Grids of at most 8192 tiles take the one-pass counting sort (csrc/gsr_binning.cu tile_count_sort), larger ones the two
    LSD radix passes; GSR_TILE_SORT=radix forces the latter.  Same point list, ranges and image, bit for bit -- from four
    tiles (every round of the counting pass is contested) to 1080p, and 1920x1088 = 8160 tiles with splats over hundreds of tiles."""
    cam = S.make_camera(W, H)
    scene = S.make_scene(P, cam, sh_degree=1, M=4, seed=11, s0_pix=s0)
    dL = S.upstream_grad(cam)
    old = os.environ.get("GSR_TILE_SORT")
    try:
        os.environ["GSR_TILE_SORT"] = "radix"
        a = Hh.run_torch(mod, cam, scene, dL)
        os.environ.pop("GSR_TILE_SORT")
        b = Hh.run_torch(mod, cam, scene, dL)
    finally:
        if old is None:
            os.environ.pop("GSR_TILE_SORT", None)
        else:
            os.environ["GSR_TILE_SORT"] = old
    assert a["num_rendered"] == b["num_rendered"] and a["num_rendered"] > 0
    assert (a["point_list"] == b["point_list"]).all()
    assert (a["ranges"] == b["ranges"]).all()
    assert (a["color"].view(np.uint32) == b["color"].view(np.uint32)).all()


def test_empty_and_invisible_inputs(mod):
    import torch
    cam = S.make_camera(64, 48)
    t = lambda a: torch.tensor(np.asarray(a), dtype=torch.float32, device="cuda")
    st = mod.GaussianRasterizationSettings(48, 64, cam["tanfovx"], cam["tanfovy"], t([0.1, 0.2, 0.3]), 1.0,
                                           t(cam["viewmatrix"]), t(cam["projmatrix"]), 0, t(cam["campos"]), False, False)
    rast = mod.GaussianRasterizer(st)
    # P == 0 (rasterize_points.cu:94): zero image, empty radii
    e3 = torch.zeros(0, 3, device="cuda")
    color, radii = rast(e3, e3, torch.zeros(0, 1, device="cuda"), shs=torch.zeros(0, 1, 3, device="cuda"),
                        scales=e3, rotations=torch.zeros(0, 4, device="cuda"))
    assert color.shape == (3, 48, 64) and float(color.abs().max()) == 0.0 and radii.numel() == 0
    # everything behind the camera: background image, zero radii, zero gradients
    scene = S.make_scene(100, cam, sh_degree=0, M=1, seed=2)
    scene["means3D"][:, 2] = -1.0
    out = Hh.run_torch(mod, cam, scene, S.upstream_grad(cam), bg=(0.1, 0.2, 0.3))
    assert out["num_rendered"] == 0 and (out["radii"] == 0).all()
    assert np.allclose(out["color"][1], 0.2)
    assert all(np.abs(g).max() == 0 for g in out["grads"].values() if g is not None)


def test_mark_visible(mod):
    import torch
    cam = S.look_at_camera(64, 64, eye=(0.5, 0, -1), target=(0, 0, 4))
    scene = S.make_scene(5000, cam, seed=9)
    t = lambda a: torch.tensor(np.asarray(a), dtype=torch.float32, device="cuda")
    st = mod.GaussianRasterizationSettings(64, 64, cam["tanfovx"], cam["tanfovy"], t([0, 0, 0]), 1.0, t(cam["viewmatrix"]),
                                           t(cam["projmatrix"]), 0, t(cam["campos"]), False, False)
    vis = mod.GaussianRasterizer(st).markVisible(t(scene["means3D"])).cpu().numpy()
    assert vis.dtype == np.bool_ and (vis == O.mark_visible(scene["means3D"], cam["viewmatrix"])).all()
    if Hh.have_ref():
        ref = Hh.ref_module()
        rv = ref.GaussianRasterizer(ref.GaussianRasterizationSettings(*st)).markVisible(t(scene["means3D"])).cpu().numpy()
        assert (vis == rv).all()


def test_tile_row_bands_reassemble_the_full_image(mod):
    """Render-only tile-row sharding: rendering each band separately and stacking them
    gives the same pixels, bit for bit, as one full render."""
    import torch
    import gsr_multi_gpu as MG
    cam = S.make_camera(250, 130)
    scene = S.make_scene(20000, cam, sh_degree=3, M=16, seed=0)
    t = lambda a: torch.tensor(np.asarray(a), dtype=torch.float32, device="cuda")
    st = mod.GaussianRasterizationSettings(130, 250, cam["tanfovx"], cam["tanfovy"], t([0.3, 0.5, 0.7]), 1.0,
                                           t(cam["viewmatrix"]), t(cam["projmatrix"]), 3, t(cam["campos"]), False, False)
    args = dict(means3D=t(scene["means3D"]), means2D=torch.zeros(20000, 3, device="cuda"), opacities=t(scene["opacities"]),
                shs=t(scene["shs"]), scales=t(scene["scales"]), rotations=t(scene["rotations"]))
    with torch.no_grad():
        full, radii = mod.GaussianRasterizer(st)(**args)
        for world in (2, 3, 8, 16):
            img = torch.zeros_like(full)
            for band in MG.split_tile_rows(MG.tile_rows(130), world):
                if band[1] == band[0]:
                    continue
                c, r = mod.GaussianRasterizer(MG.banded_settings(st, band))(**args)
                p0, p1 = MG.band_pixel_rows(band, 130)
                img[:, p0:p1] = c[:, p0:p1]
                assert (r == radii).all()
            assert (img == full).all()


def test_full_size_properties_c2(mod):
    """BASELINE configs[1] at full size (3M Gaussians, SH3, 1080p): size-independent properties, plus
    structural parity with the reference build when it is available."""
    import torch
    cam = S.make_camera(1920, 1080)
    scene = S.make_scene(3_000_000, cam, sh_degree=3, M=16, seed=0)
    dL = S.upstream_grad(cam)
    mine = Hh.run_torch(mod, cam, scene, dL, bg=(0.3, 0.5, 0.7))
    R = mine["num_rendered"]
    assert R == int(mine["tiles_touched"].astype(np.int64).sum())
    assert 3.0 <= R / 3e6 <= 5.5                              # generator calibrated to R/P ~ 4
    rg = mine["ranges"].astype(np.int64)
    touched = rg[:, 1] > rg[:, 0]
    assert (rg[touched, 1] - rg[touched, 0]).sum() == R       # ranges tile the list exactly
    order = np.argsort(rg[touched, 0])
    assert (rg[touched][order][1:, 0] == rg[touched][order][:-1, 1]).all()
    # depth-sorted inside every tile: check via float32 depths recomputed on the host
    depth = (scene["means3D"] @ cam["viewmatrix"][:3, 2] + cam["viewmatrix"][3, 2]).astype(np.float32)
    pl = mine["point_list"].astype(np.int64)
    tile_of = np.repeat(np.nonzero(touched)[0][order], (rg[touched][order][:, 1] - rg[touched][order][:, 0]))
    d = depth[pl]
    same_tile = tile_of[1:] == tile_of[:-1]
    assert (d[1:][same_tile] >= d[:-1][same_tile] - 1e-4).all()
    assert np.isfinite(mine["color"]).all() and (mine["final_T"] >= 0).all() and (mine["final_T"] <= 1).all()
    for k in ("means3D", "means2D", "shs", "opacities", "scales", "rotations"):
        assert np.isfinite(mine["grads"][k]).all()
    assert (mine["grads"]["shs"][mine["radii"] == 0] == 0).all()
    if Hh.have_ref():
        ref = Hh.run_torch(Hh.ref_module(), cam, scene, dL, bg=(0.3, 0.5, 0.7), is_ref=True)
        assert_full_parity(mine, ref, exact_blend=True)


def test_multiview_sh_gradient_kernel_equals_sum_of_single_view_gradients(mod):
    """The fused SH-gradient exchange kernel (gsr_multiview_sh_grad), with the ranks emulated on ONE GPU: three cameras'
    backward passes publish their masked dL/dRGB blocks (ordinary device tensors stand in for the peer-mapped blocks),
    then one launch over all three blocks must equal the sum of the three ordinary dL_dsh -- and the other gradients
    must be untouched by the extra output."""
    import torch
    import bench as B
    P, W, H = 30000, 320, 200
    cam0 = S.make_camera(W, H)
    scene = S.make_scene(P, cam0, sh_degree=3, M=16, seed=4)
    cams = [B.ring_camera(S, W, H, r) for r in (0, 3, 5)]
    t = lambda a, rg=False: torch.tensor(np.asarray(a), dtype=torch.float32, device="cuda").requires_grad_(rg)
    blocks, ref_sum, ref_small = [], None, None
    for deg in (3, 1):
        blocks, ref_sum = [], None
        for ci, cam in enumerate(cams):
            dL = t(S.upstream_grad(cam, seed=10 + ci))
            st = mod.GaussianRasterizationSettings(H, W, cam["tanfovx"], cam["tanfovy"], t([0.1, 0.2, 0.3]), 1.0, t(cam["viewmatrix"]),
                                                   t(cam["projmatrix"]), deg, t(cam["campos"]), False, False)
            ins = dict(means3D=t(scene["means3D"]), sh=t(scene["shs"]), opac=t(scene["opacities"]), sc=t(scene["scales"]),
                       rot=t(scene["rotations"]))
            e = torch.Tensor([])
            nr, color, radii, gb, bb, ib = mod._C.rasterize_gaussians(st.bg, ins["means3D"], e, ins["opac"], ins["sc"], ins["rot"], 1.0, e,
                                                                      st.viewmatrix, st.projmatrix, st.tanfovx, st.tanfovy, H, W, ins["sh"],
                                                                      deg, st.campos, False, False)
            bargs = (st.bg, ins["means3D"], radii, e, ins["sc"], ins["rot"], 1.0, e, st.viewmatrix, st.projmatrix, st.tanfovx, st.tanfovy,
                     dL, ins["sh"], deg, st.campos, gb, nr, bb, ib, False)
            plain = mod._C.rasterize_gaussians_backward(*bargs)
            blk = torch.zeros(4 + 3 * P, device="cuda")
            blk[:3] = st.campos
            ev = torch.cuda.Event()
            ev.record()                                   # forces the cudaEvent to exist
            # odd cameras take the early-publish path (dL/dRGB written right after the blending backward + event record)
            fused = mod._C.rasterize_gaussians_backward(*bargs, 0, 0, blk, True, int(ev.cuda_event) if ci % 2 else 0)
            for a, b, name in zip(plain, fused, ("means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations")):
                if name != "sh":
                    assert Hh.rel_l2(b.cpu().numpy(), a.cpu().numpy()) <= 1e-5, name
            blocks.append(blk)
            ref_sum = plain[5].clone() if ref_sum is None else ref_sum + plain[5]
        out = torch.full((P, 16, 3), 7.0, device="cuda")
        mod._C.multiview_sh_grad([b.data_ptr() for b in blocks], t(scene["means3D"]), out, deg)
        torch.cuda.synchronize()
        K = (deg + 1) ** 2
        assert Hh.rel_l2(out[:, :K].cpu().numpy(), ref_sum[:, :K].cpu().numpy()) <= 1e-5
        assert float(out[:, K:].abs().max()) == 0.0 if K < 16 else True


def test_chunked_backward_with_packed_small_gradients_matches_plain_backward(mod):
    """Multi-view exchange plumbing on ONE GPU: the per-Gaussian backward launched in ranges (gsr_backward_chunk_plan),
    one event per range, the 14 small floats written as (P,14) rows and scattered back by gsr_unpack_small_grads --
    every gradient must match the plain single-launch backward."""
    import torch
    P, W, H = 150000, 320, 200
    cam = S.make_camera(W, H)
    scene = S.make_scene(P, cam, sh_degree=3, M=16, seed=9, s0_pix=1.5)
    t = lambda a: torch.tensor(np.asarray(a), dtype=torch.float32, device="cuda")
    dL = t(S.upstream_grad(cam, seed=21))
    st = mod.GaussianRasterizationSettings(H, W, cam["tanfovx"], cam["tanfovy"], t([0.1, 0.2, 0.3]), 1.0, t(cam["viewmatrix"]),
                                           t(cam["projmatrix"]), 3, t(cam["campos"]), False, False)
    m3, sh, op, sc, rot = t(scene["means3D"]), t(scene["shs"]), t(scene["opacities"]), t(scene["scales"]), t(scene["rotations"])
    e = torch.Tensor([])
    nr, color, radii, gb, bb, ib = mod._C.rasterize_gaussians(st.bg, m3, e, op, sc, rot, 1.0, e, st.viewmatrix, st.projmatrix, st.tanfovx,
                                                              st.tanfovy, H, W, sh, 3, st.campos, False, False)
    bargs = (st.bg, m3, radii, e, sc, rot, 1.0, e, st.viewmatrix, st.projmatrix, st.tanfovx, st.tanfovy, dL, sh, 3, st.campos,
             gb, nr, bb, ib, False)
    plain = mod._C.rasterize_gaussians_backward(*bargs)
    bounds = [int(b) for b in mod._C.backward_chunk_plan(P)]
    assert bounds[0] == 0 and bounds[-1] == P and len(bounds) == 5 and all(b1 >= b0 for b0, b1 in zip(bounds, bounds[1:]))
    assert [int(b) for b in mod._C.backward_chunk_plan(1000)] == [0, 1000]
    blk = torch.zeros(4 + 3 * P, device="cuda")
    evs = [torch.cuda.Event() for _ in range(5)]
    for ev in evs:
        ev.record()
    pack = torch.full((14 * P,), 5.0, device="cuda")
    fused = mod._C.rasterize_gaussians_backward(*bargs, 0, 0, blk, False, int(evs[0].cuda_event), [int(ev.cuda_event) for ev in evs[1:]], pack)
    side = torch.cuda.Stream()
    g2d, gcol, gop, g3d, gcov, gsh, gsc, grot = fused
    for c in range(4):                     # consume each range as soon as its event fires, on another stream
        side.wait_event(evs[1 + c])
        with torch.cuda.stream(side):
            mod._C.unpack_small_grads(bounds[c], bounds[c + 1], pack, g3d, g2d, gop, gsc, grot)
    torch.cuda.synchronize()
    # the blending backward sums with float atomics, so two runs agree to rounding, not bitwise
    for a, b, name in zip(plain, fused, ("means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations")):
        assert Hh.rel_l2(b.cpu().numpy(), a.cpu().numpy()) <= 1e-5, name
