"""Shared helpers for the parity tests: run the CUDA path, the reference build and the CPU oracle
on the same numpy scene and return comparable numpy dicts."""
import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "gaussian-splatting_cc-comments_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

import gsr_synthetic as S  # noqa: E402
from oracle import gs_oracle as O  # noqa: E402

GRAD_KEYS = ("means3D", "means2D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")


def have_ref():
    return os.path.exists(os.path.join(ROOT, "oracle", "_ref", "diff_gaussian_rasterization", "_C.so"))


_ref_mod = None


def ref_module():
    """The UNMODIFIED reference package built by oracle/build_ref.py, under a non-clashing name."""
    global _ref_mod
    if _ref_mod is None:
        d = os.path.join(ROOT, "oracle", "_ref", "diff_gaussian_rasterization")
        spec = importlib.util.spec_from_file_location("ref_diff_gaussian_rasterization", os.path.join(d, "__init__.py"),
                                                      submodule_search_locations=[d])
        m = importlib.util.module_from_spec(spec)
        sys.modules["ref_diff_gaussian_rasterization"] = m
        spec.loader.exec_module(m)
        _ref_mod = m
    return _ref_mod


def my_module():
    import diff_gaussian_rasterization as m
    return m


def rel_l2(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    d = np.linalg.norm(a - b); n = np.linalg.norm(b)
    return d / n if n > 0 else d


def _a128(x):
    return (x + 127) // 128 * 128


def slice_ref_buffers(P, R, W, H, geom, binning, img):
    """Pull the arrays out of the reference's opaque buffers (layout: rasterizer_impl.cu:179-223)."""
    g = geom.cpu().numpy(); b = binning.cpu().numpy(); im = img.cpu().numpy()
    N = W * H
    o = 0
    depths = g[o:o + 4 * P].view(np.float32).copy(); o = _a128(o + 4 * P)
    clamped = g[o:o + 3 * P].copy().reshape(P, 3); o = _a128(o + 3 * P)
    o = _a128(o + 4 * P)                                  # internal_radii
    means2D = g[o:o + 8 * P].view(np.float32).copy().reshape(P, 2); o = _a128(o + 8 * P)
    cov3D = g[o:o + 24 * P].view(np.float32).copy().reshape(P, 6); o = _a128(o + 24 * P)
    conic_opacity = g[o:o + 16 * P].view(np.float32).copy().reshape(P, 4); o = _a128(o + 16 * P)
    rgb = g[o:o + 12 * P].view(np.float32).copy().reshape(P, 3); o = _a128(o + 12 * P)
    tiles_touched = g[o:o + 4 * P].view(np.uint32).copy()
    point_list = b[0:4 * R].view(np.uint32).copy()
    o = 0
    final_T = im[o:o + 4 * N].view(np.float32).copy().reshape(H, W); o = _a128(o + 4 * N)
    n_contrib = im[o:o + 4 * N].view(np.uint32).copy().reshape(H, W); o = _a128(o + 4 * N)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    ranges = im[o:o + 8 * tiles].view(np.uint32).copy().reshape(tiles, 2)
    return dict(depths=depths, clamped=clamped, means2D=means2D, cov3D=cov3D, conic_opacity=conic_opacity, rgb=rgb,
                tiles_touched=tiles_touched, point_list=point_list, final_T=final_T, n_contrib=n_contrib, ranges=ranges)


def run_torch(mod, cam, scene, dL=None, bg=(0, 0, 0), use="sh", cov="sr", device="cuda", debug=False, is_ref=False,
              scale_modifier=1.0):
    """Run a diff_gaussian_rasterization-compatible module.  use: 'sh' | 'colors'; cov: 'sr' | 'cov3d'."""
    import torch
    t = lambda a, rg=False: torch.tensor(np.asarray(a), dtype=torch.float32, device=device).requires_grad_(rg)
    P = scene["means3D"].shape[0]
    ins = dict(means3D=t(scene["means3D"], True), means2D=torch.zeros(P, 3, device=device, requires_grad=True),
               opacities=t(scene["opacities"], True))
    kw = {}
    if use == "sh":
        ins["shs"] = t(scene["shs"], True); kw["shs"] = ins["shs"]
    else:
        ins["colors_precomp"] = t(scene["colors_precomp"], True); kw["colors_precomp"] = ins["colors_precomp"]
    if cov == "sr":
        ins["scales"] = t(scene["scales"], True); ins["rotations"] = t(scene["rotations"], True)
        kw["scales"] = ins["scales"]; kw["rotations"] = ins["rotations"]
    else:
        ins["cov3D_precomp"] = t(scene["cov3D_precomp"], True); kw["cov3D_precomp"] = ins["cov3D_precomp"]
    settings = mod.GaussianRasterizationSettings(
        image_height=cam["H"], image_width=cam["W"], tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], bg=t(np.asarray(bg)),
        scale_modifier=scale_modifier, viewmatrix=t(cam["viewmatrix"]), projmatrix=t(cam["projmatrix"]),
        sh_degree=scene.get("sh_degree", 0), campos=t(cam["campos"]), prefiltered=False, debug=debug)
    rast = mod.GaussianRasterizer(settings)
    color, radii = rast(means3D=ins["means3D"], means2D=ins["means2D"], opacities=ins["opacities"], **kw)
    out = dict(color=color.detach().cpu().numpy(), radii=radii.detach().cpu().numpy())
    # intermediates from the saved buffers
    fn = color.grad_fn
    saved = fn.saved_tensors
    geom, binning, img = saved[-3], saved[-2], saved[-1]
    R = fn.num_rendered
    out["num_rendered"] = R
    if is_ref:
        out.update(slice_ref_buffers(P, R, cam["W"], cam["H"], geom, binning, img))
    else:
        pl, rg, nc, ft, packed, tt, cl, wm = mod._C.view_buffers(P, R, cam["W"], cam["H"], geom, binning, img)
        pk = packed.cpu().numpy()
        out.update(point_list=pl.cpu().numpy().view(np.uint32), ranges=rg.cpu().numpy().view(np.uint32),
                   n_contrib=nc.cpu().numpy().view(np.uint32), final_T=ft.cpu().numpy(), packed=pk,
                   means2D=pk[:, 0:2], conic_opacity=np.concatenate([pk[:, 2:5], pk[:, 6:7]], 1), rgb=pk[:, 7:10],
                   tiles_touched=tt.cpu().numpy().view(np.uint32), clamped_bits=cl.cpu().numpy(), warp_mask=wm.cpu().numpy())
    if dL is not None:
        color.backward(t(dL))
        out["grads"] = {k: (ins[k].grad.detach().cpu().numpy() if k in ins and ins[k].grad is not None else None)
                        for k in GRAD_KEYS}
    return out


def run_oracle(cam, scene, dL=None, bg=(0, 0, 0), use="sh", cov="sr", scale_modifier=1.0):
    kw = dict(sh_degree=scene.get("sh_degree", 0), scale_modifier=scale_modifier, bg=bg)
    if use == "sh": kw["shs"] = scene["shs"]
    else: kw["colors_precomp"] = scene["colors_precomp"]
    if cov == "sr": kw["scales"] = scene["scales"]; kw["rotations"] = scene["rotations"]
    else: kw["cov3D_precomp"] = scene["cov3D_precomp"]
    f = O.forward(cam, scene["means3D"], scene["opacities"], **kw)
    if dL is not None:
        f["grads"] = O.backward(cam, f, dL, scene["means3D"], **kw)
    return f
