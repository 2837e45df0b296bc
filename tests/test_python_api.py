"""The drop-in Python surface (DGR/diff_gaussian_rasterization/__init__.py) without a GPU: names, field
order, argument validation and the error behaviour of the reference."""
import inspect

import pytest
import torch

import helpers as Hh


@pytest.fixture(scope="module")
def mod():
    return Hh.my_module()


def test_exports(mod):
    for name in ("GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians", "_RasterizeGaussians", "_C"):
        assert hasattr(mod, name)
    for fn in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible"):
        assert hasattr(mod._C, fn)                       # DGR/ext.cpp:18-22


def test_settings_fields_in_reference_order(mod):
    assert mod.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")


def test_forward_signature(mod):
    sig = inspect.signature(mod.GaussianRasterizer.forward)
    assert list(sig.parameters) == ["self", "means3D", "means2D", "opacities", "shs", "colors_precomp", "scales",
                                    "rotations", "cov3D_precomp"]
    sig = inspect.signature(mod.rasterize_gaussians)
    assert list(sig.parameters) == ["means3D", "means2D", "sh", "colors_precomp", "opacities", "scales", "rotations",
                                    "cov3Ds_precomp", "raster_settings"]


def _rast(mod):
    s = mod.GaussianRasterizationSettings(16, 16, 1.0, 1.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 0,
                                          torch.zeros(3), False, False)
    return mod.GaussianRasterizer(s)


def test_exactly_one_colour_source(mod):
    r = _rast(mod)
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(m, m, torch.zeros(4, 1), scales=torch.ones(4, 3), rotations=torch.ones(4, 4))
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), colors_precomp=torch.zeros(4, 3), scales=torch.ones(4, 3),
          rotations=torch.ones(4, 4))


def test_exactly_one_covariance_source(mod):
    r = _rast(mod)
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 1, 3))
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), scales=torch.ones(4, 3))
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), scales=torch.ones(4, 3), rotations=torch.ones(4, 4),
          cov3D_precomp=torch.zeros(4, 6))


def test_means3d_shape_is_checked_like_the_reference(mod):
    """rasterize_points.cu:60-63: AT_ERROR unless means3D is (P,3)."""
    r = _rast(mod)
    bad = torch.zeros(4, 2)
    with pytest.raises(RuntimeError, match="means3D must have dimensions"):
        r(bad, bad, torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), scales=torch.ones(4, 3), rotations=torch.ones(4, 4))


def test_no_cpu_fallback(mod):
    """CPU tensors are refused loudly; nothing is computed on the host."""
    r = _rast(mod)
    m = torch.zeros(4, 3)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), scales=torch.ones(4, 3), rotations=torch.ones(4, 4))
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        r.markVisible(m)


def test_product_package_never_imports_the_oracle():
    import os
    import re
    pkg = os.path.join(Hh.ROOT, "gaussian-splatting_cc-comments_b200")
    for dp, _, fs in os.walk(pkg):
        if "build" in dp.split(os.sep):
            continue
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"gs_oracle|gs_next_oracle|gso_|gno_|from oracle|import oracle|cuda_emu", src), os.path.join(dp, f)
