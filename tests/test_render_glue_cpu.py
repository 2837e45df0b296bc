"""Host-side pieces next to the raw-parameter path: gsr_sh.eval_sh against vectors from the reference's own
utils/sh_utils.eval_sh (tests/golden/ref_torch_cpu.npz, made by tests/golden/make_golden_cpu.py), and the surface of the
raw entry points.  CPU only: the rasterizer itself has no CPU path."""
import os

import numpy as np
import pytest
import torch

import gsr_sh

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_torch_cpu.npz")


def test_eval_sh_matches_the_reference_function():
    z = np.load(GOLD)
    means, shs, campos = torch.tensor(z["means3D"]), torch.tensor(z["shs"]), torch.tensor(z["campos"])
    for deg in range(4):
        shs_view = shs.transpose(1, 2).view(-1, 3, 16)                        # gaussian_renderer/__init__.py:94-97
        d = means - campos.repeat(shs.shape[0], 1)
        d = d / d.norm(dim=1, keepdim=True)
        raw = gsr_sh.eval_sh(deg, shs_view, d) + 0.5
        assert float((raw - torch.tensor(z["rgb_raw_deg%d" % deg])).abs().max()) <= 1e-6
        assert float((torch.clamp_min(raw, 0.0) - torch.tensor(z["rgb_deg%d" % deg])).abs().max()) <= 1e-6
    with pytest.raises(AssertionError):
        gsr_sh.eval_sh(5, shs_view, d)


def test_raw_entry_points_exist_and_refuse_cpu_tensors():
    import diff_gaussian_rasterization as mod
    import gsr_render
    assert hasattr(mod._C, "rasterize_gaussians_raw") and hasattr(mod._C, "rasterize_gaussians_raw_backward")
    assert callable(gsr_render.render)
    P = 4
    s = mod.GaussianRasterizationSettings(image_height=16, image_width=16, tanfovx=0.5, tanfovy=0.5, bg=torch.zeros(3), scale_modifier=1.0,
                                          viewmatrix=torch.eye(4), projmatrix=torch.eye(4), sh_degree=3, campos=torch.zeros(3),
                                          prefiltered=False, debug=False)
    r = mod.GaussianRasterizer(s)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        r.forward_raw(torch.zeros(P, 3), torch.zeros(P, 3), torch.zeros(P, 1, 3), torch.zeros(P, 15, 3), torch.zeros(P, 1), torch.zeros(P, 3),
                      torch.zeros(P, 4))
