"""The C3 training-loop harness (tools/c3_trainer.py) on the CPU with a stand-in rasterizer: the loop's bookkeeping --
optimizer state edits during clone / split / prune, densification statistics, the growth schedule -- against the
behaviour of the reference's GaussianModel methods it restates (scene/gaussian_model.py:398-602).  The real rasterizer,
loss and optimizer only exist on the GPU; here they are replaced by differentiable torch stand-ins."""
import os
import sys
from collections import namedtuple

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import c3_trainer as T        # noqa: E402
import gsr_optim              # noqa: E402
import gsr_synthetic as S     # noqa: E402


class FakeRasterizerModule:
    """Same constructor / call surface as diff_gaussian_rasterization; the 'image' is a smooth function of every input."""
    GaussianRasterizationSettings = namedtuple("GaussianRasterizationSettings", "image_height image_width tanfovx tanfovy bg scale_modifier "
                                               "viewmatrix projmatrix sh_degree campos prefiltered debug")

    class GaussianRasterizer:
        def __init__(self, raster_settings):
            self.s = raster_settings

        def __call__(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None):
            P = means3D.shape[0]
            H, W = self.s.image_height, self.s.image_width
            idx = torch.arange(P, dtype=torch.float32)
            k = torch.stack([torch.sin(idx * 0.7), torch.cos(idx * 1.3)], 1)
            s = opacities.squeeze(-1) * (shs.mean(dim=(1, 2)) + scales.sum(1) + 0.1 * rotations.sum(1) + 0.01 * means3D.sum(1)
                                         + (means2D[:, :2] * k).sum(1))
            basis = torch.sin(idx[:, None] * torch.arange(3 * H * W, dtype=torch.float32)[None, :] * 0.37)
            radii = ((torch.arange(P) % 5) != 0).int() * 3
            return (s * radii.bool()) @ basis / P ** 0.5 + self.s.bg.view(3, 1, 1).expand(3, H, W).reshape(-1) * 0, radii


def _setup(P=300, opt_cls=torch.optim.Adam):
    cam = S.make_camera(8, 8)
    for k in ("viewmatrix", "projmatrix", "campos"):
        cam[k + "_t"] = torch.tensor(cam[k])
    scene = S.make_scene(P, cam, sh_degree=3, seed=0)
    g = T.Gaussians(scene, "cpu", opt_cls, gsr_optim.get_expon_lr_func)
    return cam, g


def _loss(img, gt, lam):
    return (img.reshape(-1) - gt.reshape(-1)).abs().mean()


def test_loop_grows_the_model_on_schedule_and_keeps_every_tensor_consistent():
    cam, g = _setup(300)
    events = []
    extent = float(torch.exp(g.scaling.detach()).max(dim=1).values.median()) / T.OptArgs.percent_dense
    T.train(g, [cam], [torch.zeros(3 * 8 * 8)], FakeRasterizerModule, _loss, iterations=41, densify_from=5, densify_every=10,
            target_P=900, extent=extent, bg=torch.zeros(3), timed=False,
            on_densify=lambda it, p0, p1, thr, st: events.append((it, p0, p1, st)))
    assert [e[0] for e in events] == [10, 20, 30, 40]
    assert events[0][1] == 300
    assert 850 <= g.P <= 950                                  # geometric schedule to 900, minus low-opacity prunes
    assert sum(st[0] for *_, st in events) > 0 and sum(st[1] for *_, st in events) > 0   # clone and split both ran
    P = g.P
    for n, w in zip(T.GROUPS, [(3,), (1, 3), (15, 3), (1,), (3,), (4,)]):
        p = getattr(g, n)
        assert p.shape == (P,) + w and p.requires_grad and isinstance(p, torch.nn.Parameter)
        grp = [x for x in g.optimizer.param_groups if x["name"] == n][0]
        assert grp["params"][0] is p                          # the optimizer holds the live tensor
        st = g.optimizer.state[p]
        assert st["exp_avg"].shape == p.shape and st["exp_avg_sq"].shape == p.shape
    assert g.xyz_gradient_accum.shape == (P, 1) and g.denom.shape == (P, 1) and g.max_radii2D.shape == (P,)
    assert len(g.optimizer.state) == 6                        # no stale entries for replaced parameters


def test_clone_split_prune_follow_the_reference_rules():
    cam, g = _setup(200)
    opt = g.optimizer
    for p in (g.xyz, g.f_dc, g.f_rest, g.opacity, g.scaling, g.rotation):        # give every parameter a state
        p.grad = torch.ones_like(p)
    opt.step()
    m_before = opt.state[g.xyz]["exp_avg"].clone()
    xyz0, sc0, op0 = g.xyz.detach().clone(), g.scaling.detach().clone(), g.opacity.detach().clone()
    extent = float(torch.exp(g.scaling.detach()).max(dim=1).values.median()) / T.OptArgs.percent_dense
    big = torch.exp(sc0).max(dim=1).values > T.OptArgs.percent_dense * extent
    g.xyz_gradient_accum = torch.zeros(200, 1); g.denom = torch.ones(200, 1)
    chosen = torch.zeros(200, dtype=torch.bool); chosen[::4] = True
    g.xyz_gradient_accum[chosen] = 1.0
    with torch.no_grad():
        n_clone, n_split, n_pruned = g.densify_and_prune(0.5, 0.0, extent, None, generator=torch.Generator().manual_seed(0))
    assert n_clone == int((chosen & ~big).sum()) and n_split == int((chosen & big).sum()) and n_pruned == 0
    assert g.P == 200 + n_clone + n_split                      # +1 per clone; +2 -1 per split
    # survivors keep their place and their moments; clones are exact copies with zero moments
    kept = ~(chosen & big)
    assert torch.equal(g.xyz.detach()[:int(kept.sum())], xyz0[kept])
    assert torch.equal(opt.state[g.xyz]["exp_avg"][:int(kept.sum())], m_before[kept])
    n_kept = int(kept.sum())
    assert torch.equal(g.xyz.detach()[n_kept:n_kept + n_clone], xyz0[chosen & ~big])
    assert torch.count_nonzero(opt.state[g.xyz]["exp_avg"][n_kept:]) == 0
    # split children: scale / 1.6, parent's opacity, twice
    child_scale = g.scaling.detach()[n_kept + n_clone:]
    assert torch.allclose(child_scale, torch.log(torch.exp(sc0[chosen & big]).repeat(2, 1) / 1.6), atol=1e-6)
    assert torch.equal(g.opacity.detach()[n_kept + n_clone:], op0[chosen & big].repeat(2, 1))
    # statistics are reset after a densification (densification_postfix)
    assert float(g.xyz_gradient_accum.abs().sum()) == 0.0 and float(g.denom.sum()) == 0.0


def test_densification_stats_accumulate_the_2d_gradient_norm_of_visible_points_only():
    cam, g = _setup(50)
    vs = torch.zeros(50, 3, requires_grad=True)
    vs.grad = torch.tensor([[3.0, 4.0, 99.0]]).repeat(50, 1)
    vis = torch.arange(50) % 2 == 0
    g.add_densification_stats(vs, vis)
    g.add_densification_stats(vs, vis)
    assert torch.equal(g.xyz_gradient_accum.squeeze(1), torch.where(vis, 10.0, 0.0))
    assert torch.equal(g.denom.squeeze(1), torch.where(vis, 2.0, 0.0))


def test_growth_threshold_selects_the_requested_number():
    cam, g = _setup(100)
    g.xyz_gradient_accum = torch.arange(100, dtype=torch.float32).view(100, 1); g.denom = torch.ones(100, 1)
    thr = T.growth_threshold(g, 130)
    assert int(((g.xyz_gradient_accum / g.denom).squeeze() >= thr).sum()) == 30
    assert T.growth_threshold(g, 100) == float("inf")
    assert int(((g.xyz_gradient_accum / g.denom).squeeze() >= T.growth_threshold(g, 10**6)).sum()) == 99   # capped at doubling; a Gaussian without gradient is never selected


def test_build_rotation_is_orthonormal_and_matches_the_quaternion_formula():
    q = torch.randn(20, 4)
    R = T.build_rotation(q)
    assert torch.allclose(R @ R.transpose(1, 2), torch.eye(3).expand(20, 3, 3), atol=1e-5)
    assert torch.allclose(torch.linalg.det(R), torch.ones(20), atol=1e-5)
    assert torch.allclose(T.build_rotation(torch.tensor([[1.0, 0, 0, 0]]))[0], torch.eye(3))


def test_buckets():
    table = [(100, [1.0, 2.0, 0.5, 0.5]), (100, [1.0, 2.0, 0.5, 0.5]), (400, [3.0, 6.0, 1.0, 2.0])]
    b = T.geometric_buckets(table, n=2)
    assert b[0]["iterations"] == 2 and b[0]["iter_ms"] == pytest.approx(4.0)
    assert b[-1]["iterations"] == 1 and b[-1]["iter_ms"] == pytest.approx(12.0)
