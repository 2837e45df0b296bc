#!/usr/bin/env python3
"""Golden vectors for densification, produced HERE on the CPU by the reference's own GaussianModel.

scene/gaussian_model.py hard-codes device="cuda" in the functions on this path (densification_postfix, densify_and_split,
training_setup).  This script imports the UNMODIFIED module from /root/reference and maps that device name to the CPU in
torch's factory functions for the duration of the run; plyfile and simple_knn (absent / GPU-only, not on this path) are
empty stand-in modules so that the import succeeds.  torch.normal is wrapped to RECORD the samples densify_and_split draws.
Stored per case: the model before (six parameter tensors, Adam moments, the three statistics), the arguments, the recorded
samples, and the model after.  Also one case of the per-iteration statistics update (train.py:153-158).

Run in the build container only (needs /root/reference); the fixture is committed, the GPU box never runs this.
"""
import os
import sys
import types
from argparse import ArgumentParser

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
NAMES = ["xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation"]
ATTRS = {"xyz": "_xyz", "f_dc": "_features_dc", "f_rest": "_features_rest", "opacity": "_opacity", "scaling": "_scaling", "rotation": "_rotation"}


def _cpu_device_shim():
    def wrap(fn):
        def inner(*a, **k):
            if str(k.get("device", "")).startswith("cuda"):
                k["device"] = "cpu"
            return fn(*a, **k)
        return inner
    for name in ("zeros", "ones", "empty", "tensor", "zeros_like", "rand", "randn", "full"):
        setattr(torch, name, wrap(getattr(torch, name)))


def _import_reference_model():
    ply = types.ModuleType("plyfile"); ply.PlyData = ply.PlyElement = object
    knn = types.ModuleType("simple_knn"); knn_c = types.ModuleType("simple_knn._C"); knn_c.distCUDA2 = None
    sys.modules.update({"plyfile": ply, "simple_knn": knn, "simple_knn._C": knn_c})
    sys.path.insert(0, REF)
    from scene.gaussian_model import GaussianModel      # reference module
    from arguments import OptimizationParams            # reference module
    return GaussianModel, OptimizationParams


def _state(g, with_moments):
    out = {}
    for n in NAMES:
        p = getattr(g, ATTRS[n])
        out["p_" + n] = p.detach().numpy().copy()
        if with_moments:
            st = g.optimizer.state[p]
            out["m_" + n] = st["exp_avg"].numpy().copy()
            out["v_" + n] = st["exp_avg_sq"].numpy().copy()
    out["accum"] = g.xyz_gradient_accum.numpy().copy()
    out["denom"] = g.denom.numpy().copy()
    out["max_radii2D"] = g.max_radii2D.numpy().copy()
    return out


def make_case(GaussianModel, opt, seed, P, with_moments, max_screen_size, extent):
    gen = torch.Generator().manual_seed(seed)
    rn = lambda *s: torch.randn(*s, generator=gen)
    g = GaussianModel(3)
    g._xyz = torch.nn.Parameter(rn(P, 3) * 2.0)
    g._features_dc = torch.nn.Parameter(rn(P, 1, 3))
    g._features_rest = torch.nn.Parameter(rn(P, 15, 3) * 0.1)
    g._opacity = torch.nn.Parameter(rn(P, 1) * 3.0 - 1.0)                     # sigmoid: a good share below 0.005 .. above 0.9
    g._scaling = torch.nn.Parameter(rn(P, 3) * 1.2 - 3.5)                     # exp: around percent_dense * extent and 0.1 * extent
    g._rotation = torch.nn.Parameter(rn(P, 4))
    g.max_radii2D = torch.zeros(P)
    g.spatial_lr_scale = 1.0
    g.training_setup(opt)
    if with_moments:                                                          # two optimizer steps so that both moments are populated
        for _ in range(2):
            for n in NAMES:
                getattr(g, ATTRS[n]).grad = rn(*getattr(g, ATTRS[n]).shape) * 1e-3
            g.optimizer.step()
            g.optimizer.zero_grad(set_to_none=True)
    # statistics as a few iterations would leave them: some Gaussians never seen (0 / 0 -> NaN -> 0)
    seen = torch.rand(P, generator=gen) > 0.15
    g.denom = (torch.randint(1, 40, (P, 1), generator=gen).float() * seen[:, None]).contiguous()
    g.xyz_gradient_accum = (torch.rand(P, 1, generator=gen) * 6e-4 * g.denom).contiguous()   # mean gradient around the 2e-4 threshold
    g.max_radii2D = torch.rand(P, generator=gen) * 40.0
    before = _state(g, with_moments)
    drawn = []
    real_normal = torch.normal
    def recording_normal(*a, **k):
        out = real_normal(*a, **k)
        drawn.append(out.detach().clone())
        return out
    torch.normal = recording_normal
    try:
        torch.manual_seed(seed + 1)
        g.densify_and_prune(opt.densify_grad_threshold, 0.005, extent, max_screen_size)       # train.py:160-162
    finally:
        torch.normal = real_normal
    assert len(drawn) == 1
    after = _state(g, with_moments)
    case = {"before_" + k: v for k, v in before.items()}
    case.update({"after_" + k: v for k, v in after.items()})
    case.update(samples=drawn[0].numpy(), max_grad=opt.densify_grad_threshold, min_opacity=0.005, extent=extent,
                percent_dense=g.percent_dense, max_screen_size=float(max_screen_size or 0.0), with_moments=int(with_moments),
                sampling_seed=seed + 1)
    return case


def make_stats_case(GaussianModel, seed, P):
    gen = torch.Generator().manual_seed(seed)
    g = GaussianModel(0)
    g.xyz_gradient_accum = torch.rand(P, 1, generator=gen)
    g.denom = torch.randint(0, 9, (P, 1), generator=gen).float()
    g.max_radii2D = torch.rand(P, generator=gen) * 30.0
    radii = (torch.randint(-2, 60, (P,), generator=gen) * (torch.rand(P, generator=gen) > 0.3)).int()
    vsp = torch.zeros(P, 3, requires_grad=True)
    vsp.grad = torch.randn(P, 3, generator=gen) * 1e-3
    case = dict(accum0=g.xyz_gradient_accum.numpy().copy(), denom0=g.denom.numpy().copy(), max_radii0=g.max_radii2D.numpy().copy(),
                radii=radii.numpy().copy(), vs_grad=vsp.grad.numpy().copy())
    visibility_filter = radii > 0                                                               # gaussian_renderer/__init__.py:108
    g.max_radii2D[visibility_filter] = torch.max(g.max_radii2D[visibility_filter], radii[visibility_filter])   # train.py:156
    g.add_densification_stats(vsp, visibility_filter)                                           # train.py:158
    case.update(accum1=g.xyz_gradient_accum.numpy().copy(), denom1=g.denom.numpy().copy(), max_radii1=g.max_radii2D.numpy().copy())
    return case


def main():
    torch.set_num_threads(1)
    GaussianModel, OptimizationParams = _import_reference_model()
    _cpu_device_shim()
    opt = OptimizationParams(ArgumentParser())
    out = {}
    cases = [("a", dict(seed=11, P=200, with_moments=True, max_screen_size=20, extent=3.0)),
             ("b", dict(seed=12, P=129, with_moments=False, max_screen_size=None, extent=3.0)),
             ("c", dict(seed=13, P=400, with_moments=True, max_screen_size=20, extent=1.2))]
    for tag, kw in cases:
        for k, v in make_case(GaussianModel, opt, **kw).items():
            out[f"{tag}_{k}"] = v
        print(tag, "P", kw["P"], "->", out[f"{tag}_after_p_xyz"].shape[0], "split samples", out[f"{tag}_samples"].shape[0])
    for k, v in make_stats_case(GaussianModel, 21, 500).items():
        out["stats_" + k] = v
    out["cases"] = np.array([t for t, _ in cases])
    path = os.path.join(HERE, "ref_densify_cpu.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
