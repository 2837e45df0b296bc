#!/usr/bin/env python3
"""BASELINE.json config C3 on one GPU: the training loop of tools/c3_trainer.py on a synthetic scene that densifies from
P0 to P_target Gaussians, one view per iteration at 1920x1080, timed per phase with CUDA events.  One JSON line.

    python tools/train_c3.py [--impl b200|reference] [--P0 1000000] [--target 6000000] [--iterations 660]
                             [--densify-from 60] [--densify-every 50] [--width 1920 --height 1080] [--views 8]

Both arms run the same loop, scene, cameras and target images (rendered once from a second random "teacher" scene with
this package's rasterizer); they differ in rasterizer, loss and optimizer as c3_trainer.py's header says.  Numbers printed
under a profiler are not bench values.
"""
import argparse
import json
import math
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "gaussian-splatting_cc-comments_b200"), os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

import c3_trainer as T          # noqa: E402
import gsr_synthetic as S       # noqa: E402


def torch_loss(window):
    """The reference's loss as its own composition of torch ops (utils/loss_utils.py:17-64, train.py:126-127)."""
    import torch.nn.functional as F

    def fn(img, gt, lam):
        C = img.size(-3)
        conv = lambda t: F.conv2d(t, window, padding=5, groups=C)
        mu1, mu2 = conv(img), conv(gt)
        mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
        s1, s2, s12 = conv(img * img) - mu1_sq, conv(gt * gt) - mu2_sq, conv(img * gt) - mu1_mu2
        C1, C2 = 0.01 ** 2, 0.03 ** 2
        ssim = (((2 * mu1_mu2 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s1 + s2 + C2))).mean()
        return (1.0 - lam) * torch.abs(img - gt).mean() + lam * (1.0 - ssim)
    return fn


def cameras(W, H, n):
    cams = []
    for v in range(n):
        a = 2 * math.pi * v / n
        cam = S.look_at_camera(W, H, eye=(0.5 * math.cos(a), 0.5 * math.sin(a), 0.0), target=(0.0, 0.0, 5.0))
        for k in ("viewmatrix", "projmatrix", "campos"):
            cam[k + "_t"] = torch.tensor(cam[k], device="cuda")
        cams.append(cam)
    return cams


def parse_args(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--P0", type=int, default=1_000_000)
    ap.add_argument("--target", type=int, default=6_000_000)
    ap.add_argument("--iterations", type=int, default=660)
    ap.add_argument("--densify-from", type=int, default=60)
    ap.add_argument("--densify-every", type=int, default=50)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--views", type=int, default=8)
    ap.add_argument("--no-raw", action="store_true", help="b200 arm: activations and the SH concatenation as torch ops (the reference's "
                    "route through GaussianRasterizer.forward) instead of the fused raw-parameter path")
    return ap.parse_args(argv)


def run(a):
    """One arm of the C3 loop; returns the record main() prints (bench.py embeds a compact form of it)."""
    assert torch.cuda.is_available(), "train_c3.py needs a GPU (there is no CPU path)"
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    import diff_gaussian_rasterization as mine
    import gsr_optim
    cams = cameras(a.width, a.height, a.views)
    bg = torch.zeros(3, device="cuda")
    base = S.make_camera(a.width, a.height)
    # target images: a second random scene seen from every camera (rendered once, this package's rasterizer, no gradient)
    teacher = S.make_scene(min(a.P0, 1_000_000), base, sh_degree=3, seed=1)
    tg = T.Gaussians(teacher, "cuda", torch.optim.Adam, gsr_optim.get_expon_lr_func)
    with torch.no_grad():
        targets = [T.render(c, tg, mine, bg)["render"].clamp(0, 1).contiguous() for c in cams]
    del tg
    scene = S.make_scene(a.P0, base, sh_degree=3, seed=0)
    if a.impl == "b200":
        import loss_utils
        import gsr_densify
        module, opt_cls, fused_densify = mine, gsr_optim.FusedAdam, gsr_densify
        loss_fn = lambda img, gt, lam: loss_utils.fused_l1_ssim_loss(img, gt, lam)[0]
    else:
        import helpers
        import loss_ref
        assert helpers.have_ref(), "oracle/_ref is missing: build it in the container (python oracle/build_ref.py)"
        module, opt_cls, fused_densify = helpers.ref_module(), torch.optim.Adam, None
        w1 = loss_ref.window_1d().cuda()
        loss_fn = torch_loss((w1.unsqueeze(1) @ w1.unsqueeze(0)).expand(3, 1, 11, 11).contiguous())
    g = T.Gaussians(scene, "cuda", opt_cls, gsr_optim.get_expon_lr_func, fused_densify=fused_densify)
    # clone vs split is decided by max scale against percent_dense * extent (scene/gaussian_model.py:512,545): put the
    # boundary at the median so both branches run
    extent = float(torch.exp(g.scaling.detach()).max(dim=1).values.median()) / T.OptArgs.percent_dense
    log = []
    torch.cuda.synchronize()
    t0 = time.time()
    table = T.train(g, cams, targets, module, loss_fn, a.iterations, a.densify_from, a.densify_every, a.target, extent, bg,
                    raw=(a.impl == "b200" and not a.no_raw),
                    on_densify=lambda it, p0, p1, thr, st: log.append({"iteration": it, "P_before": p0, "P_after": p1, "threshold": thr,
                                                                       "cloned": st[0], "split": st[1], "pruned": st[2]}))
    torch.cuda.synchronize()
    wall = time.time() - t0
    buckets = T.geometric_buckets(table)
    plain = [sum(r) for p, r in table]
    return ({
        "metric": "training ms/iter vs P, synthetic scene densifying %d -> %d Gaussians, SH3, %dx%d, 1 view/iter" % (a.P0, a.target, a.width, a.height),
        "impl": a.impl, "unit": "ms/iter", "higher_is_better": False, "data": "synthetic", "dtype": "f32",
        "config": {"workload": "c3-train-loop", "P0": a.P0, "P_target": a.target, "iterations": a.iterations, "views": a.views,
                   "densify_from": a.densify_from, "densify_every": a.densify_every,
                   "raw_parameter_path": bool(a.impl == "b200" and not a.no_raw),
                   "threshold": "quantile schedule (see tools/c3_trainer.py), not the reference's fixed 0.0002"},
        "value": float(np.mean(plain)), "final_P": g.P, "wall_s": wall, "buckets": buckets, "densifications": log})


def main():
    print(json.dumps(run(parse_args())))


if __name__ == "__main__":
    main()
