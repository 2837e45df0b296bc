#!/usr/bin/env python3
"""Micro-benchmark of the training loss at the image size of BASELINE.json's C2 (3 x 1080 x 1920): the reference's
composition of torch ops (utils/loss_utils.py:17-64 + train.py:126-127, restated below with the same calls, TF32 off)
against the fused kernels (loss_utils.fused_l1_ssim_loss), forward + backward, CUDA events, one JSON line."""
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gaussian-splatting_cc-comments_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import loss_utils as LU  # noqa: E402
import loss_ref as LR  # noqa: E402


def torch_composition(img, gt, win, lam=0.2):
    C = img.size(-3)
    conv = lambda t: F.conv2d(t, win, padding=5, groups=C)
    mu1, mu2 = conv(img), conv(gt)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1, s2, s12 = conv(img * img) - mu1_sq, conv(gt * gt) - mu2_sq, conv(img * gt) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim = (((2 * mu1_mu2 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s1 + s2 + C2))).mean()
    return (1.0 - lam) * torch.abs(img - gt).mean() + lam * (1.0 - ssim)


def timed(fn, steps, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def main():
    H, W = 1080, 1920
    torch.backends.cudnn.allow_tf32 = False
    g = torch.Generator().manual_seed(0)
    gt = torch.rand((3, H, W), generator=g).cuda()
    img = (gt + 0.1 * torch.randn((3, H, W), generator=g).cuda()).clamp(0, 1).requires_grad_(True)
    w1 = LR.window_1d().cuda()
    win = (w1.unsqueeze(1) @ w1.unsqueeze(0)).expand(3, 1, 11, 11).contiguous()

    def ref_step():
        img.grad = None
        torch_composition(img, gt, win).backward()

    def fused_step():
        img.grad = None
        LU.fused_l1_ssim_loss(img, gt, 0.2)[0].backward()

    ref_step(); g_ref = img.grad.clone(); l_ref = float(torch_composition(img, gt, win))
    fused_step(); g_fus = img.grad.clone(); l_fus = float(LU.fused_l1_ssim_loss(img, gt, 0.2)[0])
    ms_ref, ms_fus = timed(ref_step, 20, 5), timed(fused_step, 20, 5)
    bytes_alg = 3 * H * W * 4 * (2 + 3 + 2 + 3 + 1)     # fwd: 2 images in, 3 maps out; bwd: 2 images + 3 maps in, 1 gradient out
    print(json.dumps({"metric": "ms per L1+SSIM loss fwd+bwd, 3x1080x1920", "fused_ms": ms_fus, "torch_ops_ms": ms_ref,
                      "speedup": ms_ref / ms_fus, "loss_fused": l_fus, "loss_torch_ops": l_ref,
                      "grad_rel_l2": float((g_fus - g_ref).norm() / g_ref.norm()),
                      "algorithmic_bytes": bytes_alg, "fused_gbs": bytes_alg / ms_fus / 1e6}))


if __name__ == "__main__":
    main()
