"""Checker for the fused L1 + SSIM loss: the reference's formula (utils/loss_utils.py:17-64, train.py:126-127) restated
with plain torch ops in float64 on the CPU.  Test infrastructure only; pinned against outputs of the reference's own
functions in tests/golden/ref_loss_cpu.npz (tests/test_loss_cpu.py)."""
from math import exp

import torch
import torch.nn.functional as F


def window_1d(window_size=11, sigma=1.5):
    g = torch.tensor([exp(-(x - window_size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(window_size)], dtype=torch.float32)
    return g / g.sum()                                   # float32, as the reference builds it (loss_utils.py:23-25)


def ssim_map(img1, img2, window_size=11):
    """img: (C,H,W) or (B,C,H,W) float64.  Returns the SSIM map of the same shape (loss_utils.py:44-59)."""
    squeeze = img1.dim() == 3
    a = img1.unsqueeze(0) if squeeze else img1
    b = img2.unsqueeze(0) if squeeze else img2
    C = a.shape[1]
    w1 = window_1d(window_size)
    w2 = (w1.unsqueeze(1) @ w1.unsqueeze(0)).float()     # float32 outer product (loss_utils.py:29), then promoted
    win = w2.to(a.dtype).expand(C, 1, window_size, window_size).contiguous()
    pad = window_size // 2
    mu1 = F.conv2d(a, win, padding=pad, groups=C)
    mu2 = F.conv2d(b, win, padding=pad, groups=C)
    mu1_sq, mu2_sq, mu1_mu2 = mu1 * mu1, mu2 * mu2, mu1 * mu2
    s1 = F.conv2d(a * a, win, padding=pad, groups=C) - mu1_sq
    s2 = F.conv2d(b * b, win, padding=pad, groups=C) - mu2_sq
    s12 = F.conv2d(a * b, win, padding=pad, groups=C) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    m = ((2 * mu1_mu2 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s1 + s2 + C2))
    return m.squeeze(0) if squeeze else m


def loss_and_grad(img, gt, lambda_dssim=0.2):
    """float64 evaluation: dict(loss, l1, ssim, grad) with grad = d loss / d img."""
    a = img.detach().double().cpu().requires_grad_(True)
    b = gt.detach().double().cpu()
    l1 = (a - b).abs().mean()
    s = ssim_map(a, b).mean()
    loss = (1.0 - lambda_dssim) * l1 + lambda_dssim * (1.0 - s)
    (grad,) = torch.autograd.grad(loss, a, retain_graph=True)
    (grad_ssim,) = torch.autograd.grad(s, a)
    return {"loss": float(loss), "l1": float(l1), "ssim": float(s), "grad": grad.numpy(), "grad_ssim": grad_ssim.numpy()}
