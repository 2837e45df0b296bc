"""CPU checks for the fused L1 + SSIM loss: the float64 checker against vectors produced by the
reference's own utils/loss_utils.py, and the kernel's derivative algebra (csrc/gsr_loss.cu) against autograd."""
import os

import numpy as np
import torch
import torch.nn.functional as F

import loss_ref as LR

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_loss_cpu.npz")


def _cases():
    z = np.load(GOLD)
    n = len({k.split("_")[0] for k in z.files})
    return [{k[len("c%d_" % i):]: z[k] for k in z.files if k.startswith("c%d_" % i)} for i in range(n)]


def test_checker_reproduces_the_reference_functions():
    for c in _cases():
        r = LR.loss_and_grad(torch.tensor(c["img"]), torch.tensor(c["gt"]), float(c["lambda"]))
        assert abs(r["l1"] - float(c["l1"])) <= 2e-7 * max(1.0, abs(r["l1"]))
        assert abs(r["ssim"] - float(c["ssim"])) <= 2e-6
        assert abs(r["loss"] - float(c["loss"])) <= 2e-6
        g = c["grad"].astype(np.float64)
        assert np.linalg.norm(r["grad"] - g) <= 2e-5 * np.linalg.norm(g)


def _kernel_algebra(img, gt, coef_l1, coef_ssim):
    """What csrc/gsr_loss.cu computes, in float64 torch: three derivative maps, then three convolutions."""
    x, y = img.double(), gt.double()
    squeeze = x.dim() == 3
    if squeeze:
        x, y = x.unsqueeze(0), y.unsqueeze(0)
    C = x.shape[1]
    w1 = LR.window_1d().double()
    win = (w1.unsqueeze(1) @ w1.unsqueeze(0)).expand(C, 1, 11, 11).contiguous()
    conv = lambda t: F.conv2d(t, win, padding=5, groups=C)
    mu1, mu2, q, q2, r = conv(x), conv(y), conv(x * x), conv(y * y), conv(x * y)
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    s1, s2, s12 = q - mu1 * mu1, q2 - mu2 * mu2, r - mu1 * mu2
    A1, A2, B1, B2 = 2 * mu1 * mu2 + C1, 2 * s12 + C2, mu1 * mu1 + mu2 * mu2 + C1, s1 + s2 + C2
    inv = 1.0 / (B1 * B2)
    m = A1 * A2 * inv
    dm_ds1 = -m / B2
    dm_ds12 = 2 * A1 * inv
    dm_dp = 2 * mu2 * A2 * inv - 2 * mu1 * m / B1 + dm_ds1 * (-2 * mu1) + dm_ds12 * (-mu2)
    grad = coef_l1 * torch.sign(x - y) + coef_ssim * (conv(dm_dp) + 2 * x * conv(dm_ds1) + y * conv(dm_ds12))
    return m.mean().item(), (grad.squeeze(0) if squeeze else grad).numpy()


def test_kernel_derivative_algebra_equals_autograd():
    for c in _cases():
        img, gt, lam = torch.tensor(c["img"]), torch.tensor(c["gt"]), float(c["lambda"])
        n = img.numel()
        r = LR.loss_and_grad(img, gt, lam)
        s, g = _kernel_algebra(img, gt, (1 - lam) / n, -lam / n)
        # (the checker rounds the 2-D window to float32 as the reference does, the separable form does not: 1e-9 apart)
        assert abs(s - r["ssim"]) <= 1e-8
        assert np.linalg.norm(g - r["grad"]) <= 1e-6 * np.linalg.norm(r["grad"])
        _, gs = _kernel_algebra(img, gt, 0.0, 1.0 / n)
        assert np.linalg.norm(gs - r["grad_ssim"]) <= 1e-6 * np.linalg.norm(r["grad_ssim"])
