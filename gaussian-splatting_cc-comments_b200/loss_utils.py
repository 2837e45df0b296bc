"""Training loss of the reference (utils/loss_utils.py:17-64, train.py:126-127) on fused CUDA kernels for sm_100a.

Same names and arguments as the reference module, so `from utils.loss_utils import l1_loss, ssim` can be pointed here:

    l1_loss(network_output, gt)                        -> mean |a - b|                    (loss_utils.py:17-18)
    l2_loss(network_output, gt)                        -> mean (a - b)^2                  (loss_utils.py:20-21)
    ssim(img1, img2, window_size=11, size_average=True) -> mean SSIM, Gaussian window 11 / sigma 1.5, zero padding (:33-64)

plus the one call a training loop wants, the whole expression of train.py:126-127 in two kernels forward and one backward:

    loss, Ll1, ssim_value = fused_l1_ssim_loss(image, gt_image, lambda_dssim=0.2)

`ssim` and `fused_l1_ssim_loss` run `csrc/gsr_loss.cu` through the C ABI (include/gsr.h: gsr_l1_ssim_forward /
gsr_l1_ssim_backward).  CUDA float32 tensors only: there is no CPU path and nothing falls back to torch ops.
Gradients flow to the first image (the render); the second is the ground truth and gets none, as in train.py.
"""
import torch

from diff_gaussian_rasterization import _C


def l1_loss(network_output, gt):
    return torch.abs((network_output - gt)).mean()


def l2_loss(network_output, gt):
    return ((network_output - gt) ** 2).mean()


def _check(img1, img2):
    if not (img1.is_cuda and img2.is_cuda):
        raise RuntimeError("loss_utils: CUDA tensors only (there is no CPU path)")
    if img1.dtype != torch.float32 or img2.dtype != torch.float32:
        raise TypeError("loss_utils: float32 images")
    if img1.shape != img2.shape or img1.dim() not in (3, 4):
        raise ValueError("loss_utils: images must share a (C,H,W) or (B,C,H,W) shape")


class _FusedL1SSIM(torch.autograd.Function):
    """which = 0: the train.py loss, 1: mean |a - b|, 2: mean ssim -- all three come out of the same forward kernel."""

    @staticmethod
    def forward(ctx, img, gt, lambda_dssim, which):
        need_grad = img.requires_grad
        out3, maps = _C.l1_ssim_forward(img, gt, float(lambda_dssim), bool(need_grad))
        n = float(img.numel())
        if which == 0:
            ctx.coefs = ((1.0 - lambda_dssim) / n, -lambda_dssim / n)
        elif which == 1:
            ctx.coefs = (1.0 / n, 0.0)
        else:
            ctx.coefs = (0.0, 1.0 / n)
        ctx.save_for_backward(img, gt, maps)
        ctx.mark_non_differentiable(out3)
        return out3[which].clone(), out3

    @staticmethod
    def backward(ctx, grad_value, _grad_out3):
        img, gt, maps = ctx.saved_tensors
        if maps.numel() == 0:
            return None, None, None, None
        g = grad_value.reshape(1).to(torch.float32)
        dimg = _C.l1_ssim_backward(img, gt, maps, ctx.coefs[0], ctx.coefs[1], g)
        return dimg.view_as(img), None, None, None


def fused_l1_ssim_loss(image, gt_image, lambda_dssim=0.2):
    """(loss, Ll1, ssim) with loss = (1 - lambda_dssim) * l1_loss(image, gt) + lambda_dssim * (1 - ssim(image, gt)), train.py:126-127.
    `loss` carries the gradient; Ll1 and ssim are detached values for logging (train.py:146,209)."""
    _check(image, gt_image)
    loss, out3 = _FusedL1SSIM.apply(image, gt_image, float(lambda_dssim), 0)
    return loss, out3[1], out3[2]


def ssim(img1, img2, window_size=11, size_average=True):
    _check(img1, img2)
    if window_size != 11 or not size_average:
        raise NotImplementedError("loss_utils.ssim: the fused kernels implement the configuration the reference trains with "
                                  "(window_size=11, size_average=True)")
    value, _ = _FusedL1SSIM.apply(img1, img2, 1.0, 2)
    return value
