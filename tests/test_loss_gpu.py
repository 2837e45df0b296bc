"""GPU parity of the fused L1 + SSIM loss (csrc/gsr_loss.cu through the C ABI and loss_utils.py) against vectors from the
reference's own utils/loss_utils.py and against the float64 checker.  Tolerances: values 2e-6 absolute (means of
O(1) quantities in float32), gradient 1e-4 relative L2 (north_star's gradient bar)."""
import os

import numpy as np
import pytest
import torch

import loss_ref as LR
from helpers import rel_l2

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_loss_cpu.npz")


def _lu():
    import loss_utils
    return loss_utils


def _cases():
    z = np.load(GOLD)
    n = len({k.split("_")[0] for k in z.files})
    return [{k[len("c%d_" % i):]: z[k] for k in z.files if k.startswith("c%d_" % i)} for i in range(n)]


def test_fused_loss_matches_the_reference_vectors():
    lu = _lu()
    for c in _cases():
        img = torch.tensor(c["img"], device="cuda", requires_grad=True)
        gt = torch.tensor(c["gt"], device="cuda")
        loss, l1, s = lu.fused_l1_ssim_loss(img, gt, float(c["lambda"]))
        loss.backward()
        assert abs(float(l1) - float(c["l1"])) <= 2e-6
        assert abs(float(s) - float(c["ssim"])) <= 2e-6
        assert abs(float(loss.detach()) - float(c["loss"])) <= 2e-6
        assert rel_l2(img.grad.cpu().numpy(), c["grad"]) <= 1e-4
        assert not gt.requires_grad


def test_ssim_and_l1_functions_keep_the_reference_signature():
    lu = _lu()
    c = _cases()[0]
    img = torch.tensor(c["img"], device="cuda", requires_grad=True)
    gt = torch.tensor(c["gt"], device="cuda")
    s = lu.ssim(img, gt)                               # ssim(img1, img2, window_size=11, size_average=True)
    Ll1 = lu.l1_loss(img, gt)
    loss = (1.0 - 0.2) * Ll1 + 0.2 * (1.0 - s)         # train.py:126-127 composed by the caller
    loss.backward()
    assert abs(float(s) - float(c["ssim"])) <= 2e-6 and abs(float(loss) - float(c["loss"])) <= 2e-6
    assert rel_l2(img.grad.cpu().numpy(), c["grad"]) <= 1e-4
    with pytest.raises(RuntimeError):
        lu.ssim(img.detach().cpu(), gt.cpu())          # no CPU path
    with pytest.raises(NotImplementedError):
        lu.ssim(img, gt, window_size=7)


@pytest.mark.parametrize("shape", [(3, 1, 1), (3, 5, 300), (3, 270, 481), (1, 3, 64, 48)])
def test_fused_loss_matches_the_float64_checker_on_ragged_sizes(shape):
    lu = _lu()
    g = torch.Generator().manual_seed(11)
    gt = torch.rand(shape, generator=g)
    img = (gt + 0.2 * torch.randn(shape, generator=g)).clamp(0, 1)
    if shape[-1] * shape[-2] > 1:
        img[..., 0, 0] = gt[..., 0, 0]                 # |x - y| = 0 somewhere: the subgradient there is 0 (torch.abs)
    r = LR.loss_and_grad(img, gt, 0.2)
    a = img.cuda().requires_grad_(True)
    loss, l1, s = lu.fused_l1_ssim_loss(a, gt.cuda(), 0.2)
    (3.0 * loss).backward()                            # a non-unit upstream gradient
    assert abs(float(loss) - r["loss"]) <= 2e-6 and abs(float(l1) - r["l1"]) <= 2e-6 and abs(float(s) - r["ssim"]) <= 2e-6
    assert rel_l2(a.grad.cpu().numpy(), 3.0 * r["grad"]) <= 1e-4


def test_fused_loss_is_deterministic_and_needs_no_gradient_buffers_for_evaluation():
    lu = _lu()
    g = torch.Generator().manual_seed(5)
    gt = torch.rand((3, 200, 333), generator=g).cuda()
    img = torch.rand((3, 200, 333), generator=g).cuda()
    a = lu.fused_l1_ssim_loss(img, gt)[0]
    b = lu.fused_l1_ssim_loss(img, gt)[0]
    assert float(a) == float(b) and not a.requires_grad
    c = lu.fused_l1_ssim_loss(img.clone().requires_grad_(True), gt)[0]
    assert float(c) == float(a)
