#!/usr/bin/env python3
"""Golden vectors for the L1 + SSIM loss from the reference's OWN functions (utils/loss_utils.py, imported unchanged from
/root/reference, float32 on the CPU).  Run in the build container only; writes tests/golden/ref_loss_cpu.npz.

    python tests/golden/make_golden_loss.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from utils.loss_utils import l1_loss, ssim  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def case(seed, shape, lam):
    g = torch.Generator().manual_seed(seed)
    gt = torch.rand(shape, generator=g)
    # a render that resembles its target: target + smooth error + noise, clipped like an image
    img = (gt + 0.15 * torch.randn(shape, generator=g) * torch.rand(shape, generator=g)).clamp(0, 1)
    img.requires_grad_(True)
    Ll1 = l1_loss(img, gt)
    s = ssim(img, gt)
    loss = (1.0 - lam) * Ll1 + lam * (1.0 - s)            # train.py:126-127
    loss.backward()
    return {"img": img.detach().numpy(), "gt": gt.numpy(), "lambda": np.float32(lam), "l1": np.float32(Ll1.item()),
            "ssim": np.float32(s.item()), "loss": np.float32(loss.item()), "grad": img.grad.numpy()}


def main():
    torch.set_num_threads(1)
    out = {}
    for i, (seed, shape, lam) in enumerate([(1, (3, 37, 53), 0.2), (2, (3, 16, 16), 0.2), (3, (1, 9, 70), 0.5), (4, (2, 3, 33, 20), 0.2)]):
        for k, v in case(seed, shape, lam).items():
            out["c%d_%s" % (i, k)] = v
    np.savez_compressed(os.path.join(HERE, "ref_loss_cpu.npz"), **out)
    print("wrote", os.path.join(HERE, "ref_loss_cpu.npz"), {k: out[k] for k in out if k.endswith(("_loss", "_ssim", "_l1"))})


if __name__ == "__main__":
    main()
