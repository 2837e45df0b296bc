#!/usr/bin/env python3
"""Golden vectors from the reference's own torch code, generated in THIS container (CPU).

Imports, unchanged, from /root/reference:
  utils/sh_utils.py:57-112           eval_sh            (the SH->RGB path behind --convert_SHs_python,
                                                          gaussian_renderer/__init__.py:93-99)
  utils/general_utils.py:78-150      build_scaling_rotation / strip_symmetric
                                     (the covariance path behind --compute_cov3D_python,
                                      scene/gaussian_model.py:32-37)
The covariance helpers hard-code device="cuda"; torch.zeros is wrapped for the duration of the
call so the reference's own lines run on the CPU.

Writes tests/golden/ref_torch_cpu.npz (inputs + outputs).  Run:  python tests/golden/make_golden_cpu.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "gaussian-splatting_cc-comments_b200"))
sys.path.insert(0, "/root/reference")
import gsr_synthetic as S  # noqa: E402
from utils.sh_utils import eval_sh  # noqa: E402
from utils import general_utils as GU  # noqa: E402


def ref_covariance(scales, scale_modifier, rotations):
    orig = torch.zeros

    def cpu_zeros(*a, **k):
        k["device"] = "cpu"
        return orig(*a, **k)

    torch.zeros = cpu_zeros
    try:
        L = GU.build_scaling_rotation(scale_modifier * scales, rotations)
        return GU.strip_symmetric(L @ L.transpose(1, 2))
    finally:
        torch.zeros = orig


def main():
    torch.set_num_threads(1)
    cam = S.look_at_camera(160, 96, eye=(0.3, -0.2, -0.5), target=(0.0, 0.0, 5.0))
    scene = S.make_scene(1500, cam, sh_degree=3, M=16, seed=11)
    means = torch.from_numpy(scene["means3D"])
    shs = torch.from_numpy(scene["shs"])                     # (P, 16, 3)
    campos = torch.from_numpy(cam["campos"])
    out = dict(means3D=scene["means3D"], shs=scene["shs"], scales=scene["scales"], rotations=scene["rotations"],
               campos=cam["campos"])
    for deg in range(4):
        # gaussian_renderer/__init__.py:93-99
        shs_view = shs.transpose(1, 2).view(-1, 3, 16)
        dir_pp = means - campos.repeat(shs.shape[0], 1)
        dir_pp_normalized = dir_pp / dir_pp.norm(dim=1, keepdim=True)
        sh2rgb = eval_sh(deg, shs_view, dir_pp_normalized)
        out["rgb_deg%d" % deg] = torch.clamp_min(sh2rgb + 0.5, 0.0).numpy()
        out["rgb_raw_deg%d" % deg] = (sh2rgb + 0.5).numpy()
    for mod in (1.0, 0.7):
        out["cov3D_mod%s" % mod] = ref_covariance(torch.from_numpy(scene["scales"]), mod,
                                                  torch.from_numpy(scene["rotations"])).numpy()
    np.savez_compressed(os.path.join(HERE, "ref_torch_cpu.npz"), **out)
    print("wrote", os.path.join(HERE, "ref_torch_cpu.npz"), {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
