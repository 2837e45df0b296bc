#!/usr/bin/env python3
"""Diagnostic (GPU): how many (instance, pixel block) pairs the blending backward replays at C2, for 8x4 blocks (the
forward's warp masks as they are) and for 8x8 blocks (two 8x4 blocks stacked) -- the ratio decides whether two pixels
per lane pay off in the blending kernels.  Prints one JSON line."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "gaussian-splatting_cc-comments_b200"), os.path.join(ROOT, "tests"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)
import numpy as np
import torch
import gsr_synthetic as S
import diff_gaussian_rasterization as dgr


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 3_000_000
    W, H = 1920, 1080
    cam = S.make_camera(W, H)
    sc = S.make_scene(P, cam, sh_degree=3, M=16, seed=0)
    dev = torch.device("cuda")
    t = lambda a: torch.tensor(np.asarray(a), dtype=torch.float32, device=dev)
    st = dgr.GaussianRasterizationSettings(image_height=H, image_width=W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], bg=t([0, 0, 0]),
                                           scale_modifier=1.0, viewmatrix=t(cam["viewmatrix"]), projmatrix=t(cam["projmatrix"]),
                                           sh_degree=3, campos=t(cam["campos"]), prefiltered=False, debug=False)
    means3D = t(sc["means3D"]).requires_grad_(True)
    color, radii = dgr.GaussianRasterizer(st)(means3D=means3D, means2D=torch.zeros(P, 3, device=dev), opacities=t(sc["opacities"]),
                                              shs=t(sc["shs"]), scales=t(sc["scales"]), rotations=t(sc["rotations"]))
    fn = color.grad_fn
    geom, binning, img = fn.saved_tensors[-3:]
    R = fn.num_rendered
    views = dgr._C.view_buffers(P, R, W, H, geom, binning, img)
    wm = views[7].to(torch.int32)
    nc = views[2].to(torch.int64)
    # since v13 of the forward: one bit per 8x8 block of the tile (round 2 started with 8x4 blocks: 9 762 921 pairs at C2,
    # profiles/r02_mask_stats.json)
    pop = lambda x: sum(((x >> b) & 1) for b in range(4))
    pairs88 = int(pop(wm).sum())
    out = {"P": P, "R": R, "visible": int((radii > 0).sum()), "instances_with_any_block": int((wm != 0).sum()),
           "pairs_8x8": pairs88, "sum_n_contrib": int(nc.sum()), "pixel_slots_8x8": pairs88 * 64}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
