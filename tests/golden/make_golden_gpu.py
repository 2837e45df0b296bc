#!/usr/bin/env python3
"""Golden vectors from the UNMODIFIED reference CUDA rasterizer (oracle/_ref, built for sm_100 by
oracle/build_ref.py), generated on a B200 box:

    gpurun -- 'python tests/golden/make_golden_gpu.py'      # writes gpurun_out/golden/*.npz
    cp gpurun_out/golden/ref_cuda_*.npz tests/golden/

Each file holds the inputs (numpy, float32) and everything the reference produced for them:
radii, num_rendered, the sorted point list, tile ranges, n_contrib, final_T, the image, the
per-Gaussian intermediates sliced out of its geometry buffer, and the 6-8 gradients.
The reference ships no fixtures of its own; these pin the CPU oracle and the
CUDA path to the reference's behaviour without needing /root/reference at test time.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as Hh  # noqa: E402
from helpers import S  # noqa: E402

CASES = {
    # name: (P, W, H, degree, M, use, cov, bg, scale_modifier, camera)
    "sh3_160x96": (1500, 160, 96, 3, 16, "sh", "sr", (0.25, 0.5, 0.75), 1.0, "lookat"),
    "precomp_100x70": (800, 100, 70, 0, 1, "colors", "cov3d", (0.0, 0.0, 0.0), 1.0, "origin"),
    "sh1_mod_64x64": (600, 64, 64, 1, 4, "sh", "sr", (1.0, 1.0, 1.0), 0.6, "origin"),
}


def make_case(name):
    P, W, H, deg, M, use, cov, bg, smod, camk = CASES[name]
    cam = S.look_at_camera(W, H, eye=(0.3, -0.2, -0.5), target=(0.0, 0.0, 5.0)) if camk == "lookat" else S.make_camera(W, H)
    scene = S.make_scene(P, cam, sh_degree=deg, M=M, seed=11, s0_pix=3.0)
    if use == "colors":
        rng = np.random.default_rng(5)
        scene["colors_precomp"] = rng.uniform(0, 1, (P, 3)).astype(np.float32)
    if cov == "cov3d":
        from oracle import gs_oracle as O
        scene["cov3D_precomp"] = O.cov3d(scene["scales"], scene["rotations"], 1.0)
    dL = S.upstream_grad(cam, seed=2)
    return cam, scene, dL, dict(use=use, cov=cov, bg=bg, scale_modifier=smod)


def main():
    ref = Hh.ref_module()
    out_dir = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out_dir, exist_ok=True)
    for name in CASES:
        cam, scene, dL, kw = make_case(name)
        r = Hh.run_torch(ref, cam, scene, dL, is_ref=True, **kw)
        rec = {"in_" + k: v for k, v in scene.items() if isinstance(v, np.ndarray)}
        rec.update({"cam_" + k: np.asarray(v) for k, v in cam.items()})
        rec["dL"] = dL
        rec["bg"] = np.asarray(kw["bg"], np.float32)
        for k in ("color", "radii", "point_list", "ranges", "n_contrib", "final_T", "means2D", "conic_opacity", "rgb",
                  "depths", "cov3D", "tiles_touched", "clamped"):
            rec["out_" + k] = r[k]
        rec["out_num_rendered"] = np.asarray(r["num_rendered"])
        for k, g in r["grads"].items():
            if g is not None:
                rec["grad_" + k] = g
        np.savez_compressed(os.path.join(out_dir, "ref_cuda_%s.npz" % name), **rec)
        print(name, "R =", r["num_rendered"], "visible =", int((r["radii"] > 0).sum()))


if __name__ == "__main__":
    main()
