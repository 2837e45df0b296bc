#!/usr/bin/env python3
"""Stage-by-stage diagnostic on a GPU box: CUDA path vs reference build vs CPU oracle.
Writes gpurun_out/diag.json.  Not a test; used to localise mismatches quickly."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import helpers as Hh  # noqa: E402
from helpers import S  # noqa: E402


def cmp_stage(tag, mine, ref, res):
    vis = ref["radii"] > 0
    r = {}
    r["radii_mismatch"] = int((mine["radii"] != ref["radii"]).sum())
    r["num_rendered"] = [int(mine["num_rendered"]), int(ref["num_rendered"])]
    r["tiles_touched_mismatch"] = int((mine["tiles_touched"][vis] != ref["tiles_touched"][vis]).sum())
    r["means2D_bits_mismatch"] = int((mine["means2D"][vis].view(np.uint32) != ref["means2D"][vis].view(np.uint32)).sum())
    r["conic_bits_mismatch"] = int((mine["conic_opacity"][vis].view(np.uint32) != ref["conic_opacity"][vis].view(np.uint32)).sum())
    r["conic_maxrel"] = float(np.max(np.abs(mine["conic_opacity"][vis] - ref["conic_opacity"][vis]) /
                                     (np.abs(ref["conic_opacity"][vis]) + 1e-30))) if vis.any() else 0.0
    if "rgb" in ref and ref["rgb"] is not None:
        r["rgb_bits_mismatch"] = int((mine["rgb"][vis].view(np.uint32) != ref["rgb"][vis].view(np.uint32)).sum())
        r["rgb_maxabs"] = float(np.abs(mine["rgb"][vis] - ref["rgb"][vis]).max()) if vis.any() else 0.0
    if mine["num_rendered"] == ref["num_rendered"]:
        r["point_list_mismatch"] = int((mine["point_list"] != ref["point_list"]).sum())
        r["ranges_mismatch"] = int((mine["ranges"] != ref["ranges"]).sum())
    r["n_contrib_mismatch"] = int((mine["n_contrib"] != ref["n_contrib"]).sum())
    r["final_T_maxabs"] = float(np.abs(mine["final_T"] - ref["final_T"]).max())
    r["final_T_bits_mismatch"] = int((mine["final_T"].view(np.uint32) != ref["final_T"].view(np.uint32)).sum())
    r["color_maxabs"] = float(np.abs(mine["color"] - ref["color"]).max())
    r["color_bits_mismatch"] = int((mine["color"].view(np.uint32) != ref["color"].view(np.uint32)).sum())
    if "grads" in mine and "grads" in ref:
        r["grad_rel_l2"] = {k: (float(Hh.rel_l2(mine["grads"][k], ref["grads"][k]))
                                if mine["grads"].get(k) is not None and ref["grads"].get(k) is not None else None)
                            for k in Hh.GRAD_KEYS}
    res[tag] = r
    print(tag, json.dumps(r), flush=True)


def main():
    import torch
    mod = Hh.my_module()
    ref = Hh.ref_module() if Hh.have_ref() else None
    res = {"device": torch.cuda.get_device_name(0), "have_ref": ref is not None}
    cases = [("c1_10k_256", 10000, 256, 256, 0, 1), ("mid_100k_640x360", 100000, 640, 360, 3, 16),
             ("odd_20k_250x130_d2", 20000, 250, 130, 2, 16)]
    if len(sys.argv) > 1 and sys.argv[1] == "big":
        cases.append(("c2_3M_1080p", 3000000, 1920, 1080, 3, 16))
    for name, P, W, H, deg, M in cases:
        cam = S.make_camera(W, H)
        scene = S.make_scene(P, cam, sh_degree=deg, M=M, seed=0)
        dL = S.upstream_grad(cam)
        bg = (0.3, 0.5, 0.7)
        t0 = time.time()
        mine = Hh.run_torch(mod, cam, scene, dL, bg=bg)
        torch.cuda.synchronize()
        print(name, "mine done", round(time.time() - t0, 2), "R", mine["num_rendered"], flush=True)
        if ref is not None:
            rr = Hh.run_torch(ref, cam, scene, dL, bg=bg, is_ref=True)
            cmp_stage(name + "/vs_ref", mine, rr, res)
        if P <= 100000:
            oo = Hh.run_oracle(cam, scene, dL, bg=bg)
            oo["color"] = oo["color"]
            cmp_stage(name + "/vs_oracle", mine, oo, res)
            if ref is not None:
                cmp_stage(name + "/ref_vs_oracle", rr, oo, res)
    os.makedirs(os.path.join(Hh.ROOT, "gpurun_out"), exist_ok=True)
    json.dump(res, open(os.path.join(Hh.ROOT, "gpurun_out", "diag.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
