#!/usr/bin/env python3
"""Golden vectors for distCUDA2 from the reference's OWN simple-knn build (oracle/_ref/simple_knn, made by
oracle/build_ref.py: build_knn from the unmodified sources).  Needs a GPU: run on the B200 box
(`python tests/golden/make_golden_knn_gpu.py`, see tools/first_gpu_run_next_rows.sh); writes
gpurun_out/ref_knn_cuda.npz, to be committed as tests/golden/ref_knn_cuda.npz.  tests/test_knn_cpu.py picks the fixture up
when it exists and pins the CPU oracle to it (until then the oracle's header says "parity unpinned")."""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))


def main():
    so = os.path.join(ROOT, "oracle", "_ref", "simple_knn", "ref_simple_knn_C.so")
    if not (os.path.exists(so) and torch.cuda.is_available()):
        print("needs oracle/_ref/simple_knn/ref_simple_knn_C.so and a GPU")
        return 1
    spec = importlib.util.spec_from_file_location("ref_simple_knn_C", so)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    out = {}
    rng = np.random.default_rng(7)
    cases = {"normal_3000": rng.normal(size=(3000, 3)), "aniso_5000": rng.uniform(size=(5000, 3)) * np.array([100.0, 1.0, 0.01]),
             "two_scales_4000": np.concatenate([rng.normal(size=(2000, 3)) * 0.02 - 4.0, rng.uniform(-30, 30, size=(2000, 3))]),
             "tiny_4": rng.normal(size=(4, 3)), "tiny_3": rng.normal(size=(3, 3)), "box_edge_1025": rng.normal(size=(1025, 3))}
    cases["normal_3000"][:20] = cases["normal_3000"][20:40]          # coincident points
    for name, p in cases.items():
        p = p.astype(np.float32)
        out["pts_" + name] = p
        out["d2_" + name] = ref.distCUDA2(torch.tensor(p, device="cuda")).cpu().numpy()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    path = os.path.join(ROOT, "gpurun_out", "ref_knn_cuda.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, torch.cuda.get_device_name(0))
    return 0


if __name__ == "__main__":
    sys.exit(main())
