#!/usr/bin/env python3
"""Micro-benchmarks of the SURVEY 8(f) rows next to the rasterizer, one JSON line each (same keys as bench.py where they
apply: metric / value / unit / roofline / cpu_baseline / gpu_launches):

  adam : one optimizer step over the six parameter groups of a 3 M-Gaussian model (59 floats per Gaussian and their two
         moments) -- gsr_optim.FusedAdam (one kernel, 28 B per element) against torch.optim.Adam on the same GPU (the
         reference's optimizer, scene/gaussian_model.py:252) and the CPU oracle on a bounded sample.
  knn  : distCUDA2 on 1 M random points (the size of a large COLMAP cloud) -- this library's kernels against the reference's
         own simple-knn build (oracle/_ref/simple_knn, when present) and the CPU oracle on a bounded sample.

  densify : densify_and_prune (scene/gaussian_model.py:558-597) on a 3 M-Gaussian model with Adam state, and the per-iteration
         statistics update (train.py:153-158) -- gsr_densify against the torch statements on the same GPU and the CPU oracle.

Timing: CUDA events on the current stream, warm-up, synchronise on both sides; inputs larger than L2 (adam) or re-sorted
from scratch every step (knn).  Run on the GPU box:  python tools/bench_next.py [adam|knn|densify|all] [--P N] [--steps K]
"""
import argparse
import importlib.util
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gaussian-splatting_cc-comments_b200"))
sys.path.insert(0, ROOT)

NAMES = ["xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation"]
SHAPES = {"xyz": (3,), "f_dc": (1, 3), "f_rest": (15, 3), "opacity": (1,), "scaling": (3,), "rotation": (4,)}
LRS = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=1.25e-4, opacity=0.05, scaling=0.005, rotation=0.001)


def peak_hbm():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        return 6551.0, "fallback: B200_PROFILING.md measured copy bandwidth"


def timed(fn, steps, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def launches():
    import ctypes as C
    lib = C.CDLL(os.path.join(ROOT, "gaussian-splatting_cc-comments_b200", "libgsr_b200.so"))
    lib.gsr_launch_count.restype = C.c_ulonglong
    return int(lib.gsr_launch_count())


def bench_adam(P, steps, warmup):
    import gsr_optim
    from oracle import gs_next_oracle as O
    g = torch.Generator().manual_seed(0)

    def make(cls):
        ps = {n: torch.nn.Parameter(torch.randn(P, *SHAPES[n], generator=g).cuda()) for n in NAMES}
        opt = cls([{"params": [ps[n]], "lr": LRS[n], "name": n} for n in NAMES], lr=0.0, eps=1e-15)
        for n in NAMES:
            ps[n].grad = torch.randn(P, *SHAPES[n], generator=g).cuda() * 1e-3
        return ps, opt
    _, mine = make(gsr_optim.FusedAdam)
    _, ref = make(torch.optim.Adam)
    l0 = launches()
    ms_mine = timed(mine.step, steps, warmup)
    n_launch = (launches() - l0) / (steps + warmup)
    ms_ref = timed(ref.step, steps, warmup)
    elems = P * 59
    alg = 28 * elems                                   # 4 arrays read, 3 written, 4 bytes each
    peak, how = peak_hbm()
    # CPU baseline: the oracle (scalar C, one thread) on a bounded sample, scaled by element count
    sample = min(elems, 20_000_000)
    a = [np.random.default_rng(0).normal(size=sample).astype(np.float32) for _ in range(2)] + [np.zeros(sample, np.float32), np.zeros(sample, np.float32)]
    t = time.time(); O.adam_step(a[0], a[1], a[2], a[3], 1e-3, 1); cpu_s = (time.time() - t) * elems / sample
    print(json.dumps({
        "metric": "optimizer steps/s, Adam over the six parameter groups of a %d-Gaussian model (59 floats each)" % P,
        "value": 1000.0 / ms_mine, "unit": "steps/s", "ms_per_step": ms_mine, "higher_is_better": True, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "adam-step", "P": P, "elements": elems, "cache": "1.4 GB of state per step at 3 M: larger than L2"},
        "steps": steps, "warmup": warmup, "gpu_launches": n_launch,
        "reference_same_gpu": {"impl": "torch.optim.Adam (foreach)", "ms_per_step": ms_ref, "speedup": ms_ref / ms_mine},
        "roofline": {"bound": "hbm", "achieved": alg / ms_mine / 1e6, "peak": peak, "unit": "GB/s", "frac": alg / ms_mine / 1e6 / peak,
                     "traffic": None, "peak_source": how, "algorithmic_bytes": alg},
        "cpu_baseline": {"value": 1.0 / cpu_s, "unit": "steps/s", "cores": 1, "kind": "port",
                         "sample": "%d of %d elements, scaled by element count" % (sample, elems)}}))


def ref_knn_module():
    so = os.path.join(ROOT, "oracle", "_ref", "simple_knn", "ref_simple_knn_C.so")
    if not os.path.exists(so):
        return None
    spec = importlib.util.spec_from_file_location("ref_simple_knn_C", so)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def bench_knn(P, steps, warmup):
    from simple_knn._C import distCUDA2
    from oracle import gs_next_oracle as O
    rng = np.random.default_rng(0)
    pts_h = (rng.normal(size=(P, 3)) * np.array([3.0, 1.0, 2.0])).astype(np.float32)
    pts = torch.tensor(pts_h, device="cuda")
    l0 = launches()
    ms_mine = timed(lambda: distCUDA2(pts), steps, warmup)
    n_launch = (launches() - l0) / (steps + warmup)
    ref = ref_knn_module()
    ms_ref = timed(lambda: ref.distCUDA2(pts), max(1, steps // 4), 1) if ref is not None else None
    same = bool(torch.equal(ref.distCUDA2(pts), distCUDA2(pts))) if ref is not None else None
    # algorithmic bytes: read the points, write the means, sort (code, index) pairs in 4 passes, write + read the 16-byte records
    alg = P * (12 + 4 + 4 * 20 + 2 * 16)
    peak, how = peak_hbm()
    sample = min(P, 1_000_000)
    t = time.time(); O.knn_mean_dist2(pts_h[:sample]); cpu_s = time.time() - t
    print(json.dumps({
        "metric": "points/s, mean squared distance to the 3 nearest neighbours (distCUDA2), %d points" % P,
        "value": P / ms_mine * 1e3, "unit": "points/s", "ms_per_step": ms_mine, "higher_is_better": True, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "distCUDA2", "P": P, "distribution": "anisotropic normal"},
        "steps": steps, "warmup": warmup, "gpu_launches": n_launch,
        "reference_same_gpu": {"impl": "reference simple-knn rebuilt for sm_100 (oracle/_ref/simple_knn)", "ms_per_step": ms_ref,
                               "speedup": (ms_ref / ms_mine) if ms_ref else None, "bit_identical": same},
        "roofline": {"bound": "hbm", "achieved": alg / ms_mine / 1e6, "peak": peak, "unit": "GB/s", "frac": alg / ms_mine / 1e6 / peak,
                     "traffic": None, "peak_source": how, "algorithmic_bytes": alg,
                     "note": "the search kernel is bound by FP32 issue + shared-memory broadcasts, not HBM; this fraction is a floor"},
        "cpu_baseline": {"value": sample / cpu_s, "unit": "points/s", "cores": os.cpu_count(), "kind": "port",
                         "sample": "first %d points of the same cloud (OpenMP oracle)" % sample}}))


def bench_densify(P, steps, warmup):
    """densify_and_prune on a P-Gaussian model with populated Adam state: the two library calls (gsr_densify) against the
    three-round torch sequence (tools/c3_trainer.py, the restatement of scene/gaussian_model.py:366-597) on the same GPU.
    Every step starts from the same model (the input tensors are kept; outputs are new tensors), so the timed region is
    classify + totals read-back + the sampling call + apply + the optimizer hand-over.  Also the per-iteration statistics
    update (train.py:153-158) against its torch statements."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import c3_trainer as T
    import gsr_densify
    import gsr_optim
    import gsr_synthetic as S
    from oracle import gs_next_oracle as O
    cam = S.make_camera(64, 64)
    scene = S.make_scene(P, cam, sh_degree=3, seed=3)
    peak, how = peak_hbm()

    def fresh(fused):
        g = T.Gaussians(scene, "cuda", torch.optim.Adam, gsr_optim.get_expon_lr_func, fused_densify=fused)
        gen = torch.Generator(device="cuda").manual_seed(5)
        for n in T.GROUPS:
            getattr(g, n).grad = torch.randn(getattr(g, n).shape, device="cuda", generator=gen) * 1e-3
        g.optimizer.step()
        g.optimizer.zero_grad(set_to_none=True)
        g.denom = torch.randint(1, 40, (P, 1), device="cuda", generator=gen).float() * (torch.rand(P, 1, device="cuda", generator=gen) > 0.15)
        g.xyz_gradient_accum = torch.rand(P, 1, device="cuda", generator=gen) * 6e-4 * g.denom
        return g
    extent = float(np.median(np.asarray(scene["scales"]).max(axis=1))) / T.OptArgs.percent_dense
    out = {}
    for arm, fused in (("fused", gsr_densify), ("torch", None)):
        ms = []
        l0 = launches()
        for it in range(warmup + steps):
            g = fresh(fused)
            torch.manual_seed(77)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            g.densify_and_prune(0.0002, 0.005, extent, 20)
            e1.record()
            torch.cuda.synchronize()
            if it >= warmup:
                ms.append(e0.elapsed_time(e1))
            P_out = g.P
            del g
        out[arm] = (float(np.median(ms)), P_out, (launches() - l0) / (warmup + steps))
    ms_mine, P_out, n_launch = out["fused"]
    # algorithmic bytes: classify reads accum, denom, scaling, opacity (24 B) and writes a flag byte; apply reads the flag
    # and, per output row, reads and writes parameter + two moments (59 floats x 3 x 4 B each way; new rows read no moments)
    alg = P * 25 + P + P_out * 59 * 4 * 3 * 2
    # CPU baseline: the oracle (the reference's three rounds in scalar C) on a bounded sample, scaled by P
    sample = min(P, 1_000_000)
    rng = np.random.default_rng(0)
    W = {"xyz": 3, "f_dc": 3, "f_rest": 45, "opacity": 1, "scaling": 3, "rotation": 4}
    pc = {n: rng.normal(size=(sample, W[n])).astype(np.float32) for n in NAMES}
    pc["scaling"] = (pc["scaling"] * 1.2 - 3.5).astype(np.float32)
    mc = {n: np.zeros_like(pc[n]) for n in NAMES}
    den = rng.integers(1, 40, sample).astype(np.float32)
    acc = (rng.random(sample) * 6e-4 * den).astype(np.float32)
    _, sm = O.densify_masks(acc, den, pc["scaling"], 0.0002, 3.0, 0.01)
    smp = rng.normal(size=(2 * int(sm.sum()), 3)).astype(np.float32)
    t = time.time(); O.densify_and_prune(pc, mc, mc, acc, den, 0.0002, 0.005, 3.0, 0.01, 20.0, smp); cpu_s = (time.time() - t) * P / sample
    print(json.dumps({
        "metric": "densify_and_prune calls/s on a %d-Gaussian model with Adam state (6 tensors + 12 moments)" % P,
        "value": 1000.0 / ms_mine, "unit": "calls/s", "ms_per_step": ms_mine, "higher_is_better": True, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "densify-and-prune", "P": P, "P_out": P_out, "cache": "708 B per Gaussian: larger than L2 from 180 k Gaussians"},
        "steps": steps, "warmup": warmup, "gpu_launches": n_launch,
        "reference_same_gpu": {"impl": "torch mask-indexing / cat sequence (tools/c3_trainer.py)", "ms_per_step": out["torch"][0],
                               "P_out": out["torch"][1], "speedup": out["torch"][0] / ms_mine},
        "roofline": {"bound": "hbm", "achieved": alg / ms_mine / 1e6, "peak": peak, "unit": "GB/s", "frac": alg / ms_mine / 1e6 / peak,
                     "traffic": None, "peak_source": how, "algorithmic_bytes": alg,
                     "note": "the timed region is the whole call (two host round trips, torch.normal, output allocation), not the apply kernel alone"},
        "cpu_baseline": {"value": 1.0 / cpu_s, "unit": "calls/s", "cores": 1, "kind": "port",
                         "sample": "%d of %d Gaussians, scaled by count" % (sample, P)}}))

    # ---- per-iteration statistics ----
    class M:
        pass
    m = M()
    gen = torch.Generator(device="cuda").manual_seed(1)
    m.xyz_gradient_accum, m.denom, m.max_radii2D = torch.zeros(P, 1, device="cuda"), torch.zeros(P, 1, device="cuda"), torch.zeros(P, device="cuda")
    radii = (torch.randint(1, 60, (P,), device="cuda", generator=gen) * (torch.rand(P, device="cuda", generator=gen) > 0.25)).int()
    grad = torch.randn(P, 3, device="cuda", generator=gen) * 1e-3
    l0 = launches()
    ms_mine = timed(lambda: gsr_densify.add_densification_stats(m, grad, radii), steps, warmup)
    n_launch = (launches() - l0) / (steps + warmup)
    accum, denom, mr = torch.zeros(P, 1, device="cuda"), torch.zeros(P, 1, device="cuda"), torch.zeros(P, device="cuda")

    def torch_stats():
        vis = radii > 0
        mr[vis] = torch.max(mr[vis], radii[vis].float())
        accum[vis] += torch.norm(grad[vis, :2], dim=-1, keepdim=True)
        denom[vis] += 1
    ms_ref = timed(torch_stats, steps, warmup)
    alg = P * (4 + 8 + 6 * 4)                           # radii, two gradient components, three statistics read and written
    a0, d0, r0 = np.zeros(P, np.float32), np.zeros(P, np.float32), np.zeros(P, np.float32)
    gh, rh = grad.cpu().numpy(), radii.cpu().numpy()
    t = time.time(); O.densify_stats(gh, rh, a0, d0, r0); cpu_s = time.time() - t
    print(json.dumps({
        "metric": "densification statistics updates/s on a %d-Gaussian model (train.py:153-158)" % P,
        "value": 1000.0 / ms_mine, "unit": "updates/s", "ms_per_step": ms_mine, "higher_is_better": True, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "densify-stats", "P": P, "visible": int((radii > 0).sum())}, "steps": steps, "warmup": warmup, "gpu_launches": n_launch,
        "reference_same_gpu": {"impl": "torch boolean-mask statements of train.py:153-158", "ms_per_step": ms_ref, "speedup": ms_ref / ms_mine},
        "roofline": {"bound": "hbm", "achieved": alg / ms_mine / 1e6, "peak": peak, "unit": "GB/s", "frac": alg / ms_mine / 1e6 / peak,
                     "traffic": None, "peak_source": how, "algorithmic_bytes": alg},
        "cpu_baseline": {"value": 1.0 / cpu_s, "unit": "updates/s", "cores": 1, "kind": "port", "sample": "the full model"}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", nargs="?", default="all", choices=["adam", "knn", "densify", "all"])
    ap.add_argument("--P", type=int, default=0)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    a = ap.parse_args()
    assert torch.cuda.is_available(), "bench_next.py needs a GPU (there is no CPU path)"
    if a.what in ("adam", "all"):
        bench_adam(a.P or 3_000_000, a.steps, a.warmup)
    if a.what in ("knn", "all"):
        bench_knn(a.P or 1_000_000, a.steps, a.warmup)
    if a.what in ("densify", "all"):
        bench_densify(a.P or 3_000_000, min(a.steps, 10), a.warmup)


if __name__ == "__main__":
    main()
