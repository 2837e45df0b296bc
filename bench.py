#!/usr/bin/env python3
"""bench.py -- the rasterizer hot path on B200: forward + backward of one view per step.

Workload (BASELINE.json configs[1], "c2"): 3 000 000 synthetic Gaussians, SH degree 3 (M = 16), 1920x1080,
one camera per GPU; a step is GaussianRasterizer.forward + backward with a dense upstream gradient.
At N > 1 GPUs (configs[4]): one camera per rank on a ring, parameters replicated, and the step ends with the
all-reduce of the per-Gaussian gradients over NCCL; value = views/s over all ranks (weak scaling).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload c2|c1|c4-render]

--impl reference runs the UNMODIFIED reference CUDA rasterizer (oracle/_ref, built for sm_100 by
oracle/build_ref.py) through its own Python API on the same tensors -- the reference ships no CPU rasterizer,
so this is the comparison arm BASELINE.json's north_star names.  If oracle/_ref is absent it times the CPU
oracle port instead (bounded sample).  One JSON line on stdout (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "gaussian-splatting_cc-comments_b200")
for _p in (ROOT, PKG, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (P, W, H, sh_degree, M, backward?)
    "c1": (10_000, 256, 256, 0, 1, True),
    "c2": (3_000_000, 1920, 1080, 3, 16, True),
    "c4-render": (6_000_000, 3840, 2160, 3, 16, False),
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.proc = index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
            for line in self.proc.stdout:
                self.rows.append([x.strip() for x in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
        self.join(timeout=2)
        sm, mx, reasons = [], 0.0, set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def ring_camera(S, W, H, rank, n_views=8):
    if n_views == 1 or rank is None:
        return S.make_camera(W, H)
    import math
    a = 2 * math.pi * rank / n_views
    return S.look_at_camera(W, H, eye=(0.5 * math.cos(a), 0.5 * math.sin(a), 0.0), target=(0.0, 0.0, 5.0))


def load_module(impl):
    """The drop-in package (impl b200) or the unmodified reference package built into oracle/_ref."""
    import importlib.util
    if impl == "b200":
        import diff_gaussian_rasterization as m          # raises if the CUDA extension is not built
        return m
    d = os.path.join(ROOT, "oracle", "_ref", "diff_gaussian_rasterization")
    if not os.path.exists(os.path.join(d, "_C.so")):
        return None
    spec = importlib.util.spec_from_file_location("ref_diff_gaussian_rasterization", os.path.join(d, "__init__.py"),
                                                  submodule_search_locations=[d])
    m = importlib.util.module_from_spec(spec)
    sys.modules["ref_diff_gaussian_rasterization"] = m
    spec.loader.exec_module(m)
    return m


class GsrLib:
    """ctypes view of the measurement hooks of libgsr_b200.so (same library instance the extension links)."""

    def __init__(self):
        import ctypes as C
        self.C = C
        self.lib = C.CDLL(os.path.join(PKG, "libgsr_b200.so"))
        self.lib.gsr_launch_count.restype = C.c_ulonglong
        self.lib.gsr_stage_name.restype = C.c_char_p
        self.n = self.lib.gsr_num_stages()
        self.names = [self.lib.gsr_stage_name(i).decode() for i in range(self.n)]

    def launches(self):
        return int(self.lib.gsr_launch_count())

    def profiling(self, on):
        self.lib.gsr_set_profiling(int(on))

    def stage_ms(self):
        ms = (self.C.c_float * self.n)()
        cnt = (self.C.c_int * self.n)()
        self.lib.gsr_get_stage_ms(ms, cnt, self.n)
        return {self.names[i]: (float(ms[i]), int(cnt[i])) for i in range(self.n)}


class View:
    """One camera's rasterizer call, written the way gaussian_renderer.render() makes it
    (gaussian_renderer/__init__.py:44-116): settings -> GaussianRasterizer -> forward(...)."""

    def __init__(self, mod, cam, scene, device, bg=(0.0, 0.0, 0.0)):
        import torch
        self.torch, self.mod, self.cam, self.device = torch, mod, cam, device
        t = lambda a, rg=False: torch.tensor(np.asarray(a), dtype=torch.float32, device=device).requires_grad_(rg)
        P = scene["means3D"].shape[0]
        self.P = P
        self.params = dict(means3D=t(scene["means3D"], True), opacities=t(scene["opacities"], True), shs=t(scene["shs"], True),
                           scales=t(scene["scales"], True), rotations=t(scene["rotations"], True))
        self.means2D = torch.zeros(P, 3, device=device, requires_grad=True)
        self.sh_degree = scene["sh_degree"]
        self.bg = t(np.asarray(bg))
        # camera block: viewmatrix (16) | projmatrix (16) | campos (3), one small H2D per view in the e2e loop
        self.cam_host = torch.from_numpy(np.concatenate([cam["viewmatrix"].ravel(), cam["projmatrix"].ravel(),
                                                         cam["campos"].ravel()]).astype(np.float32)).pin_memory()
        self.cam_dev = self.cam_host.to(device)
        self.rast = self._rasterizer()

    def _rasterizer(self):
        c = self.cam_dev
        st = self.mod.GaussianRasterizationSettings(
            image_height=self.cam["H"], image_width=self.cam["W"], tanfovx=self.cam["tanfovx"], tanfovy=self.cam["tanfovy"],
            bg=self.bg, scale_modifier=1.0, viewmatrix=c[0:16].view(4, 4), projmatrix=c[16:32].view(4, 4),
            sh_degree=self.sh_degree, campos=c[32:35], prefiltered=False, debug=False)
        return self.mod.GaussianRasterizer(raster_settings=st)

    def zero_grad(self):
        for p in self.params.values():
            p.grad = None
        self.means2D.grad = None

    def forward(self):
        p = self.params
        return self.rast(means3D=p["means3D"], means2D=self.means2D, opacities=p["opacities"], shs=p["shs"],
                         scales=p["scales"], rotations=p["rotations"])

    def grads(self):
        return [self.params[k].grad for k in ("means3D", "opacities", "shs", "scales", "rotations")] + [self.means2D.grad]


def run_gpu(args):
    import torch
    import torch.distributed as dist
    import gsr_synthetic as S
    import gsr_multi_gpu as MG

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"           # keep stdout to the one JSON line
        os.environ.setdefault("TORCH_NCCL_HIGH_PRIORITY", "1")   # collectives overlap with compute kernels launched earlier
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)

    mod = load_module(args.impl)
    if mod is None:
        return run_cpu_reference(args, rank, world)
    P, W, H, deg, M, do_bwd = WORKLOADS[args.workload]
    if args.P:
        P = args.P                                                       # C3-style sweep over the model size
    cam0 = S.make_camera(W, H)
    scene = S.make_scene(P, cam0, sh_degree=deg, M=M, seed=0)           # identical on every rank (replicated parameters)
    # fwd+bwd: one camera per rank (multi-view training); render-only: ONE frame, its tile rows split over the ranks
    cam = ring_camera(S, W, H, rank if (world > 1 and do_bwd) else None)
    sharded = world > 1 and not do_bwd and args.impl == "b200"
    view = View(mod, cam, scene, device)
    dL_host = torch.from_numpy(S.upstream_grad(cam, seed=1 + rank)).pin_memory()
    dL = dL_host.to(device)
    lib = GsrLib() if args.impl == "b200" else None
    hbm_peak, peak_src, sm_max = peaks()

    # N > 1, this implementation: the fused exchange lives inside backward() (peer-mapped dL/dRGB blocks + one kernel
    # for the SH block, one NCCL all-reduce for the 14 small floats).  The reference arm all-reduces its tensors with NCCL.
    fused = world > 1 and args.impl == "b200" and do_bwd and not args.plain_allreduce and not args.no_exchange
    mv_exchange = None
    if fused:
        mv_exchange = MG.enable_multiview_training(P, device)

    def render_only():
        with torch.no_grad():
            if sharded:
                p = view.params
                img, radii, band = MG.render_tile_row_sharded(mod.GaussianRasterizer, view.rast.raster_settings, means3D=p["means3D"],
                                                              means2D=view.means2D, opacities=p["opacities"], shs=p["shs"],
                                                              scales=p["scales"], rotations=p["rotations"])
                return img
            return view.forward()[0]

    def step_device():
        if not do_bwd:
            return render_only()
        view.zero_grad()
        color, radii = view.forward()
        color.backward(dL)
        if world > 1 and not fused and not args.no_exchange:
            MG.allreduce_gradients(view.grads())
        return color

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, K):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(K):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=device)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # ---- warm-up ---------------------------------------------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        step_device()
    torch.cuda.synchronize()
    color, radii = view.forward()
    R = int(color.grad_fn.num_rendered)
    visible = int((radii > 0).sum().item())
    del color, radii

    # ---- timed region 1: inputs resident in HBM ("value") -----------------------------------------------------------
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.3)
    if lib:
        lib.profiling(True)
        l0 = lib.launches()
    total_ms = timed(step_device, args.steps)
    if mv_exchange is not None and os.environ.get("GSR_MV_TIMELINE") == "1" and rank == 0:
        print("exchange timeline (ms from backward start):", json.dumps(mv_exchange.timeline_report()), file=sys.stderr)
    stage = lib.stage_ms() if lib else {}
    launches = (lib.launches() - l0) if lib else None
    if lib:
        lib.profiling(False)

    # forward-only / backward-only split (same call sequence, events between the two halves)
    fwd_ms = bwd_ms = None
    if do_bwd:
        ef = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        fs, bs = [], []
        n_split = min(args.steps, 10)
        for _ in range(n_split):
            view.zero_grad()
            torch.cuda.synchronize()
            ef[0].record()
            color, _ = view.forward()
            ef[1].record()
            color.backward(dL)
            ef[2].record()
            torch.cuda.synchronize()
            fs.append(ef[0].elapsed_time(ef[1])); bs.append(ef[1].elapsed_time(ef[2]))
        fwd_ms, bwd_ms = float(np.median(fs)), float(np.median(bs))
        if rank == 0:
            print("split fwd ms:", ["%.3f" % x for x in fs], "bwd ms:", ["%.3f" % x for x in bs], file=sys.stderr)

    # ---- timed region 2: end to end through the public API with host buffers ("e2e") --------------------------------
    # Per step the host supplies what changes per view in train.py -- the camera (35 floats) and the per-pixel
    # upstream weights (the stand-in for the ground-truth image, 3*H*W floats) -- from pinned memory, and reads the
    # scalar loss back.  The Gaussians are the model's parameters and stay resident, as in the reference (nn.Parameters
    # on cuda, scene/gaussian_model.py:183-204).
    loss_host = torch.zeros(1).pin_memory()
    img_host = torch.zeros(3, H, W).pin_memory() if not do_bwd else None
    copy_stream = torch.cuda.Stream(device=device)
    copied = torch.cuda.Event()

    def step_e2e():
        view.cam_dev.copy_(view.cam_host, non_blocking=True)
        if not do_bwd:                       # render-only: the result a viewer needs is the frame itself
            img = render_only()
            if rank == 0:                    # the viewer sits on one host: one device-to-host copy of the frame
                img_host.copy_(img, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            return
        # the upstream image (the stand-in for the next ground-truth image) rides a copy stream while the forward runs,
        # as a training loop prefetches its next target; the previous step's readers of dL finished at its final sync
        with torch.cuda.stream(copy_stream):
            dL.copy_(dL_host, non_blocking=True)
            copied.record(copy_stream)
        view.zero_grad()
        color, radii = view.forward()
        torch.cuda.current_stream().wait_event(copied)
        if do_bwd:
            loss = (color * dL).sum()
            loss.backward()
            if world > 1 and not fused and not args.no_exchange:
                MG.allreduce_gradients(view.grads())
            loss_host.copy_(loss.detach().view(1), non_blocking=True)
        else:
            loss_host.copy_(color.sum().view(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for _ in range(2):
        step_e2e()
    e2e_ms = timed(step_e2e, args.steps)
    clocks = sampler.stop() if sampler else None

    per_step = total_ms / args.steps
    nviews = world if do_bwd else 1          # render-only: one frame per step however many GPUs share it (strong scaling)
    out = {
        "metric": "views/s, rasterizer fwd+bwd" if do_bwd else "frames/s, rasterizer forward (render only)",
        "value": nviews * 1000.0 / per_step, "unit": "views/s" if do_bwd else "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": per_step, "higher_is_better": True, "scaling": "weak" if do_bwd else "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "impl": args.impl,
        "config": {"workload": "%s: %d Gaussians, SH degree %d (M=%d), %dx%d, %s%s" % (
            args.workload, P, deg, M, W, H, "fwd+bwd, 1 view per GPU per step" if do_bwd else
            ("forward only, one frame per step" + (", tile rows sharded over the GPUs + all-gather of the bands" if sharded else
                                                  (", every rank renders the full frame (no sharding in this implementation)" if world > 1 else ""))),
            "" if (world == 1 or not do_bwd) else (", fused SH-gradient exchange over peer memory + NCCL all-reduce of the 14 small floats" if fused
                                   else ", NCCL all-reduce of per-Gaussian gradients")),
            "P": P, "width": W, "height": H, "sh_degree": deg, "num_rendered": R, "visible": visible,
            "cache": "inputs larger than L2 (%.0f MB of Gaussian parameters per step vs 126 MB L2)" % (P * (3 + 3 + 4 + 1 + 3 * M) * 4 / 1e6)},
        "fwd_ms": fwd_ms, "bwd_ms": bwd_ms,
        "e2e": {"value": nviews * 1000.0 / (e2e_ms / args.steps), "unit": "views/s" if do_bwd else "frames/s",
                "ms_per_step": e2e_ms / args.steps,
                "h2d_bytes_per_step": int(view.cam_host.numel() * 4 + (dL_host.numel() * 4 if do_bwd else 0)),
                "d2h_bytes_per_step": 4 if do_bwd else int(3 * H * W * 4),
                "what": ("GaussianRasterizer.forward + loss + backward through the drop-in Python API; camera and upstream "
                         "image copied from pinned host memory every step (the image on a copy stream beside the forward, as a "
                         "training loop prefetches its next target), scalar loss read back; parameters resident") if do_bwd else
                        "camera copied from pinned host memory, frame rendered through the drop-in Python API and copied back to "
                        "pinned host memory every step; parameters resident"},
        "gpu_launches": launches,
        "clocks": clocks,
    }
    if lib and not sharded:
        out.update(roofline_block(stage, P, R, W * H, visible, (deg + 1) ** 2, hbm_peak, peak_src, args.steps))
    if rank == 0 and world == 1 and args.impl == "b200" and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(R_full=R, P_full=P)
    if args.impl == "reference":
        out["reference_kind"] = "reference CUDA rasterizer, unmodified sources built for sm_100 (oracle/_ref), same GPU"
        out["cpu_baseline"] = {"value": out["value"], "unit": out["unit"], "cores": 0, "kind": "reference",
                               "sample": "not a CPU run: the reference ships no CPU rasterizer (SURVEY.md 8(d)); this arm is its unmodified "
                                         "CUDA rasterizer rebuilt for sm_100 on the same GPU, full workload, same harness"}
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(out), flush=True)
    return 0


def roofline_block(stage, P, R, N, P_vis, K, hbm_peak, peak_src, steps):
    """Per-stage algorithmic bytes (restated in DESIGN.md) over the CUDA-event mean of each stage."""
    tiles_bits_passes = 2
    alg = {
        "preprocess": P * (44 + 8) + P_vis * (12 * K + 40),
        "depth_sort": 4 * (8 + 16) * P,                    # 4 LSD passes: histogram read 4 B + scatter read/write 8+8 B per Gaussian... (keys 4 B re-read)
        "scan": 8 * P + 8 * P,                             # gather tiles_touched in depth order + exclusive scan
        "duplicate": 20 * P + 8 * R,                       # 4 B tile key + 4 B id per instance
        "tile_sort": tiles_bits_passes * (4 + 16) * R,     # 2 passes on 32-bit tile keys: 4 B histogram read + 16 B scatter
        "tile_ranges": 4 * R,
        "render": 40 * R + 20 * N,
        "bwd_zero": 48 * P,
        "render_bwd": 40 * R + 20 * N + 44 * R,
        "preprocess_bwd": P_vis * (24 * K + 132) + (P - P_vis) * (12 * K + 100),   # culled Gaussians still get zeros written
    }
    stages = {}
    for name, (ms, n) in stage.items():
        if ms is None or ms < 0:
            continue
        b = alg.get(name)
        stages[name] = {"ms": round(ms, 4), "samples": n, "alg_bytes": int(b) if b else None,
                        "gbs": round(b / ms / 1e6, 1) if b else None, "frac_hbm": round(b / ms / 1e6 / hbm_peak, 4) if b else None}
    single = [k for k in ("preprocess", "duplicate", "tile_ranges", "render", "render_bwd", "preprocess_bwd") if k in stages]
    dom = max(single, key=lambda k: stages[k]["ms"])
    a = stages[dom]["gbs"]
    pairs_upper = 256.0 * R
    rl = {"kernel": {"render": "render_fwd_kernel", "render_bwd": "render_bwd_kernel", "preprocess": "preprocess_fwd_kernel",
                     "preprocess_bwd": "preprocess_bwd_kernel_staged", "duplicate": "duplicate_coop_kernel", "tile_ranges": "tile_ranges4_kernel"}[dom],
          "bound": "hbm", "achieved": a, "peak": hbm_peak, "unit": "GB/s", "frac": round(a / hbm_peak, 4), "traffic": None,
          "peak_source": peak_src, "launch_ms": stages[dom]["ms"], "alg_bytes_per_launch": stages[dom]["alg_bytes"],
          "note": "the blending kernels are issue-slot/MUFU/shared-memory bound, not HBM bound (SURVEY.md 8(d)); the HBM "
                  "fraction is reported as the contract asks, the pair rate against the FP32-issue bound is in pair_rate"}
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")        # dram bytes per launch from the committed ncu --set full capture
    if os.path.exists(tpath):
        t = json.load(open(tpath))
        if t.get("P") == P and t.get("num_rendered") == R and rl["kernel"] in t.get("kernels", {}):
            rl["traffic"] = t["kernels"][rl["kernel"]]
            rl["traffic_source"] = t.get("source")
            wi = t.get("warp_instructions", {}).get(rl["kernel"])
            if wi:      # the bound that actually holds for the blending kernels: warp-instruction issue slots (4 schedulers per SM, 1 per clock)
                peak_gi = 148 * 4 * 1.965            # G warp-instructions/s at the 1965 MHz the run holds (see clocks)
                ach = wi / stages[dom]["ms"] / 1e6
                rl["issue_slots"] = {"warp_instructions_per_launch": wi, "achieved": round(ach, 1), "peak": round(peak_gi, 1),
                                     "unit": "G warp-instr/s", "frac": round(ach / peak_gi, 4),
                                     "source": "ncu smsp__inst_executed.sum of the committed capture / live launch time"}
    out = {"roofline": rl, "stages": stages}
    if dom in ("render", "render_bwd"):
        issue_peak = 148 * 128 * 1.965e9          # lane-instructions/s
        slots = 20.0 if dom == "render" else 45.0
        out["pair_rate"] = {"pairs_upper_bound_per_launch": pairs_upper, "gpairs_per_s": round(pairs_upper / stages[dom]["ms"] / 1e6, 1),
                            "issue_bound_ms_at_%d_slots_per_pair" % int(slots): round(pairs_upper * slots / issue_peak * 1e3, 3)}
    return out


def cpu_baseline(R_full, P_full, budget_s=20.0):
    """The CPU oracle port (oracle/gs_oracle.c, OpenMP over the host cores) on a bounded sample of the workload:
    the same generator at P = 100 000 / 960x540, forward + backward, scaled to the full workload by instance count."""
    import gsr_synthetic as S
    from oracle import gs_oracle as O          # the one place bench.py touches oracle/: the reported CPU baseline
    O.build()
    Ps, Ws, Hs = 100_000, 960, 540
    cam = S.make_camera(Ws, Hs)
    scene = S.make_scene(Ps, cam, sh_degree=3, M=16, seed=0)
    dL = S.upstream_grad(cam)
    kw = dict(shs=scene["shs"], scales=scene["scales"], rotations=scene["rotations"], sh_degree=3)
    t0 = time.time()
    n = 0
    while True:
        f = O.forward(cam, scene["means3D"], scene["opacities"], **kw)
        O.backward(cam, f, dL, scene["means3D"], **kw)
        n += 1
        if time.time() - t0 > budget_s or n >= 20:
            break
    dt = (time.time() - t0) / n
    Rs = int(f["num_rendered"])
    scale = max(Rs / max(R_full, 1), 1e-12)
    return {"value": (1.0 / dt) * scale, "unit": "views/s", "cores": os.cpu_count(), "kind": "port",
            "sample": "oracle/gs_oracle.c fwd+bwd, %d Gaussians SH3 at %dx%d (R=%d), %d runs of %.2f s; scaled to the full "
                      "workload by instance count R (x%.4f)" % (Ps, Ws, Hs, Rs, n, dt, scale),
            "sample_seconds_per_view": dt}


def run_cpu_reference(args, rank, world):
    """Fallback reference arm when oracle/_ref was not built: the CPU oracle port, rank 0 only."""
    if rank != 0:
        return 0
    P, W, H, deg, M, do_bwd = WORKLOADS[args.workload]
    cb = cpu_baseline(R_full=4 * P, P_full=P, budget_s=30.0)
    out = {"metric": "views/s, rasterizer fwd+bwd", "value": cb["value"], "unit": "views/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": 1000.0 / cb["value"], "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
           "config": {"workload": "%s (CPU oracle port; oracle/_ref not built)" % args.workload},
           "cpu_baseline": cb, "e2e": {"value": cb["value"], "unit": "views/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--P", type=int, default=0, help="override the number of Gaussians (model-size sweep, BASELINE configs[2])")
    ap.add_argument("--no-exchange", action="store_true", help="diagnostic: N > 1 without any gradient exchange (rank skew only)")
    ap.add_argument("--plain-allreduce", action="store_true", help="N > 1: one NCCL all-reduce of the whole gradient arena "
                    "instead of the fused SH-gradient exchange")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # convenience: relaunch under torchrun when called plainly with --gpus N
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
