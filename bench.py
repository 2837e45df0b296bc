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


class Model:
    """The Gaussians (replicated parameters), as gaussian_renderer.render() hands them to the rasterizer."""

    def __init__(self, scene, device):
        import torch
        t = lambda a, rg=False: torch.tensor(np.asarray(a), dtype=torch.float32, device=device).requires_grad_(rg)
        self.P = scene["means3D"].shape[0]
        self.sh_degree = scene["sh_degree"]
        self.params = dict(means3D=t(scene["means3D"], True), opacities=t(scene["opacities"], True), shs=t(scene["shs"], True),
                           scales=t(scene["scales"], True), rotations=t(scene["rotations"], True))

    def zero_grad(self):
        for p in self.params.values():
            p.grad = None

    def small(self):
        """The parameters whose .grad the multi-view exchange sums with one all-reduce (everything but the SH tensor)."""
        return [self.params[k] for k in ("means3D", "opacities", "scales", "rotations")]

    def grads(self):
        return [self.params[k].grad for k in ("means3D", "opacities", "shs", "scales", "rotations")]


class View:
    """One camera's rasterizer call, written the way gaussian_renderer.render() makes it
    (gaussian_renderer/__init__.py:44-116): settings -> GaussianRasterizer -> forward(...)."""

    def __init__(self, mod, cam, model, device, bg=(0.0, 0.0, 0.0)):
        import torch
        self.torch, self.mod, self.cam, self.device, self.model = torch, mod, cam, device, model
        self.P = model.P
        self.params = model.params
        self.means2D = torch.zeros(self.P, 3, device=device, requires_grad=True)     # screenspace_points of render()
        self.sh_degree = model.sh_degree
        self.bg = torch.tensor(np.asarray(bg), dtype=torch.float32, device=device)
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
        self.model.zero_grad()
        self.means2D.grad = None

    def forward(self):
        p = self.params
        return self.rast(means3D=p["means3D"], means2D=self.means2D, opacities=p["opacities"], shs=p["shs"],
                         scales=p["scales"], rotations=p["rotations"])

    def grads(self):
        return self.model.grads() + [self.means2D.grad]


def exchange_check(mod, device, rank, world, views_per_rank, P=40001, W=320, H=200):
    """Correctness of the multi-view exchange where the driver runs it: on a reduced scene every rank computes ALL
    world * views_per_rank views locally without any exchange and sums their gradients, then runs its own views through the
    fused exchange; the two must agree (relative L2 per tensor, bar 1e-5).  P is odd on purpose (block alignment).
    means2D.grad is compared per view (it is not an exchanged quantity).  Collective; returns a dict for the JSON line."""
    import torch
    import gsr_synthetic as S
    import gsr_multi_gpu as MG
    n_views = world * views_per_rank
    scene = S.make_scene(P, S.make_camera(W, H), sh_degree=3, M=16, seed=4)
    model = Model(scene, device)
    views = [View(mod, ring_camera(S, W, H, v, n_views=max(n_views, 2)), model, device, bg=(0.1, 0.2, 0.3)) for v in range(n_views)]
    dLs = [torch.from_numpy(S.upstream_grad(v.cam, seed=20 + i)).to(device) for i, v in enumerate(views)]
    names = ("means3D", "opacities", "shs", "scales", "rotations")
    MG.disable_multiview_training()
    model.zero_grad()
    own2d = {}
    for i, v in enumerate(views):
        v.means2D.grad = None
        color, _ = v.forward()
        color.backward(dLs[i])
        own2d[i] = v.means2D.grad.clone()
    expect = {k: model.params[k].grad.clone() for k in names}
    mine = MG.cameras_for_rank(n_views, rank, world)
    ex = MG.enable_multiview_training(P, device, views_per_rank=views_per_rank)
    worst = 0.0
    errs = {}
    for rep in range(2):                              # twice: the second step reuses the blocks and events of the first
        model.zero_grad()
        for i in mine:
            views[i].means2D.grad = None
            color, _ = views[i].forward()
            color.backward(dLs[i])
        ex.end_step(model.small())
        torch.cuda.synchronize()
        for k in names:
            e = float((model.params[k].grad - expect[k]).norm() / expect[k].norm().clamp_min(1e-30))
            errs[k] = max(errs.get(k, 0.0), e)
        for i in mine:
            e = float((views[i].means2D.grad - own2d[i]).norm() / own2d[i].norm().clamp_min(1e-30))
            errs["means2D_per_view"] = max(errs.get("means2D_per_view", 0.0), e)
    transport = ex.mode
    MG.disable_multiview_training()
    worst = max(errs.values())
    t = torch.tensor([worst], device=device)
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    worst = float(t.item())
    return {"max_rel_l2": worst, "ok": bool(worst <= 1e-5), "bar": 1e-5, "views": n_views, "views_per_rank": views_per_rank,
            "P": P, "image": "%dx%d" % (W, H), "transport": transport, "per_tensor_rank0": {k: float("%.3g" % v) for k, v in errs.items()},
            "what": "fused exchange vs the sum of all views computed locally without exchange, two steps, max over ranks"}


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
    model = Model(scene, device)
    # fwd+bwd at N > 1 (BASELINE configs[4]): 8 cameras per step, camera c on rank c mod N -> 8/N views per GPU per step
    # (N = 2, 4, 8; any other N: one camera per rank).  Render-only: ONE frame, its tile rows split over the ranks.
    multiview = world > 1 and do_bwd
    n_views = (8 if 8 % world == 0 and world <= 8 else world) if multiview else 1
    if args.views_per_rank:
        n_views = world * args.views_per_rank
    my_cams = MG.cameras_for_rank(n_views, rank, world) if multiview else [None]
    vpr = len(my_cams)
    sharded = world > 1 and not do_bwd and args.impl == "b200"
    views = [View(mod, ring_camera(S, W, H, c, n_views=max(n_views, 8)), model, device) for c in my_cams]
    view = views[0]
    dLs = [torch.from_numpy(S.upstream_grad(v.cam, seed=1 + (c or 0))).to(device) for v, c in zip(views, my_cams)]
    lib = GsrLib() if args.impl == "b200" else None
    hbm_peak, peak_src, sm_max = peaks()

    # N > 1, this implementation: the fused exchange lives inside backward() of each rank's last view of the step (all-gather of
    # the 12 B/Gaussian dL/dRGB blocks + one kernel for the SH block) and in end_step() (one NCCL all-reduce of the 11 small
    # floats).  The reference arm all-reduces its gradient tensors with NCCL after its local views.
    fused = multiview and args.impl == "b200" and not args.plain_allreduce and not args.no_exchange
    check = None
    if fused and not args.no_exchange_check:
        check = exchange_check(mod, device, rank, world, vpr)
        if not check["ok"]:
            if rank == 0:
                print(json.dumps({"exchange_check": check, "error": "multi-view exchange differs from the sum of the views"}), flush=True)
            dist.barrier()
            dist.destroy_process_group()
            return 3
    mv_exchange = MG.enable_multiview_training(P, device, views_per_rank=vpr) if fused else None

    def render_only():
        with torch.no_grad():
            if sharded:
                p = model.params
                img, radii, band = MG.render_tile_row_sharded(mod.GaussianRasterizer, view.rast.raster_settings, means3D=p["means3D"],
                                                              means2D=view.means2D, opacities=p["opacities"], shs=p["shs"],
                                                              scales=p["scales"], rotations=p["rotations"])
                return img
            return view.forward()[0]

    def reduce_after_views():
        if mv_exchange is not None:
            mv_exchange.end_step(model.small())
        elif multiview and not args.no_exchange:
            MG.allreduce_gradients(model.grads())

    def step_device():
        if not do_bwd:
            return render_only()
        model.zero_grad()
        for v, dL in zip(views, dLs):
            v.means2D.grad = None
            color, radii = v.forward()
            color.backward(dL)
        reduce_after_views()
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
    with torch.enable_grad():
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

    # forward-only / backward-only split (same call sequence, events between the two halves; first local view)
    fwd_ms = bwd_ms = None
    render_only_ms = None
    if do_bwd:
        ef = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        fs, bs = [], []
        n_split = min(args.steps, 10)
        if mv_exchange is not None:
            MG.disable_multiview_training()
        for _ in range(n_split):
            view.zero_grad()
            torch.cuda.synchronize()
            ef[0].record()
            color, _ = view.forward()
            ef[1].record()
            color.backward(dLs[0])
            ef[2].record()
            torch.cuda.synchronize()
            fs.append(ef[0].elapsed_time(ef[1])); bs.append(ef[1].elapsed_time(ef[2]))
        fwd_ms, bwd_ms = float(np.median(fs)), float(np.median(bs))
        # render-only frame rate (BASELINE metric "render FPS @1080p"): the forward under no_grad, same scene and camera
        ro = []
        for _ in range(n_split):
            torch.cuda.synchronize()
            ef[0].record()
            with torch.no_grad():
                view.forward()
            ef[1].record()
            torch.cuda.synchronize()
            ro.append(ef[0].elapsed_time(ef[1]))
        render_only_ms = float(np.median(ro))
        if mv_exchange is not None:
            import diff_gaussian_rasterization as dgr
            dgr.set_multiview_exchange(mv_exchange)
        if rank == 0:
            print("split fwd ms:", ["%.3f" % x for x in fs], "bwd ms:", ["%.3f" % x for x in bs], file=sys.stderr)

    # ---- timed region 2: end to end through the public API with host buffers ("e2e") --------------------------------
    # Per step and view the host supplies what changes per view in train.py -- the camera (35 floats) and the ground-truth
    # image as the data loader holds it (uint8, 3*H*W bytes; it stands in for the per-pixel upstream weights: the device
    # turns it into float weights w = u8/255 - 0.5 and the loss is sum(render * w)) -- from pinned memory, and reads the
    # scalar loss back.  The Gaussians are the model's parameters and stay resident, as in the reference (nn.Parameters on
    # cuda, scene/gaussian_model.py:183-204).
    loss_host = torch.zeros(1).pin_memory()
    img_host = torch.zeros(3, H, W).pin_memory() if not do_bwd else None
    gen = torch.Generator().manual_seed(100 + rank)
    gt_host = [torch.randint(0, 256, (3, H, W), dtype=torch.uint8, generator=gen).pin_memory() for _ in views] if do_bwd else []
    gt_dev = [torch.empty((3, H, W), dtype=torch.uint8, device=device) for _ in views] if do_bwd else []
    copy_stream = torch.cuda.Stream(device=device)
    copied = [torch.cuda.Event() for _ in views]
    h2d_ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in views]
    h2d_ms = []

    def step_e2e():
        if not do_bwd:                       # render-only: the result a viewer needs is the frame itself
            view.cam_dev.copy_(view.cam_host, non_blocking=True)
            img = render_only()
            if rank == 0:                    # the viewer sits on one host: one device-to-host copy of the frame
                img_host.copy_(img, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            return
        # the ground-truth images ride a copy stream while the forward runs, as a training loop prefetches its next
        # targets; the previous step's readers finished at its final sync
        with torch.cuda.stream(copy_stream):
            for i in range(len(views)):
                h2d_ev[i][0].record(copy_stream)
                gt_dev[i].copy_(gt_host[i], non_blocking=True)
                h2d_ev[i][1].record(copy_stream)
                copied[i].record(copy_stream)
        model.zero_grad()
        total = None
        for i, v in enumerate(views):
            v.cam_dev.copy_(v.cam_host, non_blocking=True)
            v.means2D.grad = None
            color, radii = v.forward()
            torch.cuda.current_stream().wait_event(copied[i])
            wgt = gt_dev[i].to(torch.float32).mul_(1.0 / 255.0).sub_(0.5)
            loss = (color * wgt).sum()
            loss.backward()
            total = loss.detach() if total is None else total + loss.detach()
        reduce_after_views()
        loss_host.copy_(total.view(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()
        h2d_ms.append(sum(a.elapsed_time(b) for a, b in h2d_ev))

    for _ in range(2):
        step_e2e()
    h2d_ms.clear()
    e2e_ms = timed(step_e2e, args.steps)
    clocks = sampler.stop() if sampler else None
    h2d_stat = None
    if do_bwd and h2d_ms:
        t = torch.tensor([float(np.mean(h2d_ms))], device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        h2d_stat = float(t.item())

    per_step = total_ms / args.steps
    nviews = world * vpr if do_bwd else 1    # render-only: one frame per step however many GPUs share it (strong scaling)
    value = nviews * 1000.0 / per_step
    how = ("fwd+bwd, %d view%s per GPU per step" % (vpr, "" if vpr == 1 else "s")) if do_bwd else (
        "forward only, one frame per step" + (", tile rows sharded over the GPUs + all-gather of the bands" if sharded else
                                              (", every rank renders the full frame (no sharding in this implementation)" if world > 1 else "")))
    if multiview:
        how += " (%d cameras per step)" % n_views
        how += (", fused SH-gradient exchange (12 B/Gaussian all-gather + one kernel) + one NCCL all-reduce of the 11 small floats"
                if fused else ", NCCL all-reduce of per-Gaussian gradients")
    out = {
        "metric": "views/s, rasterizer fwd+bwd" if do_bwd else "frames/s, rasterizer forward (render only)",
        "value": value, "unit": "views/s" if do_bwd else "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": per_step, "higher_is_better": True,
        # N > 1 fwd+bwd: the step's total work is fixed (8 cameras) and split over the GPUs; N = 1: one view per step
        "scaling": ("strong" if (multiview and n_views == 8) else "weak") if do_bwd else "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "impl": args.impl,
        "config": {"workload": "%s: %d Gaussians, SH degree %d (M=%d), %dx%d, %s" % (args.workload, P, deg, M, W, H, how),
                   "P": P, "width": W, "height": H, "sh_degree": deg, "num_rendered": R, "visible": visible,
                   "views_per_step": nviews, "views_per_gpu": vpr if do_bwd else None,
                   "cache": "inputs larger than L2 (%.0f MB of Gaussian parameters per step vs 126 MB L2)" % (P * (3 + 3 + 4 + 1 + 3 * M) * 4 / 1e6)},
        "fwd_ms": fwd_ms, "bwd_ms": bwd_ms,
        "ms_per_view": per_step / vpr if do_bwd else None,
        "gaussians_per_s": P * value, "instances_per_s": R * value,
        "render_only_fps": (1000.0 / render_only_ms) if render_only_ms else None, "render_only_ms": render_only_ms,
        "e2e": {"value": nviews * 1000.0 / (e2e_ms / args.steps), "unit": "views/s" if do_bwd else "frames/s",
                "ms_per_step": e2e_ms / args.steps,
                "h2d_bytes_per_step": int(sum(v.cam_host.numel() * 4 for v in views) + sum(g.numel() for g in gt_host)) if do_bwd
                else int(view.cam_host.numel() * 4),
                "d2h_bytes_per_step": 4 if do_bwd else int(3 * H * W * 4),
                "h2d_image_copy_ms_per_step_max_over_ranks": h2d_stat,
                "what": ("GaussianRasterizer.forward + loss + backward through the drop-in Python API for every local view; per view "
                         "the camera and the uint8 ground-truth image are copied from pinned host memory every step (the images on a "
                         "copy stream beside the forward, as a training loop prefetches its next targets; the device turns them into "
                         "float per-pixel weights), the scalar loss is read back; parameters resident") if do_bwd else
                        "camera copied from pinned host memory, frame rendered through the drop-in Python API and copied back to "
                        "pinned host memory every step; parameters resident"},
        "gpu_launches": launches,
        "clocks": clocks,
    }
    if check is not None:
        out["exchange_check"] = check
    if lib and not sharded:
        out.update(roofline_block(stage, P, R, W * H, visible, (deg + 1) ** 2, hbm_peak, peak_src, args.steps))
    if rank == 0 and world == 1 and args.impl == "b200" and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(R_full=R, P_full=P)
        out["cpu_torch_baseline"] = cpu_torch_baseline()
    if args.impl == "reference":
        out["reference_kind"] = "reference CUDA rasterizer, unmodified sources built for sm_100 (oracle/_ref), same GPU"
        out["cpu_baseline"] = {"value": out["value"], "unit": out["unit"], "cores": 0, "kind": "reference",
                               "sample": "not a CPU run: the reference ships no CPU rasterizer (SURVEY.md 8(d)); this arm is its unmodified "
                                         "CUDA rasterizer rebuilt for sm_100 on the same GPU, full workload, same harness"}
    # ---- sub-records for the other BASELINE configs (bounded; skipped with --no-extras) -------------------------------
    if not args.no_extras and args.workload == "c2":
        del views, view, dLs, gt_dev, gt_host, model, scene
        if mv_exchange is not None:
            MG.disable_multiview_training()
            mv_exchange = None
        torch.cuda.empty_cache()
        try:
            out["c4_render"] = c4_render_record(mod, args.impl, device, rank, world)
        except Exception as e:          # a sub-record must not take the headline down
            out["c4_render"] = {"error": repr(e)[:300]}
        if world == 1:
            try:
                out["c3"] = c3_record(args.impl)
            except Exception as e:
                out["c3"] = {"error": repr(e)[:300]}
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(out), flush=True)
    return 0


def c4_render_record(mod, impl, device, rank, world, P=6_000_000, W=3840, H=2160, steps=10, warmup=3):
    """BASELINE configs[3]: render-only 3840x2160, 6 M Gaussians SH3.  One frame per step; at N > 1 (this implementation) the
    frame's tile rows are split over the ranks with the Gaussians replicated and the bands are all-gathered.  Also times the
    unsharded frame on this GPU in the same process, so that speedup_vs_n1 needs no second run."""
    import torch
    import torch.distributed as dist
    import gsr_synthetic as S
    import gsr_multi_gpu as MG
    cam = S.make_camera(W, H)
    model = Model(S.make_scene(P, cam, sh_degree=3, M=16, seed=0), device)
    view = View(mod, cam, model, device)
    p = model.params
    kw = dict(means3D=p["means3D"], means2D=view.means2D, opacities=p["opacities"], shs=p["shs"], scales=p["scales"], rotations=p["rotations"])

    def full():
        with torch.no_grad():
            return view.forward()[0]

    def sharded():
        return MG.render_tile_row_sharded(mod.GaussianRasterizer, view.rast.raster_settings, **kw)[0]

    def timed(fn):
        for _ in range(warmup):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / steps], device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n1_ms = timed(full)
    with torch.enable_grad():
        R = int(view.forward()[0].grad_fn.num_rendered)
    rec = {"workload": "c4-render: %d Gaussians SH3, %dx%d, forward only" % (P, W, H), "num_rendered": R, "n_gpus": world,
           "n1_ms_per_frame": n1_ms, "steps": steps, "warmup": warmup}
    if world > 1 and impl == "b200":
        ms = timed(sharded)
        same = bool(torch.equal(sharded(), full()))
        rec.update({"ms_per_frame": ms, "frames_per_s": 1000.0 / ms, "speedup_vs_n1": n1_ms / ms,
                    "bands_bit_identical_to_full_frame": same,
                    "how": "tile rows split over the ranks (Gaussians replicated), bands all-gathered; n1 = the unsharded frame on one GPU in the same process"})
    else:
        rec.update({"ms_per_frame": n1_ms, "frames_per_s": 1000.0 / n1_ms, "speedup_vs_n1": 1.0})
    return rec


def c3_record(impl):
    """BASELINE configs[2]: the training loop on a synthetic scene densifying 1 M -> 6 M Gaussians, one view per iteration at
    1080p, one GPU (tools/train_c3.py; 660 iterations, densification every 50 from iteration 60).  The b200 line also runs the
    reference arm (unmodified reference rasterizer + torch loss / Adam / densification) when oracle/_ref is there."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import gc
    import train_c3
    import torch
    gc.collect()
    torch.cuda.empty_cache()          # the loop re-allocates at every densification: start from an empty caching allocator, as a
    torch.cuda.synchronize()          # training process would (measured: 5.6 instead of 4.3 ms/iter on top of the C2 / C4 leftovers)

    def arm(name, repeats=2):
        # The loop launches hundreds of small kernels per iteration; on a shared host a descheduled launch thread stretches whole
        # phases (seen: 4.1 - 7.1 ms per iteration for the same build on different boxes).  Both arms alike: the loop is run
        # `repeats` times and the fastest run is reported, every run's mean is listed.
        runs = []
        for _ in range(repeats):
            gc.collect()
            torch.cuda.empty_cache()
            runs.append(train_c3.run(train_c3.parse_args(["--impl", name])))
        torch.cuda.empty_cache()
        r = min(runs, key=lambda x: x["value"])
        return {"ms_per_iter_mean": r["value"], "ms_per_iter_runs": [round(x["value"], 4) for x in runs], "final_P": r["final_P"],
                "iterations": r["config"]["iterations"],
                "ms_per_iter_by_P": [[b["P_from"], round(b["iter_ms"], 3)] for b in r["buckets"]],
                "phases_ms_by_P": [[b["P_from"], round(b["fwd_ms"], 3), round(b["bwd_ms"], 3), round(b["densify_stats_ms"], 3),
                                    round(b["optimizer_ms"], 3)] for b in r["buckets"]]}
    rec = {"workload": "c3-train-loop: 1 M -> 6 M Gaussians SH3, 1920x1080, 1 view/iter; render + L1/SSIM loss + backward + densification "
                       "statistics / densify_and_prune + Adam", "phases": ["fwd", "bwd", "densify+stats", "optimizer"]}
    if impl == "b200":
        rec["b200"] = arm("b200")
        if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "diff_gaussian_rasterization", "_C.so")):
            rec["reference"] = arm("reference")
            rec["vs_reference"] = rec["reference"]["ms_per_iter_mean"] / rec["b200"]["ms_per_iter_mean"]
    else:
        rec["reference"] = arm("reference")
    return rec


def cpu_torch_baseline(budget_s=12.0):
    """The reference's CPU-side torch path that north_star asks to time on the box's host cores: SH colour evaluation
    (utils/sh_utils.py:57-112 eval_sh, restated in gsr_sh.eval_sh; gaussian_renderer/__init__.py:84-93 convert_SHs_python) and
    the covariance build (utils/general_utils.py:78-150 build_rotation / build_scaling_rotation / strip_symmetric via
    scene/gaussian_model.py:31-36), in torch on CPU tensors.  /root/reference does not exist on the GPU box, so these are the
    package's restatements of those functions (tests/test_oracle.py pins them to the reference's outputs)."""
    import torch
    import gsr_sh
    out = {"cores": os.cpu_count(), "torch_threads": torch.get_num_threads(), "kind": "port",
           "what": "eval_sh (degree 3) + covariance build in torch on the host cores, the work preprocess_fwd_kernel does per Gaussian"}
    for P in (10_000, 3_000_000):
        g = torch.Generator().manual_seed(0)
        xyz = torch.randn(P, 3, generator=g)
        sh = torch.randn(P, 16, 3, generator=g) * 0.2
        scaling = torch.rand(P, 3, generator=g) * 0.05 + 0.005
        rot = torch.nn.functional.normalize(torch.randn(P, 4, generator=g))
        campos = torch.tensor([0.0, 0.0, -5.0])

        def colours():
            d = xyz - campos
            d = d / d.norm(dim=1, keepdim=True)
            return torch.clamp_min(gsr_sh.eval_sh(3, sh.transpose(1, 2), d) + 0.5, 0.0)

        def covariance():
            r, x, y, z = rot[:, 0], rot[:, 1], rot[:, 2], rot[:, 3]
            Rm = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                              2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                              2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=1).view(P, 3, 3)
            L = Rm * scaling.unsqueeze(1)                      # R @ diag(s)
            cov = L @ L.transpose(1, 2)
            return torch.stack([cov[:, 0, 0], cov[:, 0, 1], cov[:, 0, 2], cov[:, 1, 1], cov[:, 1, 2], cov[:, 2, 2]], dim=1)

        rec = {}
        for name, fn in (("eval_sh_ms", colours), ("covariance_ms", covariance)):
            fn()
            t0, n = time.time(), 0
            while True:
                fn()
                n += 1
                if time.time() - t0 > budget_s / 4 or n >= 20:
                    break
            rec[name] = (time.time() - t0) / n * 1e3
        rec["gaussians_per_s"] = P / ((rec["eval_sh_ms"] + rec["covariance_ms"]) * 1e-3)
        out["P_%d" % P] = rec
    return out


def roofline_block(stage, P, R, N, P_vis, K, hbm_peak, peak_src, steps):
    """Per-stage algorithmic bytes (restated in DESIGN.md) over the CUDA-event mean of each stage."""
    tiles_bits_passes = 2
    # <= 8192 tiles: the one-pass counting tile sort (no separate ranges stage): keys read by the count pass, keys + ids
    # read and ids written by the scatter pass; else two radix passes + a pass over the sorted keys for the ranges
    counting = not (stage.get("tile_ranges") and stage["tile_ranges"][0] is not None and stage["tile_ranges"][0] >= 0)
    alg = {
        "preprocess": P * (44 + 8) + P_vis * (12 * K + 40),
        "depth_sort": 4 * (8 + 16) * P,                    # 4 LSD passes: histogram read 4 B + scatter read/write 8+8 B per Gaussian... (keys 4 B re-read)
        "scan": 8 * P + 8 * P,                             # gather tiles_touched in depth order + exclusive scan
        "duplicate": 20 * P + 8 * R,                       # 4 B tile key + 4 B id per instance
        "tile_sort": (4 + 12) * R if counting else tiles_bits_passes * (4 + 16) * R,   # radix: per pass 4 B histogram read + 16 B scatter
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
    the same generator at P = 1 000 000 at the workload's own 1920x1080 (a third of its instances), forward + backward,
    scaled to the full workload by instance count."""
    import gsr_synthetic as S
    from oracle import gs_oracle as O          # the one place bench.py touches oracle/: the reported CPU baseline
    O.build()
    Ps, Ws, Hs = 1_000_000, 1920, 1080
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
        if time.time() - t0 > budget_s or n >= 10:
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
    ap.add_argument("--no-extras", action="store_true", help="skip the c4_render / c3 sub-records (kernel iteration runs)")
    ap.add_argument("--no-exchange-check", action="store_true", help="N > 1: skip the correctness check of the exchange before the timed region")
    ap.add_argument("--views-per-rank", type=int, default=0, help="N > 1: cameras per GPU per step (default: 8 cameras per step over the GPUs)")
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
