"""Training loop for BASELINE.json's config C3 ("training loop on a synthetic scene growing 1M -> 6M Gaussians via
densification: variable P, 1 view/iter, 1 B200") -- a MEASUREMENT HARNESS, not a product component.

It restates the per-iteration work of the reference's train.py:73-175 and the model methods it calls
(scene/gaussian_model.py: activations :114-138, training_setup :230-257, add_densification_stats :599-602,
densify_and_prune / densify_and_clone / densify_and_split / prune_points / cat_tensors_to_optimizer :398-597) around
three swappable pieces, so that the SAME loop times either arm:

    arm "b200"      : this package's rasterizer (diff_gaussian_rasterization), fused L1+SSIM loss (loss_utils) and
                      one-kernel Adam (gsr_optim.FusedAdam)
    arm "reference" : the unmodified reference rasterizer (oracle/_ref), the reference's loss as its composition of torch
                      ops (utils/loss_utils.py:17-64) and torch.optim.Adam

Everything else -- activations, the SH concatenation, densification statistics, clone / split / prune, optimizer state
edits -- is the same torch code in both arms.  One deviation, stated in the output: the densification threshold is not
the fixed 0.0002 of the reference but the quantile of the accumulated gradient that makes P follow a geometric schedule
from P0 to P_target (BASELINE.json: "thresholds tuned so P reaches ~6M"); with a synthetic target image the absolute
gradient scale means nothing.

The pieces are injected, so tests/test_c3_trainer_cpu.py drives the loop on the CPU with a stand-in rasterizer.
"""
import math

import torch


def inverse_sigmoid(x):
    return torch.log(x / (1 - x))          # utils/general_utils.py:19-20


def build_rotation(r):
    """Unit quaternion (w, x, y, z) rows -> rotation matrices; utils/general_utils.py:78-101."""
    q = r / torch.norm(r, dim=1, keepdim=True)
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                     2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                     2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], dim=1)
    return R.view(-1, 3, 3)


class OptArgs:
    """arguments/__init__.py:89-107 (the values the loop uses)."""
    position_lr_init = 0.00016
    position_lr_final = 0.0000016
    position_lr_delay_mult = 0.01
    position_lr_max_steps = 30_000
    feature_lr = 0.0025
    opacity_lr = 0.05
    scaling_lr = 0.005
    rotation_lr = 0.001
    percent_dense = 0.01
    lambda_dssim = 0.2


GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


class Gaussians:
    """The six raw parameter tensors + the densification bookkeeping of GaussianModel, nothing else."""

    def __init__(self, scene, device, optimizer_cls, lr_func, opt=OptArgs, spatial_lr_scale=1.0, fused_densify=None):
        self.fused_densify = fused_densify          # the gsr_densify module (b200 arm) or None (torch ops, as the reference)
        t = lambda a: torch.as_tensor(a, dtype=torch.float32, device=device)
        shs = t(scene["shs"])
        self.xyz = torch.nn.Parameter(t(scene["means3D"]).contiguous())
        self.f_dc = torch.nn.Parameter(shs[:, :1, :].contiguous())
        self.f_rest = torch.nn.Parameter(shs[:, 1:, :].contiguous())
        self.opacity = torch.nn.Parameter(inverse_sigmoid(t(scene["opacities"]).clamp(1e-4, 1 - 1e-4)).contiguous())
        self.scaling = torch.nn.Parameter(torch.log(t(scene["scales"])).contiguous())
        self.rotation = torch.nn.Parameter(t(scene["rotations"]).contiguous())
        self.sh_degree = int(scene["sh_degree"])
        self.opt = opt
        self.device = device
        P = self.xyz.shape[0]
        self.xyz_gradient_accum = torch.zeros((P, 1), device=device)
        self.denom = torch.zeros((P, 1), device=device)
        self.max_radii2D = torch.zeros((P,), device=device)
        groups = [
            {"params": [self.xyz], "lr": opt.position_lr_init * spatial_lr_scale, "name": "xyz"},
            {"params": [self.f_dc], "lr": opt.feature_lr, "name": "f_dc"},
            {"params": [self.f_rest], "lr": opt.feature_lr / 20.0, "name": "f_rest"},
            {"params": [self.opacity], "lr": opt.opacity_lr, "name": "opacity"},
            {"params": [self.scaling], "lr": opt.scaling_lr, "name": "scaling"},
            {"params": [self.rotation], "lr": opt.rotation_lr, "name": "rotation"},
        ]
        self.optimizer = optimizer_cls(groups, lr=0.0, eps=1e-15)
        self.xyz_lr = lr_func(lr_init=opt.position_lr_init * spatial_lr_scale, lr_final=opt.position_lr_final * spatial_lr_scale,
                              lr_delay_mult=opt.position_lr_delay_mult, max_steps=opt.position_lr_max_steps)

    # the reference's attribute names (scene/gaussian_model.py:47-57), which gsr_densify reads and writes
    _xyz = property(lambda s: s.xyz, lambda s, v: setattr(s, "xyz", v))
    _features_dc = property(lambda s: s.f_dc, lambda s, v: setattr(s, "f_dc", v))
    _features_rest = property(lambda s: s.f_rest, lambda s, v: setattr(s, "f_rest", v))
    _opacity = property(lambda s: s.opacity, lambda s, v: setattr(s, "opacity", v))
    _scaling = property(lambda s: s.scaling, lambda s, v: setattr(s, "scaling", v))
    _rotation = property(lambda s: s.rotation, lambda s, v: setattr(s, "rotation", v))
    percent_dense = property(lambda s: s.opt.percent_dense)

    # activations, scene/gaussian_model.py:114-138
    @property
    def P(self): return self.xyz.shape[0]
    @property
    def get_scaling(self): return torch.exp(self.scaling)
    @property
    def get_rotation(self): return torch.nn.functional.normalize(self.rotation)
    @property
    def get_opacity(self): return torch.sigmoid(self.opacity)
    @property
    def get_features(self): return torch.cat((self.f_dc, self.f_rest), dim=1)

    def update_learning_rate(self, iteration):
        for g in self.optimizer.param_groups:
            if g["name"] == "xyz":
                g["lr"] = self.xyz_lr(iteration)
                return g["lr"]

    def _rebind(self, tensors):
        self.xyz, self.f_dc, self.f_rest = tensors["xyz"], tensors["f_dc"], tensors["f_rest"]
        self.opacity, self.scaling, self.rotation = tensors["opacity"], tensors["scaling"], tensors["rotation"]

    def _edit_optimizer(self, edit_param, edit_state):
        out = {}
        for group in self.optimizer.param_groups:
            old = group["params"][0]
            stored = self.optimizer.state.get(old, None)
            if stored is not None:
                stored["exp_avg"] = edit_state(stored["exp_avg"], group["name"])
                stored["exp_avg_sq"] = edit_state(stored["exp_avg_sq"], group["name"])
                del self.optimizer.state[old]
            group["params"][0] = torch.nn.Parameter(edit_param(old, group["name"]).requires_grad_(True))
            if stored is not None:
                self.optimizer.state[group["params"][0]] = stored
            out[group["name"]] = group["params"][0]
        return out

    def prune_points(self, mask):                                   # :398-437
        keep = ~mask
        self._rebind(self._edit_optimizer(lambda p, n: p.detach()[keep], lambda s, n: s[keep]))
        self.xyz_gradient_accum = self.xyz_gradient_accum[keep]
        self.denom = self.denom[keep]
        self.max_radii2D = self.max_radii2D[keep]

    def densification_postfix(self, new):                           # :447-495
        self._rebind(self._edit_optimizer(lambda p, n: torch.cat((p.detach(), new[n]), dim=0),
                                          lambda s, n: torch.cat((s, torch.zeros_like(new[n])), dim=0)))
        P = self.P
        self.xyz_gradient_accum = torch.zeros((P, 1), device=self.device)
        self.denom = torch.zeros((P, 1), device=self.device)
        self.max_radii2D = torch.zeros((P,), device=self.device)

    def densify_and_clone(self, grads, grad_threshold, scene_extent):   # :538-556
        sel = torch.norm(grads, dim=-1) >= grad_threshold
        sel = torch.logical_and(sel, torch.max(self.get_scaling, dim=1).values <= self.opt.percent_dense * scene_extent)
        self.densification_postfix({"xyz": self.xyz.detach()[sel], "f_dc": self.f_dc.detach()[sel], "f_rest": self.f_rest.detach()[sel],
                                    "opacity": self.opacity.detach()[sel], "scaling": self.scaling.detach()[sel],
                                    "rotation": self.rotation.detach()[sel]})
        return int(sel.sum())

    def densify_and_split(self, grads, grad_threshold, scene_extent, N=2, generator=None):   # :497-536
        n_init = self.P
        padded = torch.zeros((n_init,), device=self.device)
        padded[:grads.shape[0]] = grads.squeeze()
        sel = padded >= grad_threshold
        sel = torch.logical_and(sel, torch.max(self.get_scaling, dim=1).values > self.opt.percent_dense * scene_extent)
        stds = self.get_scaling.detach()[sel].repeat(N, 1)
        samples = torch.normal(mean=torch.zeros((stds.size(0), 3), device=self.device), std=stds, generator=generator)
        rots = build_rotation(self.rotation.detach()[sel]).repeat(N, 1, 1)
        new = {"xyz": torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + self.xyz.detach()[sel].repeat(N, 1),
               "scaling": torch.log(self.get_scaling.detach()[sel].repeat(N, 1) / (0.8 * N)),
               "rotation": self.rotation.detach()[sel].repeat(N, 1), "f_dc": self.f_dc.detach()[sel].repeat(N, 1, 1),
               "f_rest": self.f_rest.detach()[sel].repeat(N, 1, 1), "opacity": self.opacity.detach()[sel].repeat(N, 1)}
        self.densification_postfix(new)
        n_sel = int(sel.sum())
        self.prune_points(torch.cat((sel, torch.zeros(N * n_sel, device=self.device, dtype=torch.bool))))
        return n_sel

    def densify_and_prune(self, max_grad, min_opacity, extent, max_screen_size, generator=None):   # :558-597
        if self.fused_densify is not None:
            P0, P1, n_clone_kept, n_split = self.fused_densify.densify_and_prune(self, max_grad, min_opacity, extent, max_screen_size,
                                                                                 generator=generator)
            return n_clone_kept, n_split, None
        grads = self.xyz_gradient_accum / self.denom
        grads[grads.isnan()] = 0.0
        n_clone = self.densify_and_clone(grads, max_grad, extent)
        n_split = self.densify_and_split(grads, max_grad, extent, generator=generator)
        prune = (self.get_opacity < min_opacity).squeeze()
        if max_screen_size:
            prune = torch.logical_or(torch.logical_or(prune, self.max_radii2D > max_screen_size),
                                     self.get_scaling.max(dim=1).values > 0.1 * extent)
        n_pruned = int(prune.sum())
        self.prune_points(prune)
        return n_clone, n_split, n_pruned

    def add_densification_stats(self, viewspace_points, update_filter):   # :599-602
        self.xyz_gradient_accum[update_filter] += torch.norm(viewspace_points.grad[update_filter, :2], dim=-1, keepdim=True)
        self.denom[update_filter] += 1


def render(cam, g, rasterizer_module, bg, scaling_modifier=1.0, raw=False):
    """gaussian_renderer/__init__.py:19-127 for the default pipeline (SH evaluated and covariance built by the rasterizer).
    raw=True (this package's rasterizer only): the model's RAW tensors go to GaussianRasterizer.forward_raw, which applies
    sigmoid / exp / normalize and reads the split SH storage inside the per-Gaussian kernels -- no torch
    activation kernels, no torch.cat, and autograd has nothing to run after the rasterizer's backward."""
    screenspace_points = torch.zeros_like(g.xyz, requires_grad=True) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    settings = rasterizer_module.GaussianRasterizationSettings(
        image_height=int(cam["H"]), image_width=int(cam["W"]), tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], bg=bg,
        scale_modifier=scaling_modifier, viewmatrix=cam["viewmatrix_t"], projmatrix=cam["projmatrix_t"], sh_degree=g.sh_degree,
        campos=cam["campos_t"], prefiltered=False, debug=False)
    rasterizer = rasterizer_module.GaussianRasterizer(raster_settings=settings)
    if raw:
        image, radii = rasterizer.forward_raw(g.xyz, screenspace_points, g.f_dc, g.f_rest, g.opacity, g.scaling, g.rotation)
        return {"render": image, "viewspace_points": screenspace_points, "visibility_filter": radii > 0, "radii": radii}
    image, radii = rasterizer(means3D=g.xyz, means2D=screenspace_points, shs=g.get_features, colors_precomp=None,
                              opacities=g.get_opacity, scales=g.get_scaling, rotations=g.get_rotation, cov3D_precomp=None)
    return {"render": image, "viewspace_points": screenspace_points, "visibility_filter": radii > 0, "radii": radii}


def growth_threshold(g, target_P):
    """The accumulated-gradient value above which enough Gaussians are selected for P to reach target_P at this
    densification (each selected Gaussian nets +1: a clone adds one, a split adds two and removes the original)."""
    grads = (g.xyz_gradient_accum / g.denom).squeeze(-1)
    grads[grads.isnan()] = 0.0
    want = max(0, min(int(target_P) - g.P, g.P))
    if want == 0:
        return float("inf")
    return float(torch.topk(grads, want, sorted=True).values[-1].clamp_min(1e-30))


class Phase:
    """Five CUDA events per iteration, read back after the run (no synchronisation inside the loop); no-op on the CPU."""

    def __init__(self, enabled):
        self.enabled = enabled
        self.rows = []

    def start(self):
        self.cur = []
        self.mark()

    def mark(self):
        if self.enabled:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self.cur.append(e)

    def finish(self, P):
        if self.enabled:
            self.rows.append((P, self.cur))

    def table(self):
        out = []
        for P, ev in self.rows:
            out.append((P, [ev[i].elapsed_time(ev[i + 1]) for i in range(len(ev) - 1)]))
        return out


def train(g, cams, targets, rasterizer_module, loss_fn, iterations, densify_from, densify_every, target_P, extent, bg,
          timed=True, seed=0, on_densify=None, raw=False):
    """The loop of train.py:73-175 without logging / saving / GUI.  loss_fn(image, gt, lambda_dssim) -> loss tensor.
    Returns the per-iteration phase table [(P, [render+loss fwd ms, backward ms, stats+densify ms, optimizer ms])]."""
    gen = torch.Generator(device=g.device).manual_seed(seed)
    n_events = max(1, (iterations - 1 - densify_from) // densify_every)
    P0 = g.P
    ph = Phase(timed and g.device != "cpu" and torch.cuda.is_available())
    event = 0
    for iteration in range(1, iterations + 1):
        g.update_learning_rate(iteration)
        cam = cams[(iteration - 1) % len(cams)]
        gt = targets[(iteration - 1) % len(cams)]
        ph.start()
        pkg = render(cam, g, rasterizer_module, bg, raw=raw)
        loss = loss_fn(pkg["render"], gt, g.opt.lambda_dssim)
        ph.mark()
        loss.backward()
        ph.mark()
        with torch.no_grad():
            vis, radii = pkg["visibility_filter"], pkg["radii"]
            if g.fused_densify is not None:
                g.fused_densify.add_densification_stats(g, pkg["viewspace_points"], radii)
            else:
                g.max_radii2D[vis] = torch.max(g.max_radii2D[vis], radii[vis].to(g.max_radii2D.dtype))
                g.add_densification_stats(pkg["viewspace_points"], vis)
            P_before = g.P
            if iteration > densify_from and iteration % densify_every == 0 and iteration < iterations:
                event += 1
                goal = int(round(P0 * (target_P / P0) ** (min(event, n_events) / n_events)))
                thr = growth_threshold(g, goal)
                stats = g.densify_and_prune(thr, 0.005, extent, None, generator=gen)
                if on_densify is not None:
                    on_densify(iteration, P_before, g.P, thr, stats)
                ph.mark()
                # the gradients belong to the old parameter tensors: as in the reference, this iteration's step sees none
                g.optimizer.zero_grad(set_to_none=True)
            else:
                ph.mark()
            g.optimizer.step()
            g.optimizer.zero_grad(set_to_none=True)
        ph.mark()
        ph.finish(P_before)
    return ph.table()


def geometric_buckets(table, n=6):
    """Mean phase times per bucket of P (geometric edges between the smallest and largest P seen)."""
    if not table:
        return []
    Ps = [p for p, _ in table]
    lo, hi = min(Ps), max(Ps)
    edges = [lo * (hi / lo) ** (i / n) for i in range(n + 1)] if hi > lo else [lo, hi]
    out = []
    for i in range(len(edges) - 1):
        rows = [r for p, r in table if edges[i] <= p < edges[i + 1] or (i == len(edges) - 2 and p == edges[-1])]
        if not rows:
            continue
        k = len(rows[0])
        mean = [sum(r[j] for r in rows) / len(rows) for j in range(k)]
        out.append({"P_from": int(edges[i]), "P_to": int(edges[i + 1]), "iterations": len(rows), "fwd_ms": mean[0], "bwd_ms": mean[1],
                    "densify_stats_ms": mean[2], "optimizer_ms": mean[3], "iter_ms": sum(mean)})
    return out
