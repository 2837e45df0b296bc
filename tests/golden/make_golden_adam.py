#!/usr/bin/env python3
"""Golden vectors for the Adam step, produced HERE (CPU) by the optimizer the reference uses.

The reference's optimizer is `torch.optim.Adam(l, lr=0.0, eps=1e-15)` over six named groups with the learning rates of
arguments.OptimizationParams, and the xyz rate follows utils.general_utils.get_expon_lr_func
(scene/gaussian_model.py:243-275).  This script imports those two reference modules from /root/reference, builds that
optimizer on CPU tensors shaped like a small model, runs a few iterations with fixed gradients (some rows zero, like
Gaussians that were not visible) and one densification-style state edit (moments extended with zeros,
scene/gaussian_model.py:447-471), and stores inputs and every intermediate state in tests/golden/ref_adam_cpu.npz.

Run in the build container only (needs /root/reference); the fixture is committed, the GPU box never runs this.
"""
import os
import sys
from argparse import ArgumentParser

import numpy as np
import torch

REF = "/root/reference"
sys.path.insert(0, REF)
from arguments import OptimizationParams            # noqa: E402  (reference module)
from utils.general_utils import get_expon_lr_func   # noqa: E402  (reference module)

HERE = os.path.dirname(os.path.abspath(__file__))
NAMES = ["xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation"]
SHAPES = {"xyz": (3,), "f_dc": (1, 3), "f_rest": (15, 3), "opacity": (1,), "scaling": (3,), "rotation": (4,)}


def main():
    torch.manual_seed(20260101)
    torch.set_num_threads(1)
    opt_args = OptimizationParams(ArgumentParser())
    spatial_lr_scale = 4.7
    P, P_new, steps = 101, 13, 5
    params = {n: torch.nn.Parameter(torch.randn(P, *SHAPES[n])) for n in NAMES}
    lrs = {"xyz": opt_args.position_lr_init * spatial_lr_scale, "f_dc": opt_args.feature_lr, "f_rest": opt_args.feature_lr / 20.0,
           "opacity": opt_args.opacity_lr, "scaling": opt_args.scaling_lr, "rotation": opt_args.rotation_lr}
    groups = [{"params": [params[n]], "lr": lrs[n], "name": n} for n in NAMES]        # scene/gaussian_model.py:243-250
    optimizer = torch.optim.Adam(groups, lr=0.0, eps=1e-15)                          # :252
    sched = get_expon_lr_func(lr_init=opt_args.position_lr_init * spatial_lr_scale, lr_final=opt_args.position_lr_final * spatial_lr_scale,
                              lr_delay_mult=opt_args.position_lr_delay_mult, max_steps=opt_args.position_lr_max_steps)
    out = {"names": np.array(NAMES), "steps": steps, "P": P, "P_new": P_new, "spatial_lr_scale": spatial_lr_scale,
           "position_lr_init": opt_args.position_lr_init, "position_lr_final": opt_args.position_lr_final,
           "position_lr_delay_mult": opt_args.position_lr_delay_mult, "position_lr_max_steps": opt_args.position_lr_max_steps}
    for n in NAMES:
        out["init_" + n] = params[n].detach().numpy().copy()
        out["lr_" + n] = lrs[n]
    iteration = 0
    for it in range(steps):
        iteration = 1 + it * 700                     # spread over the schedule so the xyz rate visibly changes
        if it == 3:
            # densification_postfix-style edit: parameters and moments grow by P_new rows (cat_tensors_to_optimizer, :447-471)
            for group in optimizer.param_groups:
                n = group["name"]
                ext = torch.randn(P_new, *SHAPES[n])
                out["ext_" + n] = ext.numpy().copy()
                st = optimizer.state.get(group["params"][0], None)
                st["exp_avg"] = torch.cat((st["exp_avg"], torch.zeros_like(ext)), dim=0)
                st["exp_avg_sq"] = torch.cat((st["exp_avg_sq"], torch.zeros_like(ext)), dim=0)
                del optimizer.state[group["params"][0]]
                group["params"][0] = torch.nn.Parameter(torch.cat((group["params"][0].detach(), ext), dim=0).requires_grad_(True))
                optimizer.state[group["params"][0]] = st
        for group in optimizer.param_groups:          # update_learning_rate, :259-275
            if group["name"] == "xyz":
                group["lr"] = sched(iteration)
                out["xyz_lr_step%d" % it] = group["lr"]
        out["iteration_step%d" % it] = iteration
        for group in optimizer.param_groups:
            p = group["params"][0]
            g = torch.randn_like(p) * (10.0 ** torch.randint(-6, 1, (p.shape[0],) + (1,) * (p.dim() - 1)).float())
            rows = torch.rand(p.shape[0]) < 0.3       # invisible Gaussians: exactly zero gradient
            g[rows] = 0.0
            p.grad = g
            out["grad_%s_step%d" % (group["name"], it)] = g.numpy().copy()
        optimizer.step()
        optimizer.zero_grad(set_to_none=True)
        for group in optimizer.param_groups:
            p = group["params"][0]
            st = optimizer.state[p]
            n = group["name"]
            out["param_%s_step%d" % (n, it)] = p.detach().numpy().copy()
            out["exp_avg_%s_step%d" % (n, it)] = st["exp_avg"].numpy().copy()
            out["exp_avg_sq_%s_step%d" % (n, it)] = st["exp_avg_sq"].numpy().copy()
            out["state_step_%s_step%d" % (n, it)] = float(st["step"])
    path = os.path.join(HERE, "ref_adam_cpu.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, "%.1f kB" % (os.path.getsize(path) / 1e3), "torch", torch.__version__)


if __name__ == "__main__":
    main()
