"""The optimizer step of the reference's training loop on ONE fused sm_100a kernel.

The reference builds `torch.optim.Adam(l, lr=0.0, eps=1e-15)` over six named groups (scene/gaussian_model.py:243-252) and
calls `optimizer.step()` / `optimizer.zero_grad(set_to_none=True)` once per iteration (train.py:173-175).  `FusedAdam` is a
`torch.optim.Optimizer` with the same constructor arguments, the same `param_groups` and the same per-parameter `state`
entries (`step`, `exp_avg`, `exp_avg_sq`), so the code that edits the optimizer during densification
(`replace_tensor_to_optimizer`, `_prune_optimizer`, `cat_tensors_to_optimizer`, scene/gaussian_model.py:366-471) and
`state_dict()` / `load_state_dict()` (checkpoints, train.py:180, scene/gaussian_model.py:74-107) work on it unchanged.
Its `step()` updates every group with one launch of `csrc/gsr_adam.cu` through the C ABI (`gsr_adam_step`, include/gsr.h)
instead of torch's seven multi-tensor passes.  CUDA float32 parameters only: there is no CPU path.

Also here: `get_expon_lr_func` (utils/general_utils.py:29-71), the schedule of the xyz learning rate, and
`gaussian_param_groups`, the group list of GaussianModel.training_setup.
"""
import math

import torch

from diff_gaussian_rasterization import _C


def get_expon_lr_func(lr_init, lr_final, lr_delay_steps=0, lr_delay_mult=1.0, max_steps=1000000):
    """Log-linear interpolation from lr_init (step 0) to lr_final (step max_steps) with an optional sine-eased warm-up
    that starts at lr_init * lr_delay_mult; same arguments and values as utils/general_utils.py:29-71."""
    def helper(step):
        if step < 0 or (lr_init == 0.0 and lr_final == 0.0):
            return 0.0
        if lr_delay_steps > 0:
            delay_rate = lr_delay_mult + (1 - lr_delay_mult) * math.sin(0.5 * math.pi * min(max(step / lr_delay_steps, 0.0), 1.0))
        else:
            delay_rate = 1.0
        t = min(max(step / max_steps, 0.0), 1.0)
        return delay_rate * math.exp(math.log(lr_init) * (1 - t) + math.log(lr_final) * t)
    return helper


def gaussian_param_groups(xyz, features_dc, features_rest, opacity, scaling, rotation, training_args, spatial_lr_scale):
    """The six parameter groups of GaussianModel.training_setup (scene/gaussian_model.py:243-250)."""
    return [
        {"params": [xyz], "lr": training_args.position_lr_init * spatial_lr_scale, "name": "xyz"},
        {"params": [features_dc], "lr": training_args.feature_lr, "name": "f_dc"},
        {"params": [features_rest], "lr": training_args.feature_lr / 20.0, "name": "f_rest"},
        {"params": [opacity], "lr": training_args.opacity_lr, "name": "opacity"},
        {"params": [scaling], "lr": training_args.scaling_lr, "name": "scaling"},
        {"params": [rotation], "lr": training_args.rotation_lr, "name": "rotation"},
    ]


class FusedAdam(torch.optim.Optimizer):
    """torch.optim.Adam(params, lr, betas, eps) -- the configuration the reference uses -- with a one-kernel step().

    Not supported (the reference does not use them; asking for them raises): weight_decay, amsgrad, maximize, sparse or
    non-float32 gradients.  `step(visible=mask)` restricts the update to the Gaussians where the (P,) bool mask is set,
    leaving the others' parameters AND moments untouched -- an extension, not what the reference does (torch decays the
    moments of Gaussians with zero gradient as well); every parameter's first dimension must then be P.
    """

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False, *, maximize=False):
        if not 0.0 <= lr:
            raise ValueError(f"Invalid learning rate: {lr}")
        if not 0.0 <= eps:
            raise ValueError(f"Invalid epsilon value: {eps}")
        if not 0.0 <= betas[0] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 0: {betas[0]}")
        if not 0.0 <= betas[1] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 1: {betas[1]}")
        if weight_decay != 0 or amsgrad or maximize:
            raise NotImplementedError("FusedAdam implements the configuration the reference trains with: "
                                      "weight_decay=0, amsgrad=False, maximize=False")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=0, amsgrad=False, maximize=False))

    def _init_state(self, p):
        """State of a parameter that has not been stepped yet, laid out as torch.optim.Adam's (torch/optim/adam.py)."""
        state = self.state[p]
        if len(state) == 0:
            state["step"] = torch.tensor(0.0, dtype=torch.float32)
            state["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            state["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return state

    @torch.no_grad()
    def step(self, closure=None, visible=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        # one launch per (device, betas, eps) combination: for the reference's optimizer that is one launch in total
        batches = {}
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda:
                    raise RuntimeError("FusedAdam: CUDA parameters only (there is no CPU path)")
                if p.grad.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                if p.dtype != torch.float32 or p.grad.dtype != torch.float32:
                    raise TypeError("FusedAdam: float32 parameters and gradients")
                if not p.is_contiguous():
                    raise RuntimeError("FusedAdam: parameters must be contiguous (they are updated in place)")
                state = self._init_state(p)
                state["step"] += 1
                for k in ("exp_avg", "exp_avg_sq"):
                    if not state[k].is_contiguous():
                        state[k] = state[k].contiguous()
                b = batches.setdefault((p.device, beta1, beta2, group["eps"]), ([], [], [], [], [], []))
                b[0].append(p)
                b[1].append(p.grad.contiguous())
                b[2].append(state["exp_avg"])
                b[3].append(state["exp_avg_sq"])
                b[4].append(float(group["lr"]))
                b[5].append(int(state["step"]))
        for (dev, beta1, beta2, eps), b in batches.items():
            vis = None
            if visible is not None:
                vis = visible.to(device=dev).contiguous()
                if vis.dtype not in (torch.bool, torch.uint8):
                    raise TypeError("FusedAdam.step: `visible` must be a bool or uint8 mask")
            _C.adam_step(b[0], b[1], b[2], b[3], b[4], b[5], float(beta1), float(beta2), float(eps), vis)
        return loss
