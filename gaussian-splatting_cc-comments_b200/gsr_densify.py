"""Adaptive density control of the reference's GaussianModel on sm_100a kernels (second half).

Two drop-ins for what train.py does to the model between the backward pass and the optimizer step:

  add_densification_stats(model, viewspace_point_tensor, radii)
      train.py:153-158 -- `max_radii2D[vis] = max(max_radii2D[vis], radii[vis])` and
      GaussianModel.add_densification_stats (scene/gaussian_model.py:599-602) -- as ONE kernel launch, without the
      nonzero() synchronisations of boolean-mask indexing.
  densify_and_prune(model, max_grad, min_opacity, extent, max_screen_size)
      GaussianModel.densify_and_prune (scene/gaussian_model.py:558-597) with everything it calls (:366-556): two library
      calls (`gsr_densify_classify`, `gsr_densify_apply`, include/gsr.h) move every surviving row of the six parameter
      tensors and their Adam moments once, into the reference's final order, instead of three rounds of mask-indexing
      and torch.cat.  The samples of densify_and_split are drawn by the SAME torch.normal call on the same shapes, so
      the global generator advances as it does in the reference.

`model` is anything with the reference's attributes: `_xyz _features_dc _features_rest _opacity _scaling _rotation`,
`optimizer` (torch.optim.Adam or gsr_optim.FusedAdam, groups named xyz / f_dc / f_rest / opacity / scaling / rotation),
`xyz_gradient_accum`, `denom`, `max_radii2D`, `percent_dense`.  CUDA float32 only: there is no CPU path (the extension
rejects host tensors).
"""
import torch
from torch import nn

from diff_gaussian_rasterization import _C

_ATTR = {"xyz": "_xyz", "f_dc": "_features_dc", "f_rest": "_features_rest", "opacity": "_opacity", "scaling": "_scaling",
         "rotation": "_rotation"}
SPLIT_SELECTED = 16       # GSR_DENSIFY_SPLIT_SEL (include/gsr.h)


def add_densification_stats(model, viewspace_point_tensor, radii):
    """Updates model.max_radii2D, model.xyz_gradient_accum and model.denom in place for the Gaussians with radii > 0."""
    grad = viewspace_point_tensor.grad if isinstance(viewspace_point_tensor, torch.Tensor) and viewspace_point_tensor.grad is not None \
        else viewspace_point_tensor
    _C.densify_stats(grad.contiguous(), radii.to(torch.int32).contiguous(), model.xyz_gradient_accum, model.denom, model.max_radii2D)


def _groups_by_name(optimizer):
    groups = {}
    for group in optimizer.param_groups:
        assert len(group["params"]) == 1, "one tensor per parameter group, as GaussianModel.training_setup builds them"
        groups[group["name"]] = group
    missing = [n for n in _ATTR if n not in groups]
    if missing:
        raise ValueError("gsr_densify: optimizer has no parameter group named " + ", ".join(missing))
    return groups


def densify_and_prune(model, max_grad, min_opacity, extent, max_screen_size, N=2, generator=None):
    """In place on `model` and its optimizer; returns (P_before, P_after, n_cloned_kept, n_split).
    Same arguments as GaussianModel.densify_and_prune; N is densify_and_split's N."""
    groups = _groups_by_name(model.optimizer)
    names = [g["name"] for g in model.optimizer.param_groups if g["name"] in _ATTR]
    params = [groups[n]["params"][0] for n in names]
    P = params[0].shape[0]
    states = [model.optimizer.state.get(p, None) for p in params]
    states = [s if s else None for s in states]                     # an empty defaultdict entry is "no state yet"
    screen = float(max_screen_size) if max_screen_size else 0.0
    hyper = (float(max_grad), float(min_opacity), float(extent), float(model.percent_dense), screen, int(N))

    scaling = params[names.index("scaling")].detach()
    opacity = params[names.index("opacity")].detach()
    flags, totals_dev, scratch = _C.densify_classify(model.xyz_gradient_accum.contiguous(), model.denom.contiguous(), scaling.contiguous(),
                                                     opacity.contiguous(), *hyper)
    totals = [int(x) for x in totals_dev.tolist()]                  # the one host synchronisation of the whole operation
    P_out, n_split = totals[0], totals[1]

    # densify_and_split's own sampling call (scene/gaussian_model.py:515-517), same shapes, same generator use
    if n_split > 0:
        split = (flags & SPLIT_SELECTED) != 0
        stds = torch.exp(scaling)[split].repeat(N, 1)
        means = torch.zeros((stds.size(0), 3), device=stds.device)
        samples = torch.normal(mean=means, std=stds, generator=generator)
    else:
        samples = torch.zeros((0, 3), device=scaling.device)

    new_p, new_m, new_v = _C.densify_apply(
        flags, scratch, totals, samples.contiguous(), [p.detach().contiguous() for p in params],
        [s["exp_avg"].contiguous() if s is not None else None for s in states],
        [s["exp_avg_sq"].contiguous() if s is not None else None for s in states],
        names.index("xyz"), names.index("scaling"), names.index("rotation"), *hyper)

    # hand the new tensors to the optimizer and the model the way _prune_optimizer / cat_tensors_to_optimizer do
    for i, n in enumerate(names):
        group, old = groups[n], params[i]
        fresh = nn.Parameter(new_p[i].requires_grad_(True))
        if states[i] is not None:
            stored = states[i]
            stored["exp_avg"], stored["exp_avg_sq"] = new_m[i], new_v[i]
            del model.optimizer.state[old]
            group["params"][0] = fresh
            model.optimizer.state[fresh] = stored
        else:
            group["params"][0] = fresh
        setattr(model, _ATTR[n], fresh)
    dev = new_p[0].device
    model.xyz_gradient_accum = torch.zeros((P_out, 1), device=dev)   # densification_postfix (:493-495), then indexed by the prunes
    model.denom = torch.zeros((P_out, 1), device=dev)
    model.max_radii2D = torch.zeros((P_out,), device=dev)
    return P, P_out, totals[3], n_split
