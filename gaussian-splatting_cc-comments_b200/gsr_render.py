"""`render()` of the reference (gaussian_renderer/__init__.py:19-127) with the activations fused into the rasterizer.

Same signature and result dictionary, so `from gaussian_renderer import render` can be pointed here:

    render(viewpoint_camera, pc, pipe, bg_color, scaling_modifier=1.0, override_color=None)
        -> {"render", "viewspace_points", "visibility_filter", "radii"}

The reference computes `pc.get_opacity` (sigmoid), `pc.get_scaling` (exp), `pc.get_rotation` (normalize) and
`pc.get_features` (cat of `_features_dc` / `_features_rest`) with separate torch kernels on every call and autograd
runs their backward after the rasterizer's; here the rasterizer reads `pc._opacity`, `pc._scaling`, `pc._rotation`,
`pc._features_dc`, `pc._features_rest` directly (GaussianRasterizer.forward_raw) and returns the gradients w.r.t. them.
When a pipeline option asks for something the fused path does not cover (`pipe.convert_SHs_python`,
`pipe.compute_cov3D_python`, `override_color`, or a model whose activations are not the reference's defaults), the call
takes the reference's route through `GaussianRasterizer.forward` -- still this package's CUDA kernels; there is no CPU path.
"""
import math

import torch

from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer


def _default_activations(pc):
    """The fused kernels hard-wire exp / sigmoid / normalize (GaussianModel.setup_functions, scene/gaussian_model.py:27-42)."""
    wired = (("scaling_activation", torch.exp), ("opacity_activation", torch.sigmoid), ("rotation_activation", torch.nn.functional.normalize))
    return all(getattr(pc, name, fn) is fn for name, fn in wired)


def _settings_for(view, pc, pipe, bg_color, scaling_modifier):
    return GaussianRasterizationSettings(
        image_height=int(view.image_height), image_width=int(view.image_width),
        tanfovx=math.tan(0.5 * view.FoVx), tanfovy=math.tan(0.5 * view.FoVy),
        bg=bg_color, scale_modifier=scaling_modifier,
        viewmatrix=view.world_view_transform, projmatrix=view.full_proj_transform,
        sh_degree=pc.active_sh_degree, campos=view.camera_center, prefiltered=False, debug=pipe.debug)


def _activated_inputs(view, pc, pipe, scaling_modifier, override_color):
    """Keyword arguments of GaussianRasterizer.forward for the pipeline options the fused path does not cover
    (gaussian_renderer/__init__.py:69-101): python-side covariance, python-side SH evaluation, a colour override."""
    kw = {"opacities": pc.get_opacity}
    if pipe.compute_cov3D_python:
        kw["cov3D_precomp"] = pc.get_covariance(scaling_modifier)
    else:
        kw["scales"], kw["rotations"] = pc.get_scaling, pc.get_rotation
    if override_color is not None:
        kw["colors_precomp"] = override_color
    elif pipe.convert_SHs_python:
        from gsr_sh import eval_sh
        coeffs = pc.get_features.transpose(1, 2).view(-1, 3, (pc.max_sh_degree + 1) ** 2)
        towards = pc.get_xyz - view.camera_center.repeat(pc.get_features.shape[0], 1)
        towards = towards / towards.norm(dim=1, keepdim=True)
        kw["colors_precomp"] = torch.clamp_min(eval_sh(pc.active_sh_degree, coeffs, towards) + 0.5, 0.0)
    else:
        kw["shs"] = pc.get_features
    return kw


def render(viewpoint_camera, pc, pipe, bg_color, scaling_modifier=1.0, override_color=None):
    xyz = pc.get_xyz
    # the tensor whose .grad carries the screen-space gradient the densification statistics read (train.py:157-158)
    viewspace = torch.zeros_like(xyz, dtype=xyz.dtype, requires_grad=True, device=xyz.device) + 0
    try:
        viewspace.retain_grad()
    except Exception:
        pass
    rasterizer = GaussianRasterizer(raster_settings=_settings_for(viewpoint_camera, pc, pipe, bg_color, scaling_modifier))
    fused = (override_color is None and not pipe.convert_SHs_python and not pipe.compute_cov3D_python and _default_activations(pc))
    if fused:
        image, radii = rasterizer.forward_raw(xyz, viewspace, pc._features_dc, pc._features_rest, pc._opacity, pc._scaling, pc._rotation)
    else:
        image, radii = rasterizer(means3D=xyz, means2D=viewspace, **_activated_inputs(viewpoint_camera, pc, pipe, scaling_modifier, override_color))
    return {"render": image, "viewspace_points": viewspace, "visibility_filter": radii > 0, "radii": radii}
