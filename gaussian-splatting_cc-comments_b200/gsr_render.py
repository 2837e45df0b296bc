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
    return (getattr(pc, "scaling_activation", torch.exp) is torch.exp and getattr(pc, "opacity_activation", torch.sigmoid) is torch.sigmoid
            and getattr(pc, "rotation_activation", torch.nn.functional.normalize) is torch.nn.functional.normalize)


def render(viewpoint_camera, pc, pipe, bg_color, scaling_modifier=1.0, override_color=None):
    screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype, requires_grad=True, device=pc.get_xyz.device) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    tanfovx = math.tan(viewpoint_camera.FoVx * 0.5)
    tanfovy = math.tan(viewpoint_camera.FoVy * 0.5)
    raster_settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width), tanfovx=tanfovx, tanfovy=tanfovy,
        bg=bg_color, scale_modifier=scaling_modifier, viewmatrix=viewpoint_camera.world_view_transform,
        projmatrix=viewpoint_camera.full_proj_transform, sh_degree=pc.active_sh_degree, campos=viewpoint_camera.camera_center,
        prefiltered=False, debug=pipe.debug)
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)
    means3D = pc.get_xyz
    means2D = screenspace_points
    fused = (override_color is None and not pipe.convert_SHs_python and not pipe.compute_cov3D_python and _default_activations(pc))
    if fused:
        rendered_image, radii = rasterizer.forward_raw(means3D, means2D, pc._features_dc, pc._features_rest, pc._opacity, pc._scaling,
                                                       pc._rotation)
    else:
        scales = rotations = cov3D_precomp = None
        if pipe.compute_cov3D_python:
            cov3D_precomp = pc.get_covariance(scaling_modifier)
        else:
            scales, rotations = pc.get_scaling, pc.get_rotation
        shs = colors_precomp = None
        if override_color is None:
            if pipe.convert_SHs_python:
                from gsr_sh import eval_sh
                shs_view = pc.get_features.transpose(1, 2).view(-1, 3, (pc.max_sh_degree + 1) ** 2)
                dir_pp = pc.get_xyz - viewpoint_camera.camera_center.repeat(pc.get_features.shape[0], 1)
                dir_pp_normalized = dir_pp / dir_pp.norm(dim=1, keepdim=True)
                colors_precomp = torch.clamp_min(eval_sh(pc.active_sh_degree, shs_view, dir_pp_normalized) + 0.5, 0.0)
            else:
                shs = pc.get_features
        else:
            colors_precomp = override_color
        rendered_image, radii = rasterizer(means3D=means3D, means2D=means2D, shs=shs, colors_precomp=colors_precomp,
                                           opacities=pc.get_opacity, scales=scales, rotations=rotations, cov3D_precomp=cov3D_precomp)
    return {"render": rendered_image, "viewspace_points": screenspace_points, "visibility_filter": radii > 0, "radii": radii}
