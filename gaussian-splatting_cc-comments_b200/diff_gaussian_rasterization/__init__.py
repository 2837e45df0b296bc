"""Drop-in `diff_gaussian_rasterization` backed by the sm_100a rasterizer in ../csrc.

Mirrors the public surface of the reference package
(DGR/diff_gaussian_rasterization/__init__.py: GaussianRasterizationSettings :168-180,
GaussianRasterizer :182-258, rasterize_gaussians :22-45, _RasterizeGaussians :47-166) so that
gaussian_renderer.render() and train.py run unchanged.  There is no CPU path: importing this
package without the built CUDA extension raises.
"""
from typing import NamedTuple
import os

import torch
import torch.nn as nn

_here = os.path.dirname(os.path.abspath(__file__))
if not os.path.exists(os.path.join(_here, "_C.so")):
    raise ImportError(
        "diff_gaussian_rasterization._C is not built: run `python __graft_entry__.py build` "
        "(or gaussian-splatting_cc-comments_b200/build.py).  There is no CPU fallback.")
from . import _C  # noqa: E402


def cpu_deep_copy_tuple(input_tuple):
    return tuple(x.cpu().clone() if isinstance(x, torch.Tensor) else x for x in input_tuple)


# Optional multi-view gradient exchange (one camera per GPU): an object with `.backward_extra(P, campos)` and
# `.exchange(...)`, installed by gsr_multi_gpu.enable_multiview_training().  None = single-GPU behaviour, exactly the
# reference's.
_multiview = None


def set_multiview_exchange(obj):
    global _multiview
    _multiview = obj


def _band(raster_settings):
    """Optional tile-row band (begin, end) carried by settings that have one (multi-GPU render)."""
    return getattr(raster_settings, "tile_rows", None) or (0, 0)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                        cov3Ds_precomp, raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales,
                                     rotations, cov3Ds_precomp, raster_settings)


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                cov3Ds_precomp, raster_settings):
        rs = raster_settings
        args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier,
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
                rs.image_height, rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        band = _band(rs)
        if rs.debug:
            cpu_args = cpu_deep_copy_tuple(args)  # before anything can corrupt them
            try:
                out = _C.rasterize_gaussians(*args, *band)
            except Exception:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise
        else:
            out = _C.rasterize_gaussians(*args, *band)
        num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer = out
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
                              geomBuffer, binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii)
        return color, radii

    @staticmethod
    def backward(ctx, grad_out_color, _):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
         geomBuffer, binningBuffer, imgBuffer) = ctx.saved_tensors
        args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier,
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color,
                sh, rs.sh_degree, rs.campos, geomBuffer, ctx.num_rendered, binningBuffer, imgBuffer, rs.debug)
        band = _band(rs)
        mv = _multiview if (_multiview is not None and sh.numel() > 0) else None
        if mv is not None:   # publish masked dL/dRGB for the fused SH-gradient exchange instead of writing dL_dsh
            band = (*band, *mv.backward_extra(means3D.shape[0], rs.campos))
        if rs.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                grads = _C.rasterize_gaussians_backward(*args, *band)
            except Exception:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise
        else:
            grads = _C.rasterize_gaussians_backward(*args, *band)
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp,
         grad_sh, grad_scales, grad_rotations) = grads
        if mv is not None:   # the SH gradient of the step's last local view is the sum over all views of all ranks
            grad_sh = mv.exchange(_C, means3D, rs.sh_degree, grad_sh,
                                  (grad_means3D, grad_means2D, grad_opacities, grad_scales, grad_rotations))
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities,
                grad_scales, grad_rotations, grad_cov3Ds_precomp, None)


# ---- raw-parameter path (extension) ---------------------------------------------------------
def rasterize_gaussians_raw(means3D, means2D, features_dc, features_rest, raw_opacities, raw_scales, raw_rotations,
                            raster_settings):
    return _RasterizeGaussiansRaw.apply(means3D, means2D, features_dc, features_rest, raw_opacities, raw_scales,
                                        raw_rotations, raster_settings)


class _RasterizeGaussiansRaw(torch.autograd.Function):
    """_RasterizeGaussians on the model's RAW tensors: sigmoid / exp / normalize and the SH concatenation that
    gaussian_renderer/__init__.py:69-101 performs with torch kernels before the rasterizer (and autograd after it) happen
    inside the per-Gaussian kernels.  Gradients come back w.r.t. the raw tensors."""

    @staticmethod
    def forward(ctx, means3D, means2D, features_dc, features_rest, raw_opacities, raw_scales, raw_rotations, raster_settings):
        rs = raster_settings
        if _band(rs) != (0, 0) or _multiview is not None:
            raise Exception("the raw-parameter path is single-GPU, full-image (use GaussianRasterizer.forward for bands / multi-view)")
        out = _C.rasterize_gaussians_raw(rs.bg, means3D, features_dc, features_rest, raw_opacities, raw_scales, raw_rotations,
                                         rs.scale_modifier, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height,
                                         rs.image_width, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer = out
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(means3D, features_dc, features_rest, raw_scales, raw_rotations, radii, geomBuffer, binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii)
        return color, radii

    @staticmethod
    def backward(ctx, grad_out_color, _):
        rs = ctx.raster_settings
        means3D, features_dc, features_rest, raw_scales, raw_rotations, radii, geomBuffer, binningBuffer, imgBuffer = ctx.saved_tensors
        (grad_means2D, grad_opacities, grad_means3D, grad_dc, grad_rest, grad_scales, grad_rotations) = _C.rasterize_gaussians_raw_backward(
            rs.bg, means3D, radii, features_dc, features_rest, raw_scales, raw_rotations, rs.scale_modifier, rs.viewmatrix, rs.projmatrix,
            rs.tanfovx, rs.tanfovy, grad_out_color, rs.sh_degree, rs.campos, geomBuffer, ctx.num_rendered, binningBuffer, imgBuffer, rs.debug)
        return grad_means3D, grad_means2D, grad_dc, grad_rest, grad_opacities, grad_scales, grad_rotations, None


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        empty = torch.Tensor([])
        shs = empty if shs is None else shs
        colors_precomp = empty if colors_precomp is None else colors_precomp
        scales = empty if scales is None else scales
        rotations = empty if rotations is None else rotations
        cov3D_precomp = empty if cov3D_precomp is None else cov3D_precomp
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, self.raster_settings)

    def forward_raw(self, means3D, means2D, features_dc, features_rest, raw_opacities, raw_scales, raw_rotations):
        """Extension: the model's raw tensors (`_features_dc`, `_features_rest`, `_opacity`, `_scaling`, `_rotation` of
        scene/gaussian_model.py:44-50) instead of activated ones; same (color, radii).  See gsr_render.render."""
        return rasterize_gaussians_raw(means3D, means2D, features_dc, features_rest, raw_opacities, raw_scales, raw_rotations,
                                       self.raster_settings)
