"""Spherical-harmonics colour in torch: `eval_sh` of the reference (utils/sh_utils.py:57-112), the path behind
`--convert_SHs_python` (gaussian_renderer/__init__.py:93-99).  Same name, arguments and result; written as a basis
matrix times the coefficients instead of the reference's term-by-term accumulation.  Device-agnostic torch code, used by
gsr_render.render only when that pipeline option is set (the default path evaluates SH inside the rasterizer)."""
import torch

C0 = 0.28209479177387814
C1 = 0.4886025119029199
C2 = (1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396)
C3 = (-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658, 1.445305721320277,
      -0.5900435899266435)


def sh_basis(deg, dirs):
    """(..., 3) unit directions -> (..., (deg+1)^2) real SH basis values with the reference's signs and constants."""
    assert 0 <= deg <= 3
    x, y, z = dirs[..., 0], dirs[..., 1], dirs[..., 2]
    b = [torch.full_like(x, C0)]
    if deg > 0:
        b += [-C1 * y, C1 * z, -C1 * x]
    if deg > 1:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        b += [C2[0] * xy, C2[1] * yz, C2[2] * (2.0 * zz - xx - yy), C2[3] * xz, C2[4] * (xx - yy)]
        if deg > 2:
            b += [C3[0] * y * (3 * xx - yy), C3[1] * xy * z, C3[2] * y * (4 * zz - xx - yy), C3[3] * z * (2 * zz - 3 * xx - 3 * yy),
                  C3[4] * x * (4 * zz - xx - yy), C3[5] * z * (xx - yy), C3[6] * x * (xx - 3 * yy)]
    return torch.stack(b, dim=-1)


def eval_sh(deg, sh, dirs):
    """sh: (..., C, (max_deg+1)^2) coefficients, dirs: (..., 3) unit directions -> (..., C) colours (before the +0.5 shift)."""
    assert deg <= 4 and deg >= 0
    if deg > 3:
        raise NotImplementedError("SH degree 4 is not used by the rasterizer path (max_sh_degree = 3)")
    K = (deg + 1) ** 2
    assert sh.shape[-1] >= K
    return (sh[..., :K] * sh_basis(deg, dirs).unsqueeze(-2)).sum(dim=-1)
