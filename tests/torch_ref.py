"""Independent float64 torch-autograd model of the rasterizer's math (TEST INFRASTRUCTURE).

Written from the formulas, not from the kernels: pinhole EWA projection (forward.cu:84-140), cov3D
from scale/rotation (forward.cu:146-180), SH colour (forward.cu:21-81), front-to-back alpha blending
with the reference's thresholds (forward.cu:425-468).  The discrete decisions (visibility, per-tile
order, which pairs are skipped, where a pixel terminates) are recomputed here in float64; gradients
come from autograd, so this checks the hand-derived backward of both the oracle and the CUDA path.
Only usable on small scenes (dense per-tile alpha matrices).
"""
import numpy as np
import torch

C0 = 0.28209479177387814
C1 = 0.4886025119029199
C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
      1.445305721320277, -0.5900435899266435]


def sh_color(deg, sh, dirs):
    x, y, z = dirs[:, 0:1], dirs[:, 1:2], dirs[:, 2:3]
    res = C0 * sh[:, 0]
    if deg > 0:
        res = res - C1 * y * sh[:, 1] + C1 * z * sh[:, 2] - C1 * x * sh[:, 3]
    if deg > 1:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        res = (res + C2[0] * xy * sh[:, 4] + C2[1] * yz * sh[:, 5] + C2[2] * (2 * zz - xx - yy) * sh[:, 6] +
               C2[3] * xz * sh[:, 7] + C2[4] * (xx - yy) * sh[:, 8])
    if deg > 2:
        res = (res + C3[0] * y * (3 * xx - yy) * sh[:, 9] + C3[1] * xy * z * sh[:, 10] +
               C3[2] * y * (4 * zz - xx - yy) * sh[:, 11] + C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[:, 12] +
               C3[4] * x * (4 * zz - xx - yy) * sh[:, 13] + C3[5] * z * (xx - yy) * sh[:, 14] +
               C3[6] * x * (xx - 3 * yy) * sh[:, 15])
    return res + 0.5


def cov3d_from(scales, rots, mod):
    r, x, y, z = rots[:, 0], rots[:, 1], rots[:, 2], rots[:, 3]        # NOT normalised (forward.cu:155)
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                     2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                     2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).view(-1, 3, 3)
    L = R * (mod * scales)[:, None, :]          # R @ diag(s)
    return L @ L.transpose(1, 2)


def render(cam, scene, bg, use="sh", cov="sr", scale_modifier=1.0, dL=None):
    """Returns dict(color (3,H,W) float64 numpy, radii-free) and, with dL, gradients as numpy float64."""
    dt = torch.float64
    t = lambda a: torch.tensor(np.asarray(a), dtype=dt)
    W, H = cam["W"], cam["H"]
    ins = {"means3D": t(scene["means3D"]).requires_grad_(True), "opacities": t(scene["opacities"]).requires_grad_(True)}
    if use == "sh":
        ins["shs"] = t(scene["shs"]).requires_grad_(True)
    else:
        ins["colors_precomp"] = t(scene["colors_precomp"]).requires_grad_(True)
    if cov == "sr":
        ins["scales"] = t(scene["scales"]).requires_grad_(True)
        ins["rotations"] = t(scene["rotations"]).requires_grad_(True)
    else:
        ins["cov3D_precomp"] = t(scene["cov3D_precomp"]).requires_grad_(True)
    V = t(cam["viewmatrix"]); Pm = t(cam["projmatrix"]); campos = t(cam["campos"])
    tanx, tany = cam["tanfovx"], cam["tanfovy"]
    fx, fy = W / (2.0 * tanx), H / (2.0 * tany)
    m = ins["means3D"]
    P = m.shape[0]
    m1 = torch.cat([m, torch.ones(P, 1, dtype=dt)], 1)
    tv = m1 @ V                      # row-vector convention: matrices are stored transposed
    hom = m1 @ Pm
    pw = 1.0 / (hom[:, 3] + 1e-7)
    ndc = hom[:, :2] * pw[:, None]
    px = ((ndc[:, 0] + 1.0) * W - 1.0) * 0.5
    py = ((ndc[:, 1] + 1.0) * H - 1.0) * 0.5
    means2D = torch.zeros(P, 2, dtype=dt, requires_grad=True)      # screen-space offset (gradient probe)
    # gradient wrt means2D in the reference is d/d(ndc-scaled pixel) * 0.5*W: probe in pixel units, rescale later
    px = px + means2D[:, 0]
    py = py + means2D[:, 1]
    if cov == "sr":
        Sig = cov3d_from(ins["scales"], ins["rotations"], scale_modifier)
    else:
        c = ins["cov3D_precomp"]
        Sig = torch.stack([c[:, 0], c[:, 1], c[:, 2], c[:, 1], c[:, 3], c[:, 4], c[:, 2], c[:, 4], c[:, 5]], 1).view(-1, 3, 3)
    tz = tv[:, 2]
    limx, limy = 1.3 * tanx, 1.3 * tany
    # The reference's backward treats a clamped t.x / t.y as a constant (x_grad_mul = 0 and the t.z
    # derivative taken at fixed t.x, backward.cu:177-178,262-265); mirror that convention.
    rx, ry = (tv[:, 0] / tz).detach(), (tv[:, 1] / tz).detach()
    txc = torch.where((rx < -limx) | (rx > limx), (torch.clamp(rx, -limx, limx) * tz).detach(), tv[:, 0])
    tyc = torch.where((ry < -limy) | (ry > limy), (torch.clamp(ry, -limy, limy) * tz).detach(), tv[:, 1])
    z0 = torch.zeros_like(tz)
    J = torch.stack([fx / tz, z0, -(fx * txc) / (tz * tz), z0, fy / tz, -(fy * tyc) / (tz * tz)], 1).view(-1, 2, 3)
    Wm = V[:3, :3].t()               # world->view rotation
    T = J @ Wm
    cov2 = T @ Sig @ T.transpose(1, 2)
    a = cov2[:, 0, 0] + 0.3
    b = cov2[:, 0, 1]
    c = cov2[:, 1, 1] + 0.3
    det = a * c - b * b
    conA, conB, conC = c / det, -b / det, a / det
    mid = 0.5 * (a + c)
    lam = mid + torch.sqrt(torch.clamp(mid * mid - det, min=0.1))
    radius = torch.ceil(3.0 * torch.sqrt(lam)).detach()
    visible = (tz.detach() > 0.2) & (det.detach() != 0)
    if use == "sh":
        d = m - campos[None]
        d = d / d.norm(dim=1, keepdim=True)
        col = torch.clamp_min(sh_color(scene["sh_degree"], ins["shs"], d), 0.0)
    else:
        col = ins["colors_precomp"]
    op = ins["opacities"][:, 0]
    bgt = t(bg)
    gx, gy = (W + 15) // 16, (H + 15) // 16
    pxd, pyd, rd, zd = px.detach().numpy(), py.detach().numpy(), radius.numpy(), tz.detach().numpy()
    out = torch.zeros(3, H, W, dtype=dt)
    for ty in range(gy):
        for tx in range(gx):
            x0, y0 = tx * 16, ty * 16
            # rect test of getRect (auxiliary.h:48-58)
            rx0 = np.clip(((pxd - rd) / 16).astype(np.int64), 0, gx); rx1 = np.clip(((pxd + rd + 15) / 16).astype(np.int64), 0, gx)
            ry0 = np.clip(((pyd - rd) / 16).astype(np.int64), 0, gy); ry1 = np.clip(((pyd + rd + 15) / 16).astype(np.int64), 0, gy)
            sel = visible.numpy() & (rx0 <= tx) & (tx < rx1) & (ry0 <= ty) & (ty < ry1)
            ids = np.nonzero(sel)[0]
            if ids.size == 0:
                xs = slice(x0, min(x0 + 16, W)); ys = slice(y0, min(y0 + 16, H))
                out[:, ys, xs] = bgt[:, None, None]
                continue
            # depth order as float32 bits, ties by index (stable)
            ids = ids[np.argsort(zd[ids].astype(np.float32), kind="stable")]
            idt = torch.from_numpy(ids)
            xs = torch.arange(x0, min(x0 + 16, W), dtype=dt); ys = torch.arange(y0, min(y0 + 16, H), dtype=dt)
            PX, PY = torch.meshgrid(xs, ys, indexing="xy")           # (h, w)
            dx = px[idt][:, None, None] - PX[None]
            dy = py[idt][:, None, None] - PY[None]
            power = -0.5 * (conA[idt][:, None, None] * dx * dx + conC[idt][:, None, None] * dy * dy) - conB[idt][:, None, None] * dx * dy
            # the reference's backward differentiates through the 0.99 cap as if it were the identity
            # (backward.cu:529,577-598 use G and opacity regardless of the min): straight-through here too
            araw = op[idt][:, None, None] * torch.exp(power)
            alpha = araw + (torch.clamp(araw, max=0.99) - araw).detach()
            keep = (power.detach() <= 0) & (alpha.detach() >= 1.0 / 255.0)
            alpha = torch.where(keep, alpha, torch.zeros_like(alpha))
            Tcum = torch.cumprod(1.0 - alpha, 0)                       # T after each Gaussian
            Tbefore = torch.cat([torch.ones_like(Tcum[:1]), Tcum[:-1]], 0)
            # termination: first j with T_after < 1e-4 is NOT blended, nor anything behind it
            stop = (Tcum.detach() < 1e-4) & keep
            stopped = torch.cumsum(stop.to(torch.int64), 0) > 0
            alpha_eff = torch.where(stopped, torch.zeros_like(alpha), alpha)
            w = alpha_eff * Tbefore
            Cpix = (w[:, None] * col[idt][:, :, None, None]).sum(0)
            Tfinal = torch.prod(1.0 - alpha_eff, 0)
            out[:, int(ys[0]):int(ys[-1]) + 1, int(xs[0]):int(xs[-1]) + 1] = Cpix + Tfinal[None] * bgt[:, None, None]
    res = {"color": out.detach().numpy(), "radii_f": (radius * visible).numpy()}
    if dL is not None:
        out.backward(t(dL))
        g = {k: (v.grad.numpy() if v.grad is not None else None) for k, v in ins.items()}
        m2 = np.zeros((P, 3))
        # reference: dL_dmean2D is wrt NDC-like units, = pixel-gradient * 0.5*W (backward.cu:587-588)
        m2[:, 0] = means2D.grad[:, 0].numpy() * 0.5 * W
        m2[:, 1] = means2D.grad[:, 1].numpy() * 0.5 * H
        g["means2D"] = m2
        res["grads"] = g
    return res
