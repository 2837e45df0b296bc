"""Seeded synthetic scenes and cameras for parity tests and benchmarks.

Camera matrices follow the reference's own formulas: getWorld2View2 / getProjectionMatrix
(utils/graphics_utils.py:38-75) and Camera.__init__ (scene/cameras.py:57-61): matrices are stored
transposed (row-vector convention), full_proj = world_view @ projection, campos = inv(world_view)[3,:3].
Everything is generated on the CPU with numpy (float32), so the same bits feed the CPU oracle, the
reference build and the CUDA path.
"""
import math
import numpy as np

# median pixel sigma calibrated so that num_rendered / P ~= 4 (see calibrate() below)
S0_PIX = {(256, 256): 1.95, (1920, 1080): 1.95, (3840, 2160): 1.95}


def make_camera(W, H, fx=None, R=None, T=None, znear=0.01, zfar=100.0):
    fx = 0.9 * W if fx is None else fx
    fy = fx
    fovx = 2.0 * math.atan(W / (2.0 * fx))
    fovy = 2.0 * math.atan(H / (2.0 * fy))
    R = np.eye(3) if R is None else np.asarray(R, dtype=np.float64)
    T = np.zeros(3) if T is None else np.asarray(T, dtype=np.float64)
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = R.transpose()
    Rt[:3, 3] = T
    Rt[3, 3] = 1.0
    Rt = np.linalg.inv(np.linalg.inv(Rt))          # getWorld2View2 with translate=0, scale=1
    world_view = np.float32(Rt).transpose().copy()
    tx, ty = math.tan(fovx / 2), math.tan(fovy / 2)
    top, right = ty * znear, tx * znear
    Pj = np.zeros((4, 4), dtype=np.float32)
    Pj[0, 0] = 2.0 * znear / (2 * right)
    Pj[1, 1] = 2.0 * znear / (2 * top)
    Pj[3, 2] = 1.0
    Pj[2, 2] = zfar / (zfar - znear)
    Pj[2, 3] = -(zfar * znear) / (zfar - znear)
    proj = Pj.transpose().copy()
    full = (world_view @ proj).astype(np.float32)
    campos = np.linalg.inv(world_view)[3, :3].astype(np.float32)
    return dict(W=W, H=H, tanfovx=math.tan(fovx * 0.5), tanfovy=math.tan(fovy * 0.5), fx=fx,
                viewmatrix=np.ascontiguousarray(world_view), projmatrix=np.ascontiguousarray(full),
                campos=np.ascontiguousarray(campos))


def look_at_camera(W, H, eye, target, up=(0.0, -1.0, 0.0), fx=None):
    """COLMAP-style camera (x right, y down, z forward) at `eye` looking at `target`."""
    eye = np.asarray(eye, dtype=np.float64); target = np.asarray(target, dtype=np.float64)
    z = target - eye; z /= np.linalg.norm(z)
    x = np.cross(-np.asarray(up, dtype=np.float64), z); x /= np.linalg.norm(x)
    y = np.cross(z, x)
    Rc2w = np.stack([x, y, z], axis=1)               # camera-to-world rotation
    T = -Rc2w.transpose() @ eye                      # world-to-camera translation
    return make_camera(W, H, fx=fx, R=Rc2w, T=T)


def make_scene(P, cam, sh_degree=3, M=None, seed=0, s0_pix=None, near_frac=0.02):
    """Random Gaussians in front of `cam` (assumed at the origin looking down +z for placement)."""
    rng = np.random.default_rng(seed)
    W, H = cam["W"], cam["H"]
    M = (sh_degree + 1) ** 2 if M is None else M
    s0 = S0_PIX.get((W, H), 1.95) if s0_pix is None else s0_pix
    z = rng.uniform(1.0, 9.0, P)
    nn = int(P * near_frac)
    if nn:
        z[:nn] = rng.uniform(-1.0, 0.2, nn)
    x = z * cam["tanfovx"] * rng.uniform(-1.15, 1.15, P)
    y = z * cam["tanfovy"] * rng.uniform(-1.15, 1.15, P)
    means = np.stack([x, y, z], 1).astype(np.float32)
    perm = rng.permutation(P)
    means = means[perm]
    zz = np.abs(means[:, 2:3]).astype(np.float64) + 1e-3
    scales = (zz / cam["fx"]) * np.exp(rng.normal(math.log(s0), 0.6, (P, 3)))
    scales = scales.astype(np.float32)
    q = rng.normal(0, 1, (P, 4))
    rots = (q / np.linalg.norm(q, axis=1, keepdims=True)).astype(np.float32)
    opac = (1.0 / (1.0 + np.exp(-rng.normal(0, 2, (P, 1))))).astype(np.float32)
    shs = rng.normal(0, 0.05, (P, M, 3))
    shs[:, 0, :] = (rng.uniform(0, 1, (P, 3)) - 0.5) / 0.28209479
    shs = shs.astype(np.float32)
    return dict(means3D=means, scales=scales, rotations=rots, opacities=opac, shs=shs,
                sh_degree=sh_degree, M=M)


def upstream_grad(cam, seed=1):
    rng = np.random.default_rng(seed)
    return rng.uniform(-1, 1, (3, cam["H"], cam["W"])).astype(np.float32)
