"""ctypes/numpy front end of oracle/gs_oracle.c -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.
The product path (gaussian-splatting_cc-comments_b200/) never does.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "gs_oracle.c")
_SO = os.path.join(_HERE, "libgs_oracle.so")


def build(force=False):
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(_SRC):
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-mfma", "-fopenmp", "-fPIC", "-shared",
                               "-o", _SO, _SRC, "-lm"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.gso_bin.restype = C.c_int64
        _lib.gso_higher_msb.restype = C.c_uint32
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def mark_visible(means3D, viewmatrix):
    m = _f(means3D); P = m.shape[0]
    out = np.zeros(P, np.uint8)
    lib().gso_mark_visible(C.c_int(P), _p(m), _p(_f(viewmatrix)), _p(out))
    return out.astype(bool)


def forward(cam, means3D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
            cov3D_precomp=None, sh_degree=0, scale_modifier=1.0, bg=(0, 0, 0), render=True):
    """Full forward.  Returns a dict with every intermediate the parity tests compare."""
    L = lib()
    means3D = _f(means3D); P = means3D.shape[0]
    W, H = cam["W"], cam["H"]
    shs = _f(shs); colors_precomp = _f(colors_precomp); scales = _f(scales); rotations = _f(rotations)
    cov3D_precomp = _f(cov3D_precomp); opacities = _f(opacities)
    M = 0 if shs is None else shs.shape[1]
    V, Pm, cp = _f(cam["viewmatrix"]), _f(cam["projmatrix"]), _f(cam["campos"])
    radii = np.zeros(P, np.int32); xy = np.zeros((P, 2), np.float32); depths = np.zeros(P, np.float32)
    cov3D = np.zeros((P, 6), np.float32); rgb = np.zeros((P, 3), np.float32)
    conic_opacity = np.zeros((P, 4), np.float32); tiles = np.zeros(P, np.uint32); clamped = np.zeros((P, 3), np.uint8)
    L.gso_preprocess(C.c_int(P), C.c_int(sh_degree), C.c_int(M), _p(means3D), _p(scales), C.c_float(scale_modifier),
                     _p(rotations), _p(opacities), _p(shs), _p(cov3D_precomp), _p(colors_precomp), _p(V), _p(Pm), _p(cp),
                     C.c_int(W), C.c_int(H), C.c_float(cam["tanfovx"]), C.c_float(cam["tanfovy"]),
                     _p(radii), _p(xy), _p(depths), _p(cov3D), _p(rgb), _p(conic_opacity), _p(tiles), _p(clamped))
    gx, gy = (W + 15) // 16, (H + 15) // 16
    R = int(L.gso_bin(C.c_int(P), C.c_int(W), C.c_int(H), _p(xy), _p(depths), _p(radii), _p(tiles), None, None, None))
    point_list = np.zeros(max(R, 1), np.uint32); keys = np.zeros(max(R, 1), np.uint64)
    ranges = np.zeros((gx * gy, 2), np.uint32)
    L.gso_bin(C.c_int(P), C.c_int(W), C.c_int(H), _p(xy), _p(depths), _p(radii), _p(tiles), _p(point_list), _p(keys), _p(ranges))
    out = dict(radii=radii, means2D=xy, depths=depths, cov3D=cov3D, rgb=rgb, conic_opacity=conic_opacity,
               tiles_touched=tiles, clamped=clamped, num_rendered=R, point_list=point_list[:R], keys=keys[:R],
               ranges=ranges)
    if render:
        feats = colors_precomp if colors_precomp is not None else rgb
        bg = _f(np.asarray(bg))
        color = np.zeros((3, H, W), np.float32); final_T = np.zeros((H, W), np.float32); n_contrib = np.zeros((H, W), np.uint32)
        L.gso_render(C.c_int(W), C.c_int(H), _p(ranges), _p(point_list), _p(xy), _p(feats), _p(conic_opacity), _p(bg),
                     _p(color), _p(final_T), _p(n_contrib))
        out.update(color=color, final_T=final_T, n_contrib=n_contrib)
    return out


def backward(cam, fwd, dL_dpix, means3D, shs=None, colors_precomp=None, scales=None, rotations=None,
             cov3D_precomp=None, sh_degree=0, scale_modifier=1.0, bg=(0, 0, 0)):
    """Backward given the dict returned by forward().  Returns the reference's 8 gradients + dL_dconic."""
    L = lib()
    means3D = _f(means3D); P = means3D.shape[0]
    W, H = cam["W"], cam["H"]
    shs = _f(shs); colors_precomp = _f(colors_precomp); scales = _f(scales); rotations = _f(rotations)
    cov3D_precomp = _f(cov3D_precomp)
    M = 0 if shs is None else shs.shape[1]
    V, Pm, cp = _f(cam["viewmatrix"]), _f(cam["projmatrix"]), _f(cam["campos"])
    feats = colors_precomp if colors_precomp is not None else fwd["rgb"]
    cov3D = cov3D_precomp if cov3D_precomp is not None else fwd["cov3D"]
    bg = _f(np.asarray(bg)); dL_dpix = _f(dL_dpix)
    d_mean2D = np.zeros((P, 3), np.float32); d_conic = np.zeros((P, 4), np.float32)
    d_opac = np.zeros((P, 1), np.float32); d_col = np.zeros((P, 3), np.float32)
    pl = np.ascontiguousarray(fwd["point_list"]) if fwd["num_rendered"] else np.zeros(1, np.uint32)
    L.gso_render_backward(C.c_int(P), C.c_int(W), C.c_int(H), _p(fwd["ranges"]), _p(pl), _p(bg), _p(fwd["means2D"]),
                          _p(fwd["conic_opacity"]), _p(feats), _p(fwd["final_T"]), _p(fwd["n_contrib"]), _p(dL_dpix),
                          _p(d_mean2D), _p(d_conic), _p(d_opac), _p(d_col))
    d_mean3D = np.zeros((P, 3), np.float32); d_cov3D = np.zeros((P, 6), np.float32)
    d_sh = np.zeros((P, M, 3), np.float32); d_scale = np.zeros((P, 3), np.float32); d_rot = np.zeros((P, 4), np.float32)
    L.gso_preprocess_backward(C.c_int(P), C.c_int(sh_degree), C.c_int(M), _p(means3D), _p(fwd["radii"]), _p(shs),
                              _p(fwd["clamped"]), _p(scales), _p(rotations), C.c_float(scale_modifier), _p(_f(cov3D)),
                              _p(V), _p(Pm), C.c_int(W), C.c_int(H), C.c_float(cam["tanfovx"]), C.c_float(cam["tanfovy"]),
                              _p(cp), _p(d_mean2D), _p(d_conic), _p(d_mean3D), _p(d_col), _p(d_cov3D), _p(d_sh),
                              _p(d_scale), _p(d_rot))
    return dict(means2D=d_mean2D, colors_precomp=d_col, opacities=d_opac, means3D=d_mean3D, cov3D_precomp=d_cov3D,
                shs=d_sh, scales=d_scale, rotations=d_rot, conic=d_conic)


def sh_to_rgb(means3D, campos, shs, sh_degree):
    m = _f(means3D); s = _f(shs); P, M = s.shape[0], s.shape[1]
    rgb = np.zeros((P, 3), np.float32); cl = np.zeros((P, 3), np.uint8)
    lib().gso_sh_to_rgb(C.c_int(P), C.c_int(sh_degree), C.c_int(M), _p(m), _p(_f(campos)), _p(s), _p(rgb), _p(cl))
    return rgb, cl.astype(bool)


def cov3d(scales, rotations, scale_modifier=1.0):
    s = _f(scales); r = _f(rotations); P = s.shape[0]
    out = np.zeros((P, 6), np.float32)
    lib().gso_cov3d(C.c_int(P), _p(s), C.c_float(scale_modifier), _p(r), _p(out))
    return out
