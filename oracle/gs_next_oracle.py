"""ctypes/numpy front end of oracle/gs_next_oracle.c (Adam step, 3-NN mean squared distance) -- TEST INFRASTRUCTURE ONLY.

Only tests/ and the cpu_baseline leg of the bench tools may import this module.  The product path never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "gs_next_oracle.c")
_SO = os.path.join(_HERE, "libgs_next_oracle.so")
_lib = None


def build(force=False):
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(_SRC):
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-mfma", "-fopenmp", "-fPIC", "-shared", "-o", _SO, _SRC, "-lm"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.gno_adam_step.restype = None
        _lib.gno_adam_step.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_double, C.c_double,
                                       C.c_double, C.c_double, C.c_longlong]
        _lib.gno_knn_mean_d2.restype = C.c_int
        _lib.gno_knn_mean_d2.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
    return _lib


def adam_step(param, grad, exp_avg, exp_avg_sq, lr, step, beta1=0.9, beta2=0.999, eps=1e-15):
    """In-place torch.optim.Adam update of float32 numpy arrays (param, exp_avg, exp_avg_sq are modified); `step` is the
    count after this update (>= 1)."""
    for a in (param, exp_avg, exp_avg_sq):
        assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"] and a.flags["WRITEABLE"]
    g = np.ascontiguousarray(grad, dtype=np.float32)
    assert g.size == param.size == exp_avg.size == exp_avg_sq.size
    lib().gno_adam_step(param.ctypes.data, g.ctypes.data, exp_avg.ctypes.data, exp_avg_sq.ctypes.data, param.size, float(lr),
                        float(beta1), float(beta2), float(eps), int(step))


def knn_mean_dist2(points):
    """distCUDA2: (P,3) float32 -> (P,) mean squared distance to the three nearest other points."""
    p = np.ascontiguousarray(points, dtype=np.float32)
    assert p.ndim == 2 and p.shape[1] == 3
    out = np.empty(p.shape[0], np.float32)
    if lib().gno_knn_mean_d2(p.shape[0], p.ctypes.data, out.ctypes.data) != 0:
        raise MemoryError("gno_knn_mean_d2")
    return out


# ---- densification (gno_densify_*): the reference's clone / split / prune rounds on numpy arrays ---------------------------
GROUP_ORDER = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")     # optimizer group order, scene/gaussian_model.py:243-250


def densify_masks(accum, denom, scaling, max_grad, extent, percent_dense):
    """(clone_mask, split_mask) over the P original Gaussians, as densify_and_clone / densify_and_split select them."""
    a, d, s = (np.ascontiguousarray(x, dtype=np.float32) for x in (accum, denom, scaling))
    P = s.shape[0]
    cm, sm = np.zeros(P, np.uint8), np.zeros(P, np.uint8)
    L = lib()
    L.gno_densify_masks.restype = C.c_int
    L.gno_densify_masks.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_void_p]
    L.gno_densify_masks(P, a.ctypes.data, d.ctypes.data, s.ctypes.data, float(max_grad), float(extent), float(percent_dense),
                        cm.ctypes.data, sm.ctypes.data)
    return cm.astype(bool), sm.astype(bool)


def densify_and_prune(params, exp_avg, exp_avg_sq, accum, denom, max_grad, min_opacity, extent, percent_dense, max_screen_size,
                      samples, N=2):
    """params / exp_avg / exp_avg_sq: dicts name -> (P, ...) float32 over GROUP_ORDER (exp_avg, exp_avg_sq may be None).
    samples: (N * n_split, 3).  Returns (params', exp_avg', exp_avg_sq') with the reference's row order."""
    names = [n for n in GROUP_ORDER if n in params]
    P = params["xyz"].shape[0]
    flat = lambda a: np.ascontiguousarray(a, dtype=np.float32).reshape(P, -1) if P else np.zeros((0, int(np.prod(a.shape[1:]))), np.float32)
    p = [flat(params[n]) for n in names]
    widths = (C.c_int * len(names))(*[a.shape[1] for a in p])
    has_state = exp_avg is not None
    m = [flat(exp_avg[n]) for n in names] if has_state else None
    v = [flat(exp_avg_sq[n]) for n in names] if has_state else None
    a, d = (np.ascontiguousarray(x, dtype=np.float32).reshape(-1) for x in (accum, denom))
    cm, sm = densify_masks(a, d, p[names.index("scaling")], max_grad, extent, percent_dense)
    smp = np.ascontiguousarray(samples, dtype=np.float32).reshape(-1, 3)
    assert smp.shape[0] == N * int(sm.sum()), (smp.shape, int(sm.sum()))
    cap = P + int(cm.sum()) + N * int(sm.sum()) + 1
    ptrs = lambda arrs: (C.c_void_p * len(arrs))(*[x.ctypes.data for x in arrs])
    po = [np.zeros((cap, x.shape[1]), np.float32) for x in p]
    mo = [np.zeros((cap, x.shape[1]), np.float32) for x in p] if has_state else None
    vo = [np.zeros((cap, x.shape[1]), np.float32) for x in p] if has_state else None
    L = lib()
    L.gno_densify_and_prune.restype = C.c_int
    L.gno_densify_and_prune.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int, C.c_void_p,
                                        C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    n = L.gno_densify_and_prune(P, len(names), widths, ptrs(p), ptrs(m) if has_state else None, ptrs(v) if has_state else None,
                                a.ctypes.data, d.ctypes.data, float(max_grad), float(min_opacity), float(extent), float(percent_dense),
                                float(max_screen_size or 0.0), int(N), smp.ctypes.data, names.index("xyz"), names.index("scaling"),
                                names.index("rotation"), names.index("opacity"), cap, ptrs(po), ptrs(mo) if has_state else None,
                                ptrs(vo) if has_state else None)
    assert n >= 0
    shape = lambda name, x: x[:n].reshape((n,) + tuple(params[name].shape[1:]))
    out = lambda arrs: {nm: shape(nm, x) for nm, x in zip(names, arrs)}
    return out(po), (out(mo) if has_state else None), (out(vo) if has_state else None)


def densify_stats(vs_grad, radii, accum, denom, max_radii2D=None):
    """In place on float32 arrays accum, denom, max_radii2D (P,)."""
    g = np.ascontiguousarray(vs_grad, dtype=np.float32)
    r = np.ascontiguousarray(radii, dtype=np.int32)
    for x in (accum, denom) + ((max_radii2D,) if max_radii2D is not None else ()):
        assert x.dtype == np.float32 and x.flags["C_CONTIGUOUS"]
    L = lib()
    L.gno_densify_stats.restype = None
    L.gno_densify_stats.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.gno_densify_stats(r.shape[0], g.ctypes.data, r.ctypes.data, accum.ctypes.data, denom.ctypes.data,
                        max_radii2D.ctypes.data if max_radii2D is not None else None)
