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
