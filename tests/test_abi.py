"""C-ABI library: loads without a GPU, exports every symbol include/gsr.h declares, and its host-side
logic (sizes, argument validation, error text) behaves.  No compute calls here."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "gaussian-splatting_cc-comments_b200", "libgsr_b200.so")
HDR = os.path.join(ROOT, "include", "gsr.h")


def declared_functions():
    src = open(HDR).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"\b(gsr_[a-z0-9_]+)\s*\(", src)
    typedef_fn = set(re.findall(r"\(\*\s*(gsr_[a-z0-9_]+)\s*\)", src))
    return sorted(set(names) - typedef_fn)


@pytest.fixture(scope="module")
def lib():
    return C.CDLL(LIB)


def test_header_declares_the_path():
    fns = declared_functions()
    for must in ("gsr_forward", "gsr_backward", "gsr_mark_visible", "gsr_required_bytes", "gsr_last_error"):
        assert must in fns


def test_every_declared_symbol_is_exported(lib):
    for name in declared_functions():
        assert hasattr(lib, name), "libgsr_b200.so does not export " + name


def test_version(lib):
    assert lib.gsr_version() == 1


def test_required_bytes_monotone_and_aligned(lib):
    lib.gsr_required_bytes.restype = C.c_size_t
    lib.gsr_required_bytes.argtypes = [C.c_int, C.c_size_t, C.c_size_t]
    prev = 0
    for n in (0, 1, 1000, 10**6, 3 * 10**6):
        b = lib.gsr_required_bytes(0, n, 0)
        assert b >= prev and b % 256 == 0
        prev = b
    # geometry state: 48 B record + sort/scan arrays; stays well under the reference's 79 B + CUB temp x2
    per = lib.gsr_required_bytes(0, 3 * 10**6, 0) / 3e6
    assert 48 <= per <= 100
    per_inst = lib.gsr_required_bytes(1, 12 * 10**6, 0) / 12e6
    assert 16 <= per_inst <= 23                      # 2 x (tile key + id) ping-pong = 16 B / instance + look-back tables
    img = lib.gsr_required_bytes(2, 1920 * 1080, 120 * 68)
    assert img >= 1920 * 1080 * 8 + 120 * 68 * 8
    assert lib.gsr_required_bytes(7, 10, 0) == 0


def test_scratch_bytes(lib):
    lib.gsr_backward_scratch_bytes.restype = C.c_size_t
    assert lib.gsr_backward_scratch_bytes(0) >= 0
    assert lib.gsr_backward_scratch_bytes(1000) >= 1000 * 48


def test_stage_names(lib):
    lib.gsr_stage_name.restype = C.c_char_p
    n = lib.gsr_num_stages()
    names = [lib.gsr_stage_name(i).decode() for i in range(n)]
    assert names[0] == "preprocess" and "render" in names and "render_bwd" in names and "preprocess_bwd" in names
    assert lib.gsr_stage_name(99) == b""


class FwdArgs(C.Structure):
    _fields_ = [("P", C.c_int), ("D", C.c_int), ("M", C.c_int), ("background", C.c_void_p), ("width", C.c_int),
                ("height", C.c_int), ("means3D", C.c_void_p), ("shs", C.c_void_p), ("colors_precomp", C.c_void_p),
                ("opacities", C.c_void_p), ("scales", C.c_void_p), ("scale_modifier", C.c_float),
                ("rotations", C.c_void_p), ("cov3D_precomp", C.c_void_p), ("viewmatrix", C.c_void_p),
                ("projmatrix", C.c_void_p), ("cam_pos", C.c_void_p), ("tan_fovx", C.c_float), ("tan_fovy", C.c_float),
                ("prefiltered", C.c_int), ("out_color", C.c_void_p), ("radii", C.c_void_p), ("debug", C.c_int),
                ("tile_row_begin", C.c_int), ("tile_row_end", C.c_int),
                ("raw_activations", C.c_int), ("features_dc", C.c_void_p), ("features_rest", C.c_void_p)]


class BwdArgs(C.Structure):
    _fields_ = [("P", C.c_int), ("D", C.c_int), ("M", C.c_int), ("R", C.c_int), ("background", C.c_void_p), ("width", C.c_int),
                ("height", C.c_int), ("means3D", C.c_void_p), ("shs", C.c_void_p), ("colors_precomp", C.c_void_p), ("scales", C.c_void_p),
                ("scale_modifier", C.c_float), ("rotations", C.c_void_p), ("cov3D_precomp", C.c_void_p), ("viewmatrix", C.c_void_p),
                ("projmatrix", C.c_void_p), ("cam_pos", C.c_void_p), ("tan_fovx", C.c_float), ("tan_fovy", C.c_float), ("radii", C.c_void_p),
                ("geom_buffer", C.c_void_p), ("binning_buffer", C.c_void_p), ("image_buffer", C.c_void_p), ("dL_dpix", C.c_void_p),
                ("dL_dmean2D", C.c_void_p), ("dL_dconic", C.c_void_p), ("dL_dopacity", C.c_void_p), ("dL_dcolor", C.c_void_p),
                ("dL_dmean3D", C.c_void_p), ("dL_dcov3D", C.c_void_p), ("dL_dsh", C.c_void_p), ("dL_dscale", C.c_void_p), ("dL_drot", C.c_void_p),
                ("grad_scratch", C.c_void_p), ("debug", C.c_int), ("tile_row_begin", C.c_int), ("tile_row_end", C.c_int),
                ("dL_drgb", C.c_void_p), ("skip_dL_dsh", C.c_int), ("drgb_ready_event", C.c_void_p), ("chunk_done_events", C.c_void_p),
                ("n_chunk_events", C.c_int), ("small_pack", C.c_void_p), ("raw_activations", C.c_int), ("features_dc", C.c_void_p),
                ("features_rest", C.c_void_p), ("dL_dfeatures_dc", C.c_void_p), ("dL_dfeatures_rest", C.c_void_p)]


RESIZE = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_size_t)


def _call_forward(lib, args):
    cb = RESIZE(lambda user, n: None)
    nr = C.c_int(-5)
    lib.gsr_last_error.restype = C.c_char_p
    rc = lib.gsr_forward(C.byref(args), cb, None, cb, None, cb, None, None, C.byref(nr))
    return rc, nr.value, lib.gsr_last_error().decode()


def test_forward_rejects_bad_arguments_without_touching_the_gpu(lib):
    a = FwdArgs()
    a.P, a.width, a.height = -1, 16, 16
    rc, nr, msg = _call_forward(lib, a)
    assert rc == 1 and "bad sizes" in msg and nr == 0
    a.P, a.width = 10, 0
    rc, _, _ = _call_forward(lib, a)
    assert rc == 1
    a.width = 16
    rc, _, msg = _call_forward(lib, a)           # required pointers missing
    assert rc == 1 and "pointer" in msg
    rc = lib.gsr_forward(None, None, None, None, None, None, None, None, None)
    assert rc == 1


def test_forward_with_no_gaussians_is_a_no_op(lib):
    """P == 0 returns before any launch (rasterize_points.cu:94)."""
    a = FwdArgs()
    a.P, a.width, a.height = 0, 32, 32
    rc, nr, _ = _call_forward(lib, a)
    assert rc == 0 and nr == 0


def test_mark_visible_argument_checks(lib):
    assert lib.gsr_mark_visible(0, None, None, None, None, None) == 0
    assert lib.gsr_mark_visible(-1, None, None, None, None, None) == 1
    assert lib.gsr_mark_visible(5, None, None, None, None, None) == 1


def test_backward_argument_checks(lib):
    assert lib.gsr_backward(None, None) == 1


def _fill(args, names, value=0x1000):
    for n in names:
        setattr(args, n, value)


def test_raw_parameter_path_argument_checks(lib):
    """Split SH storage and raw activations are validated on the host before any launch."""
    lib.gsr_last_error.restype = C.c_char_p
    a = FwdArgs()
    a.P, a.width, a.height, a.M, a.D = 10, 16, 16, 16, 3
    _fill(a, ["means3D", "opacities", "viewmatrix", "projmatrix", "cam_pos", "background", "out_color", "scales", "rotations"])
    a.features_dc = 0x1000                       # features_rest missing although M > 1
    rc, _, msg = _call_forward(lib, a)
    assert rc == 1 and "split SH" in msg
    a.features_rest = 0x1000
    a.shs = 0x1000                               # both layouts at once
    rc, _, msg = _call_forward(lib, a)
    assert rc == 1 and "split SH" in msg
    a.shs = None
    a.D = 4                                      # degree does not fit
    rc, _, msg = _call_forward(lib, a)
    assert rc == 1 and "degree" in msg

    b = BwdArgs()
    b.P, b.width, b.height, b.M, b.D, b.R = 10, 16, 16, 16, 3, 0
    _fill(b, ["means3D", "radii", "geom_buffer", "binning_buffer", "image_buffer", "dL_dpix", "viewmatrix", "projmatrix", "cam_pos",
              "background", "grad_scratch", "dL_dmean2D", "dL_dopacity", "dL_dcolor", "dL_dmean3D", "dL_dcov3D", "dL_dscale", "dL_drot",
              "scales", "rotations"])
    assert lib.gsr_backward(C.byref(b), None) == 1 and b"output pointer" in lib.gsr_last_error()      # M > 0 and no dL_dsh
    b.features_dc = b.features_rest = 0x1000
    assert lib.gsr_backward(C.byref(b), None) == 1 and b"split SH" in lib.gsr_last_error()            # gradients of the split tensors missing
    b.dL_dfeatures_dc = b.dL_dfeatures_rest = 0x1000
    b.M = 25
    assert lib.gsr_backward(C.byref(b), None) == 1 and b"split SH" in lib.gsr_last_error()            # M <= 16
    b.M = 16
    b.rotations = 0x1004
    assert lib.gsr_backward(C.byref(b), None) == 1 and b"aligned" in lib.gsr_last_error()
    b.rotations = 0x1000
    b.small_pack = 0x1000
    assert lib.gsr_backward(C.byref(b), None) == 1 and b"multi-view" in lib.gsr_last_error()
    b.small_pack = None
    b.features_dc = b.features_rest = b.dL_dfeatures_dc = b.dL_dfeatures_rest = None
    b.dL_dsh = 0x1000
    b.raw_activations = 1
    b.dL_drgb = 0x1000
    assert lib.gsr_backward(C.byref(b), None) == 1 and b"multi-view" in lib.gsr_last_error()


def test_backward_chunk_plan_covers_every_gaussian_once(lib):
    """gsr_backward_chunk_plan is host arithmetic: ranges are contiguous, aligned to 256 Gaussians and cover [0, P)."""
    b = (C.c_int * 5)()
    for P in (0, 1, 1000, 65535, 65536, 100000, 3000000, 6000001):
        n = lib.gsr_backward_chunk_plan(P, b)
        assert 1 <= n <= 4
        bounds = [b[i] for i in range(n + 1)]
        assert bounds[0] == 0 and bounds[-1] == P
        assert all(x <= y for x, y in zip(bounds, bounds[1:]))
        assert all(x % 256 == 0 for x in bounds[:-1])
        if P >= 65536:
            assert n == 4 and 0 < bounds[1] <= P // 8       # a small head range, then the rest
    assert lib.gsr_backward_chunk_plan(10, None) == 0


def test_exchange_entry_points_check_their_arguments(lib):
    lib.gsr_multimem_allreduce.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_int, C.c_int, C.c_void_p]
    assert lib.gsr_multimem_allreduce(None, 0, 16, 0, 2, None) == 1            # no multicast address
    assert lib.gsr_multimem_allreduce(C.c_void_p(0x1000), 2, 16, 0, 2, None) == 1   # range start not 16-byte aligned
    assert lib.gsr_multimem_allreduce(C.c_void_p(0x1000), 0, 16, 2, 2, None) == 1   # rank outside the world
    lib.gsr_multimem_broadcast.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    assert lib.gsr_multimem_broadcast(None, None, 0, None) == 0                    # nothing to send
    assert lib.gsr_multimem_broadcast(C.c_void_p(0x1000), C.c_void_p(0x2000), 6, None) == 1   # not a multiple of 4 floats
    assert lib.gsr_multimem_broadcast(C.c_void_p(0x1004), C.c_void_p(0x2000), 8, None) == 1   # destination not 16-byte aligned
    lib.gsr_unpack_small_grads.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 7
    assert lib.gsr_unpack_small_grads(0, 0, None, None, None, None, None, None, None) == 0    # empty range: no-op
    assert lib.gsr_unpack_small_grads(0, 8, None, None, None, None, None, None, None) == 1
    assert lib.gsr_unpack_small_grads(8, 0, None, None, None, None, None, None, None) == 1


class AdamGroup(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p), ("n", C.c_size_t),
                ("lr", C.c_double), ("step", C.c_longlong), ("row_width", C.c_int)]


def test_adam_and_knn_entry_points_check_their_arguments(lib):
    lib.gsr_adam_step.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_void_p]
    lib.gsr_last_error.restype = C.c_char_p
    assert lib.gsr_adam_step(None, 0, 0.9, 0.999, 1e-15, None, None) == 0                 # no groups: no-op
    assert lib.gsr_adam_step(None, 2, 0.9, 0.999, 1e-15, None, None) == 1
    g = (AdamGroup * 1)()
    g[0].n = 0
    assert lib.gsr_adam_step(g, 1, 0.9, 0.999, 1e-15, None, None) == 0                    # empty group: skipped, no launch
    assert lib.gsr_adam_step(g, 1, 1.0, 0.999, 1e-15, None, None) == 1                    # beta1 outside [0,1), as torch rejects it
    assert lib.gsr_adam_step(g, 1, 0.9, 0.999, -1.0, None, None) == 1
    g[0].n = 16
    assert lib.gsr_adam_step(g, 1, 0.9, 0.999, 1e-15, None, None) == 1                    # null tensors
    assert b"null tensor" in lib.gsr_last_error()
    g[0].param = g[0].grad = g[0].exp_avg = g[0].exp_avg_sq = 0x1000
    g[0].step = 0
    assert lib.gsr_adam_step(g, 1, 0.9, 0.999, 1e-15, None, None) == 1                    # step counts from 1
    assert b"step" in lib.gsr_last_error()
    g[0].step = 1
    g[0].lr = -1.0
    assert lib.gsr_adam_step(g, 1, 0.9, 0.999, 1e-15, None, None) == 1
    g[0].lr = 0.1
    g[0].row_width = 0
    assert lib.gsr_adam_step(g, 1, 0.9, 0.999, 1e-15, C.c_void_p(0x2000), None) == 1     # mask needs a row width

    lib.gsr_knn_scratch_bytes.restype = C.c_size_t
    lib.gsr_knn_scratch_bytes.argtypes = [C.c_int]
    assert lib.gsr_knn_scratch_bytes(0) == 0
    per = lib.gsr_knn_scratch_bytes(10**6) / 1e6
    assert 32 <= per <= 48                         # 2 x (code + index) ping-pong + 16 B sorted record + tables and boxes
    assert lib.gsr_knn_scratch_bytes(10**6) % 256 == 0
    lib.gsr_knn_mean_dist2.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    assert lib.gsr_knn_mean_dist2(0, None, None, None, None) == 0                         # nothing to do
    assert lib.gsr_knn_mean_dist2(-1, None, None, None, None) == 1
    assert lib.gsr_knn_mean_dist2(5, None, None, None, None) == 1
    assert lib.gsr_knn_mean_dist2(5, C.c_void_p(0x1000), C.c_void_p(0x2000), C.c_void_p(0x3010), None) == 1   # scratch alignment


class DensifyParams(C.Structure):
    _fields_ = [("max_grad", C.c_double), ("min_opacity", C.c_double), ("extent", C.c_double), ("percent_dense", C.c_double),
                ("max_screen_size", C.c_double), ("n_split_copies", C.c_int)]


class DensifyGroup(C.Structure):
    _fields_ = [("param", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p), ("param_out", C.c_void_p),
                ("exp_avg_out", C.c_void_p), ("exp_avg_sq_out", C.c_void_p), ("row_width", C.c_int)]


def densify_argtypes(lib):
    lib.gsr_densify_scratch_bytes.restype = C.c_size_t
    lib.gsr_densify_scratch_bytes.argtypes = [C.c_int]
    lib.gsr_densify_stats.argtypes = [C.c_int] + [C.c_void_p] * 6
    lib.gsr_densify_classify.argtypes = [C.c_int] + [C.c_void_p] * 9
    lib.gsr_densify_apply.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                      C.c_int, C.c_int, C.c_int, C.c_void_p]


def test_densify_entry_points_check_their_arguments(lib):
    densify_argtypes(lib)
    A = lambda x: C.c_void_p(x)
    assert lib.gsr_densify_scratch_bytes(0) == 0 and lib.gsr_densify_scratch_bytes(1000) >= 4 * 4 * 4
    assert lib.gsr_densify_scratch_bytes(1000) % 256 == 0
    assert lib.gsr_densify_stats(0, None, None, None, None, None, None) == 0
    assert lib.gsr_densify_stats(-1, None, None, None, None, None, None) == 1
    assert lib.gsr_densify_stats(4, A(0x100), None, A(0x100), A(0x100), None, None) == 1          # radii missing
    ok = DensifyParams(0.0002, 0.005, 3.0, 0.01, 20.0, 2)
    assert lib.gsr_densify_classify(5, A(0x100), A(0x100), A(0x100), A(0x100), None, A(0x100), A(0x100), A(0x100), None) == 1
    bad = DensifyParams(0.0002, 0.005, 3.0, 0.01, 20.0, 0)
    assert lib.gsr_densify_classify(5, A(0x100), A(0x100), A(0x100), A(0x100), C.byref(bad), A(0x100), A(0x100), A(0x100), None) == 1
    assert b"n_split_copies" in lib.gsr_last_error()
    assert lib.gsr_densify_classify(5, A(0x100), A(0x100), A(0x100), A(0x100), C.byref(ok), A(0x100), A(0x100), A(0x110), None) == 1
    assert b"aligned" in lib.gsr_last_error()
    assert lib.gsr_densify_classify(5, None, A(0x100), A(0x100), A(0x100), C.byref(ok), A(0x100), A(0x100), A(0x100), None) == 1
    # apply: totals must be consistent with one another and with P; roles must point at groups of the right width
    groups = (DensifyGroup * 3)()
    for g, w in zip(groups, (3, 3, 4)):
        g.param, g.param_out, g.row_width = 0x1000, 0x2000, w
    tot = (C.c_int * 5)(7, 1, 4, 1, 1)
    args = lambda t=tot, grp=groups, n=3, roles=(0, 1, 2), smp=A(0x3000): (10, A(0x100), A(0x100), t, C.byref(ok), smp, grp, n, *roles, None)
    assert lib.gsr_densify_apply(0, None, None, None, C.byref(ok), None, None, 0, 0, 0, 0, None) == 0
    assert lib.gsr_densify_apply(*args(t=(C.c_int * 5)(8, 1, 4, 1, 1))) == 1 and b"totals" in lib.gsr_last_error()
    assert lib.gsr_densify_apply(*args(t=(C.c_int * 5)(7, 0, 4, 1, 1))) == 1                      # kept children without split parents
    assert lib.gsr_densify_apply(*args(roles=(0, 1, 1))) == 1 and b"row widths" in lib.gsr_last_error()
    assert lib.gsr_densify_apply(*args(roles=(0, 0, 2))) == 1
    assert lib.gsr_densify_apply(*args(smp=None)) == 1 and b"samples" in lib.gsr_last_error()
    assert lib.gsr_densify_apply(*args(n=9)) == 1
    groups[1].exp_avg = 0x4000                                                                      # a moment without its pair
    assert lib.gsr_densify_apply(*args()) == 1 and b"pairs" in lib.gsr_last_error()
    groups[1].exp_avg = None
    groups[2].param_out = groups[2].param                                                           # in place is not supported
    assert lib.gsr_densify_apply(*args()) == 1 and b"alias" in lib.gsr_last_error()


def test_launch_counter_starts_at_zero_on_cpu(lib):
    lib.gsr_launch_count.restype = C.c_ulonglong
    assert lib.gsr_launch_count() == 0


def test_library_has_no_cpu_fallback_symbols():
    """The product library must not link the oracle."""
    import subprocess
    out = subprocess.run(["nm", "-D", LIB], capture_output=True, text=True).stdout
    assert "gso_" not in out and "gno_" not in out and "emu_" not in out
    ext = os.path.join(ROOT, "gaussian-splatting_cc-comments_b200", "diff_gaussian_rasterization", "_C.so")
    out = subprocess.run(["nm", "-D", ext], capture_output=True, text=True).stdout
    assert "gso_" not in out and "gno_" not in out
