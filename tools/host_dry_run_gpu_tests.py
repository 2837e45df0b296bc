#!/usr/bin/env python3
"""Dry run, on the HOST, of the Python side of the `first_gpu_run` GPU tests (tests/test_zz_next_rows_gpu.py).

TEST INFRASTRUCTURE, not a product path and not a parity claim: those tests exist to run the real kernels on a B200.
This script only makes sure their Python (fixtures, model set-up, optimizer surgery, comparisons, launch counting) is
free of bugs before GPU minutes are spent on them: inside this process the device name "cuda" is mapped to the CPU in
torch's factory functions, and the torch extension's entry points for the Adam step, distCUDA2 and densification are replaced
by ctypes doubles over the host-emulated C ABI (tests/cuda_emu/emu_abi.cpp -- the product's own kernel sources run by
emu.h).  The rasterizer-facing tests (raw-parameter path) and distCUDA2's device checks need the real binding and are
not covered.

    python tools/host_dry_run_gpu_tests.py [P of the torch-sequence comparison, default 20000]
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "gaussian-splatting_cc-comments_b200"), os.path.join(ROOT, "tests"),
                os.path.join(ROOT, "tests", "cuda_emu"), os.path.join(ROOT, "tools")]
import torch          # noqa: E402
import torch.optim    # noqa: E402


def _to_cpu(fn):
    def inner(*a, **k):
        if str(k.get("device", "")).startswith("cuda"):
            k["device"] = "cpu"
        return fn(*a, **k)
    return inner


def install_device_shim():
    for name in ("zeros", "ones", "empty", "tensor", "rand", "randn", "randint", "full"):
        setattr(torch, name, _to_cpu(getattr(torch, name)))
    as_tensor = _to_cpu(torch.as_tensor)
    torch.as_tensor = lambda *a, **k: as_tensor(*a, **k).clone()      # a device copy never aliases the host array
    torch.Tensor.cuda = lambda self, *a, **k: self.clone()
    torch.Tensor.is_cuda = property(lambda self: True)                 # the drop-ins' own "CUDA only" guards (dry run only)
    real = torch.Generator

    class HostGenerator(real):
        def __new__(cls, device="cpu"):
            return real("cpu")
    torch.Generator = HostGenerator


def main():
    import build_emu
    import c3_trainer  # noqa: F401  (imported before the shim replaces torch.Generator)
    lib = C.CDLL(build_emu.build("emu_abi"))
    lib.gsr_last_error.restype = C.c_char_p
    lib.gsr_launch_count.restype = C.c_ulonglong
    install_device_shim()

    import test_densify_cpu as TD
    from test_abi import AdamGroup
    TD.densify_argtypes(lib)
    lib.gsr_adam_step.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_void_p]

    def adam_step(params, grads, exp_avgs, exp_avg_sqs, lrs, steps, beta1, beta2, eps, visible=None):
        groups = (AdamGroup * len(params))()
        for g, p, gr, m, v, lr, st in zip(groups, params, grads, exp_avgs, exp_avg_sqs, lrs, steps):
            g.param, g.grad, g.exp_avg, g.exp_avg_sq = p.data_ptr(), gr.data_ptr(), m.data_ptr(), v.data_ptr()
            g.n, g.lr, g.step = p.numel(), lr, st
            g.row_width = p.numel() // visible.numel() if visible is not None and visible.numel() else 1
        vis = visible.to(torch.uint8).contiguous() if visible is not None else None
        if lib.gsr_adam_step(groups, len(params), beta1, beta2, eps, C.c_void_p(vis.data_ptr()) if vis is not None else None, None) != 0:
            raise RuntimeError(lib.gsr_last_error().decode())

    lib.gsr_knn_scratch_bytes.restype = C.c_size_t
    lib.gsr_knn_mean_dist2.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]

    def dist_cuda2(points):
        pts = points.contiguous()
        P = pts.shape[0]
        out = torch.zeros(P)
        raw = torch.zeros(lib.gsr_knn_scratch_bytes(P) + 512, dtype=torch.uint8)
        scratch = raw.data_ptr() + (-raw.data_ptr()) % 256
        if P and lib.gsr_knn_mean_dist2(P, C.c_void_p(pts.data_ptr()), C.c_void_p(out.data_ptr()), C.c_void_p(scratch), None) != 0:
            raise RuntimeError(lib.gsr_last_error().decode())
        return out

    from diff_gaussian_rasterization import _C
    import gsr_densify
    _C.adam_step = adam_step
    _C.dist_cuda2 = dist_cuda2
    gsr_densify._C = TD._EmulatedExtension(lib)
    import test_zz_next_rows_gpu as G
    G._launches = lambda: int(lib.gsr_launch_count())
    G._ref_knn = lambda: None                                          # the reference's simple-knn build is CUDA only

    class MonkeyPatch:
        def __init__(self): self.saved = []
        def setattr(self, obj, name, val): self.saved.append((obj, name, getattr(obj, name))); setattr(obj, name, val)
        def undo(self):
            for obj, name, val in reversed(self.saved): setattr(obj, name, val)

    def run(name, *args):
        mp = MonkeyPatch()
        fn = getattr(G, name)
        try:
            fn(*args, mp) if "monkeypatch" in fn.__code__.co_varnames[:fn.__code__.co_argcount] else fn(*args)
        finally:
            mp.undo()
        print("ok ", name, args, flush=True)

    run("test_fused_adam_follows_the_golden_trajectory")
    for P in (1, 1000, 20_003):
        run("test_fused_adam_equals_torch_adam_on_the_gpu_and_the_oracle", P)
    run("test_fused_adam_unaligned_views_and_visibility_mask")
    for P, seed in ((1, 0), (2, 1), (3, 2), (4, 3), (255, 4), (256, 5), (257, 6), (5000, 7)):
        run("test_dist_cuda2_is_bit_exact_against_the_oracle_and_the_reference_build", P, seed)
    for tag in "abc":
        run("test_densify_and_prune_reproduces_the_reference_models_fixture", tag)
    run("test_densify_and_prune_equals_the_torch_sequence_on_the_gpu_with_the_same_seed", int(sys.argv[1]) if len(sys.argv) > 1 else 20000, 20)
    run("test_densification_statistics_kernel_against_torch_and_the_fixture")
    print("dry run complete: the Python side of these GPU tests is consistent (the kernels themselves still need the B200)")


if __name__ == "__main__":
    main()
