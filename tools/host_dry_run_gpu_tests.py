#!/usr/bin/env python3
"""Dry run, on the HOST, of the Python side of the `first_gpu_run` GPU tests (tests/test_zz_next_rows_gpu.py).

TEST INFRASTRUCTURE, not a product path and not a parity claim: those tests exist to run the real kernels on a B200.
This script only makes sure their Python (fixtures, model set-up, optimizer surgery, comparisons, launch counting) is
free of bugs before GPU minutes are spent on them: inside this process the device name "cuda" is mapped to the CPU in
torch's factory functions, and `diff_gaussian_rasterization._C` is a HOST build of the product's own binding
(csrc/torch_binding.cpp with the device guard / current stream as no-ops and `is_cuda` read as `is_cpu`,
tests/cuda_emu/build_emu.py: build_binding) linked against the host-emulated C ABI (tests/cuda_emu/emu_abi.cpp -- the
product's own kernel sources run by emu.h).  So the Python API, the binding's tensor / arena arithmetic, the C ABI's
orchestration and the kernels' logic all run; what does not run is the hardware.  Tests whose point is a device check
(host tensors must be rejected) or a size that the emulation cannot finish in minutes are left out.

    python tools/host_dry_run_gpu_tests.py --scripts --smoke   (__graft_entry__.smoke() only, one minute)
    python tools/host_dry_run_gpu_tests.py --scripts      (bench_next.py and train_c3.py --impl b200 at toy sizes)
    python tools/host_dry_run_gpu_tests.py [P of the torch-sequence comparison, default 20000] [substring of the test names to run]
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
    import importlib.util
    import build_emu
    import c3_trainer  # noqa: F401  (imported before the shim replaces torch.Generator)
    spec = importlib.util.spec_from_file_location("_C_host", build_emu.build_binding())
    host = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(host)
    sys.modules["diff_gaussian_rasterization._C"] = host          # every `from diff_gaussian_rasterization import _C` gets it
    lib = C.CDLL(build_emu.build("emu_abi"))
    lib.gsr_launch_count.restype = C.c_ulonglong
    install_device_shim()

    import diff_gaussian_rasterization
    assert diff_gaussian_rasterization._C is host
    import test_zz_next_rows_gpu as G
    G._launches = lambda: int(lib.gsr_launch_count())
    G._ref_knn = lambda: None                                      # the reference's simple-knn build is CUDA only

    class MonkeyPatch:
        def __init__(self): self.saved = []
        def setattr(self, obj, name, val): self.saved.append((obj, name, getattr(obj, name))); setattr(obj, name, val)
        def undo(self):
            for obj, name, val in reversed(self.saved): setattr(obj, name, val)

    count = [0]

    only = sys.argv[2] if len(sys.argv) > 2 else ""

    def run(name, *args):
        if only not in name:
            return
        mp = MonkeyPatch()
        fn = getattr(G, name)
        try:
            fn(*args, mp) if "monkeypatch" in fn.__code__.co_varnames[:fn.__code__.co_argcount] else fn(*args)
        finally:
            mp.undo()
        count[0] += 1
        print("ok ", name, args, flush=True)

    run("test_fused_adam_follows_the_golden_trajectory")
    for P in (1, 1000, 20_003):
        run("test_fused_adam_equals_torch_adam_on_the_gpu_and_the_oracle", P)
    run("test_fused_adam_unaligned_views_and_visibility_mask")
    for P, seed in ((1, 0), (2, 1), (3, 2), (4, 3), (255, 4), (256, 5), (257, 6), (5000, 7)):
        run("test_dist_cuda2_is_bit_exact_against_the_oracle_and_the_reference_build", P, seed)
    for P, deg, M in ((1, 3, 16), (777, 1, 16), (300, 0, 16), (2000, 2, 9), (900, 0, 1), (129, 1, 4)):
        run("test_raw_parameter_path_equals_torch_activations_plus_the_rasterizer", P, deg, M)
    run("test_fused_render_is_a_drop_in_for_the_reference_render")
    for tag in "abc":
        run("test_densify_and_prune_reproduces_the_reference_models_fixture", tag)
    run("test_densify_and_prune_equals_the_torch_sequence_on_the_gpu_with_the_same_seed", int(sys.argv[1]) if len(sys.argv) > 1 else 20000, 20)
    run("test_densification_statistics_kernel_against_torch_and_the_fixture")
    print("dry run complete: %d test invocations; the Python side, the binding and the kernels' logic are consistent "
          "(the kernels themselves still need the B200)" % count[0])


def dry_run_scripts():
    """The measurement scripts of round 2's first GPU call (tools/first_gpu_run_next_rows.sh) at toy sizes: bench_next.py (Adam,
    distCUDA2, densification) and the b200 arm of train_c3.py (render -> fused loss -> backward -> statistics -> densify ->
    FusedAdam, for a few iterations).  The printed numbers mean nothing; the point is that every line of them executes."""
    import importlib.util
    import time
    import build_emu
    import c3_trainer  # noqa: F401
    spec = importlib.util.spec_from_file_location("_C_host", build_emu.build_binding())
    host = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(host)
    sys.modules["diff_gaussian_rasterization._C"] = host
    install_device_shim()

    class HostEvent:
        def __init__(self, enable_timing=False): self.t = 0.0
        def record(self, stream=None): self.t = time.time()
        def elapsed_time(self, other): return (other.t - self.t) * 1e3
        def synchronize(self): pass
    torch.cuda.Event = HostEvent
    torch.cuda.synchronize = lambda *a, **k: None
    torch.cuda.is_available = lambda: True
    torch.cuda.is_current_stream_capturing = lambda: False
    sys.modules["torch.cuda.graphs"].is_current_stream_capturing = lambda: False

    if "--smoke" in sys.argv:
        import __graft_entry__
        __graft_entry__.smoke()                                    # the driver's round-end smoke test, same code path
        print("ok  __graft_entry__.smoke()", flush=True)
        return
    import bench_next
    bench_next.ref_knn_module = lambda: None                      # the reference's simple-knn build is CUDA only
    sys.argv = ["bench_next.py", "all", "--P", "3000", "--steps", "2", "--warmup", "1"]
    bench_next.main()
    print("ok  bench_next.py all", flush=True)
    import train_c3
    sys.argv = ["train_c3.py", "--impl", "b200", "--P0", "400", "--target", "900", "--iterations", "25", "--densify-from", "5",
                "--densify-every", "10", "--width", "64", "--height", "48", "--views", "2"]
    train_c3.main()
    print("ok  train_c3.py --impl b200", flush=True)
    print("script dry run complete")


if __name__ == "__main__":
    dry_run_scripts() if "--scripts" in sys.argv else main()
