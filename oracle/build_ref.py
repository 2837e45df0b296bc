#!/usr/bin/env python3
"""Build the UNMODIFIED reference rasterizer (DGR) for sm_100 into oracle/_ref/.

TEST / BENCH INFRASTRUCTURE ONLY -- nothing in the product path imports this.

The reference's CUDA sources are compiled where they lie under
/root/reference/submodules/diff-gaussian-rasterization (no copy of the sources
enters this repository; oracle/_ref/ is git-ignored but travels to the GPU box
with the snapshot).  Flags are the stock torch.utils.cpp_extension flags with
TORCH_CUDA_ARCH_LIST=10.0 plus `-include cstdint` (rasterizer_impl.h uses
std::uintptr_t without including <cstdint>, which gcc 13 no longer forgives).
No -use_fast_math, matching the reference's setup.py:29.

Result:  oracle/_ref/diff_gaussian_rasterization/{__init__.py,_C.so}
         (the same layout `pip install --target` would give).

build_knn() does the same for the reference's simple-knn extension (submodules/simple-knn: spatial.cu, simple_knn.cu,
ext.cpp; stock flags, `-include cfloat` for FLT_MAX under gcc 13) -> oracle/_ref/simple_knn/ref_simple_knn_C.so, the checker of
gsr_knn_mean_dist2 on the GPU box.
"""
import os, shutil, sys, glob

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/submodules/diff-gaussian-rasterization"
OUT = os.path.join(HERE, "_ref")
PKG = os.path.join(OUT, "diff_gaussian_rasterization")


def build(verbose=False):
    if not os.path.isdir(REF):
        return False  # GPU box: only the prebuilt files exist
    os.makedirs(PKG, exist_ok=True)
    so = os.path.join(PKG, "_C.so")
    srcs = [os.path.join(REF, p) for p in (
        "cuda_rasterizer/rasterizer_impl.cu", "cuda_rasterizer/forward.cu",
        "cuda_rasterizer/backward.cu", "rasterize_points.cu", "ext.cpp")]
    newest = max(os.path.getmtime(s) for s in srcs)
    if os.path.exists(so) and os.path.getmtime(so) >= newest:
        return True
    os.environ["TORCH_CUDA_ARCH_LIST"] = "10.0"
    os.environ.setdefault("MAX_JOBS", "8")
    from torch.utils.cpp_extension import load
    bdir = os.path.join(OUT, "build")
    os.makedirs(bdir, exist_ok=True)
    load(name="_C", sources=srcs,
         extra_include_paths=[os.path.join(REF, "third_party/glm")],
         extra_cflags=["-O3"], extra_cuda_cflags=["-include", "cstdint"],
         build_directory=bdir, verbose=verbose, is_python_module=False)
    built = glob.glob(os.path.join(bdir, "_C*.so"))
    assert built, "reference build produced no .so"
    shutil.copy2(built[0], so)
    # the reference's Python wrapper, installed next to its extension (pip-install equivalent)
    shutil.copy2(os.path.join(REF, "diff_gaussian_rasterization/__init__.py"),
                 os.path.join(PKG, "__init__.py"))
    return True


REF_KNN = "/root/reference/submodules/simple-knn"
KNN_MODULE = "ref_simple_knn_C"
PKG_KNN = os.path.join(OUT, "simple_knn")


def build_knn(verbose=False):
    if not os.path.isdir(REF_KNN):
        return False  # GPU box: only the prebuilt files exist
    os.makedirs(PKG_KNN, exist_ok=True)
    so = os.path.join(PKG_KNN, KNN_MODULE + ".so")
    srcs = [os.path.join(REF_KNN, p) for p in ("spatial.cu", "simple_knn.cu", "ext.cpp")]
    if os.path.exists(so) and os.path.getmtime(so) >= max(os.path.getmtime(s) for s in srcs):
        return True
    os.environ["TORCH_CUDA_ARCH_LIST"] = "10.0"
    os.environ.setdefault("MAX_JOBS", "8")
    from torch.utils.cpp_extension import load
    bdir = os.path.join(OUT, "build_knn")
    os.makedirs(bdir, exist_ok=True)
    # Its own module name: torch renames a second load(name="_C") in one process to "_C_v1", and the file then exports
    # PyInit__C_v1 while being imported as "_C" (round 1's ten XFAILs).  The exported init symbol is checked below.
    load(name=KNN_MODULE, sources=srcs, extra_cflags=["-O3"], extra_cuda_cflags=["-include", "cfloat", "-include", "cstdint"],
         build_directory=bdir, verbose=verbose, is_python_module=False)
    built = glob.glob(os.path.join(bdir, KNN_MODULE + "*.so"))
    assert built, "reference simple-knn build produced no .so"
    shutil.copy2(built[0], so)
    import subprocess
    syms = subprocess.run(["nm", "-D", "--defined-only", so], capture_output=True, text=True).stdout
    assert ("PyInit_" + KNN_MODULE) in syms.split(), "reference simple-knn module does not export PyInit_" + KNN_MODULE
    with open(os.path.join(PKG_KNN, "__init__.py"), "w"):
        pass   # the reference's package directory holds only the compiled module (submodules/simple-knn/simple_knn/.gitkeep)
    return True


if __name__ == "__main__":
    ok = build(verbose="-v" in sys.argv)
    print("reference build:", "ok" if ok else "skipped (no /root/reference)")
    ok = build_knn(verbose="-v" in sys.argv)
    print("reference simple-knn build:", "ok" if ok else "skipped (no /root/reference)")
