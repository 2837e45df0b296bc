#!/usr/bin/env python3
"""In-tree build of the sm_100a rasterizer.

  libgsr_b200.so                         -- the C-ABI library (include/gsr.h), pure CUDA runtime, no torch
  diff_gaussian_rasterization/_C.so      -- the torch extension module over it (same surface as DGR/ext.cpp)

Both are built with explicit nvcc / g++ commands (no JIT cache: the .so files travel with the
repository snapshot to the GPU box).  Nothing is rebuilt when the outputs are newer than the sources.
"""
import os, subprocess, sys, sysconfig, glob
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libgsr_b200.so")
EXT = os.path.join(HERE, "diff_gaussian_rasterization", "_C.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
CU = ["gsr_api.cu", "gsr_preprocess.cu", "gsr_binning.cu", "gsr_render.cu", "gsr_render_bwd.cu", "gsr_preprocess_bwd.cu", "gsr_exchange.cu", "gsr_loss.cu", "gsr_adam.cu", "gsr_knn.cu", "gsr_densify.cu"]


def _newer(out, deps):
    return os.path.exists(out) and all(os.path.getmtime(out) >= os.path.getmtime(d) for d in deps)


def _run(cmd, verbose):
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("build step failed: " + " ".join(cmd))
    if verbose and r.stderr.strip():
        print(r.stderr)


def build(verbose=False, force=False):
    os.makedirs(OBJ, exist_ok=True)
    headers = glob.glob(os.path.join(CSRC, "*.cuh")) + [os.path.join(ROOT, "include", "gsr.h")]
    jobs = []
    objs = []
    for f in CU:
        src = os.path.join(CSRC, f)
        obj = os.path.join(OBJ, f.replace(".cu", ".o"))
        objs.append(obj)
        if force or not _newer(obj, [src] + headers):
            jobs.append([NVCC, *ARCH, "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-c", src, "-o", obj])
    with ThreadPoolExecutor(max_workers=8) as ex:
        list(ex.map(lambda c: _run(c, verbose), jobs))
    if force or jobs or not _newer(LIB, objs):
        _run([NVCC, *ARCH, "-shared", "-o", LIB, *objs, "-cudart", "shared"], verbose)

    bsrc = os.path.join(CSRC, "torch_binding.cpp")
    if force or not _newer(EXT, [bsrc, LIB] + headers):
        import torch
        from torch.utils.cpp_extension import include_paths, library_paths
        inc = include_paths("cuda") if "cuda" in include_paths.__code__.co_varnames[:1] else include_paths()
        try:
            inc = include_paths("cuda")
        except TypeError:
            inc = include_paths(cuda=True)
        try:
            libp = library_paths("cuda")
        except TypeError:
            libp = library_paths(cuda=True)
        cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-DTORCH_EXTENSION_NAME=_C", "-DTORCH_API_INCLUDE_EXTENSION_H",
               "-D_GLIBCXX_USE_CXX11_ABI=%d" % int(torch._C._GLIBCXX_USE_CXX11_ABI)]
        cmd += ["-I" + p for p in inc] + ["-I" + sysconfig.get_paths()["include"], "-I/usr/local/cuda/include"]
        cmd += [bsrc, "-o", EXT, "-L" + HERE, "-lgsr_b200", "-Wl,-rpath,$ORIGIN/.."]
        cmd += ["-L" + p for p in libp] + ["-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python",
                                           "-L/usr/local/cuda/lib64", "-lcudart"]
        _run(cmd, verbose)
    return LIB, EXT


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
