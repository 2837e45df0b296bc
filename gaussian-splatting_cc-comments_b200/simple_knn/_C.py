"""`simple_knn._C` of the reference (submodules/simple-knn/ext.cpp:15-17) on this library's kernels.

    distCUDA2(points)  ->  (P,) float32: mean squared distance from every point to its three nearest other points

Same name, argument and result as the reference's pybind function (spatial.cu:15-26); used by
GaussianModel.create_from_pcd as `torch.clamp_min(distCUDA2(points), 0.0000001)` (scene/gaussian_model.py:207).
CUDA float32 only: there is no CPU path and nothing falls back to torch ops.
"""
from diff_gaussian_rasterization import _C as _gsr


def distCUDA2(points):
    return _gsr.dist_cuda2(points)
