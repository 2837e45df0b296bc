"""Edges of the hot path that the parity cases do not reach ("untested edges"):

* `prefiltered=True` (reference: forward.cu:204-214 via auxiliary.h:144-175): same result as False when every Gaussian passes,
  and the trap when one does not;
* the `debug=True` FAILURE paths of the Python layer (DGR/diff_gaussian_rasterization/__init__.py:87-94,141-148): a failing
  forward writes snapshot_fw.dump, a failing backward writes snapshot_bw.dump, both re-raise;
* the C ABI driven by a host that is NOT torch: gsr_forward / gsr_backward called through ctypes with memory from cudaMalloc
  and a plain C callback for the three growable buffers -- the binding INTEGRATION.md section 2 describes.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

import helpers as Hh
from helpers import S, ROOT

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mod():
    return Hh.my_module()


def _settings(mod, cam, t, sh_degree=3, prefiltered=False, debug=False):
    return mod.GaussianRasterizationSettings(
        image_height=cam["H"], image_width=cam["W"], tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], bg=t([0.1, 0.2, 0.3]),
        scale_modifier=1.0, viewmatrix=t(cam["viewmatrix"]), projmatrix=t(cam["projmatrix"]), sh_degree=sh_degree,
        campos=t(cam["campos"]), prefiltered=prefiltered, debug=debug)


def _inputs(scene, t):
    P = scene["means3D"].shape[0]
    import torch
    return dict(means3D=t(scene["means3D"]).requires_grad_(True), means2D=torch.zeros(P, 3, device="cuda", requires_grad=True),
                opacities=t(scene["opacities"]), shs=t(scene["shs"]), scales=t(scene["scales"]), rotations=t(scene["rotations"]))


def test_prefiltered_true_gives_the_same_image_when_every_gaussian_is_in_front_of_the_camera(mod):
    import torch
    cam = S.make_camera(160, 96)
    scene = S.make_scene(3000, cam, sh_degree=3, M=16, seed=11, near_frac=0.0)
    assert (scene["means3D"] @ cam["viewmatrix"].reshape(4, 4)[:3, 2] + cam["viewmatrix"].reshape(4, 4)[3, 2] > 0.2).all()
    t = lambda a: torch.tensor(np.asarray(a), dtype=torch.float32, device="cuda")
    a, ra = mod.GaussianRasterizer(_settings(mod, cam, t))(**_inputs(scene, t))
    b, rb = mod.GaussianRasterizer(_settings(mod, cam, t, prefiltered=True))(**_inputs(scene, t))
    assert torch.equal(a, b) and torch.equal(ra, rb) and int((ra > 0).sum()) > 1000


_FAIL_FORWARD = r'''
import os, sys
import numpy as np, torch
sys.path[:0] = [%(tests)r, %(pkg)r, %(root)r]
import helpers as Hh
from helpers import S
mod = Hh.my_module()
cam = S.make_camera(64, 48)
scene = S.make_scene(200, cam, sh_degree=0, M=1, seed=2)
scene["means3D"][7, 2] = -1.0                      # behind the camera: culled, which prefiltered=True forbids (forward.cu:204-214)
t = lambda a: torch.tensor(np.asarray(a), dtype=torch.float32, device="cuda")
st = mod.GaussianRasterizationSettings(48, 64, cam["tanfovx"], cam["tanfovy"], t([0, 0, 0]), 1.0, t(cam["viewmatrix"]),
                                       t(cam["projmatrix"]), 0, t(cam["campos"]), True, True)
P = 200
try:
    mod.GaussianRasterizer(st)(means3D=t(scene["means3D"]), means2D=torch.zeros(P, 3, device="cuda"), opacities=t(scene["opacities"]),
                               shs=t(scene["shs"]), scales=t(scene["scales"]), rotations=t(scene["rotations"]))
except Exception as e:
    print("RAISED", type(e).__name__)
    sys.stdout.flush()
    os._exit(7)                                    # the context is dead after a trap: leave without touching CUDA again
print("NO ERROR")
'''


def test_debug_forward_failure_writes_snapshot_fw_and_raises(tmp_path):
    """prefiltered=True with a Gaussian behind the camera traps in the per-Gaussian kernel; with debug=True the library
    synchronises after the stage, reports it, and the Python layer dumps its arguments before re-raising."""
    code = _FAIL_FORWARD % dict(tests=os.path.join(ROOT, "tests"), pkg=os.path.join(ROOT, "gaussian-splatting_cc-comments_b200"), root=ROOT)
    r = subprocess.run([sys.executable, "-c", code], cwd=str(tmp_path), capture_output=True, text=True, timeout=300)
    assert r.returncode == 7 and "RAISED" in r.stdout, (r.returncode, r.stdout[-500:], r.stderr[-1500:])
    assert "snapshot_fw.dump" in r.stdout
    import torch
    dump = torch.load(str(tmp_path / "snapshot_fw.dump"))
    assert len(dump) == 19 and dump[1].shape == (200, 3) and dump[1].device.type == "cpu"      # the 19-argument pack, on the CPU


def test_debug_backward_failure_writes_snapshot_bw_and_raises(mod, tmp_path, monkeypatch):
    import torch
    monkeypatch.chdir(tmp_path)
    cam = S.make_camera(64, 48)
    scene = S.make_scene(300, cam, sh_degree=3, M=16, seed=5)
    t = lambda a: torch.tensor(np.asarray(a), dtype=torch.float32, device="cuda")
    ins = _inputs(scene, t)
    color, _ = mod.GaussianRasterizer(_settings(mod, cam, t, debug=True))(**ins)
    color.grad_fn.num_rendered = -3                # what a corrupted context would hand to the backward: rejected by gsr_backward
    with pytest.raises(Exception):
        color.sum().backward()
    dump = torch.load(str(tmp_path / "snapshot_bw.dump"))
    assert len(dump) == 21 and dump[17] == -3 and dump[1].device.type == "cpu"
    # and the library is still usable afterwards (an argument error is not a sticky CUDA error)
    color2, _ = mod.GaussianRasterizer(_settings(mod, cam, t))(**_inputs(scene, t))
    assert torch.isfinite(color2).all()


# ------------------------------------------------------------------------------------------------ a host that is not torch
class _Cuda:
    """cudaMalloc / cudaMemcpy through ctypes on the CUDA runtime libgsr_b200.so itself links (no torch allocations)."""

    def __init__(self):
        out = subprocess.run(["ldd", os.path.join(ROOT, "gaussian-splatting_cc-comments_b200", "libgsr_b200.so")], capture_output=True, text=True).stdout
        path = [ln.split("=>")[1].split("(")[0].strip() for ln in out.splitlines() if "libcudart" in ln and "=>" in ln]
        self.rt = C.CDLL(path[0] if path and os.path.exists(path[0]) else "libcudart.so.12")
        self.rt.cudaMalloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
        self.rt.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
        self.rt.cudaFree.argtypes = [C.c_void_p]
        self.live = []

    def malloc(self, n):
        p = C.c_void_p()
        assert self.rt.cudaMalloc(C.byref(p), max(int(n), 16)) == 0
        self.live.append(p.value)
        return p.value

    def put(self, a):
        a = np.ascontiguousarray(a)
        p = self.malloc(a.nbytes)
        assert self.rt.cudaMemcpy(p, a.ctypes.data, a.nbytes, 1) == 0
        return p

    def get(self, p, shape, dtype=np.float32):
        a = np.empty(shape, dtype)
        assert self.rt.cudaDeviceSynchronize() == 0
        assert self.rt.cudaMemcpy(a.ctypes.data, p, a.nbytes, 2) == 0
        return a

    def free_all(self):
        for p in self.live:
            self.rt.cudaFree(p)
        self.live = []


def test_c_abi_from_a_non_torch_host_with_cudamalloc_and_a_c_callback(mod):
    """gsr_forward / gsr_backward (include/gsr.h) with raw device pointers, the three buffers handed out by a C callback that
    calls cudaMalloc, the legacy default stream (NULL); results against the torch binding on the same scene."""
    from test_abi import FwdArgs, BwdArgs, RESIZE
    lib = C.CDLL(os.path.join(ROOT, "gaussian-splatting_cc-comments_b200", "libgsr_b200.so"))
    lib.gsr_last_error.restype = C.c_char_p
    lib.gsr_backward_scratch_bytes.restype = C.c_size_t
    cu = _Cuda()
    cam = S.make_camera(160, 96)
    P, M = 2500, 16
    scene = S.make_scene(P, cam, sh_degree=3, M=M, seed=21)
    dL = S.upstream_grad(cam, seed=3)
    want = Hh.run_torch(mod, cam, scene, dL, bg=(0.1, 0.2, 0.3))
    W, H = cam["W"], cam["H"]
    f32 = lambda a: np.asarray(a, np.float32)
    d = {k: cu.put(f32(scene[k])) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    bg, vm, pm, cp = cu.put(f32([0.1, 0.2, 0.3])), cu.put(f32(cam["viewmatrix"])), cu.put(f32(cam["projmatrix"])), cu.put(f32(cam["campos"]))
    out_color, radii = cu.malloc(3 * H * W * 4), cu.malloc(P * 4)
    handed = {}

    def resize(user, n):                               # resizeFunctional of rasterize_points.cu:28-36, without torch
        handed[user] = cu.malloc(n)
        return handed[user]
    cb = RESIZE(resize)
    a = FwdArgs()
    a.P, a.D, a.M, a.width, a.height = P, 3, M, W, H
    a.background, a.means3D, a.shs, a.opacities, a.scales, a.rotations = bg, d["means3D"], d["shs"], d["opacities"], d["scales"], d["rotations"]
    a.scale_modifier, a.viewmatrix, a.projmatrix, a.cam_pos = 1.0, vm, pm, cp
    a.tan_fovx, a.tan_fovy, a.out_color, a.radii = cam["tanfovx"], cam["tanfovy"], out_color, radii
    nr = C.c_int(0)
    rc = lib.gsr_forward(C.byref(a), cb, C.c_void_p(1), cb, C.c_void_p(2), cb, C.c_void_p(3), None, C.byref(nr))
    assert rc == 0, lib.gsr_last_error()
    assert nr.value == want["num_rendered"] and set(handed) == {1, 2, 3}
    assert np.array_equal(cu.get(radii, (P,), np.int32), want["radii"])
    assert np.array_equal(cu.get(out_color, (3, H, W)).view(np.uint32), want["color"].view(np.uint32))      # bit-identical image

    b = BwdArgs()
    b.P, b.D, b.M, b.R, b.width, b.height = P, 3, M, nr.value, W, H
    b.background, b.means3D, b.shs, b.scales, b.rotations, b.scale_modifier = bg, d["means3D"], d["shs"], d["scales"], d["rotations"], 1.0
    b.viewmatrix, b.projmatrix, b.cam_pos, b.tan_fovx, b.tan_fovy, b.radii = vm, pm, cp, cam["tanfovx"], cam["tanfovy"], radii
    b.geom_buffer, b.binning_buffer, b.image_buffer, b.dL_dpix = handed[1], handed[2], handed[3], cu.put(f32(dL))
    widths = dict(dL_dmean2D=3, dL_dconic=4, dL_dopacity=1, dL_dcolor=3, dL_dmean3D=3, dL_dcov3D=6, dL_dsh=3 * M, dL_dscale=3, dL_drot=4)
    for k, w in widths.items():
        setattr(b, k, cu.malloc(P * w * 4))             # uninitialised on purpose: every element is written
    b.grad_scratch = cu.malloc(lib.gsr_backward_scratch_bytes(P))
    rc = lib.gsr_backward(C.byref(b), None)
    assert rc == 0, lib.gsr_last_error()
    got = dict(means3D=cu.get(b.dL_dmean3D, (P, 3)), means2D=cu.get(b.dL_dmean2D, (P, 3)), shs=cu.get(b.dL_dsh, (P, M, 3)),
               opacities=cu.get(b.dL_dopacity, (P, 1)), scales=cu.get(b.dL_dscale, (P, 3)), rotations=cu.get(b.dL_drot, (P, 4)))
    for k, g in got.items():
        assert Hh.rel_l2(g, want["grads"][k]) <= 1e-5, k      # same kernels; only the order of the float reductions differs
    cu.free_all()
