"""Adam step, CPU side: the oracle against vectors produced by torch.optim.Adam configured as the
reference configures it (tests/golden/ref_adam_cpu.npz, made by tests/golden/make_golden_adam.py from the reference's
arguments.OptimizationParams and get_expon_lr_func), and the host logic of gsr_optim.FusedAdam."""
import os

import numpy as np
import pytest
import torch

import gsr_optim
from oracle import gs_next_oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_adam_cpu.npz")
SHAPES = {"xyz": (3,), "f_dc": (1, 3), "f_rest": (15, 3), "opacity": (1,), "scaling": (3,), "rotation": (4,)}


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def test_oracle_reproduces_torch_adam_over_the_golden_trajectory(gold):
    """Five steps, a changing xyz learning rate, zero-gradient rows, and a densification-style growth of the moments
    before step 3.  Bar: every parameter and moment within 2 ulp-of-the-largest (1e-7 relative to the tensor's max);
    torch's CPU kernels contract a few multiply-adds differently from its CUDA functors, which the oracle follows."""
    steps = int(gold["steps"])
    exact = total = 0
    for n in gold["names"]:
        p = gold["init_" + n].copy()
        m = np.zeros_like(p)
        v = np.zeros_like(p)
        for it in range(steps):
            if it == 3:
                ext = gold["ext_" + n]
                p = np.concatenate([p, ext]); m = np.concatenate([m, np.zeros_like(ext)]); v = np.concatenate([v, np.zeros_like(ext)])
            lr = float(gold["xyz_lr_step%d" % it]) if n == "xyz" else float(gold["lr_" + n])
            O.adam_step(p, gold["grad_%s_step%d" % (n, it)], m, v, lr, it + 1)
            assert float(gold["state_step_%s_step%d" % (n, it)]) == it + 1
            for a, k in ((p, "param"), (m, "exp_avg"), (v, "exp_avg_sq")):
                r = gold["%s_%s_step%d" % (k, n, it)]
                assert a.shape == r.shape
                assert np.abs(a - r).max() <= 1e-7 * max(float(np.abs(r).max()), 1e-30), (n, it, k)
                exact += int((a == r).sum()); total += a.size
    assert exact / total > 0.99            # measured 0.9987: the rest differ in the last bit


def test_first_step_moves_every_touched_parameter_by_lr(gold):
    """Property of Adam with eps = 1e-15: after step 1, |delta| = lr where the gradient is non-zero, 0 where it is zero."""
    rng = np.random.default_rng(0)
    p = rng.normal(size=1000).astype(np.float32)
    g = rng.normal(size=1000).astype(np.float32) * 10.0 ** rng.integers(-6, 1, size=1000).astype(np.float32)
    g[::7] = 0
    p0 = p.copy(); m = np.zeros_like(p); v = np.zeros_like(p)
    O.adam_step(p, g, m, v, 0.01, 1)
    d = p - p0
    assert np.all(d[::7] == 0)
    nz = g != 0
    assert np.allclose(np.abs(d[nz]), 0.01, rtol=1e-4) and np.all(np.sign(d[nz]) == -np.sign(g[nz]))


def test_expon_lr_func_matches_the_reference_function(gold):
    f = gsr_optim.get_expon_lr_func(lr_init=float(gold["position_lr_init"]) * float(gold["spatial_lr_scale"]),
                                    lr_final=float(gold["position_lr_final"]) * float(gold["spatial_lr_scale"]),
                                    lr_delay_mult=float(gold["position_lr_delay_mult"]), max_steps=int(gold["position_lr_max_steps"]))
    for it in range(int(gold["steps"])):
        assert f(int(gold["iteration_step%d" % it])) == pytest.approx(float(gold["xyz_lr_step%d" % it]), rel=1e-12)
    assert f(-1) == 0.0
    assert gsr_optim.get_expon_lr_func(0.0, 0.0)(10) == 0.0
    g = gsr_optim.get_expon_lr_func(1e-2, 1e-4, lr_delay_steps=100, lr_delay_mult=0.1, max_steps=1000)
    assert g(0) == pytest.approx(1e-3) and g(1000) == pytest.approx(1e-4) and g(5000) == pytest.approx(1e-4)


class _Args:
    position_lr_init = 0.00016
    feature_lr = 0.0025
    opacity_lr = 0.05
    scaling_lr = 0.005
    rotation_lr = 0.001


def _optimizer(P=7):
    params = {n: torch.nn.Parameter(torch.randn(P, *s)) for n, s in SHAPES.items()}
    groups = gsr_optim.gaussian_param_groups(params["xyz"], params["f_dc"], params["f_rest"], params["opacity"], params["scaling"],
                                             params["rotation"], _Args, spatial_lr_scale=2.0)
    return params, gsr_optim.FusedAdam(groups, lr=0.0, eps=1e-15)


def test_fused_adam_has_torch_adams_surface():
    params, opt = _optimizer()
    assert isinstance(opt, torch.optim.Optimizer)
    assert [g["name"] for g in opt.param_groups] == ["xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation"]
    assert opt.param_groups[0]["lr"] == pytest.approx(0.00032) and opt.param_groups[2]["lr"] == pytest.approx(0.0025 / 20)
    assert all(g["eps"] == 1e-15 and g["betas"] == (0.9, 0.999) for g in opt.param_groups)
    ref = torch.optim.Adam([{"params": [torch.nn.Parameter(torch.zeros(1))], "lr": 0.1, "name": "xyz"}], lr=0.0, eps=1e-15)
    for k in ("lr", "betas", "eps", "weight_decay", "amsgrad", "maximize"):
        assert k in opt.param_groups[0] and k in ref.param_groups[0]
    for bad in (dict(weight_decay=0.1), dict(amsgrad=True), dict(maximize=True)):
        with pytest.raises(NotImplementedError):
            gsr_optim.FusedAdam([torch.nn.Parameter(torch.zeros(1))], **bad)
    with pytest.raises(ValueError):
        gsr_optim.FusedAdam([torch.nn.Parameter(torch.zeros(1))], betas=(1.0, 0.999))


def test_there_is_no_cpu_path():
    params, opt = _optimizer()
    for p in params.values():
        p.grad = torch.ones_like(p)
    with pytest.raises(RuntimeError, match="no CPU path"):
        opt.step()
    opt.zero_grad(set_to_none=True)
    opt.step()                                 # nothing has a gradient: nothing to do, no error (as torch)


def test_state_layout_lets_the_references_densification_code_edit_it():
    """cat_tensors_to_optimizer / _prune_optimizer (scene/gaussian_model.py:398-471) index optimizer.state by parameter
    and touch "exp_avg" / "exp_avg_sq"; state_dict round-trips for checkpoints (train.py:180)."""
    params, opt = _optimizer(P=7)
    for p in params.values():
        st = opt._init_state(p)
        assert set(st) == {"step", "exp_avg", "exp_avg_sq"} and st["exp_avg"].shape == p.shape and float(st["step"]) == 0.0
    # the reference's growth edit, verbatim in behaviour
    for group in opt.param_groups:
        ext = torch.randn(3, *SHAPES[group["name"]])
        stored = opt.state.get(group["params"][0], None)
        assert stored is not None
        stored["exp_avg"] = torch.cat((stored["exp_avg"], torch.zeros_like(ext)), dim=0)
        stored["exp_avg_sq"] = torch.cat((stored["exp_avg_sq"], torch.zeros_like(ext)), dim=0)
        del opt.state[group["params"][0]]
        group["params"][0] = torch.nn.Parameter(torch.cat((group["params"][0], ext), dim=0).requires_grad_(True))
        opt.state[group["params"][0]] = stored
    # the reference's prune edit
    mask = torch.tensor([True, False] * 5)
    for group in opt.param_groups:
        stored = opt.state.get(group["params"][0], None)
        stored["exp_avg"] = stored["exp_avg"][mask]
        stored["exp_avg_sq"] = stored["exp_avg_sq"][mask]
        del opt.state[group["params"][0]]
        group["params"][0] = torch.nn.Parameter(group["params"][0][mask].requires_grad_(True))
        opt.state[group["params"][0]] = stored
        assert opt.state[group["params"][0]]["exp_avg"].shape[0] == 5
    sd = opt.state_dict()
    params2, opt2 = _optimizer(P=5)
    opt2.load_state_dict(sd)
    assert [g["name"] for g in opt2.param_groups] == [g["name"] for g in opt.param_groups]
    assert all(opt2.state[g["params"][0]]["exp_avg"].shape[0] == 5 for g in opt2.param_groups)
