"""The host-buffer entry points (fes_*_host_*: CPU tensors in, CPU tensors out, computed on the GPU) on a B200:
bit-identical to the device-pointer path, repeated calls reuse the library-owned staging pool, the phase timing is
reported, and a multi-pass `fes.shoot` on host tensors (staged on the device once) equals the device result."""
import numpy as np
import pytest
import torch

from cocofest_b200 import _lib, fes
from tests.problems import random_problem, to_lib

pytestmark = pytest.mark.gpu


def _cpu(*ts):
    return [None if t is None else t.cpu().pin_memory() for t in ts]


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_ivp_host_equals_device_and_reports_phases(dtype):
    dev = torch.device("cuda", 0)
    pb = random_problem("hmed2018_with_fatigue", B=5000, seed=4, shared_ctrl=False)
    params, x0, win, ctrl = to_lib(pb, dev, dtype)
    kw = dict(n_steps=10, n_shooting=pb["ns"], dirs=fes.FES_DIR_X0, trace="nodes", sens_trace=True, trace_batch=64)
    d = fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], ctrl=ctrl, **kw)
    hp, hx, hw, hc = _cpu(params, x0, win, ctrl)
    h = None
    for _ in range(3):  # the second and third call reuse the pool and the pinned result buffers
        h = fes.ivp(pb["kind"], pb["fat"], hp, hx, hw, pb["dt"], ctrl=hc, out=h, **kw)
        assert h.x_final.device.type == "cpu" and h.x_final.is_pinned()
        for a, b in ((h.x_final, d.x_final), (h.trace, d.trace), (h.sens_final, d.sens_final), (h.sens_trace, d.sens_trace)):
            assert torch.equal(a, b.cpu())
        t = fes.last_host_timing()
        assert t["kernels_ms"] > 0 and t["h2d_ms"] > 0 and t["d2h_ms"] > 0
        assert t["total_ms"] >= t["h2d_ms"] + t["kernels_ms"] + t["d2h_ms"] - 1e-3
    assert _lib.lib().fes_host_pool_trim() == 0
    again = fes.ivp(pb["kind"], pb["fat"], hp, hx, hw, pb["dt"], ctrl=hc, **kw)   # the pool grows back on demand
    assert torch.equal(again.x_final, h.x_final)


def test_shoot_on_host_tensors_runs_every_pass_on_the_device():
    dev = torch.device("cuda", 0)
    pb = random_problem("hmed2018_with_fatigue", B=300, N=12, ns=6, seed=6, shared_ctrl=False)
    params, x0, win, ctrl = to_lib(pb, dev)
    rng = np.random.default_rng(1)
    X = (x0[None] * torch.tensor(rng.uniform(0.9, 1.1, (pb["ns"] + 1, pb["nx"], pb["B"])), device=dev)).contiguous()
    want = ("x", "u", "p", "pf", "w")   # 1 + 2 + 1 + 3 passes
    d = fes.shoot(pb["kind"], pb["fat"], params, X, win, pb["dt"], n_steps=5, ctrl=ctrl, want=want, endpoints=True)
    n0 = _lib.lib().fes_launch_count()
    hp, hX, hw, hc = _cpu(params, X, win, ctrl)
    h = fes.shoot(pb["kind"], pb["fat"], hp, hX, hw, pb["dt"], n_steps=5, ctrl=hc, want=want, endpoints=True)
    # 7 passes; the window tables are rebuilt only when their layout changes (Hmed: never after the first pass)
    assert _lib.lib().fes_launch_count() - n0 == 7 + 1
    for a, b in ((h.G, d.G), (h.Phi, d.Phi), (h.Jx, d.Jx), (h.Ju, d.Ju), (h.Jp, d.Jp), (h.Jw, d.Jw)):
        assert a.device.type == "cpu" and torch.equal(a, b.cpu())
    # a single pass goes through fes_shoot_host_* itself
    h1 = fes.shoot(pb["kind"], pb["fat"], hp, hX, hw, pb["dt"], n_steps=5, ctrl=hc, want=("x",))
    assert torch.equal(h1.Jx, d.Jx.cpu()) and torch.equal(h1.G, d.G.cpu())
    assert fes.last_host_timing()["kernels_ms"] > 0


def test_shoot_with_force_factor_on_host_tensors():
    """One pass (Ding2007) goes through fes_shoot_host_* with the extended control rows; three passes (Hmed2018, N = 12) are
    staged on the device once."""
    dev = torch.device("cuda", 0)
    for name, passes in (("ding2007_with_fatigue", 1), ("hmed2018_with_fatigue", 3)):
        pb = random_problem(name, B=200, N=12, ns=5, seed=8, shared_ctrl=False)
        params, x0, win, ctrl = to_lib(pb, dev)
        rng = np.random.default_rng(2)
        X = (x0[None] * torch.tensor(rng.uniform(0.9, 1.1, (pb["ns"] + 1, pb["nx"], pb["B"])), device=dev)).contiguous()
        fc = torch.tensor(rng.uniform(0.5, 1.2, (pb["ns"], pb["B"])), device=dev)
        d = fes.shoot(pb["kind"], pb["fat"], params, X, win, pb["dt"], n_steps=5, ctrl=ctrl, endpoints=True, force_factor=fc)
        n0 = _lib.lib().fes_launch_count()
        hp, hX, hw, hc, hf = _cpu(params, X, win, ctrl, fc)
        h = fes.shoot(pb["kind"], pb["fat"], hp, hX, hw, pb["dt"], n_steps=5, ctrl=hc, endpoints=True, force_factor=hf)
        assert _lib.lib().fes_launch_count() - n0 == passes + 1
        for a, b in ((h.G, d.G), (h.Phi, d.Phi), (h.Jx, d.Jx), (h.Ju, d.Ju), (h.Jc, d.Jc)):
            assert a.device.type == "cpu" and torch.equal(a, b.cpu())


def test_identify_host_equals_device():
    dev = torch.device("cuda", 0)
    from cocofest_b200 import protocols

    proto, p, x0, ctrl = protocols.hmed2018_candidates(2000, device=dev)
    win = torch.tensor(proto["win"], device=dev)
    truth = torch.tensor(fes.default_params(fes.FES_HMED2018), device=dev)
    F = fes.ivp(fes.FES_HMED2018, False, truth, x0[:, :1].contiguous(), win, proto["dt"], ctrl=ctrl, n_shooting=proto["ns"], trace="nodes")
    target = F.trace[:, 1, 0].contiguous()
    d = fes.identify(fes.FES_HMED2018, p, win, proto["dt"], target, ctrl=ctrl, max_iter=3)
    hp, hw, ht, hc = _cpu(p, win, target, ctrl)
    h = fes.identify(fes.FES_HMED2018, hp, hw, proto["dt"], ht, ctrl=hc, max_iter=3)
    for a, b in ((h.params, d.params), (h.cost, d.cost), (h.grad, d.grad), (h.hess, d.hess), (h.evaluations, d.evaluations)):
        assert torch.equal(a, b.cpu())


@pytest.mark.parametrize("name", ["ding2007_with_fatigue", "hmed2018"])
def test_second_order_host_equals_device(name):
    """fes_shoot_hess_host_f64 (CPU tensors in and out) against fes_shoot_hess_f64 on device tensors: bit-identical, with the
    phase record of the host path."""
    dev = torch.device("cuda", 0)
    pb = random_problem(name, B=300, N=5, ns=6, seed=6, shared_ctrl=False)
    params, x0, win, ctrl = to_lib(pb, dev)
    ns, nx, B = pb["ns"], x0.shape[0], x0.shape[-1]
    X = (x0[None] * torch.linspace(1.0, 0.9, ns + 1, dtype=torch.float64, device=dev)[:, None, None]).contiguous()
    X[:, 0] += 0.1
    X[:, 1] += 20.0
    lam = torch.rand((ns, nx, B), dtype=torch.float64, device=dev)
    d = fes.shoot_hessian(pb["kind"], pb["fat"], params, X, win, pb["dt"], n_steps=4, ctrl=ctrl, multipliers=lam)
    hp, hX, hw, hc, hl = _cpu(params, X, win, ctrl, lam)
    h = fes.shoot_hessian(pb["kind"], pb["fat"], hp, hX, hw, pb["dt"], n_steps=4, ctrl=hc, multipliers=hl)
    assert h.D2.device.type == "cpu" and h.H.is_pinned()
    assert torch.equal(h.D2, d.D2.cpu()) and torch.equal(h.H, d.H.cpu())
    t = fes.last_host_timing()
    assert t["kernels_ms"] > 0 and t["h2d_ms"] > 0 and t["d2h_ms"] > 0
