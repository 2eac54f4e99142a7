"""Run-time behaviour of the boundary: NaN propagates silently and stays in its own trajectory (as in the reference,
SURVEY.md section 8b), the device-pointer entry points honour the caller's stream, and -- because the library
allocates nothing and keeps no state -- a call can be captured in a CUDA graph and replayed on new inputs."""
import numpy as np
import pytest
import torch

from cocofest_b200 import fes
from tests.conftest import BACKENDS
from tests.problems import random_problem, to_lib


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_nan_stays_in_its_own_trajectory(backend):
    pb = random_problem("ding2003_with_fatigue", B=70, seed=2)
    params, x0, win, ctrl = to_lib(pb, backend)
    ok = fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], n_steps=5, n_shooting=pb["ns"], dirs=fes.FES_DIR_X0)
    x0n = x0.clone()
    x0n[1, 3] = float("nan")
    x0n[0, 66] = float("inf")
    bad = fes.ivp(pb["kind"], pb["fat"], params, x0n, win, pb["dt"], n_steps=5, n_shooting=pb["ns"], dirs=fes.FES_DIR_X0)
    keep = torch.ones(70, dtype=torch.bool)
    keep[[3, 66]] = False
    assert torch.equal(bad.x_final[:, keep], ok.x_final[:, keep])
    assert torch.equal(bad.sens_final[..., keep], ok.sens_final[..., keep])
    assert not torch.isfinite(bad.x_final[:, 3]).all() and not torch.isfinite(bad.x_final[:, 66]).all()


@pytest.mark.gpu
def test_calls_follow_the_current_stream():
    dev = torch.device("cuda", 0)
    pb = random_problem("ding2007_with_fatigue", B=4096, seed=7, shared_ctrl=False)
    params, x0, win, ctrl = to_lib(pb, dev)
    ref = fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], n_steps=10, ctrl=ctrl, n_shooting=pb["ns"], dirs=fes.FES_DIR_X0)
    torch.cuda.synchronize()
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):
        x0s = x0 * 1.0  # produced on the side stream: the kernel must be ordered after it
        out = fes.ivp(pb["kind"], pb["fat"], params, x0s, win, pb["dt"], n_steps=10, ctrl=ctrl, n_shooting=pb["ns"],
                      dirs=fes.FES_DIR_X0)
    side.synchronize()
    assert torch.equal(out.x_final, ref.x_final) and torch.equal(out.sens_final, ref.sens_final)


@pytest.mark.gpu
def test_shoot_is_capturable_in_a_cuda_graph():
    """One NMPC-style evaluation (defects + d Phi/d x, d Phi/d u) captured once, replayed on new node states."""
    dev = torch.device("cuda", 0)
    pb = random_problem("ding2007_with_fatigue", B=2048, N=10, ns=33, freq=33.0, seed=9, shared_ctrl=False)
    params, x0, win, ctrl = to_lib(pb, dev)
    rng = np.random.default_rng(0)

    def nodes():
        X = torch.empty((pb["ns"] + 1, 5, pb["B"]), dtype=torch.float64, device=dev)
        X[:] = x0[None]
        return X * torch.tensor(rng.uniform(0.9, 1.1, X.shape), device=dev)

    X_static = nodes()
    call = lambda: fes.shoot(pb["kind"], pb["fat"], params, X_static, win, pb["dt"], n_steps=10, ctrl=ctrl, want=("x", "u"))
    warm = torch.cuda.Stream(device=dev)
    warm.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(warm):
        call()
    torch.cuda.current_stream(dev).wait_stream(warm)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        res = call()
    for _ in range(2):
        Xn = nodes()
        X_static.copy_(Xn)
        graph.replay()
        torch.cuda.synchronize()
        eager = fes.shoot(pb["kind"], pb["fat"], params, Xn, win, pb["dt"], n_steps=10, ctrl=ctrl, want=("x", "u"))
        assert torch.equal(res.G, eager.G) and torch.equal(res.Jx, eager.Jx) and torch.equal(res.Ju, eager.Ju)
