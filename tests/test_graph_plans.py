"""IvpPlan / ShootPlan calls captured in CUDA graphs (fes.GraphedPlan): same bits as the direct calls after the inputs
are rewritten in place, every pass of a multi-pass plan inside the one graph, and the single-trajectory case timed both ways
(measured on the B200: 41.0 us either way -- the call is bound by its chain of dependent steps, not by the launch)."""
import time

import numpy as np
import pytest
import torch

from cocofest_b200 import fes
from tests.problems import random_problem, to_lib

pytestmark = pytest.mark.gpu


def test_graphed_shoot_plan_replays_every_pass_with_new_inputs():
    dev = torch.device("cuda", 0)
    pb = random_problem("hmed2018_with_fatigue", B=40, N=12, ns=6, seed=51, shared_ctrl=False)
    params, x0, win, ctrl = to_lib(pb, dev)
    X = (x0[None] * torch.ones((pb["ns"] + 1, 1, 1), dtype=torch.float64, device=dev)).contiguous()
    plan = fes.ShootPlan(pb["kind"], pb["fat"], params, win, pb["dt"], pb["ns"], pb["B"], ctrl=ctrl, want=("x", "u", "p"), endpoints=True)
    graphed = plan.graph(X, ctrl, params)          # 1 + 2 + 1 passes in one graph
    for rep in range(3):
        Xn = (X * (1.0 - 0.03 * rep)).contiguous()
        cn = (ctrl * (1.0 + 0.01 * rep)).contiguous()
        ref = fes.shoot(pb["kind"], pb["fat"], params, Xn, win, pb["dt"], ctrl=cn, want=("x", "u", "p"), endpoints=True)
        graphed.inputs[0].copy_(Xn)                # the graph holds the pointers of X / ctrl: new values go INTO them
        graphed.inputs[1].copy_(cn)
        got = graphed.replay()
        torch.cuda.synchronize()
        for a, b in ((got.G, ref.G), (got.Phi, ref.Phi), (got.Jx, ref.Jx), (got.Ju, ref.Ju), (got.Jp, ref.Jp)):
            assert torch.equal(a, b)


def test_graphed_ivp_plan_and_the_launch_bound_case():
    dev = torch.device("cuda", 0)
    pb = random_problem("ding2007_with_fatigue", B=1, N=6, ns=10, seed=52)
    params, x0, win, ctrl = to_lib(pb, dev)
    plan = fes.IvpPlan(pb["kind"], pb["fat"], params, x0, win, pb["dt"], ctrl=ctrl, n_shooting=pb["ns"], dirs=fes.FES_DIR_X0, trace="nodes")
    graphed = plan.graph()
    direct = plan()
    keep = (direct.x_final.clone(), direct.sens_final.clone(), direct.trace.clone())
    x0.mul_(1.01)
    ref = fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], ctrl=ctrl, n_shooting=pb["ns"], dirs=fes.FES_DIR_X0, trace="nodes")
    got = graphed()
    torch.cuda.synchronize()
    assert torch.equal(got.x_final, ref.x_final) and torch.equal(got.sens_final, ref.sens_final) and torch.equal(got.trace, ref.trace)
    assert not torch.equal(got.x_final, keep[0])

    def per_call(fn, n=300):
        for _ in range(20):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        torch.cuda.synchronize()
        return 1e6 * (time.perf_counter() - t0) / n

    us_plan, us_graph = per_call(plan), per_call(graphed)
    print(f"single trajectory, 100 RK4 steps with d x/d x0: {us_plan:.1f} us per IvpPlan call, {us_graph:.1f} us per graph replay")
    assert us_graph < 1.5 * us_plan    # never worse by much; the numbers go to DESIGN.md from the GPU log
