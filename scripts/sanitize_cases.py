#!/usr/bin/env python
"""Small invocations of every kernel family for compute-sanitizer (memcheck / racecheck / synccheck): partial last
blocks with shadow lanes, node chunks staged through shared memory between block-wide barriers, per-trajectory and
shared tables, the identification kernel's shared-memory scratch, the host-buffer entry points.

    compute-sanitizer --tool racecheck --error-exitcode 3 python scripts/sanitize_cases.py
"""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

from cocofest_b200 import fes, protocols  # noqa: E402
from tests.problems import MODELS, random_problem, to_lib  # noqa: E402

dev = torch.device("cuda", 0)
n = 0
for name in MODELS:
    for dtype in (torch.float64, torch.float32):
        # B = 70: one full block + 6 live lanes and 58 shadow lanes; ns = 40 with N = 12 -> several shared-memory chunks for Hmed
        pb = random_problem(name, B=70, N=12, ns=40, seed=1, shared_ctrl=True)
        params, x0, win, ctrl = to_lib(pb, dev, dtype)
        for dirs in (0, fes.FES_DIR_X0, fes.FES_DIR_P):
            fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], n_steps=4, ctrl=ctrl, n_shooting=pb["ns"], dirs=dirs,
                    trace="nodes", sens_trace=bool(dirs), trace_batch=9)
            n += 1
        X = (x0[None] * torch.ones((pb["ns"] + 1, 1, 1), dtype=dtype, device=dev)).contiguous()
        fes.shoot(pb["kind"], pb["fat"], params, X, win, pb["dt"], n_steps=4, ctrl=ctrl, want=("x", "u", "p", "w"))
        n += 1
    pb = random_problem(name, B=33, N=5, ns=6, seed=2, per_traj_win=True, shared_ctrl=False)
    params, x0, win, ctrl = to_lib(pb, dev)
    fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], n_steps=3, ctrl=ctrl, n_shooting=pb["ns"], dirs=fes.FES_DIR_X0)
    n += 1
# long shared table: 2 s of the fatigue protocol = 66 nodes, chunked staging with barriers inside the node loop
proto = protocols.fatigue_protocol(final_time=2)
p, x0 = protocols.fatigue_population(200, device=dev)
fes.ivp(fes.FES_DING2003, True, p, x0, torch.tensor(proto["win"], device=dev), proto["dt"], n_shooting=proto["ns"],
        dirs=fes.FES_DIR_X0, trace="nodes", trace_batch=64)
# identification (shared-memory scratch columns), device and host entry points
proto, p, x0, ctrl = protocols.hmed2018_candidates(150, device=dev)
win = torch.tensor(proto["win"], device=dev)
tgt = fes.ivp(fes.FES_HMED2018, False, torch.tensor(fes.default_params(fes.FES_HMED2018), device=dev), x0[:, :1].contiguous(), win,
              proto["dt"], ctrl=ctrl, n_shooting=proto["ns"], trace="nodes").trace[:, 1, 0].contiguous()
for it in (0, 3):
    fes.identify(fes.FES_HMED2018, p, win, proto["dt"], tgt, ctrl=ctrl, max_iter=it)
    fes.identify(fes.FES_HMED2018, p.cpu(), win.cpu(), proto["dt"], tgt.cpu(), ctrl=ctrl.cpu(), max_iter=it)
    n += 2
torch.cuda.synchronize()
print(f"sanitize_cases: {n + 1} calls completed, launches {fes._lib.lib().fes_launch_count()}")
