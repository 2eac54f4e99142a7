#!/usr/bin/env python
"""Prints the actual errors of the bench kernel (Ding2003 + fatigue, d x/d x0) against the fp64 C oracle at several horizons
of the fatigue protocol, in fp64 and in fp32 mode: how much of the 1e-9 / 1e-4 budget a build really uses.  Every state
and every (row, column) sensitivity entry is measured on ITS OWN scale (the largest value it takes over the batch).
FES_B200_LIB selects the build.  Test tooling: imports oracle/.

    python scripts/parity_margin.py [--seconds 20,100,300] [--batch 16]
"""
import argparse
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cocofest_b200 import fes, protocols  # noqa: E402
from oracle import fes_oracle_c as foc  # noqa: E402
from tests.problems import entry_err  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--seconds", default="20,100,300")
ap.add_argument("--batch", type=int, default=16)
a = ap.parse_args()
dev = torch.device("cuda", 0)
for seconds in [int(s) for s in a.seconds.split(",")]:
    proto = protocols.fatigue_protocol(final_time=seconds)
    params, x0 = protocols.fatigue_population(a.batch, device=dev)
    win = torch.tensor(proto["win"], device=dev)
    for dtype, name in ((torch.float64, "f64"), (torch.float32, "f32")):
        p, x = params.to(dtype), x0.to(dtype)
        res = fes.ivp(fes.FES_DING2003, True, p, x, win, proto["dt"], n_steps=10, n_shooting=proto["ns"], dirs=fes.FES_DIR_X0,
                      trace="nodes", dtype=dtype)
        torch.cuda.synchronize()
        # the oracle starts from the same (rounded) inputs the library was given
        ref = foc.ivp(0, True, 4, proto["N"], proto["ns"], 10, proto["dt"], 0.020, 1.04, p.double().cpu().numpy().T.copy(),
                      x.double().cpu().numpy().T.copy(), proto["win"][None], None, seed_x0=True, trace_B=a.batch)
        err_x = entry_err(res.x_final.double().cpu().numpy().T, ref["final"], keep=(1,), tiny=1e-200)
        err_s = entry_err(np.transpose(res.sens_final.double().cpu().numpy(), (2, 0, 1)), ref["sens_final"], keep=(1, 2), tiny=1e-200)
        tr = np.transpose(res.trace.double().cpu().numpy(), (2, 0, 1))
        on = [k for k in range(17, proto["ns"] + 1, 66)]       # mid on-phase nodes: every state is O(its scale) there
        err_t = entry_err(tr[:, on], ref["trace"][:, on], keep=(2,), tiny=1e-200)
        print(f"{name} seconds={seconds} batch={a.batch} entry-wise rel err: final state {err_x:.2e}  states at on-phase nodes {err_t:.2e}  "
              f"d x(T)/d x0 {err_s:.2e}", flush=True)
