#!/usr/bin/env python
"""Prints the actual relative errors of the bench kernel (Ding2003 + fatigue, d x/d x0) against the C oracle on the
first `--seconds` of the fatigue protocol: how much of the 1e-9 budget a build really uses (FES_B200_LIB selects
the build).  Test tooling: imports oracle/."""
import argparse
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cocofest_b200 import fes, protocols  # noqa: E402
from oracle import fes_oracle_c as foc  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--seconds", type=int, default=20)
ap.add_argument("--batch", type=int, default=64)
a = ap.parse_args()
dev = torch.device("cuda", 0)
proto = protocols.fatigue_protocol(final_time=a.seconds)
params, x0 = protocols.fatigue_population(a.batch, device=dev)
win = torch.tensor(proto["win"], device=dev)
res = fes.ivp(fes.FES_DING2003, True, params, x0, win, proto["dt"], n_steps=10, n_shooting=proto["ns"], dirs=fes.FES_DIR_X0,
              trace="nodes")
torch.cuda.synchronize()
ref = foc.ivp(0, True, 4, proto["N"], proto["ns"], 10, proto["dt"], 0.020, 1.04, params.cpu().numpy().T.copy(),
              x0.cpu().numpy().T.copy(), proto["win"][None], None, seed_x0=True, trace_B=a.batch)
xf = res.x_final.cpu().numpy().T
err_x = np.max(np.abs(xf - ref["final"]) / np.maximum(np.abs(ref["final"]), 1e-12 * np.abs(ref["final"]).max()))
sf = np.transpose(res.sens_final.cpu().numpy(), (2, 0, 1))
err_s = np.max(np.abs(sf - ref["sens_final"])) / np.max(np.abs(ref["sens_final"]))
tr = np.transpose(res.trace.cpu().numpy(), (2, 0, 1))
err_t = np.max(np.abs(tr - ref["trace"]) / np.maximum(np.abs(ref["trace"]), 1e-9 * np.abs(ref["trace"]).max(axis=(0, 1))))
print(f"seconds={a.seconds} batch={a.batch} rel_err final={err_x:.3e} trace={err_t:.3e} sens={err_s:.3e}")
