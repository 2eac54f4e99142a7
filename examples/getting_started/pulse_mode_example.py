"""Single pulses, doublets and triplets of the same 10 Hz train (reference: examples/getting_started/pulse_mode_example.py),
integrated with IvpFes on the GPU -- and then the three modes for a population of muscles in ONE launch each, which is what
the batched path adds.  Plotting is left out (matplotlib is not part of the accelerated path).

    python examples/getting_started/pulse_mode_example.py
"""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))

from cocofest_b200 import IvpFes, ModelMaker, fes  # noqa: E402

final_time = 1
base_train = [k / 10 for k in range(10)]
peaks = {}
for mode in ("single", "doublet", "triplet"):
    # a fresh model per mode: IvpFes expands model.stim_time in place for doublets / triplets (as the reference does)
    model = ModelMaker.create_model("ding2003_with_fatigue", stim_time=list(base_train))
    ivp = IvpFes({"model": model, "pulse_mode": mode}, {"final_time": final_time})
    result, time = ivp.integrate()
    force = result["F"][0]
    peaks[mode] = float(force.max())
    print(f"{mode:8s}: {len(ivp.stim_time):2d} pulses, n_shooting = {ivp.n_shooting:4d}, {force.shape[0]:5d} points, "
          f"peak force {force.max():7.2f} N, force at t = 1 s {force[-1]:7.2f} N")

    # the same protocol for 50 000 muscles whose parameters scatter +-20 % around the defaults
    B = 50_000
    gen = torch.Generator().manual_seed(0)
    params = torch.tensor(model.param_vector())[:, None] * (0.8 + 0.4 * torch.rand((fes.FES_NP, B), generator=gen, dtype=torch.float64))
    batch = ivp.integrate_batch(params=params, trace="nodes", trace_batch=1)
    f_end = batch.x_final[1]
    print(f"          population of {B}: force at t = 1 s {f_end.mean().item():7.2f} +- {f_end.std().item():.2f} N")

assert peaks["single"] < peaks["doublet"] < peaks["triplet"], "more pulses per burst, more force"
