"""The reference's examples/getting_started/model_integration.py with the import line changed: the stimulation is
on for 1 s at 33 Hz and off for 1 s, the model is integrated with IvpFes and the force decreases with fatigue.
Only ``from cocofest import ...`` / ``from bioptim import OdeSolver`` became ``from cocofest_b200 import ...``
(plotting left out: matplotlib is not part of the accelerated path).  Needs a CUDA device.

Second half: the same protocol for a POPULATION of muscles in one call -- what the batched path is for."""
import numpy as np

from cocofest_b200 import (
    IvpFes,
    DingModelFrequencyWithFatigue,
    DingModelPulseIntensityFrequencyWithFatigue,
    DingModelPulseWidthFrequencyWithFatigue,
    OdeSolver,
)

# --- Set stimulation time apparition --- #
final_time = 3
stim_time = [val for start in range(0, final_time, 2) for val in np.linspace(start, start + 1, 34)[:-1]]

# --- Set Ding2003 model --- #
# model = DingModelFrequencyWithFatigue(stim_time=stim_time, sum_stim_truncation=10)
# fes_parameters = {"model": model}

# --- Set Ding2007 model --- #
# pulse_width = np.random.uniform(0.00019, 0.0006, len(stim_time)).tolist()
# model = DingModelPulseWidthFrequencyWithFatigue(stim_time=stim_time, sum_stim_truncation=10)
# fes_parameters = {"model": model, "pulse_width": pulse_width}

# --- Set Hmed2018 model --- #
pulse_intensity = np.random.randint(20, 130, len(stim_time)).tolist()
model = DingModelPulseIntensityFrequencyWithFatigue(stim_time=stim_time, sum_stim_truncation=10)
fes_parameters = {"model": model, "pulse_intensity": pulse_intensity}

# --- Build ivp --- #
ivp_parameters = {"final_time": final_time, "ode_solver": OdeSolver.RK1(n_integration_steps=10)}
ivp = IvpFes(fes_parameters=fes_parameters, ivp_parameters=ivp_parameters)

result, time = ivp.integrate()
result["time"] = time
print(f"single muscle: {result['F'].shape[1]} points, peak force {result['F'].max():.2f} N, "
      f"A {result['A'][0, 0]:.1f} -> {result['A'][0, -1]:.1f}")

# --- The same train for 100 000 muscles whose parameters scatter +-20 % around the defaults, with d x(T)/d x0 --- #
import torch
from cocofest_b200 import fes

B = 100_000
gen = torch.Generator().manual_seed(0)
params = torch.tensor(model.param_vector())[:, None] * (0.8 + 0.4 * torch.rand((fes.FES_NP, B), generator=gen, dtype=torch.float64))
batch = ivp.integrate_batch(params=params, dirs=fes.FES_DIR_X0, trace="nodes", trace_batch=4)
print(f"population: final force {batch.x_final[1].mean().item():.2f} +- {batch.x_final[1].std().item():.2f} N over {B} muscles, "
      f"sensitivity block {tuple(batch.sens_final.shape)}")
