"""Synthetic stimulation protocols and input generators of the BASELINE.json configurations
(SURVEY.md section 8d).  Host-side numpy/torch only; seeds fixed so that every run and the CPU baseline
see the same inputs."""
from __future__ import annotations

import torch

from . import fes
from ._lib import FES_DING2003, FES_DING2007, FES_HMED2018, PARAM_INDEX


def fatigue_protocol(final_time=300, frequency=33, on_s=1, period_s=2, truncation=10):
    """Config 2: `frequency` Hz pulse trains, `on_s` s on / (period_s - on_s) s off for `final_time` s
    (reference examples/getting_started/model_integration.py:19-20).  Returns dict(win, ns, dt, N, stim)."""
    stim = [s + k / frequency for s in range(0, int(final_time), period_s) for k in range(int(frequency * on_s))]
    ns = int(round(final_time * frequency))  # every pulse sits on the k/frequency grid
    win, idx = fes.stim_windows(stim, ns, final_time, truncation)
    return dict(win=win, ns=ns, dt=final_time / ns, N=truncation, stim=stim, stim_idx=idx, final_time=final_time)


def fatigue_population(batch, seed=0, kind=FES_DING2003, device="cpu", offset=0):
    """Per-trajectory parameters uniform(0.8, 1.2) x default for (a_rest, km_rest, tau1_rest, tau2, alpha_a,
    alpha_tau1, alpha_km, tau_fat) and the matching rest states.  `offset` shifts the seed per rank so
    that shards of a multi-GPU run are distinct populations.  Returns (params [14,B], x0 [5,B]) float64."""
    g = torch.Generator(device="cpu").manual_seed(seed + offset)
    p = torch.tensor(fes.default_params(kind))[:, None].repeat(1, batch)
    p[:8] *= 0.8 + 0.4 * torch.rand((8, batch), generator=g, dtype=torch.float64)
    x0 = torch.zeros((5, batch), dtype=torch.float64)
    x0[2], x0[3], x0[4] = p[PARAM_INDEX["a_rest"]], p[PARAM_INDEX["tau1_rest"]], p[PARAM_INDEX["km_rest"]]
    return p.to(device), x0.to(device)


def train_protocol(n_pulses=10, frequency=10, final_time=2.0, truncation=10):
    """Configs 3 / 4: `n_pulses` at `frequency` Hz from t = 0, observed until final_time
    (examples/identification/ding2007_model_identification.py:24-31)."""
    stim = [k / frequency for k in range(n_pulses)]
    ns = fes.prepare_n_shooting(stim, final_time)
    win, idx = fes.stim_windows(stim, ns, final_time, truncation)
    return dict(win=win, ns=ns, dt=final_time / ns, N=truncation, stim=stim, stim_idx=idx, final_time=final_time)


def ding2007_sweep(batch, seed=1, seed_pw=2, device="cpu"):
    """Config 3: parameter sets (a_scale, km_rest, tau1_rest, pd0, pdt) uniform in their ranges; shared
    pulse widths ~ U(2e-4, 6e-4) per pulse.  Returns (proto, params [14,B], x0 [2,B], ctrl [ns,1])."""
    proto = train_protocol()
    g = torch.Generator().manual_seed(seed)
    u = torch.rand((5, batch), generator=g, dtype=torch.float64)
    p = torch.tensor(fes.default_params(FES_DING2007))[:, None].repeat(1, batch)
    lo_hi = {"a_scale": (2000, 8000), "km_rest": (0.05, 0.4), "tau1_rest": (0.02, 0.1), "pd0": (1e-4, 2e-4), "pdt": (1e-4, 3e-4)}
    for r, (k, (lo, hi)) in enumerate(lo_hi.items()):
        p[PARAM_INDEX[k]] = lo + (hi - lo) * u[r]
    pw = 2e-4 + 4e-4 * torch.rand(len(proto["stim"]), generator=torch.Generator().manual_seed(seed_pw), dtype=torch.float64)
    # pulse width must exceed pd0 for A_eff > 0: the sweep keeps pd0 <= 2e-4 <= pw
    ctrl = torch.stack([pw[proto["stim_idx"][n][-1]] for n in range(proto["ns"])])[:, None]
    x0 = torch.zeros((2, batch), dtype=torch.float64)
    return proto, p.to(device), x0.to(device), ctrl.to(device)


HMED_BOUNDS = {  # identification bounds of hmed2018_force_parameter_identification.py:103-160
    "a_rest": (1, 10000), "km_rest": (0.001, 1), "tau1_rest": (0.0001, 1), "tau2": (0.0001, 1),
    "ar": (0.01, 1), "bs": (0.001, 0.1), "Is": (1, 150), "cr": (0.01, 2),
}


def hmed2018_candidates(batch, seed=3, seed_i=4, device="cpu", spread=0.3):
    """Config 4: candidate vectors of the 8 identifiable parameters.  Drawn uniform within +-spread of the
    defaults, clipped to the identification bounds (the full bounds span five decades and mostly give
    saturated, uninformative responses); shared intensities ~ randint(30, 130) per pulse."""
    proto = train_protocol()
    g = torch.Generator().manual_seed(seed)
    p = torch.tensor(fes.default_params(FES_HMED2018))[:, None].repeat(1, batch)
    for k, (lo, hi) in HMED_BOUNDS.items():
        i = PARAM_INDEX[k]
        p[i] = (p[i] * (1 - spread + 2 * spread * torch.rand(batch, generator=g, dtype=torch.float64))).clamp(lo, hi)
    inten = torch.randint(30, 130, (len(proto["stim"]),), generator=torch.Generator().manual_seed(seed_i)).to(torch.float64)
    N, ns = proto["N"], proto["ns"]
    ctrl = torch.full((ns, N), 50.0, dtype=torch.float64)
    for n in range(ns):
        idx = proto["stim_idx"][n]
        for s, src in enumerate(idx):
            ctrl[n, N - len(idx) + s] = inten[src]
    x0 = torch.zeros((2, batch), dtype=torch.float64)
    return proto, p.to(device), x0.to(device), ctrl.to(device)


def nmpc_scenarios(batch, seed=5, device="cpu", frequency=33, truncation=10):
    """Config 5: one cyclic NMPC horizon (33 pulses at 33 Hz over 1 s, previous cycle's last 10 pulses as
    previous_stim) for `batch` scenarios of DingModelPulseWidthFrequencyWithFatigue: decision states around
    rest +- 10 %, F in [0, 300], pulse widths in [pd0, 6e-4].
    Returns (proto, params [14], X [ns+1,5,B], U [ns,1,B])."""
    stim = [k / frequency for k in range(frequency)]
    prev = [-(truncation - k) / frequency for k in range(truncation)]
    ns = frequency
    win, idx = fes.stim_windows(stim, ns, 1.0, truncation, previous_time=prev)
    proto = dict(win=win, ns=ns, dt=1.0 / ns, N=truncation, stim=stim, stim_idx=idx, final_time=1.0)
    p = torch.tensor(fes.default_params(FES_DING2007))
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.rand(s, generator=g, dtype=torch.float64)
    X = torch.empty((ns + 1, 5, batch), dtype=torch.float64)
    X[:, 0] = 0.05 + 0.9 * r(ns + 1, batch)
    X[:, 1] = 300.0 * r(ns + 1, batch)
    X[:, 2] = p[PARAM_INDEX["a_scale"]] * (0.9 + 0.2 * r(ns + 1, batch))
    X[:, 3] = p[PARAM_INDEX["tau1_rest"]] * (0.9 + 0.2 * r(ns + 1, batch))
    X[:, 4] = p[PARAM_INDEX["km_rest"]] * (0.9 + 0.2 * r(ns + 1, batch))
    pd0 = float(p[PARAM_INDEX["pd0"]])
    U = pd0 + (6e-4 - pd0) * r(ns, 1, batch)
    return proto, p.to(device), X.to(device), U.to(device)
