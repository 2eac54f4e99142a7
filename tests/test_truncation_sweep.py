"""The truncation sweep (examples/sensitivity/truncation_sweep.py; reference examples/sensitivity/truncation/
sensitivity_analysis.py:55-132) batches trajectories of DIFFERENT truncations into one launch by left-padding the shorter
windows with the reference's pad time.  Pinned here: a padded window gives the same bits as the short one, both match
the oracle driven with the short window, the ABI accepts the sweep's long windows, and the example's grouping covers
every (frequency, truncation) point of the reference's loop nest."""
import importlib.util
from pathlib import Path

import numpy as np
import pytest
import torch

from cocofest_b200 import fes
from cocofest_b200._lib import FesLibraryError
from oracle import fes_oracle as fo
from oracle import fes_oracle_c as foc
from tests.conftest import BACKENDS, ROOT

spec = importlib.util.spec_from_file_location("truncation_sweep", ROOT / "examples" / "sensitivity" / "truncation_sweep.py")
sweep = importlib.util.module_from_spec(spec)
spec.loader.exec_module(sweep)


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("mode", ["single", "doublet", "triplet"])
def test_padded_window_equals_the_shorter_truncation_bit_for_bit(backend, mode):
    freq, T, S = 12, 1.0, 10
    pts = [q for q in sweep.sweep_points(mode, freq, freq, T)]          # truncations 1..12 of one pulse train
    ns = pts[0][3]
    assert all(q[3] == ns for q in pts)
    p = fes.default_params(fes.FES_DING2003)
    rest = fo.rest_state(fo.DING2003, True, p)
    win = sweep.padded_windows(pts, T)                                   # [ns+1, 12, 12]
    assert (win[:, :-1, 0] == fes.PAD_TIME).all() and win[-1, -1, 0] > 0
    x0 = torch.tensor(np.tile(rest[:, None], (1, len(pts))), device=backend)
    batch = fes.ivp(fes.FES_DING2003, True, torch.tensor(p, device=backend), x0, torch.tensor(win, device=backend), T / ns,
                    n_steps=S, n_shooting=ns, trace="nodes")
    for b, (_, j, stim, _) in enumerate(pts):
        short, _ = fes.stim_windows(stim, ns, T, j)
        one = fes.ivp(fes.FES_DING2003, True, torch.tensor(p, device=backend), x0[:, :1].contiguous(),
                      torch.tensor(short, device=backend), T / ns, n_steps=S, n_shooting=ns, trace="nodes")
        assert torch.equal(one.x_final[:, 0], batch.x_final[:, b]) and torch.equal(one.trace[:, :, 0], batch.trace[:, :, b]), j
        ref = foc.ivp(fo.DING2003, True, 4, j, ns, S, T / ns, 0.020, fo.R0_KM_RELATIONSHIP, p[None], rest[None], short[None])
        got = batch.x_final[:, b].cpu().numpy()
        assert np.max(np.abs(got - ref["final"][0]) / np.abs(ref["final"][0])) < 1e-9, j
    # the truncation matters: the shortest and the longest window end at different forces
    assert abs(float(batch.x_final[1, 0] - batch.x_final[1, -1])) > 1e-3 * abs(float(batch.x_final[1, -1]))


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_windows_longer_than_64_pulses(backend):
    """The sweep reaches truncation 100 (x 3 pulses in triplet mode).  Models that only need the two window sums accept up
    to 1024 slots; per-slot tables (Hmed2018, pulse-time columns) keep the limit of 64."""
    T, freq, N = 1.0, 100, 100
    stim = list(np.round([k / freq for k in range(freq)], 3))
    ns = fes.prepare_n_shooting(stim, T)
    win, _ = fes.stim_windows(stim, ns, T, N)
    p = fes.default_params(fes.FES_DING2003)
    rest = fo.rest_state(fo.DING2003, True, p)
    res = fes.ivp(fes.FES_DING2003, True, torch.tensor(p, device=backend), torch.tensor(rest[:, None].copy(), device=backend),
                  torch.tensor(win, device=backend), T / ns, n_steps=4, n_shooting=ns, dirs=fes.FES_DIR_X0)
    ref = foc.ivp(fo.DING2003, True, 4, N, ns, 4, T / ns, 0.020, fo.R0_KM_RELATIONSHIP, p[None], rest[None], win[None], seed_x0=True)
    assert np.max(np.abs(res.x_final[:, 0].cpu().numpy() - ref["final"][0]) / np.abs(ref["final"][0])) < 1e-9
    ph = fes.default_params(fes.FES_HMED2018)
    with pytest.raises(FesLibraryError, match="size"):
        fes.ivp(fes.FES_HMED2018, False, torch.tensor(ph, device=backend), torch.zeros((2, 1), dtype=torch.float64, device=backend),
                torch.tensor(win, device=backend), T / ns, n_steps=4, n_shooting=ns,
                ctrl=torch.full((ns, N), 50.0, dtype=torch.float64, device=backend))
    X = torch.tensor(np.tile(rest[None, :, None], (ns + 1, 1, 1)), device=backend)
    with pytest.raises(FesLibraryError, match="size"):
        fes.shoot(fes.FES_DING2003, True, torch.tensor(p, device=backend), X, torch.tensor(win, device=backend), T / ns, n_steps=4,
                  want=("w",))


def test_sweep_grouping_covers_the_reference_loop_nest():
    for mode, coef in sweep.COEFFICIENT.items():
        pts = sweep.sweep_points(mode, 1, 30)
        assert len(pts) == 30 * 31 // 2                                       # sum_{i=1..30} i truncations
        assert [(i, j) for i, j, _, _ in pts] == [(i, j) for i in range(1, 31) for j in range(1, i + 1)]
        for i, j, stim, ns in pts:
            assert len(stim) == i * coef and 1000 % ns == 0                   # millisecond grid: n_shooting divides 1000
    # the example end to end on the host build of the kernels is covered by the padded-window test; its GPU run time is
    # recorded under profiles/ (r02_truncation_sweep.txt)


@pytest.mark.gpu
def test_sweep_example_runs_and_matches_point_by_point():
    out, spent = sweep.run_mode("doublet", max_stim=12)
    assert len(out["parameter_list"]) == 78 and spent > 0
    i, j = 7, 3
    stim = fes.pulse_mode_times(list(np.round([k / i for k in range(i)], 3)), "doublet")
    ns = fes.prepare_n_shooting(stim, 1.0)
    win, _ = fes.stim_windows(stim, ns, 1.0, j)
    p = fes.default_params(fes.FES_DING2003)
    rest = fo.rest_state(fo.DING2003, True, p)
    ref = foc.ivp(fo.DING2003, True, 4, j, ns, 10, 1.0 / ns, 0.020, fo.R0_KM_RELATIONSHIP, p[None], rest[None], win[None])
    assert abs(out["force_total_results"][i - 1][j - 1] - ref["final"][0, 1]) < 1e-9 * abs(ref["final"][0, 1])
    assert abs(out["km_total_results"][i - 1][j - 1] - ref["final"][0, 4]) < 1e-9 * abs(ref["final"][0, 4])
