"""Cyclic-NMPC window carry-over and warm-start shifting for B concurrent scenarios (FesNmpcBatch), against the
reference semantics of FesNmpc.update_stim (optimization/fes_nmpc.py:69-85) restated with the oracle's window
builder, and closed-loop consistency: integrating cycle after cycle with carried windows and states equals one
long integration of the concatenated train."""
import numpy as np
import pytest
import torch

import cocofest_b200 as cf
from cocofest_b200 import fes
from cocofest_b200.ocp_fes import FesNmpcBatch
from oracle import fes_oracle as fo
from tests.conftest import BACKENDS


def _model(n_cycles=2, freq=33, N=10):
    stim = [k / freq for k in range(freq * n_cycles)]
    return cf.DingModelPulseWidthFrequencyWithFatigue(stim_time=stim, sum_stim_truncation=N), stim


def test_update_stim_matches_reference_semantics():
    model, stim = _model()
    nm = FesNmpcBatch(model, 1.0, n_cycles_simultaneous=2, n_cycles_to_advance=1, device="cpu")
    assert (nm.n_shooting, nm.cycle_len) == (66, 33)
    # first horizon: padded window (ding2003.py:394-397)
    assert float(nm.shooting.window[0, 0]) == -10000000 and float(nm.shooting.window[0, -1]) == 0.0
    carried = nm.update_stim()
    # fes_nmpc.py:71-73 with the horizon advanced by one cycle: the N pulses preceding the new start, shifted
    expect = [t - 1.0 for t in stim if t - 1.0 < 0][-10:]
    np.testing.assert_allclose(carried, expect, atol=1e-12)
    win_ref, _ = fo.stim_windows(stim, 66, 2.0, 10, previous_time=expect)
    np.testing.assert_allclose(nm.shooting.window.numpy(), win_ref, atol=1e-12)
    # periodic train: every later horizon sees the same window
    w1 = nm.shooting.window.clone()
    nm.update_stim()
    np.testing.assert_allclose(nm.shooting.window.numpy(), w1.numpy(), atol=1e-12)
    assert nm.horizon == 2


def test_advance_shifts_frames_and_repeats_last_cycle():
    model, _ = _model(n_cycles=3)
    nm = FesNmpcBatch(model, 1.0, n_cycles_simultaneous=3, n_cycles_to_advance=1, device="cpu")
    ns = nm.n_shooting
    X = torch.arange(ns + 1, dtype=torch.float64)[:, None, None].repeat(1, 5, 4)
    U = torch.arange(ns, dtype=torch.float64)[:, None, None].repeat(1, 1, 4)
    x0, Xn, Un = nm.advance(X, U)
    assert Xn.shape == X.shape and Un.shape == U.shape
    assert float(x0[0, 0]) == 33.0
    np.testing.assert_array_equal(Xn[: ns - 32, 0, 0].numpy(), np.arange(33, ns + 1))
    np.testing.assert_array_equal(Xn[ns - 32:, 0, 0].numpy(), np.arange(ns - 32, ns + 1))  # last cycle again
    np.testing.assert_array_equal(Un[: ns - 33, 0, 0].numpy(), np.arange(33, ns))
    np.testing.assert_array_equal(Un[ns - 33:, 0, 0].numpy(), np.arange(ns - 33, ns))
    with pytest.raises(ValueError):
        FesNmpcBatch(model, 1.0, n_cycles_simultaneous=3, n_cycles_to_advance=4, device="cpu")


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_closed_loop_equals_one_long_integration(backend):
    """Three horizons of one cycle each, states and windows carried, against the same 3 s train integrated at once:
    the carried window reproduces the long train's window from the second cycle on (N pulses of memory)."""
    B, N, freq = 6, 10, 33
    model, stim1 = _model(n_cycles=1)
    nm = FesNmpcBatch(model, 1.0, device=backend, ode_solver=cf.OdeSolver.RK4(5))
    rng = np.random.default_rng(0)
    p = torch.tensor(fes.default_params(fes.FES_DING2007), device=backend)
    U = torch.tensor(rng.uniform(2e-4, 6e-4, (3, 33, 1, B)), device=backend)
    x = torch.zeros((5, B), dtype=torch.float64, device=backend)
    x[2], x[3], x[4] = p[0], p[2], p[1]
    x_start = x.clone()
    ends = []
    for c in range(3):
        sh = nm.shooting
        res = fes.ivp(fes.FES_DING2007, True, p, x, sh.window, sh.dt, n_steps=5, ctrl=U[c], n_shooting=33, trace="nodes")
        # the integrated nodes are defect-free for the horizon's own evaluator
        ev = nm.evaluate(res.trace, U[c], want=())
        assert float(ev.G.abs().max()) < 1e-9 * float(res.trace.abs().max())
        x = res.x_final
        ends.append(x.clone())
        nm.update_stim()
    stim3 = [k / freq for k in range(3 * freq)]
    win3, _ = fes.stim_windows(stim3, 99, 3.0, N)
    long = fes.ivp(fes.FES_DING2007, True, p, x_start, torch.tensor(win3, device=backend), 1.0 / freq, n_steps=5,
                   ctrl=U.reshape(99, 1, B), n_shooting=99, trace="nodes")
    for c in range(3):
        got, ref = ends[c].cpu().numpy(), long.trace[33 * (c + 1)].cpu().numpy()
        assert np.max(np.abs(got - ref) / np.maximum(np.abs(ref), 1e-12)) < 1e-9


def test_update_stim_partial_advance_shift_and_kept_attributes():
    """n_cycles_simultaneous = 3 > n_cycles_to_advance = 1: the reference (fes_nmpc.py:72) shifts the last N pulses by the
    whole phase time (3 s), the product by the time the window really advances (1 s) -- pinned here, with the customised
    model attributes the next horizon's model must keep."""
    # a train that is NOT the same in every cycle (the third cycle stimulates at 11 Hz): on a periodic train the two rules
    # pick different pulses with identical times
    stim = [k / 33 for k in range(66)] + [2 + k / 11 for k in range(11)]
    model = cf.DingModelPulseWidthFrequencyWithFatigue(stim_time=stim, sum_stim_truncation=10)
    model.tauc, model.r0_km_relationship, model.a_scale = 0.013, 1.1, 4000.0
    model._muscle_name = "biceps"
    nm = FesNmpcBatch(model, 1.0, n_cycles_simultaneous=3, n_cycles_to_advance=1, device="cpu")
    carried = nm.update_stim()
    np.testing.assert_allclose(carried, [t - 1.0 for t in stim if t < 1.0][-10:], atol=1e-12)     # the last N pulses BEFORE t = 1 s
    reference_rule = [t - 3.0 for t in stim[-10:]]                                                # fes_nmpc.py:71-72
    assert not np.allclose(carried, reference_rule)
    assert np.allclose(np.array(carried) - np.round(np.array(carried) * 33) / 33, 0, atol=1e-12)  # still on the 33 Hz grid
    nxt = nm.model
    assert (nxt.tauc, nxt.r0_km_relationship, nxt.a_scale, nxt.muscle_name) == (0.013, 1.1, 4000.0, "biceps")
    assert type(nxt) is type(model) and nxt is not model and model.previous_stim["time"] != nxt.previous_stim["time"]
    # a full advance is where the two rules agree
    again = cf.DingModelPulseWidthFrequencyWithFatigue(stim_time=stim, sum_stim_truncation=10)
    full = FesNmpcBatch(again, 1.0, n_cycles_simultaneous=3, n_cycles_to_advance=3, device="cpu")
    np.testing.assert_allclose(full.update_stim(), reference_rule, atol=1e-12)


def test_update_stim_carries_pulse_intensities_for_hmed():
    stim = [k / 20 for k in range(20)]
    model = cf.DingModelPulseIntensityFrequencyWithFatigue(stim_time=stim, sum_stim_truncation=6,
                                                           previous_stim={"time": [-0.1, -0.05], "pulse_intensity": [40.0, 45.0]})
    nm = FesNmpcBatch(model, 1.0, device="cpu")
    with pytest.raises(ValueError, match="pulse_intensity"):
        nm.update_stim()
    inten = [50.0 + k for k in range(20)]
    carried = nm.update_stim(pulse_intensity=inten)
    np.testing.assert_allclose(carried, [t - 1.0 for t in stim[-6:]], atol=1e-12)
    assert nm.model.previous_stim["pulse_intensity"] == inten[-6:]
    # a horizon shorter than the window: the model's own previous pulses are carried with their intensities
    short = cf.DingModelPulseIntensityFrequencyWithFatigue(stim_time=[0.0, 0.5], sum_stim_truncation=4,
                                                           previous_stim={"time": [-0.3, -0.2], "pulse_intensity": [40.0, 45.0]})
    nm2 = FesNmpcBatch(short, 1.0, device="cpu")
    nm2.update_stim(pulse_intensity=[60.0, 70.0])
    np.testing.assert_allclose(nm2.model.previous_stim["time"], [-1.3, -1.2, -1.0, -0.5], atol=1e-12)
    assert nm2.model.previous_stim["pulse_intensity"] == [40.0, 45.0, 60.0, 70.0]
