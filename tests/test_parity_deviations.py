"""The places where the product deliberately does NOT reproduce a line of the reference, each pinned so that the choice is
explicit (docstring of cocofest_b200/identification.py).  The NMPC carry-over shift under a partial
advance is the fourth one; tests/test_nmpc_batch.py::test_update_stim_partial_advance_shift_and_kept_attributes pins it."""
import numpy as np
import pytest
import torch

import cocofest_b200 as cf
from cocofest_b200 import fes
from cocofest_b200.identification import DingModelPulseIntensityFrequencyForceParameterIdentification as HmedId
from oracle import fes_oracle as fo
from oracle import fes_oracle_c as foc
from tests.conftest import BACKENDS


def test_hmed_identification_controls_follow_the_window_slots():
    """Deviation 1: slot j of node n carries the intensity of the pulse that sits in slot j of node n's window.  The
    reference's formula (fes_identification_ocp.py:332-343) reads control_value[stim_idx[j-i][j-i]]: for node i = 1, slot
    j = 3 that is the pulse list of node 2 at position 2, which does not exist yet (IndexError on the reference's own
    example train or, through Python's negative indices, some other pulse) -- and on equal intensities both rules give
    the same matrix."""
    stim = [k / 10 for k in range(10)]
    N, T = 4, 2.0
    ns = fes.prepare_n_shooting(stim, T)
    win, idx = fes.stim_windows(stim, ns, T, N)
    inten = [30.0 + 7 * k for k in range(10)]
    u = HmedId._controls(None, inten, idx, ns, N)
    for n in range(ns):
        for s in range(N):
            if win[n, s] > fes.PAD_TIME:   # a real pulse: its own intensity, whatever node looks at it
                assert u[n, s] == inten[stim.index(win[n, s])], (n, s)
    # the reference's literal index formula on the same data pairs real pulses with other pulses' intensities
    try:
        lit = np.array([[inten[idx[j - i][j - i]] if i < j + 1 else inten[idx[i][0]] for i in range(ns)] for j in range(N)]).T
    except IndexError:
        lit = None   # (it also runs off the index lists for other trains)
    if lit is not None:
        wrong = [(n, s) for n in range(ns) for s in range(N) if win[n, s] > fes.PAD_TIME and lit[n, s] != inten[stim.index(win[n, s])]]
        assert wrong, "the literal formula would have to pair every pulse with its own intensity to be kept"
    const = HmedId._controls(None, [55.0] * 10, idx, ns, N)
    assert (const == 55.0).all()


def test_pulse_on_a_node_belongs_to_that_node():
    """Deviation 2: with dt = 0.3 / 3 the product 1 * dt is 0.09999999999999999 < 0.1: the reference's raw `<=`
    (ding2003.py:416) would start the pulse at 0.1 one node late, yet its golden vectors (tests/shard1/test_ivp.py) are
    reproduced only with the pulse in node 1's window -- which the grid-exact assignment gives."""
    dt = 0.3 / 3
    assert 1 * dt < 0.1                                  # the float trap itself
    win, idx = fes.stim_windows([0, 0.1, 0.2], 3, 0.3, 3)
    assert win[1, -1] == 0.1 and idx[1] == [0, 1]        # node 1 sees the pulse at 0.1
    assert win[2, -1] == 0.2
    ow, oi = fo.stim_windows([0, 0.1, 0.2], 3, 0.3, 3)   # the oracle (pinned to the goldens) agrees
    np.testing.assert_array_equal(win, ow)
    assert idx == oi


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_ivp_fes_scalar_pulse_intensity_is_the_intensity(backend):
    """Deviation 3: IvpFes with a scalar pulse_intensity integrates with THAT intensity in every window slot (the
    reference reads self.pulse_width at ivp_fes.py:339, an unrelated attribute): same trajectory as the one-element list,
    which the reference handles correctly, and as the oracle driven with the scalar."""
    runs = []
    for pi in (50, [50]):
        model = cf.DingModelPulseIntensityFrequency(stim_time=[0, 0.1, 0.2], sum_stim_truncation=3)
        ivp = cf.IvpFes({"model": model, "pulse_intensity": pi},
                        {"final_time": 0.3, "ode_solver": cf.OdeSolver.RK4(n_integration_steps=10)}, device=backend)
        res, _ = ivp.integrate()
        runs.append(np.asarray(res["F"]).ravel())
    np.testing.assert_array_equal(runs[0], runs[1])
    p = fes.default_params(fes.FES_HMED2018)
    win, _ = fes.stim_windows([0, 0.1, 0.2], 3, 0.3, 3)
    ref = foc.ivp(fo.HMED2018, False, 4, 3, 3, 10, 0.1, 0.020, 1.04, p[None], np.zeros((1, 2)), win[None],
                  np.full((1, 3, 3), 50.0), trace_B=1, trace_every=1)
    np.testing.assert_allclose(runs[0], ref["trace"][0, :, 1], rtol=1e-9, atol=1e-12)


def test_identification_bounds_are_validated_and_nonpositive_boxes_sample_uniformly(tmp_path):
    """a min_bound <= 0 used to turn every start but the first into NaN (log of a non-positive number)."""
    import pickle

    path = str(tmp_path / "d.pkl")
    with open(path, "wb") as f:
        pickle.dump({"time": list(np.linspace(0, 1, 11)), "force": [0.0] * 11, "stim_time": [0.0, 0.5]}, f)
    mk = lambda settings: cf.DingModelFrequencyForceParameterIdentification(
        model=cf.DingModelFrequency(), data_path=[path], identification_method="full", final_time=1.0,
        key_parameter_to_identify=["a_rest", "tau2"], additional_key_settings=settings, n_starts=16, device="cpu")
    with pytest.raises(ValueError, match="min_bound"):
        mk({"a_rest": {"min_bound": 5000, "max_bound": 100}})
    with pytest.raises(ValueError, match="initial_guess"):
        mk({"a_rest": {"initial_guess": 1000000}})
    ident = mk({"tau2": {"min_bound": -0.5, "max_bound": 1, "initial_guess": 0.5}})
    P = ident._starting_points(fes.default_params(fes.FES_DING2003))
    assert bool(torch.isfinite(P).all())
    t2 = P[fes.PARAM_INDEX["tau2"]]
    assert float(t2[0]) == 0.5 and float(t2.min()) >= -0.5 and float(t2.max()) <= 1.0 and float(t2.min()) < 0.0
    a = P[fes.PARAM_INDEX["a_rest"]]
    assert float(a.min()) >= 1 and float(a.max()) <= 10000 and float(a[1:].median()) < 2000   # log-uniform: median ~ 100
