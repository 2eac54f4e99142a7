"""The reference-named identification classes end to end, as examples/identification/*_model_identification.py use
them: simulate with IvpFes, write the reference's pickle format, identify, compare with the generating parameters.
`emul` runs on the CPU CI (host build of the kernel source), `cuda` is the product on a B200."""
import pickle

import numpy as np
import pytest

import cocofest_b200 as cf
from cocofest_b200 import identification as ident
from tests.conftest import BACKENDS


def _simulate(tmp_path, model, device, final_time=2, name="sim.pkl", **fes_extra):
    ivp = cf.IvpFes({"model": model, **fes_extra}, {"final_time": final_time}, device=device)
    result, time = ivp.integrate()
    path = str(tmp_path / name)
    data = {"time": time, "force": result["F"][0], "stim_time": model.stim_time}
    for k in ("pulse_width", "pulse_intensity"):
        if k in fes_extra:
            data[k] = fes_extra[k]
    with open(path, "wb") as f:
        pickle.dump(data, f)
    return path


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_ding2003_example(backend, tmp_path):
    """examples/identification/ding2003_model_identification.py (noise-free)."""
    stim_time = list(np.linspace(0, 1, 11)[:-1])
    path = _simulate(tmp_path, cf.DingModelFrequency(stim_time=stim_time, sum_stim_truncation=10), backend)
    ocp = cf.DingModelFrequencyForceParameterIdentification(
        model=cf.DingModelFrequency(stim_time=stim_time, sum_stim_truncation=10), final_time=2, data_path=[path],
        identification_method="full", double_step_identification=False,
        key_parameter_to_identify=["a_rest", "km_rest", "tau1_rest", "tau2"], additional_key_settings={},
        n_starts=8, device=backend)
    got = ocp.force_model_identification()
    truth = cf.DingModelFrequency()
    for key in ("a_rest", "km_rest", "tau1_rest", "tau2"):
        assert got[key] == pytest.approx(getattr(truth, key), rel=1e-5), key
        assert getattr(ocp.model, key) == got[key]
    res = ocp.force_identification_result
    np.testing.assert_allclose(res.force, res.force_data, atol=1e-6 * np.max(res.force_data))
    assert res.all_costs.shape[0] == 8


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_ding2007_example_subset_with_fixed_parameters(backend, tmp_path):
    """examples/identification/ding2007_model_identification.py: identify a subset, the others stay at the model's
    values; the pulse widths come from the pickle."""
    stim_time = list(np.linspace(0, 1, 11)[:-1])
    pw = [0.0003, 0.00035, 0.0004, 0.00045, 0.0005, 0.00055, 0.0006, 0.00025, 0.0002, 0.00038]
    path = _simulate(tmp_path, cf.DingModelPulseWidthFrequency(stim_time=stim_time, sum_stim_truncation=10), backend,
                     pulse_width=pw)
    model = cf.DingModelPulseWidthFrequency(stim_time=stim_time, sum_stim_truncation=10)
    ocp = cf.DingModelPulseWidthFrequencyForceParameterIdentification(
        model=model, final_time=2, data_path=[path], identification_method="full",
        key_parameter_to_identify=["tau1_rest", "tau2", "km_rest", "a_scale", "pd0", "pdt"],
        additional_key_settings={"a_scale": {"initial_guess": 4000}}, n_starts=32, device=backend)
    got = ocp.force_model_identification()
    truth = cf.DingModelPulseWidthFrequency()
    res = ocp.force_identification_result
    assert res.cost < 1e-10 * float(np.sum(res.force_data ** 2) * (2 / (len(res.time) - 1)))
    for key in ("tau1_rest", "km_rest", "tau2"):
        assert got[key] == pytest.approx(getattr(truth, key), rel=1e-3), key
    # a_scale, pd0 and pdt only enter through a_scale (1 - exp(-(pw - pd0)/pdt)): the fitted force curve, not the
    # triple itself, is what the data pin when few distinct widths are used
    np.testing.assert_allclose(res.force, res.force_data, atol=1e-5 * np.max(res.force_data))


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_hmed2018_example(backend, tmp_path):
    """final_time = 1 s so that n_shooting = n_stim = N: the only case where IvpFes' literal control index
    formula (ivp_fes.py:326-342) puts every intensity in the slot of its own pulse."""
    stim_time = list(np.linspace(0, 1, 11)[:-1])
    pi = [30.0, 45.0, 60.0, 75.0, 90.0, 100.0, 110.0, 120.0, 70.0, 50.0]
    path = _simulate(tmp_path, cf.DingModelPulseIntensityFrequency(stim_time=stim_time, sum_stim_truncation=10), backend,
                     final_time=1, pulse_intensity=pi)
    ocp = cf.DingModelPulseIntensityFrequencyForceParameterIdentification(
        model=cf.DingModelPulseIntensityFrequency(stim_time=stim_time, sum_stim_truncation=10), final_time=1,
        data_path=[path], identification_method="full", key_parameter_to_identify=["a_rest", "km_rest", "tau1_rest", "tau2"],
        additional_key_settings={}, n_starts=8, device=backend)
    got = ocp.force_model_identification()
    truth = cf.DingModelPulseIntensityFrequency()
    for key in ("a_rest", "km_rest", "tau1_rest", "tau2"):
        assert got[key] == pytest.approx(getattr(truth, key), rel=1e-5), key
    assert ocp.model.ar == truth.ar  # not identified: untouched


def test_data_extraction_and_input_checks(tmp_path):
    """identification_method.py:5-78, 253-278 and the constructor checks (ding2003_...:177-275)."""
    for i, t0 in enumerate((0.0, 0.5)):
        with open(tmp_path / f"d{i}.pkl", "wb") as f:
            pickle.dump({"time": [[t0 + 0.0, t0 + 0.5, t0 + 1.0]], "force": [[0.0, 1.0 + i, 2.0]],
                         "stim_time": [t0, t0 + 0.5]}, f)
    paths = [str(tmp_path / "d0.pkl"), str(tmp_path / "d1.pkl")]
    time, stim, force, disc = ident.full_data_extraction(paths)
    assert force == [0.0, 1.0, 2.0, 0.0, 2.0, 2.0]
    assert time == [0.0, 0.5, 1.0, 1.0, 1.5, 2.0]
    # second file: rebased to its first pulse, then shifted by the previous file's last time TWICE (reference :52-59)
    assert disc == [2] and stim == [0.0, 0.5, 2.0, 2.5]
    assert ident.force_at_node_in_ocp([0, 1, 2], [0, 10, 20], 4, 2) == [0.0, 5.0, 10.0, 15.0, 20.0]
    with pytest.raises(NotImplementedError):
        ident.sparse_data_extraction(paths)
    model = cf.DingModelFrequency(stim_time=[0, 0.5], sum_stim_truncation=10)
    kw = dict(model=model, data_path=paths, key_parameter_to_identify=["a_rest"], additional_key_settings={}, device="cpu")
    with pytest.raises(ValueError, match="fatigue"):
        cf.DingModelFrequencyForceParameterIdentification(**{**kw, "model": cf.DingModelFrequencyWithFatigue(stim_time=[0, 0.5])})
    with pytest.raises(ValueError, match="identification method"):
        cf.DingModelFrequencyForceParameterIdentification(**kw, identification_method="partial")
    with pytest.raises(TypeError, match="bool"):
        cf.DingModelFrequencyForceParameterIdentification(**kw, double_step_identification="no")
    with pytest.raises(ValueError, match="key_parameter_to_identify"):
        cf.DingModelFrequencyForceParameterIdentification(**{**kw, "key_parameter_to_identify": ["pd0"]})
    with pytest.raises(TypeError, match="pkl"):
        cf.DingModelFrequencyForceParameterIdentification(**{**kw, "data_path": ["x.csv"]})
    with pytest.raises(TypeError):
        cf.DingModelFrequencyForceParameterIdentification(**{**kw, "additional_key_settings": {"a_rest": {"initial_guess": "1"}}})


def test_pickle_formats_round_trip(tmp_path):
    """identification pickles and the SolutionToPickle dictionary (result/pickle.py:25-36)."""
    from cocofest_b200 import pickle_io as pio

    t = np.linspace(0, 1, 11)
    for i in range(3):
        pio.save_identification_data(str(tmp_path / f"s{i}.pkl"), t, (i + 1) * t, [0.0, 0.5], pulse_width=[3e-4, 4e-4])
    d = pio.load_identification_data(str(tmp_path / "s1.pkl"))
    assert set(d) == {"time", "force", "stim_time", "pulse_width"}
    target, stim = pio.batch_targets([str(tmp_path / f"s{i}.pkl") for i in range(3)], 4, 1.0)
    assert target.shape == (5, 3) and stim == [0.0, 0.5]
    np.testing.assert_allclose(target[:, 2], 3 * np.linspace(0, 1, 5))
    with open(tmp_path / "bad.pkl", "wb") as f:
        pickle.dump({"time": [0]}, f)
    with pytest.raises(KeyError):
        pio.load_identification_data(str(tmp_path / "bad.pkl"))
    sol = {"time": np.array([0.0, 0.1, 0.1, 0.2]), "states": {"F": np.array([[0.0, 1.0, 1.0, 2.0]]), "Cn": np.array([0.0, 0.5, 0.5, 0.7])},
           "parameters": {"a_rest": 3009.0}, "parameters_bounds": {"a_rest": (1, 10000)}, "time_to_optimize": 0.01}
    pio.SolutionToPickle(sol, "sol.pkl", str(tmp_path) + "/").pickle()
    with open(tmp_path / "sol.pkl", "rb") as f:
        out = pickle.load(f)
    assert list(out) == ["time", "states", "control", "parameters", "parameters_bounds", "time_to_optimize", "bio_model_path"]
    np.testing.assert_array_equal(out["time"], [0.0, 0.1, 0.2])
    np.testing.assert_array_equal(out["states"]["F"], [0.0, 1.0, 2.0])
