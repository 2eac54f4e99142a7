"""The example scripts run as a user would run them (GPU box only: they integrate on the device)."""
import subprocess
import sys

import pytest

from tests.conftest import ROOT


@pytest.mark.gpu
def test_model_integration_example_runs():
    r = subprocess.run([sys.executable, str(ROOT / "examples" / "getting_started" / "model_integration.py")],
                       capture_output=True, text=True, cwd=str(ROOT), env={**__import__("os").environ, "PYTHONPATH": str(ROOT)})
    assert r.returncode == 0, r.stderr[-2000:]
    assert "single muscle: 991 points" in r.stdout and "population: final force" in r.stdout


def _run(*parts, args=()):
    return subprocess.run([sys.executable, str(ROOT.joinpath("examples", *parts)), *args], capture_output=True, text=True,
                          cwd=str(ROOT), env={**__import__("os").environ, "PYTHONPATH": str(ROOT)})


@pytest.mark.gpu
def test_pulse_mode_example_runs():
    r = _run("getting_started", "pulse_mode_example.py")
    assert r.returncode == 0, r.stderr[-2000:]
    assert "single  : 10 pulses, n_shooting =   10" in r.stdout and "triplet : 30 pulses" in r.stdout


@pytest.mark.gpu
def test_identification_example_runs_for_the_three_models():
    r = _run("identification", "model_identification.py", args=("--n-starts", "64"))
    assert r.returncode == 0, r.stderr[-2000:]
    for name in ("ding2003:", "ding2007:", "hmed2018:"):
        assert name in r.stdout


@pytest.mark.gpu
def test_nmpc_batch_example_runs():
    r = _run("getting_started", "nmpc_cyclic_batch.py", args=("--scenarios", "2048", "--cycles", "2"))
    assert r.returncode == 0, r.stderr[-2000:]
    assert "cycle 1: 2048 scenarios x 33 intervals" in r.stdout


@pytest.mark.gpu
def test_truncation_sweep_example_runs(tmp_path):
    r = _run("sensitivity", "truncation_sweep.py", args=("--max-stim", "15", "--modes", "single,triplet", "--out-dir", str(tmp_path)))
    assert r.returncode == 0, r.stderr[-2000:]
    assert "total: 240 points" in r.stdout and (tmp_path / "truncation_triplet.pkl").exists()


@pytest.mark.gpu
def test_force_factor_example_runs():
    r = _run("dynamics", "force_length_velocity_factor.py", args=("--scenarios", "4096"))
    assert r.returncode == 0, r.stderr[-2000:]
    assert "with    relationships" in r.stdout and "through Jc" in r.stdout


@pytest.mark.gpu
def test_node_shooting_comparison_example_runs():
    r = _run("sensitivity", "node_shooting_comparison.py", args=("--starts", "1024"))
    assert r.returncode == 0, r.stderr[-2000:]
    assert "observed order of RK4" in r.stdout and "errors agree within a factor" in r.stdout
