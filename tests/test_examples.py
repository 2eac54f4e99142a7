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
