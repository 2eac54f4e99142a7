"""The drop-in boundary used from plain C (tests/c_abi/abi_client.c): libfes_b200.so is self-contained (no torch,
no Python), refuses bad arguments with negative codes, fails loudly without a device, and on a B200 reproduces the
reference's own golden force trajectory (tests/shard1/test_ivp.py:28-44) through fes_ivp_host_f64."""
import json
import subprocess
from pathlib import Path

import numpy as np
import pytest
import torch

from tests.conftest import ROOT

SRC = ROOT / "tests" / "c_abi" / "abi_client.c"
EXE = ROOT / "tests" / "c_abi" / "abi_client"
LIBDIR = ROOT / "cocofest_b200" / "csrc"
GOLD = json.loads((Path(__file__).parent / "golden" / "ivp_golden.json").read_text())


@pytest.fixture(scope="module")
def client():
    assert (LIBDIR / "libfes_b200.so").exists(), "build first: python -c 'import __graft_entry__ as g; g.build()'"
    subprocess.run(["gcc", "-O1", "-Wall", "-Wextra", "-Werror", "-std=c11", f"-I{ROOT / 'include'}", "-o", str(EXE), str(SRC),
                    f"-L{LIBDIR}", "-lfes_b200", f"-Wl,-rpath,{LIBDIR}"], check=True)
    return str(EXE)


def test_library_is_self_contained_and_refuses_bad_arguments(client):
    deps = subprocess.run(["ldd", str(LIBDIR / "libfes_b200.so")], capture_output=True, text=True, check=True).stdout
    assert "torch" not in deps and "python" not in deps.lower()
    r = subprocess.run([client, "--errors"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "refused 13 of 13" in r.stdout


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the behaviour WITHOUT a device")
def test_plain_c_call_fails_loudly_without_a_device(client):
    r = subprocess.run([client], capture_output=True, text=True)
    assert r.returncode == 10 and "fes_ivp_host_f64 ->" in r.stderr and r.stdout == ""


@pytest.mark.gpu
def test_plain_c_call_reproduces_the_reference_golden(client):
    r = subprocess.run([client], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    force = np.array([float(x) for x in r.stdout.split()])
    np.testing.assert_almost_equal(force, np.array(GOLD["F"]["ding2003_with_fatigue"]), decimal=GOLD["decimal"])


@pytest.mark.gpu
def test_plain_c_call_with_force_factor_matches_the_python_surface(client):
    """fes_shoot_host_f64 with FES_DIR_X0 | FES_DIR_C from C: end points and dPhi/dfactor are those of fes.shoot(force_factor=)."""
    from cocofest_b200 import fes

    r = subprocess.run([client, "--factor"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    vals = np.array([float(x) for x in r.stdout.split()]).reshape(2, 3, 5)
    dev = torch.device("cuda", 0)
    pad = float(fes.PAD_TIME)
    win = torch.tensor([[pad, pad, 0.0], [pad, 0.0, 0.1], [0.0, 0.1, 0.2]], dtype=torch.float64, device=dev)
    p = torch.tensor(fes.default_params(fes.FES_DING2003), device=dev)
    x0 = torch.tensor([0, 0, 3009, 0.050957, 0.103], dtype=torch.float64, device=dev)
    X = x0[None, :, None].repeat(4, 1, 1).contiguous()
    res = fes.shoot(fes.FES_DING2003, True, p, X, win, 0.3 / 3, n_steps=10, endpoints=True,   # the client's dt = 0.3 / 3
                    force_factor=torch.tensor([1.0, 0.8, 1.15], dtype=torch.float64, device=dev))
    assert np.array_equal(vals[0], res.Phi[:, :, 0].cpu().numpy())
    assert np.array_equal(vals[1], res.Jc[:, :, 0].cpu().numpy())


@pytest.mark.gpu
def test_plain_c_call_of_the_second_order_entry_point_matches_the_python_surface(client):
    """fes_shoot_hess_host_f64 from C: the contracted blocks are those of fes.shoot_hessian(multipliers=1)."""
    from cocofest_b200 import fes

    r = subprocess.run([client, "--hess"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    vals = np.array([float(x) for x in r.stdout.split()]).reshape(3, 15)
    dev = torch.device("cuda", 0)
    pad = float(fes.PAD_TIME)
    win = torch.tensor([[pad, pad, 0.0], [pad, 0.0, 0.1], [0.0, 0.1, 0.2]], dtype=torch.float64, device=dev)
    p = torch.tensor(fes.default_params(fes.FES_DING2003), device=dev)
    X = torch.tensor([0, 0, 3009, 0.050957, 0.103], dtype=torch.float64, device=dev)[None, :, None].repeat(4, 1, 1).contiguous()
    X[:3, 0, 0] = 0.2
    X[:3, 1, 0] = torch.tensor([30.0, 40.0, 50.0], dtype=torch.float64, device=dev)
    lam = torch.ones((3, 5, 1), dtype=torch.float64, device=dev)
    res = fes.shoot_hessian(fes.FES_DING2003, True, p, X, win, 0.3 / 3, n_steps=10, multipliers=lam, second=False)
    assert np.array_equal(vals, res.H[:, :, 0].cpu().numpy())
    assert np.abs(vals).max() > 0
