"""pytest configuration: the ``gpu`` marker and shared helpers.

``-m "not gpu"`` runs in the build container (no GPU): oracle vs golden vectors, host logic, C-ABI
symbol checks, host emulation of the kernel math, gloo world_size-2 sharding tests.
``-m gpu`` runs on a B200: the parity tests proper, all through the C-ABI library.
"""
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


# ---- which library executes the kernels in a test -----------------------------------------------------
# "cuda": the product, libfes_b200.so on a B200 (tests marked gpu).
# "emul": tests/host_emul/libfes_emul.so, a HOST BUILD OF THE SAME KERNEL SOURCE (fes_core.cuh) used only so
#         that the GPU-less CI can check the kernel arithmetic against the oracle.  It is injected here,
#         explicitly; nothing in cocofest_b200 knows about it.
import ctypes
import subprocess

BACKENDS = [pytest.param("emul", id="emul"), pytest.param("cuda", marks=pytest.mark.gpu, id="cuda")]


def _load_emul():
    from cocofest_b200 import _lib

    d = ROOT / "tests" / "host_emul"
    subprocess.run(["make", "-C", str(d)], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return _lib._bind(ctypes.CDLL(str(d / "libfes_emul.so")))


@pytest.fixture
def backend(request, monkeypatch):
    """Yields the torch device the test should put its tensors on."""
    import torch
    from cocofest_b200 import _lib

    kind = request.param
    if kind == "emul":
        monkeypatch.setattr(_lib, "_LIB", _load_emul())
        yield torch.device("cpu")
    else:
        monkeypatch.setattr(_lib, "_LIB", None)
        _lib.lib()
        yield torch.device("cuda")
