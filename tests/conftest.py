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
