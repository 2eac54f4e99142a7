"""fes_shoot_static_f64: the interval kernel with its tables in the library's static device array (include/fes_b200.h).

Collected last on purpose.  Called through the C ABI with ctypes, compared with
fes_shoot_f64 on the same inputs: the same kernels run on the same tables, so the outputs must be IDENTICAL bits -- the
tolerance the parity tests use between two runs of one kernel (tests/test_parity_shoot.py, ShootPlan vs one-shot).  Several
calls in a row with different models and table layouts: the array keeps the previous call's tables, a call that read anything
it had not rebuilt would differ.
"""
import ctypes

import pytest
import torch

from cocofest_b200 import _lib, fes
from tests.conftest import BACKENDS
from tests.problems import random_problem, to_lib

CASES = [  # (model, batch, window length, intervals, seed): Ding2007 keeps two sums per node, Hmed2018 per-slot weights
    ("ding2007_with_fatigue", 300, 6, 7, 11),
    ("hmed2018_with_fatigue", 200, 8, 9, 12),
    ("ding2007_with_fatigue", 70, 10, 33, 13),
    ("ding2003_with_fatigue", 129, 5, 4, 14),
]


def _shoot_through(entry, pb, dev, with_workspace):
    """One call of `entry` (a C-ABI symbol) on the problem; returns (G, Phi, J)."""
    L = _lib.lib()
    params, x0, win, ctrl = to_lib(pb, dev)
    ns, nx, B = pb["ns"], pb["nx"], pb["B"]
    X = (x0[None] * torch.linspace(1.0, 0.7, ns + 1, dtype=torch.float64, device=dev)[:, None, None]).contiguous()
    dirs = fes.FES_DIR_X0 | (fes.FES_DIR_U if pb["nu"] else 0)
    nd = L.fes_num_dirs(pb["kind"], int(pb["fat"]), dirs)
    assert nd > 0
    flags = _lib.FES_WIN_SHARED | (_lib.FES_CTRL_SHARED if ctrl is not None else 0)
    desc = fes._desc(pb["kind"], pb["fat"], 4, pb["N"], ns, 10, dirs, flags, B, pb["dt"], 0.0, None, fes.R0_KM_RELATIONSHIP)
    G = torch.full((ns, nx, B), float("nan"), dtype=torch.float64, device=dev)
    Phi = torch.full((ns, nx, B), float("nan"), dtype=torch.float64, device=dev)
    J = torch.full((ns, nx, nd, B), float("nan"), dtype=torch.float64, device=dev)
    buf = _lib.FesShootBuffers(fes._ptr(params), fes._ptr(X), fes._ptr(win), fes._ptr(ctrl), fes._ptr(G), fes._ptr(Phi),
                               fes._ptr(J), nd, 0, None, 0)
    ws = None
    if with_workspace:
        wbytes = L.fes_workspace_bytes(ctypes.byref(desc), 0)
        assert 0 < wbytes <= _lib.FES_STATIC_SCRATCH_BYTES
        ws = torch.empty(int(wbytes), dtype=torch.uint8, device=dev)
        buf.workspace, buf.workspace_bytes = ws.data_ptr(), wbytes
    rc = getattr(L, entry)(ctypes.byref(desc), ctypes.byref(buf), ctypes.c_void_p(fes._stream_ptr(dev)))
    assert rc == 0, (entry, rc, L.fes_strerror(rc))
    if dev.type == "cuda":
        torch.cuda.synchronize(dev)
    del ws
    return G, Phi, J


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_static_scratch_entry_point_matches_fes_shoot_on_calls_in_a_row(backend):
    dev = backend
    if dev.type == "cuda":
        dev = torch.device("cuda", torch.cuda.current_device())
    free = []
    for rep in range(2):
        for name, B, N, ns, seed in CASES:
            pb = random_problem(name, B=B, N=N, ns=ns, seed=seed + rep)
            ref = _shoot_through("fes_shoot_f64", pb, dev, with_workspace=True)
            got = _shoot_through("fes_shoot_static_f64", pb, dev, with_workspace=False)
            for a, b, what in zip(got, ref, ("G", "Phi", "J")):
                assert not bool(torch.isnan(a).any()), (name, what, "an output element was not written")
                assert torch.equal(a, b), (name, rep, what, float((a - b).abs().max()))
            if dev.type == "cuda":
                free.append(torch.cuda.mem_get_info(dev)[0])
    # the same through the public API (which allocates its own workspace): one more independent reference
    name, B, N, ns, seed = CASES[0]
    pb = random_problem(name, B=B, N=N, ns=ns, seed=seed)
    params, x0, win, ctrl = to_lib(pb, dev)
    X = (x0[None] * torch.linspace(1.0, 0.7, ns + 1, dtype=torch.float64, device=dev)[:, None, None]).contiguous()
    api = fes.shoot(pb["kind"], pb["fat"], params, X, win, pb["dt"], n_steps=10, ctrl=ctrl, want=("x", "u"), endpoints=True)
    G, Phi, J = _shoot_through("fes_shoot_static_f64", pb, dev, with_workspace=False)
    assert torch.equal(G, api.G) and torch.equal(Phi, api.Phi)
    assert torch.equal(J[:, :, :pb["nx"]], api.Jx) and torch.equal(J[:, :, pb["nx"]:], api.Ju)
    if dev.type == "cuda":
        # nothing was taken from the device by the calls themselves: the free memory after the later calls is what it was
        # after the first ones (torch's caching allocator serves the test's own tensors from blocks it already holds)
        assert min(free[len(CASES):]) >= min(free[:len(CASES)]) - (8 << 20), free


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_static_scratch_entry_point_rejects_what_does_not_fit(backend):
    L = _lib.lib()
    # per-scenario windows of a large batch: 33 nodes x 2 sums x 2^20 scenarios x 8 B = 554 MB of tables
    desc = fes._desc(fes.FES_DING2007, True, 4, 10, 33, 10, fes.FES_DIR_X0 | fes.FES_DIR_U, 0, 1 << 20, 1 / 33, 0.0, None,
                     fes.R0_KM_RELATIONSHIP)
    assert L.fes_workspace_bytes(ctypes.byref(desc), 0) > _lib.FES_STATIC_SCRATCH_BYTES
    buf = _lib.FesShootBuffers(None, None, None, None, None, None, None, 0, 0, None, 0)
    rc = L.fes_shoot_static_f64(ctypes.byref(desc), ctypes.byref(buf), None)
    assert rc == -8
    assert b"FES_STATIC_SCRATCH_BYTES" in L.fes_strerror(rc)
    with pytest.raises(_lib.FesLibraryError, match="FES_STATIC_SCRATCH_BYTES"):
        _lib.check(rc, "fes_shoot_static_f64")
    # a descriptor that is wrong anyway is refused with its own code, as by fes_shoot_f64
    desc.n_steps = 0
    assert L.fes_shoot_static_f64(ctypes.byref(desc), ctypes.byref(buf), None) == L.fes_shoot_f64(ctypes.byref(desc), ctypes.byref(buf), None) < 0
