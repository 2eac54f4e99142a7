"""ctypes loader for oracle/libfes_oracle.so (the C restatement)  --  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
Layouts are the oracle's own (batch OUTERMOST, row-major), see fes_oracle.c.
"""
import ctypes
import os
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB = None


def build(force: bool = False) -> Path:
    so = _HERE / "libfes_oracle.so"
    if force or not so.exists() or so.stat().st_mtime < (_HERE / "fes_oracle.c").stat().st_mtime:
        env = dict(os.environ)
        env.pop("CC", None)
        subprocess.run(["make", "-C", str(_HERE), "clean", "all"], check=True, env=env,
                       stdout=subprocess.DEVNULL)
    return so


def _bind(cdll):
    cdll.fes_oracle_ivp.restype = ctypes.c_int
    cdll.fes_oracle_shoot.restype = ctypes.c_int
    cdll.fes_oracle_shoot_fc.restype = ctypes.c_int
    cdll.fes_oracle_max_threads.restype = ctypes.c_int
    return cdll


def lib():
    global _LIB
    if _LIB is None:
        _LIB = _bind(ctypes.CDLL(str(build())))
    return _LIB


TIMED_CFLAGS = ["-O3", "-march=native", "-fopenmp", "-fPIC", "-std=c11"]
_TIMED = None


def timed_lib():
    """The SAME source compiled for speed (-O3 -march=native, fp contraction allowed) on the machine that runs the
    benchmark: bench.py's timed CPU baseline.  The parity oracle stays the strict -O2 -ffp-contract=off build of
    `lib()`.  Built into oracle/_timed/ (git-ignored); returns (cdll, flags)."""
    global _TIMED
    if _TIMED is None:
        out = _HERE / "_timed"
        out.mkdir(exist_ok=True)
        so = out / "libfes_oracle_timed.so"
        cc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"   # the image's CC wrapper has no libgomp.spec
        subprocess.run([cc, *TIMED_CFLAGS, "-shared", "-o", str(so), str(_HERE / "fes_oracle.c"), "-lm"], check=True)
        _TIMED = _bind(ctypes.CDLL(str(so)))
    return _TIMED, " ".join(TIMED_CFLAGS)


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _c(a, dtype=np.float64):
    return None if a is None else np.ascontiguousarray(a, dtype=dtype)


def max_threads() -> int:
    return int(lib().fes_oracle_max_threads())


def ivp(kind, fatigue, scheme, N, ns, S, dt, tauc, r0_km, params, x0, win, u=None, seed_x0=False, pdir=(),
        trace_B=0, trace_every=None, sens_trace=False, n_threads=0, cdll=None):
    """params [B|1, 14]; x0 [B, nx]; win [B|1, ns+1, N]; u [B|1, ns, nu] -> dict(final, trace, sens_final, sens_trace).
    cdll: another build of the same source (timed_lib), default the strict parity build."""
    nx = 5 if fatigue else 2
    x0 = _c(x0).reshape(-1, nx)
    B = x0.shape[0]
    params = _c(params).reshape(-1, 14)
    win = _c(win).reshape(-1, ns + 1, N)
    nu = 0
    if u is not None:
        u = _c(u)
        nu = u.shape[-1]
        u = u.reshape(-1, ns, nu)
    pdir_a = np.asarray(list(pdir), dtype=np.int32)
    nd = (nx if seed_x0 else 0) + len(pdir_a)
    final = np.empty((B, nx))
    trace = None
    if trace_B:
        trace_every = S if trace_every is None else trace_every
        npts = ns * S + 1 if trace_every == 1 else ns + 1
        trace = np.empty((trace_B, npts, nx))
    else:
        trace_every = S
    sfin = np.empty((B, nx, nd)) if nd else None
    strace = np.empty((trace_B, ns + 1, nx, nd)) if (nd and sens_trace and trace_B) else None
    rc = (cdll or lib()).fes_oracle_ivp(
        ctypes.c_int(kind), ctypes.c_int(int(fatigue)), ctypes.c_int(scheme), ctypes.c_int(N), ctypes.c_int(ns),
        ctypes.c_int(S), ctypes.c_double(dt), ctypes.c_double(tauc), ctypes.c_double(r0_km), ctypes.c_longlong(B),
        _p(params), ctypes.c_int(int(params.shape[0] == 1 and B != 1 or params.shape[0] == 1)),
        _p(x0), _p(win), ctypes.c_int(int(win.shape[0] == 1)), _p(u),
        ctypes.c_int(int(u is not None and u.shape[0] == 1)), ctypes.c_int(nu),
        ctypes.c_int(int(seed_x0)), ctypes.c_int(len(pdir_a)), _p(pdir_a),
        _p(final), _p(trace), ctypes.c_longlong(trace_B), ctypes.c_int(trace_every),
        _p(sfin), _p(strace), ctypes.c_int(n_threads))
    if rc != 0:
        raise RuntimeError(f"fes_oracle_ivp rc={rc}")
    return {"final": final, "trace": trace, "sens_final": sfin, "sens_trace": strace}


def shoot(kind, fatigue, scheme, N, ns, S, dt, t_start, tauc, r0_km, params, X, win, U=None, pdir=(),
          want_x=True, want_u=False, want_w=False, n_threads=0, force_factor=None):
    """X [B, ns+1, nx]; win [B|1, ns+1, N]; U [B|1, ns, nu] -> dict(G, Jx, Ju, Jp, Jw[, Jc]).
    force_factor [B|1, ns]: fl*fv + fp of f_dot_fun (ding2003.py:309-311) per interval; adds Jc [B, ns, nx]."""
    nx = 5 if fatigue else 2
    X = _c(X).reshape(-1, ns + 1, nx)
    B = X.shape[0]
    params = _c(params).reshape(-1, 14)
    win = _c(win).reshape(-1, ns + 1, N)
    nu = 0
    if U is not None:
        U = _c(U)
        nu = U.shape[-1]
        U = U.reshape(-1, ns, nu)
    pdir_a = np.asarray(list(pdir), dtype=np.int32)
    G = np.empty((B, ns, nx))
    Jx = np.empty((B, ns, nx, nx)) if want_x else None
    Ju = np.empty((B, ns, nx, nu)) if (want_u and nu) else None
    Jp = np.empty((B, ns, nx, len(pdir_a))) if len(pdir_a) else None
    Jw = np.empty((B, ns, nx, N)) if want_w else None
    args = [
        ctypes.c_int(kind), ctypes.c_int(int(fatigue)), ctypes.c_int(scheme), ctypes.c_int(N), ctypes.c_int(ns),
        ctypes.c_int(S), ctypes.c_double(dt), ctypes.c_double(t_start), ctypes.c_double(tauc),
        ctypes.c_double(r0_km), ctypes.c_longlong(B), _p(params), ctypes.c_int(int(params.shape[0] == 1)),
        _p(X), _p(win), ctypes.c_int(int(win.shape[0] == 1)), _p(U),
        ctypes.c_int(int(U is not None and U.shape[0] == 1)), ctypes.c_int(nu),
        ctypes.c_int(len(pdir_a)), _p(pdir_a), _p(G), _p(Jx), _p(Ju), _p(Jp), _p(Jw), ctypes.c_int(n_threads)]
    out = {"G": G, "Jx": Jx, "Ju": Ju, "Jp": Jp, "Jw": Jw}
    if force_factor is None:
        rc = lib().fes_oracle_shoot(*args)
    else:
        fc = _c(force_factor).reshape(-1, ns)
        out["Jc"] = np.empty((B, ns, nx))
        rc = lib().fes_oracle_shoot_fc(*args, _p(fc), ctypes.c_int(int(fc.shape[0] == 1)), _p(out["Jc"]))
    if rc != 0:
        raise RuntimeError(f"fes_oracle_shoot rc={rc}")
    return out
