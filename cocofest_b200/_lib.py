"""ctypes binding of ``libfes_b200.so`` (C ABI declared in ``include/fes_b200.h``).

The library is the product: there is no Python or CPU implementation of the integration behind it.
``lib()`` raises ``FesLibraryError`` if the shared object has not been built
(``python -c "import __graft_entry__ as g; g.build()"`` or ``make -C cocofest_b200/csrc``).
"""
import ctypes
from ctypes import POINTER, Structure, c_char_p, c_double, c_int, c_int32, c_int64, c_void_p
from pathlib import Path

import os

CSRC = Path(__file__).resolve().parent / "csrc"
# FES_B200_LIB points at another build of the SAME library (launch-shape sweeps, scripts/build_variants.sh)
LIB_PATH = Path(os.environ["FES_B200_LIB"]) if os.environ.get("FES_B200_LIB") else CSRC / "libfes_b200.so"

FES_ABI_VERSION = 1
FES_DING2003, FES_DING2007, FES_HMED2018 = 0, 1, 2
FES_NP = 14
FES_DIR_X0, FES_DIR_U, FES_DIR_P, FES_DIR_W, FES_DIR_PF = 1, 2, 4, 8, 16
FES_DIR_C = 32  # shoot: the force factor fl * fv + fp of the interval as an input row of ctrl and a Jacobian column
FES_TILE = 5
FES_PARAMS_SHARED, FES_WIN_SHARED, FES_CTRL_SHARED, FES_TABLES_READY, FES_TARGET_SHARED = 1, 2, 4, 8, 16
FES_STATIC_SCRATCH_BYTES = 4 * 1024 * 1024  # fes_shoot_static_f64: size of the library-owned static device scratch

PARAM_NAMES = ("a_rest", "km_rest", "tau1_rest", "tau2", "alpha_a", "alpha_tau1", "alpha_km", "tau_fat",
               "pd0", "pdt", "ar", "bs", "Is", "cr")
PARAM_INDEX = {n: i for i, n in enumerate(PARAM_NAMES)}
PARAM_INDEX["a_scale"] = 0  # ding2007: a_scale takes the a_rest slot (ding2007_with_fatigue.py:108,211)


class FesLibraryError(RuntimeError):
    pass


class FesDesc(Structure):
    _fields_ = [
        ("abi_version", c_int32), ("model_kind", c_int32), ("fatigue", c_int32), ("scheme", c_int32),
        ("n_trunc", c_int32), ("n_shooting", c_int32), ("n_steps", c_int32), ("dirs", c_int32),
        ("flags", c_int32), ("tile_offset", c_int32), ("trace_every", c_int32), ("reserved0", c_int32),
        ("batch", c_int64), ("trace_batch", c_int64),
        ("dt", c_double), ("t_start", c_double), ("tauc", c_double), ("r0_km", c_double),
    ]


class FesIvpBuffers(Structure):
    _fields_ = [
        ("params", c_void_p), ("x0", c_void_p), ("stim_win", c_void_p), ("ctrl", c_void_p),
        ("x_final", c_void_p), ("trace", c_void_p), ("sens_final", c_void_p), ("sens_trace", c_void_p),
        ("workspace", c_void_p), ("workspace_bytes", c_int64),
    ]


class FesShootBuffers(Structure):
    _fields_ = [
        ("params", c_void_p), ("X", c_void_p), ("stim_win", c_void_p), ("ctrl", c_void_p),
        ("G", c_void_p), ("Phi", c_void_p), ("J", c_void_p), ("j_cols", c_int32), ("j_col0", c_int32),
        ("workspace", c_void_p), ("workspace_bytes", c_int64),
    ]


class FesIdentBuffers(Structure):
    _fields_ = [
        ("params", c_void_p), ("x0", c_void_p), ("stim_win", c_void_p), ("ctrl", c_void_p), ("target", c_void_p),
        ("weight", c_void_p), ("lower", c_void_p), ("upper", c_void_p), ("params_out", c_void_p), ("cost", c_void_p),
        ("grad", c_void_p), ("hess", c_void_p), ("iters", c_void_p), ("workspace", c_void_p), ("workspace_bytes", c_int64),
    ]


class FesHessBuffers(Structure):
    _fields_ = [
        ("params", c_void_p), ("X", c_void_p), ("stim_win", c_void_p), ("ctrl", c_void_p), ("multipliers", c_void_p),
        ("D2", c_void_p), ("H", c_void_p),
    ]


class FesIdentOptions(Structure):
    _fields_ = [("max_iter", c_int32), ("p_mask", c_int32), ("lambda0", c_double), ("ftol", c_double), ("xtol", c_double)]


# every symbol include/fes_b200.h declares, with its prototype
PROTOTYPES = {
    "fes_version": (c_int, []),
    "fes_strerror": (c_char_p, [c_int]),
    "fes_num_states": (c_int, [c_int]),
    "fes_num_controls": (c_int, [c_int, c_int]),
    "fes_num_dirs": (c_int, [c_int, c_int, c_int]),
    "fes_workspace_bytes": (c_int64, [POINTER(FesDesc), c_int]),
    "fes_ivp_f64": (c_int, [POINTER(FesDesc), POINTER(FesIvpBuffers), c_void_p]),
    "fes_ivp_f32": (c_int, [POINTER(FesDesc), POINTER(FesIvpBuffers), c_void_p]),
    "fes_shoot_f64": (c_int, [POINTER(FesDesc), POINTER(FesShootBuffers), c_void_p]),
    "fes_shoot_f32": (c_int, [POINTER(FesDesc), POINTER(FesShootBuffers), c_void_p]),
    "fes_shoot_static_f64": (c_int, [POINTER(FesDesc), POINTER(FesShootBuffers), c_void_p]),
    "fes_shoot_hess_num_dirs": (c_int, [c_int, c_int, c_int, c_int]),
    "fes_shoot_hess_f64": (c_int, [POINTER(FesDesc), POINTER(FesHessBuffers), c_void_p]),
    "fes_shoot_hess_host_f64": (c_int, [POINTER(FesDesc), POINTER(FesHessBuffers), c_void_p]),
    "fes_ivp_host_f64": (c_int, [POINTER(FesDesc), POINTER(FesIvpBuffers), c_void_p]),
    "fes_ivp_host_f32": (c_int, [POINTER(FesDesc), POINTER(FesIvpBuffers), c_void_p]),
    "fes_shoot_host_f64": (c_int, [POINTER(FesDesc), POINTER(FesShootBuffers), c_void_p]),
    "fes_shoot_host_f32": (c_int, [POINTER(FesDesc), POINTER(FesShootBuffers), c_void_p]),
    "fes_identify_f64": (c_int, [POINTER(FesDesc), POINTER(FesIdentBuffers), POINTER(FesIdentOptions), c_void_p]),
    "fes_identify_f32": (c_int, [POINTER(FesDesc), POINTER(FesIdentBuffers), POINTER(FesIdentOptions), c_void_p]),
    "fes_identify_host_f64": (c_int, [POINTER(FesDesc), POINTER(FesIdentBuffers), POINTER(FesIdentOptions), c_void_p]),
    "fes_identify_host_f32": (c_int, [POINTER(FesDesc), POINTER(FesIdentBuffers), POINTER(FesIdentOptions), c_void_p]),
    "fes_probe_fma_peak": (c_int, [c_int, c_int, c_void_p, POINTER(c_double)]),
    "fes_launch_count": (c_int64, []),
    "fes_last_host_timing": (c_int, [POINTER(c_double)]),
    "fes_host_pool_trim": (c_int, []),
}

_LIB = None


def _bind(cdll, required=PROTOTYPES):
    for name, (res, args) in required.items():
        try:
            fn = getattr(cdll, name)
        except AttributeError as e:
            raise FesLibraryError(f"{cdll._name} does not export {name}") from e
        fn.restype = res
        fn.argtypes = args
    return cdll


def lib():
    """The loaded CUDA library.  Fails loudly when it is missing: there is no fallback."""
    global _LIB
    if _LIB is None:
        if not LIB_PATH.exists():
            raise FesLibraryError(
                f"{LIB_PATH} not found: the sm_100a CUDA library must be built first "
                "(make -C cocofest_b200/csrc, or __graft_entry__.build()); cocofest_b200 has no CPU fallback")
        _LIB = _bind(ctypes.CDLL(str(LIB_PATH)))
        if _LIB.fes_version() != FES_ABI_VERSION:
            raise FesLibraryError("libfes_b200.so ABI version mismatch")
    return _LIB


def is_host_emulation() -> bool:
    """True when the loaded library is the test suite's host build of the kernel source (tests/host_emul), which the
    tests inject explicitly; the product library does not export this symbol."""
    try:
        return bool(lib().fes_is_host_emulation())
    except AttributeError:
        return False


def check(rc: int, what: str):
    if rc != 0:
        msg = lib().fes_strerror(rc)
        raise FesLibraryError(f"{what} failed with code {rc}: {msg.decode() if msg else '?'}")
