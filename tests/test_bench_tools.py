"""The measurement tooling that bench.py relies on, checked without a GPU: the static SASS counts the Makefile writes next to
the library describe THAT library (sha256), hold a record for the kernel of every benchmark configuration under the name
bench.py looks it up by, and the recorded headline count agrees with what DESIGN.md section 4 states."""
import hashlib
import json

import pytest
import torch

from cocofest_b200 import _lib, fes
from scripts import bench_configs as bc

COUNTS = _lib.LIB_PATH.parent / "sass_counts.json"


@pytest.mark.skipif(not COUNTS.exists(), reason="sass_counts.json is written by make -C cocofest_b200/csrc (cuobjdump needed)")
def test_static_counts_describe_the_built_library_and_cover_the_bench_kernels():
    data = json.loads(COUNTS.read_text())
    assert data["lib_sha256"] == hashlib.sha256(_lib.LIB_PATH.read_bytes()).hexdigest()
    kernels, src = bc.load_sass_counts()
    assert kernels is not None, src
    wanted = [
        bc.kernel_name("ivp", "f64", fes.FES_DING2003, True, fes.FES_DIR_X0),               # headline
        bc.kernel_name("ivp", "f32", fes.FES_DING2003, True, fes.FES_DIR_X0),
        bc.kernel_name("ivp", "f64", fes.FES_DING2003, True, 0),                             # configs 1, 2fwd
        bc.kernel_name("ivp", "f64", fes.FES_DING2003, True, fes.FES_DIR_P | fes.FES_DIR_PF),
        bc.kernel_name("ivp", "f64", fes.FES_DING2007, False, fes.FES_DIR_P),                # config 3
        bc.kernel_name("ivp", "f64", fes.FES_HMED2018, False, fes.FES_DIR_P),                # config 4
        bc.kernel_name("ident", "f64", fes.FES_HMED2018),                                    # config 4b / 4c
        bc.kernel_name("ident", "f32", fes.FES_HMED2018),
        bc.kernel_name("shoot", "f64", fes.FES_DING2007, True, fes.FES_DIR_X0 | fes.FES_DIR_U),   # config 5
        bc.kernel_name("shoot", "f64", fes.FES_DING2007, True, fes.FES_DIR_P | fes.FES_DIR_PF),   # config 5pf
    ]
    for name in wanted:
        assert name in kernels, name
        rec = kernels[name]
        inst = rec["fp64_inst"] if "double" in name else rec["fp32_inst"]
        assert inst > 50 and rec["mufu"] >= 4 and rec["registers"] <= 255, (name, rec)
    head = kernels[wanted[0]]
    assert head["fp64_inst"] == 247 and head["registers"] <= 164 and head["stack_bytes"] == 0
    assert head["steps_in_path"] == 2, "the trajectory kernels run RK4 steps in pairs (FES_IVP_UNROLL)"
    # the RK4 step of every kernel stays free of local-memory traffic except where DESIGN.md says so (LamCache, Hmed2018)
    dense = kernels[wanted[3]]
    assert dense["fp64_inst"] < 450, "the parameter columns of a fatigue model go through the 2 x 2 step propagator"


def test_counts_of_another_build_are_refused(tmp_path, monkeypatch):
    fake = tmp_path / "libfes_b200.so"
    fake.write_bytes(b"not the library")
    (tmp_path / "sass_counts.json").write_text(json.dumps({"lib_sha256": "0" * 64, "kernels": {}}))
    monkeypatch.setattr(_lib, "LIB_PATH", fake)
    kernels, reason = bc.load_sass_counts()
    assert kernels is None and "another build" in reason
    (tmp_path / "sass_counts.json").unlink()
    kernels, reason = bc.load_sass_counts()
    assert kernels is None and "not found" in reason


def test_flop_count_of_the_bench_kernel_is_the_documented_350():
    import bench

    assert bench.flop_alg_per_step() == 350 and bench.flop_survey_per_step(5, 5) == 978
