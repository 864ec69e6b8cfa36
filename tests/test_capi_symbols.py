"""The drop-in boundary without a GPU: libcfdnn_b200.so loads, exports every function include/cfdnn_b200.h declares, the ctypes
binding declares every one of them, and — there being no CPU path — a compute entry fails loudly when no device is
visible.  No kernel is launched here."""
import os
import re

import pytest

import refharness as R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "cfdnn_b200.h")
ERR_NO_DEVICE = 3       # CFDNN_ERR_NO_DEVICE


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)                 # comments out
    names = re.findall(r"\b(cfdnn_[a-z0-9_]+)\s*\(", text)
    # typedef'd struct tags are not followed by '(' so everything found is a function declaration
    return sorted(set(names))


# The product library is inspected in a child process: the CPU test double exports the same C names, and two libraries
# with the same global symbols in one process would interpose each other.
PROBE = r"""
import ctypes as C, json, sys
sys.path.insert(0, sys.argv[1])
import refharness as R
lib = R.capi.load()
names = json.loads(sys.argv[2])
out = {"missing": [n for n in names if not hasattr(lib, n)], "version": lib.cfdnn_version(), "devices": lib.cfdnn_device_count()}
if out["devices"] == 0:
    md = R.capi.MeshDesc(Nx=8, Ny=8, Nz=8, Ng=1, dx=0.1, dy=0.1, dz=0.1, Nz_global=8)
    cfg = R.capi.Config()
    lib.cfdnn_config_default(C.byref(cfg))
    h = C.c_void_p()
    out["mg_create"] = lib.cfdnn_mg_create(C.byref(md), C.byref(h))
    out["mg_handle"] = bool(h.value)
    out["error"] = lib.cfdnn_last_error().decode()
    out["solver_create"] = lib.cfdnn_solver_create(C.byref(md), C.byref(cfg), C.byref(h))
    out["fft_create"] = lib.cfdnn_fft_create(C.byref(md), C.byref(h))
print(json.dumps(out))
"""


@pytest.fixture(scope="module")
def probe():
    import json
    import subprocess
    import sys
    if not os.path.exists(R.capi.DEFAULT_LIB):
        pytest.fail("cfd-nn_b200/libcfdnn_b200.so is missing: python -c 'import __graft_entry__ as g; g.build()'")
    r = subprocess.run([sys.executable, "-c", PROBE, os.path.join(ROOT, "tests"), json.dumps(declared_functions())],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]                        # AttributeError if a bound symbol is not exported
    return json.loads(r.stdout.strip().splitlines()[-1])


def test_header_declares_the_entry_points_the_docs_name():
    names = declared_functions()
    assert len(names) >= 60
    for must in ("cfdnn_solver_create", "cfdnn_solver_step", "cfdnn_solver_run", "cfdnn_solver_compute_adaptive_dt",
                 "cfdnn_mg_solve", "cfdnn_mg_solve_dev", "cfdnn_nn_update", "cfdnn_nn_update_dev", "cfdnn_fft_solve_dev",
                 "cfdnn_les_update", "cfdnn_solver_set_rans_model", "cfdnn_solver_comm_init", "cfdnn_adaptive_dt_noalloc"):
        assert must in names, must


def test_library_exports_every_declared_function(probe):
    assert not probe["missing"], f"declared in include/cfdnn_b200.h but not exported by the library: {probe['missing']}"
    assert isinstance(probe["version"], int)


def test_binding_declares_every_function_of_the_header():
    unbound = [n for n in declared_functions() if n not in R.capi._SIGS]
    assert not unbound, f"no ctypes prototype in cfd-nn_b200/_capi.py for: {unbound}"
    stale = [n for n in R.capi._SIGS if n not in declared_functions()]
    assert not stale, f"bound but not declared in the header: {stale}"


def test_no_cpu_fallback_when_no_device_is_visible(probe):
    if probe["devices"] > 0:
        pytest.skip("a CUDA device is visible: the failure mode of a GPU-less host cannot be shown here")
    assert probe["mg_create"] == ERR_NO_DEVICE and not probe["mg_handle"]
    assert "device" in probe["error"].lower()
    assert probe["solver_create"] == ERR_NO_DEVICE and probe["fft_create"] == ERR_NO_DEVICE
