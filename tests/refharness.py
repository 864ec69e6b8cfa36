"""Test helpers: drive the oracle (oracle/_ref/ref_harness = the UNMODIFIED reference CPU build)
and the backend under test through the C ABI on identical inputs.  TEST INFRASTRUCTURE."""
import ctypes as C
import importlib.util
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
DEVSIM_LIB = os.path.join(ROOT, "tests", "devsim", "libcfdnn_devsim.so")


def _load_pkg():
    if "cfdnn_b200" in sys.modules:
        return sys.modules["cfdnn_b200"]
    spec = importlib.util.spec_from_file_location(
        "cfdnn_b200", os.path.join(ROOT, "cfd-nn_b200", "__init__.py"),
        submodule_search_locations=[os.path.join(ROOT, "cfd-nn_b200")])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["cfdnn_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


pkg = _load_pkg()
capi = pkg._capi

backend = pkg.backend
VBC, PBC, TI, SCHEME = backend.VBC, backend.PBC, backend.TI, backend.SCHEME
shapes, interior, rel_l2, FIELD_ID = backend.shapes, backend.interior, backend.rel_l2, backend.FIELD_ID
Backend, tgv_init = backend.Backend, backend.tgv_init


def have_oracle():
    return os.path.exists(HARNESS)


def run_ref(mode, **kw):
    """Run the reference harness; returns (meta dict, out prefix)."""
    args = [HARNESS, f"mode={mode}"] + [f"{k}={v}" for k, v in kw.items()]
    o = subprocess.run(args, capture_output=True, text=True)
    if o.returncode != 0:
        raise RuntimeError(f"ref_harness failed ({o.returncode}): {o.stderr[-2000:]}")
    return json.loads(o.stdout.strip().splitlines()[-1])


def load_state(prefix, Nx, Ny, Nz, Ng=1):
    sh = shapes(Nx, Ny, Nz, Ng)
    out = {}
    for name, key in [("u", "u"), ("v", "v"), ("w", "w"), ("p", "c"), ("pcorr", "c"), ("div", "c"),
                      ("rhs", "c"), ("nut", "c"), ("k", "c"), ("omega", "c")]:
        path = f"{prefix}_{name}.f64"
        if os.path.exists(path):
            out[name] = np.fromfile(path).reshape(sh[key])
    return out


def load_mesh_arrays(pre):
    """y-metric arrays of a stretched mesh as dumped by ref_harness (dump_mesh)"""
    return {k: np.ascontiguousarray(np.fromfile(f"{pre}_{k}.f64")) for k in ("dyv", "dyc", "aS", "aN", "aP", "yc", "yf")}


