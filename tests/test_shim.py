"""The C++ drop-in binding (include/cfdnn_b200_shim.hpp): a reference RANSSolver /
MultigridPoissonSolver and the B200 shim classes side by side in ONE C++ program
(tests/shim/shim_parity.cpp), same Mesh / Config / VelocityBC / VectorField objects."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "tests", "shim", "_build")


def _ensure_built():
    exe = os.path.join(BUILD, "shim_parity_sim")
    if not os.path.exists(exe):
        if not os.path.isdir("/root/reference/include"):
            pytest.skip("shim parity program is built where /root/reference exists")
        subprocess.run(["bash", os.path.join(ROOT, "tests", "shim", "build.sh")], check=True)


def _run(exe, *args, env=None):
    r = subprocess.run([os.path.join(BUILD, exe), *map(str, args)], capture_output=True, text=True, timeout=900,
                       env=dict(os.environ, **(env or {})))
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    return r.returncode, json.loads(line)


def _weights_dir(td):
    import nnmodels as M
    return M.write_text_dir("tbnn_channel_caseholdout", os.path.join(td, "m"))


def _check_closure(d):
    assert d["les_wale"]["u"] <= 1e-10 and d["les_wale"]["nu_t"] <= 1e-6, d["les_wale"]
    assert d["sst"]["u"] <= 1e-10 and d["sst"]["k"] <= 1e-10 and d["sst"]["nu_t"] <= 1e-9, d["sst"]     # native SST k-omega
    c = d["closure"]
    assert c and c["ok"], c
    assert c["interface_ok"] and c["views_ok"]
    for k in ("b200_solver_b200_model", "ref_solver_b200_model", "b200_solver_ref_model_bridged"):
        assert c[k]["nu_t"] <= 1e-6 and c[k]["u"] <= 1e-9, (k, c[k])
    assert c["device_view_vs_host_update"] <= 1e-12


def test_shim_matches_reference_classes_on_test_double(oracle, sim, tmp_path):
    _ensure_built()
    # plain one-sweep smoother on the CPU test double (the blocked kernels run as ~500 fibers per CTA there: slow)
    rc, d = _run("shim_parity_sim", 16, 2, _weights_dir(str(tmp_path)), env={"CFDNN_MG_TB": "0"})
    assert rc == 0 and d["ok"], d
    _check_closure(d)
    assert d["cycles_ref"] == d["cycles_b200"] and d["mg_cycles_ref"] == d["mg_cycles_b200"]
    assert d["u_rel_l2"] <= 1e-10 and d["p_rel_l2"] <= 1e-10 and d["mg_p_rel_l2"] <= 1e-10
    assert d["invalid_argument"]


@pytest.mark.gpu
def test_shim_matches_reference_classes_on_gpu(gpu):
    if not os.path.exists(os.path.join(BUILD, "shim_parity_gpu")):
        pytest.fail("tests/shim/_build/shim_parity_gpu missing (built by __graft_entry__.build())")
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        rc, d = _run("shim_parity_gpu", 32, 3, _weights_dir(td))
    assert rc == 0 and d["ok"], d
    _check_closure(d)
    assert d["u_rel_l2"] <= 1e-10 and d["p_rel_l2"] <= 1e-10 and d["mg_p_rel_l2"] <= 1e-10
