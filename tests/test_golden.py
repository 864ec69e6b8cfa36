"""Parity against COMMITTED reference outputs (tests/golden/reference_fields.npz, written by
tests/golden/make_step_golden.py from the unmodified reference CPU build).  No oracle binary is needed at
run time: the CPU test double is checked in the `not gpu` suite, the product library on the B200."""
import os

import numpy as np
import pytest

import cases

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_fields.npz"))
STEP = [c for c in cases.STEP_CASES if f"step_{c[0]}_u" in GOLD.files]
MG = [c for c in cases.MG_CASES if f"mg_{c[0]}_p" in GOLD.files]


def _check_step(B, case, tol):
    meta, got = cases.run_step_backend(B, case)
    ref = {n: GOLD[f"step_{case[0]}_{n}"] for n in cases.FIELDS}
    if len(GOLD[f"step_{case[0]}_cycles"]):
        assert list(meta["cycles"]) == list(GOLD[f"step_{case[0]}_cycles"])      # same MG iteration counts
    assert meta["residuals"] == pytest.approx(list(GOLD[f"step_{case[0]}_residuals"]), rel=1e-9)
    cases.compare_states(got, ref, tol, case[0])                                  # BASELINE gate: 1e-10 after one step


def _check_mg(B, case, tol):
    st, p = cases.run_mg_backend(B, case)
    pref = GOLD[f"mg_{case[0]}_p"]
    assert st["cycles"] == int(GOLD[f"mg_{case[0]}_cycles"][0])
    assert np.linalg.norm(p - pref) <= tol * np.linalg.norm(pref)


def test_golden_file_is_complete():
    assert len(STEP) == 4 and len(MG) == 3


@pytest.mark.parametrize("case", STEP, ids=[c[0] for c in STEP])
def test_step_matches_golden_on_test_double(sim, monkeypatch, case):
    monkeypatch.setenv("CFDNN_MG_TB", "0")        # plain smoother on the fiber-based test double (speed)
    _check_step(sim, case, 1e-10)


@pytest.mark.parametrize("case", MG, ids=[c[0] for c in MG])
def test_mg_matches_golden_on_test_double(sim, monkeypatch, case):
    monkeypatch.setenv("CFDNN_MG_TB", "0")
    _check_mg(sim, case, 1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize("case", STEP, ids=[c[0] for c in STEP])
def test_step_matches_golden_on_gpu(gpu, case):
    _check_step(gpu, case, 1e-10)


@pytest.mark.gpu
@pytest.mark.parametrize("case", MG, ids=[c[0] for c in MG])
def test_mg_matches_golden_on_gpu(gpu, case):
    _check_mg(gpu, case, 1e-12)
