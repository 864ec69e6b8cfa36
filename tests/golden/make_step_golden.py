#!/usr/bin/env python
"""tests/golden/make_step_golden.py — regenerate reference_fields.npz from the UNMODIFIED reference CPU build
(oracle/_ref/ref_harness, built by oracle/Makefile from /root/reference).  Run in the build container:
    python tests/golden/make_step_golden.py
The fixtures pin the parity tests to reference outputs even where the oracle binary is not available:
  step_*  : ghost-inclusive fields after the case's steps (tests/cases.py STEP_CASES entries named below)
  mg_*    : pressure returned by MultigridPoissonSolver::solve() + cycle count (MG_CASES entries)
Inputs are rebuilt by the tests from tests/cases.py (step_init / mg_inputs are pure functions of N)."""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import cases  # noqa: E402

STEP = ("tgv16_rk3_skew_fixed8", "tgv16_euler_conv", "channel16_rk3_walls_y", "box16_euler_walls_all")
MG = ("per32_conv1e8", "mixed_dirichlet_fixed4", "neumann_16x32x16_conv")

out = {}
for c in cases.STEP_CASES:
    if c[0] in STEP:
        with tempfile.TemporaryDirectory() as td:
            meta, ref = cases.run_step_ref(c, td)
        for n in cases.FIELDS:
            out[f"step_{c[0]}_{n}"] = ref[n]
        out[f"step_{c[0]}_cycles"] = np.array(meta["cycles"], dtype=np.int64)
        out[f"step_{c[0]}_residuals"] = np.array(meta["residuals"])
for c in cases.MG_CASES:
    if c[0] in MG:
        with tempfile.TemporaryDirectory() as td:
            meta, p = cases.run_mg_ref(c, td)
        out[f"mg_{c[0]}_p"] = p
        out[f"mg_{c[0]}_cycles"] = np.array([meta["cycles"]], dtype=np.int64)
np.savez_compressed(os.path.join(HERE, "reference_fields.npz"), **out)
print("wrote", len(out), "arrays")
