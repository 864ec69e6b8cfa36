#!/usr/bin/env python
"""scripts/mgbench.py — time fixed-cycle multigrid solves on an N^3 periodic box through the C ABI.
usage: python scripts/mgbench.py N [cycles] [reps] [bc]   (env CFDNN_MG_TB selects the smoother variant)
Prints ms per V-cycle (CUDA events inside the library) and the algorithmic GB/s (500 B/cell/V-cycle)."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import refharness as R  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
cycles = int(sys.argv[2]) if len(sys.argv) > 2 else 4
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
bc = sys.argv[4] if len(sys.argv) > 4 else "PPPPPP"
B = R.Backend()
capi = R.capi
md = B.mesh(N, N, N)
h = C.c_void_p()
B.check(B.lib.cfdnn_mg_create(C.byref(md), C.byref(h)))
B.check(B.lib.cfdnn_mg_set_bc(h, (C.c_int * 6)(*[R.PBC[c] for c in bc])))
rng = np.random.default_rng(0)
rhs = np.zeros((N + 2, N + 2, N + 2))
r = rng.standard_normal((N, N, N)); r -= r.mean()
rhs[1:-1, 1:-1, 1:-1] = r
cfg = capi.PoissonCfg()
B.lib.cfdnn_poisson_cfg_default(C.byref(cfg))
cfg.fixed_cycles = cycles
st = capi.PoissonStats()
best = 1e30
for it in range(reps):
    p = np.zeros_like(rhs)
    B.check(B.lib.cfdnn_mg_solve(h, capi.dptr(rhs), capi.dptr(p), C.byref(cfg), C.byref(st)))
    ms, n = C.c_double(), C.c_int()
    B.lib.cfdnn_mg_last_timing(h, C.byref(ms), C.byref(n))
    best = min(best, ms.value / max(n.value, 1))
print(f"N={N} bc={bc} tb={os.environ.get('CFDNN_MG_TB', 'default')} vcycle_ms={best:.4f} "
      f"algorithmic_GBs={500.0 * N ** 3 / (best * 1e-3) / 1e9:.1f} checksum={float(np.abs(p).sum()):.12e}")
B.lib.cfdnn_mg_destroy(h)
