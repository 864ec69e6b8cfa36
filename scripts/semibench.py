#!/usr/bin/env python
"""scripts/semibench.py — time fixed-cycle semi-coarsening MG solves on a y-stretched channel box
(BASELINE config 3 geometry: 4pi x [-1,1] x 2pi, tanh beta = 2) through the C ABI.
usage: python scripts/semibench.py Nx Ny Nz [cycles] [reps]
The mesh metric arrays come from the oracle binary (oracle/_ref/ref_harness, mode=mesh)."""
import ctypes as C
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402
import refharness as R  # noqa: E402

Nx, Ny, Nz = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (192, 96, 192)
cycles = int(sys.argv[4]) if len(sys.argv) > 4 else 2
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 2
B = R.Backend()
capi = R.capi
with tempfile.TemporaryDirectory() as td:
    R.run_ref("mesh", Nx=Nx, Ny=Ny, Nz=Nz, out=os.path.join(td, "m"), **cases.CHANNEL_BOX)
    arrs = R.load_mesh_arrays(os.path.join(td, "m_mesh"))
md = B.stretched_mesh(Nx, Ny, Nz, arrs, Lx=cases.CHANNEL_BOX["x1"], Lz=cases.CHANNEL_BOX["z1"])
h = C.c_void_p()
B.check(B.lib.cfdnn_mg_create(C.byref(md), C.byref(h)))
B.check(B.lib.cfdnn_mg_set_bc(h, (C.c_int * 6)(*[R.PBC[c] for c in "PPNNPP"])))
rhs, p0 = cases.mg_inputs_smooth((Nx, Ny, Nz))
cfg = capi.PoissonCfg()
B.lib.cfdnn_poisson_cfg_default(C.byref(cfg))
cfg.fixed_cycles = cycles
st = capi.PoissonStats()
best = 1e30
for it in range(reps):
    p = p0.copy()
    B.check(B.lib.cfdnn_mg_solve(h, capi.dptr(rhs), capi.dptr(p), C.byref(cfg), C.byref(st)))
    ms, n = C.c_double(), C.c_int()
    B.lib.cfdnn_mg_last_timing(h, C.byref(ms), C.byref(n))
    best = min(best, ms.value / max(n.value, 1))
print(f"semibench {Nx}x{Ny}x{Nz} cycles={cycles} vcycle_ms={best:.3f} Mcells/s-per-vcycle={Nx * Ny * Nz / best / 1e3:.1f} "
      f"checksum={float(np.abs(p).sum()):.12e}")
B.lib.cfdnn_mg_destroy(h)
