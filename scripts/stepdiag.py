#!/usr/bin/env python
"""scripts/stepdiag.py N nsteps [dt] — per-step V-cycle counts / residual / step ms of the TGV bench workload."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import refharness as R  # noqa: E402

N = int(sys.argv[1]); ns = int(sys.argv[2]); dt = float(sys.argv[3]) if len(sys.argv) > 3 else 0.01
B = R.Backend()
mesh = B.mesh(N, N, N)
cfg = B.config(nu=0.01, dt=dt, time_integrator=R.TI["rk3"], convective_scheme=R.SCHEME["skew"], adaptive_dt=0,
               poisson_fixed_cycles=0, poisson_tol_rhs=1e-8, poisson_max_vcycles=50)
h = B.solver(mesh, cfg, "PPPPPP")
u, v, w = R.tgv_init(N, N, N)
for n, a in (("u", u), ("v", v), ("w", w)):
    B.upload(h, n, a)
B.check(B.lib.cfdnn_solver_initialize(h))
seen = 0
for s in range(ns):
    r = B.step(h)
    a, b, c = C.c_double(), C.c_double(), C.c_int()
    B.lib.cfdnn_solver_last_timing(h, C.byref(a), C.byref(b), C.byref(c))
    cyc = B.cycles(h)
    print(f"step {s + 1:3d} residual={r:.6e} ms={a.value:8.2f} vcycles={cyc[seen:]}", flush=True)
    seen = len(cyc)
B.lib.cfdnn_solver_destroy(h)
