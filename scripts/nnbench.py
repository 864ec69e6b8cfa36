#!/usr/bin/env python
"""scripts/nnbench.py — time the NN-TBNN closure update on an Nx x Ny x Nz channel through the C ABI.
usage: python scripts/nnbench.py [Nx Ny Nz] [reps]   (env CFDNN_NN_SIMT=1 selects the scalar kernel)
Prints kernel ms (CUDA events inside the library), Mcells/s and FP64 TFLOP/s of the MLP
(17 536 FLOP per cell for the 5-64-64-64-4 TBNN)."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import nncases as NC  # noqa: E402
import nnmodels as M  # noqa: E402
import refharness as R  # noqa: E402

Nx, Ny, Nz = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (256, 256, 256)
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
preset = os.environ.get("NN_PRESET", "tbnn_channel_caseholdout")
B = R.Backend()
h = M.create_preset(B, preset)
B.check(B.lib.cfdnn_nn_set_params(h, NC.NU, 1.0, 1.0, 1e-10, 1.0))
W, _, _, _ = M.load_preset(preset)
flop_per_cell = sum(2 * w.shape[0] * w.shape[1] for w in W)
sh = R.shapes(Nx, Ny, Nz)
u2, v2, k2, o2 = NC.channel_fields(Nx, Ny, seed=1, k_val=0.01, om_val=10.0)
u3 = np.zeros(sh["u"]); v3 = np.zeros(sh["v"]); k3 = np.zeros(sh["c"]); o3 = np.zeros(sh["c"])
u3[1:-1], v3[1:-1], k3[1:-1], o3[1:-1] = u2[0], v2[0], k2[0], o2[0]
best = 1e30
for it in range(reps):
    got = NC.backend_update(B, h, Nx, Ny, u3, v3, k3, o3, Nz=Nz, want_tau=True)
    ms = C.c_double()
    B.lib.cfdnn_nn_last_timing(h, C.byref(ms))
    best = min(best, ms.value)
cells = Nx * Ny * Nz
tc = B.lib.cfdnn_nn_uses_tensor_cores(h)
print(f"nnbench {preset} {Nx}x{Ny}x{Nz} tensor_cores={tc} kernel_ms={best:.3f} Mcells/s={cells / best / 1e3:.1f} "
      f"mlp_TFLOPs={cells * flop_per_cell / (best * 1e-3) / 1e12:.2f} nut_checksum={float(got['nut'].sum()):.12e}")
B.lib.cfdnn_nn_destroy(h)
