#!/usr/bin/env python
"""scripts/fftbench.py — time the FFT Poisson solver (cfdnn_fft_solve_dev) on device-resident fields.
usage: python scripts/fftbench.py Nx Ny Nz [reps]   (env CFDNN_FFT_LX / CFDNN_FFT_LZ: lines per CTA of the x / z transforms)
Prints ms per solve (CUDA events on the launching stream, after warm-up) and GB/s at 100 B per cell."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

pkg = bench.load_package()
K = pkg.backend
B = K.Backend()
Nx, Ny, Nz = (int(a) for a in sys.argv[1:4])
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 10
md = B.mesh(Nx, Ny, Nz, Lx=4 * np.pi, Ly=2.0, Lz=2 * np.pi)
h = C.c_void_p()
B.check(B.lib.cfdnn_fft_create(C.byref(md), C.byref(h)))
B.check(B.lib.cfdnn_fft_set_bc(h, (C.c_int * 6)(*[K.PBC[c] for c in "PPNNPP"])))
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
rhs = torch.zeros((Nz + 2, Ny + 2, Nx + 2), dtype=torch.float64, device=dev)
rhs[1:-1, 1:-1, 1:-1] = torch.randn((Nz, Ny, Nx), dtype=torch.float64, device=dev, generator=g)
p = torch.zeros_like(rhs)
st = torch.cuda.current_stream().cuda_stream
it = C.c_int()


def solve():
    B.check(B.lib.cfdnn_fft_solve_dev(h, C.c_void_p(rhs.data_ptr()), C.c_void_p(p.data_ptr()), C.c_void_p(st), C.byref(it)))


for _ in range(3):
    solve()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    solve()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
# residual of the discrete system on the device (periodic x/z, Neumann y): max |L p - (f - mean f)|
dx, dy, dz = md.dx, md.dy, md.dz
P = p
lap = ((P[1:-1, 1:-1, 2:] - 2 * P[1:-1, 1:-1, 1:-1] + P[1:-1, 1:-1, :-2]) / dx ** 2 +
       (P[1:-1, 2:, 1:-1] - 2 * P[1:-1, 1:-1, 1:-1] + P[1:-1, :-2, 1:-1]) / dy ** 2 +
       (P[2:, 1:-1, 1:-1] - 2 * P[1:-1, 1:-1, 1:-1] + P[:-2, 1:-1, 1:-1]) / dz ** 2)
f = rhs[1:-1, 1:-1, 1:-1]
res = float((lap - (f - f.mean())).abs().max() / f.abs().max())
cells = Nx * Ny * Nz
print(f"fft {Nx}x{Ny}x{Nz} LX={os.environ.get('CFDNN_FFT_LX', 'auto')} LZ={os.environ.get('CFDNN_FFT_LZ', 'auto')} "
      f"ms_per_solve={ms:.4f} algorithmic_GBs={cells * 100.0 / ms / 1e6:.1f} rel_residual={res:.2e}")
B.lib.cfdnn_fft_destroy(h)
