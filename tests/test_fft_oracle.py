"""Pins oracle/fft_poisson.py (numpy restatement of the reference's GPU-only FFT Poisson solver) to the reference
itself: the reference's multigrid (oracle/_ref, unmodified CPU build) solves the same discrete system, so the two
solutions must agree up to the gauge constant and the multigrid's convergence tolerance."""
import os
import sys
import tempfile

import numpy as np
import pytest

import cases
import refharness as R

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "oracle"))
import fft_poisson as F  # noqa: E402

PIN_CASES = [
    ("uniform_32x16x32", (32, 16, 32), dict(x1=4 * np.pi, y0=-1.0, y1=1.0, z1=2 * np.pi)),
    ("ragged_24x12x20", (24, 12, 20), dict(x1=2 * np.pi, y0=0.0, y1=2.0, z1=np.pi)),
    ("stretched_32x24x16", (32, 24, 16), dict(**cases.CHANNEL_BOX)),
]


@pytest.mark.parametrize("case", PIN_CASES, ids=[c[0] for c in PIN_CASES])
def test_restatement_agrees_with_reference_multigrid(oracle, case):
    name, N, box = case
    Nx, Ny, Nz = N
    rhs, _ = cases.mg_inputs_smooth(N) if "stretch_beta" in box else cases.mg_inputs(N)
    with tempfile.TemporaryDirectory() as td:
        rhs.tofile(os.path.join(td, "rhs.f64"))
        R.run_ref("mg", Nx=Nx, Ny=Ny, Nz=Nz, bc="PPNNPP", fixed=0, maxv=200, tol_rhs=1e-13, tol_rel=0, tol=0, tol_abs=0,
                  rhs=os.path.join(td, "rhs.f64"), out=os.path.join(td, "ref"), **box)
        p_mg = np.fromfile(os.path.join(td, "ref_p.f64")).reshape(rhs.shape)[1:-1, 1:-1, 1:-1]
        arrs = R.load_mesh_arrays(os.path.join(td, "ref_mesh")) if "stretch_beta" in box else None
    dx, dz, dy = box["x1"] / Nx, box["z1"] / Nz, (box["y1"] - box["y0"]) / Ny
    f = rhs[1:-1, 1:-1, 1:-1]
    if arrs is None:
        p = F.solve(f, dx, dz, dy=dy)
        res = F.laplacian(p, dx, dz, dy=dy) - f
        w = np.full(Ny, dy)
    else:
        aS, aN, w = arrs["aS"][1:-1], arrs["aN"][1:-1], arrs["dyv"][1:-1]
        p = F.solve(f, dx, dz, dyv=w, aS=aS, aN=aN)
        res = F.laplacian(p, dx, dz, aS=aS, aN=aN) - f
    # (i) the restatement solves the discrete system (f has zero volume-weighted mean up to round-off)
    fm = (f * w[None, :, None]).sum() / (Nx * Nz * w.sum())
    assert np.abs(res + fm).max() <= 1e-10 * np.abs(f).max()
    # (ii) gauge: the zero mode vanishes on the first row
    assert abs(p[:, 0, :].mean()) <= 1e-12 * np.abs(p).max()
    # (iii) same solution as the reference's multigrid, up to the constant
    a = p - (p * w[None, :, None]).sum() / (Nx * Nz * w.sum())
    b = p_mg - (p_mg * w[None, :, None]).sum() / (Nx * Nz * w.sum())
    assert R.rel_l2(a, b) <= 1e-8, R.rel_l2(a, b)


def test_ghost_fill_matches_the_reference_order():
    p = np.random.default_rng(3).standard_normal((4, 3, 6))
    g = F.with_ghosts(p)
    assert np.array_equal(g[1:-1, 1:-1, 0], p[:, :, -1]) and np.array_equal(g[1:-1, 1:-1, -1], p[:, :, 0])
    assert np.array_equal(g[1:-1, 0, 1:-1], p[:, 0, :]) and np.array_equal(g[1:-1, -1, 1:-1], p[:, -1, :])
    assert np.array_equal(g[0], g[-2]) and np.array_equal(g[-1], g[1])
