"""oracle/fft_poisson.py — TEST INFRASTRUCTURE, not product.

CPU restatement (numpy) of the reference's FFT-hybrid Poisson solver for channels (periodic x/z, walls y):
`FFTPoissonSolver::solve_device`, /root/reference/src/poisson_solver_fft.cpp:960-1098.

Why a restatement: the whole reference file is under `#ifdef USE_GPU_OFFLOAD` and needs cuFFT, cuSPARSE and OpenMP
target offload — it has no CPU path and cannot be built in this container, so `oracle/_ref` does not contain it.
PARITY PIN: not against FFT outputs of the reference (none can be produced here) but against the reference's OWN
multigrid solver (`oracle/_ref`, `mode=mg`), which solves the same discrete system — the 7-point Laplacian with
periodic x/z and Neumann y, `yLap_aS/aN` on stretched meshes: `tests/test_fft_oracle.py` requires the two
solutions to agree (mean removed) to the multigrid's convergence tolerance and checks the discrete residual of
this restatement directly.  The reference's own FFT-vs-MG test uses a 0.1 relative gate
(tests/test_fft_unified.cpp:136-138).

Only tests/ may import this module.
"""
import numpy as np


def eigenvalues(N, h):
    """lambda(theta) = (2 - 2 cos theta) / h^2, second-order Laplacian symbol (poisson_solver_fft.cpp:494-504)"""
    return (2.0 - 2.0 * np.cos(2.0 * np.pi * np.arange(N) / N)) / (h * h)


def tridiagonal_coeffs(Ny, dy, aS=None, aN=None):
    """compute_tridiagonal_coeffs (:509-547): aS/aN = 1/dy^2 (uniform) or yLap_aS/aN[j + Ng] (stretched), the wall
    rows lose the coefficient that points through the wall (Neumann)"""
    if aS is None:
        lo = np.full(Ny, 1.0 / (dy * dy))
        up = np.full(Ny, 1.0 / (dy * dy))
    else:
        lo = np.array(aS, dtype=np.float64).copy()
        up = np.array(aN, dtype=np.float64).copy()
    lo[0] = 0.0
    up[Ny - 1] = 0.0
    return lo, up, -(lo + up)


def solve(rhs, dx, dz, dy=None, dyv=None, aS=None, aN=None):
    """rhs: interior array [Nz][Ny][Nx].  Returns p (interior) with the reference's gauge: the (kx, kz) = (0, 0) mode
    is pinned to 0 at j = 0 (:617-627, :668-670).

    Steps of solve_device: volume-weighted mean subtraction (:53-89, :930-938: sum f dyv / (Nx Nz Ly)), forward 2-D FFT
    over (x, z) for every y (:1001), one tridiagonal system in y per mode with diagonal -(aS + aN) - (lambda_x +
    lambda_z) (:606-641), inverse FFT and 1 / (Nx Nz) normalisation (:1085, :948)."""
    Nz, Ny, Nx = rhs.shape
    w = np.full(Ny, dy) if dyv is None else np.asarray(dyv, dtype=np.float64)
    mean = float((rhs * w[None, :, None]).sum() / (Nx * Nz * w.sum()))
    f = rhs - mean
    F = np.fft.fft2(f, axes=(0, 2))                     # unnormalised forward transform, like cufftExecD2Z
    lo, up, base = tridiagonal_coeffs(Ny, dy, aS, aN)
    lam = eigenvalues(Nz, dz)[:, None] + eigenvalues(Nx, dx)[None, :]        # [kz][kx]
    # Thomas algorithm per mode, vectorised over the modes (:1027-1075)
    c = np.zeros((Ny, Nz, Nx))
    d = np.zeros((Ny, Nz, Nx), dtype=np.complex128)
    for j in range(Ny):
        diag = base[j] - lam
        lower = np.full_like(lam, lo[j] if j > 0 else 0.0)
        upper = np.full_like(lam, up[j] if j < Ny - 1 else 0.0)
        r = F[:, j, :].copy()
        if j == 0:                                      # pinned row of the zero mode
            diag = diag.copy(); diag[0, 0] = 1.0
            upper[0, 0] = 0.0
            r[0, 0] = 0.0
            c[j] = upper / diag
            d[j] = r / diag
        else:
            den = diag - lower * c[j - 1]
            c[j] = upper / den
            d[j] = (r - lower * d[j - 1]) / den
    P = np.zeros_like(F)
    P[:, Ny - 1, :] = d[Ny - 1]
    for j in range(Ny - 2, -1, -1):
        P[:, j, :] = d[j] - c[j] * P[:, j + 1, :]
    p = np.fft.ifft2(P, axes=(0, 2))                    # includes the 1 / (Nx Nz) normalisation
    return np.ascontiguousarray(p.real)


def with_ghosts(p):
    """unpack + apply_bc_device (:806-866): periodic x, Neumann (copy) y, periodic z, in that order; Ng = 1"""
    Nz, Ny, Nx = p.shape
    g = np.zeros((Nz + 2, Ny + 2, Nx + 2))
    g[1:-1, 1:-1, 1:-1] = p
    g[1:-1, 1:-1, 0] = g[1:-1, 1:-1, Nx]
    g[1:-1, 1:-1, Nx + 1] = g[1:-1, 1:-1, 1]
    g[1:-1, 0, :] = g[1:-1, 1, :]
    g[1:-1, Ny + 1, :] = g[1:-1, Ny, :]
    g[0] = g[Nz]
    g[Nz + 1] = g[1]
    return g


def laplacian(p, dx, dz, dy=None, aS=None, aN=None):
    """the discrete operator the solve inverts (interior p, periodic x/z, Neumann y) — for residual checks"""
    Nz, Ny, Nx = p.shape
    lo, up, _ = tridiagonal_coeffs(Ny, dy, aS, aN)
    L = (np.roll(p, 1, 2) + np.roll(p, -1, 2) - 2 * p) / (dx * dx) + (np.roll(p, 1, 0) + np.roll(p, -1, 0) - 2 * p) / (dz * dz)
    ps = np.concatenate([p[:, :1], p[:, :-1]], axis=1)
    pn = np.concatenate([p[:, 1:], p[:, -1:]], axis=1)
    L += lo[None, :, None] * (ps - p) + up[None, :, None] * (pn - p)
    return L
