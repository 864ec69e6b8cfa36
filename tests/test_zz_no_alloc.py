"""cfdnn_adaptive_dt_noalloc — the allocation-free variant of compute_adaptive_dt (static __device__ scratch, see
include/cfdnn_b200.h and cfd-nn_b200/csrc/noalloc.cu) against the existing entry point
cfdnn_solver_compute_adaptive_dt, through the C ABI, on calls in a row that share the scratch array.

Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

import cases
import refharness as R

pytestmark = pytest.mark.gpu

K = R.backend
MAX_NY = 4096        # CFDNN_NOALLOC_MAX_NY (include/cfdnn_b200.h)
ERR_INVALID = 1      # CFDNN_ERR_INVALID
TOL = 1e-12          # the tolerance of the existing adaptive-dt tests (cases.check_adaptive_dt)


def _dev(B, h, name):
    p, n = C.c_void_p(), C.c_size_t()
    B.check(B.lib.cfdnn_solver_device_ptr(h, K.FIELD_ID[name], C.byref(p), C.byref(n)))
    assert p.value
    return p


def _both(B, h, md, cfg, with_nut):
    ref = C.c_double()
    B.check(B.lib.cfdnn_solver_compute_adaptive_dt(h, C.byref(ref)))
    got = C.c_double(-1.0)
    B.check(B.lib.cfdnn_adaptive_dt_noalloc(C.byref(md), C.byref(cfg), _dev(B, h, "u"), _dev(B, h, "v"), _dev(B, h, "w"),
                                            _dev(B, h, "nut") if with_nut else None, None, C.byref(got)))
    return got.value, ref.value


def _solver(B, N, bc, stretched, les=False, nu=0.01):
    Nx, Ny, Nz = N
    cfg = B.config(nu=nu, dt=0.002, time_integrator=K.TI["rk3"], convective_scheme=K.SCHEME["skew"], adaptive_dt=1,
                   CFL_max=0.4, poisson_fixed_cycles=4)
    if stretched:
        arrs = K.stretched_y_metrics(Ny, -1.0, 1.0, 2.0)
        md = B.stretched_mesh(Nx, Ny, Nz, arrs, Lx=4 * np.pi, Lz=2 * np.pi, y_lo=-1.0, y_hi=1.0)
    else:
        md = B.mesh(Nx, Ny, Nz)
    h = B.solver(md, cfg, bc, (1.0, 0.0, 0.0) if "W" in bc else (0.0, 0.0, 0.0))
    u, v, w = cases.step_init(N, "tgv")
    rng = np.random.default_rng(Nx * 1000 + Ny)
    u = u + 0.05 * rng.standard_normal(u.shape)
    v = v + 0.05 * rng.standard_normal(v.shape)
    w = w + 0.05 * rng.standard_normal(w.shape)
    B.upload(h, "u", u); B.upload(h, "v", v); B.upload(h, "w", w)
    if les:
        B.check(B.lib.cfdnn_solver_set_les_model(h, K.LES_MODEL["smagorinsky"], -1.0))
    B.check(B.lib.cfdnn_solver_initialize(h))
    return h, md, cfg


def test_noalloc_adaptive_dt_matches_existing_entry_point(gpu):
    B = gpu
    seen = []
    # (grid, bc, stretched y, LES closure attached): sizes change between the calls, so every call finds the previous
    # call's partials / maxima / dyv in the shared static array
    for N, bc, stretched, les in (((32, 24, 16), "PPPPPP", False, False),
                                  ((16, 40, 12), "PPWWPP", True, False),
                                  ((24, 16, 20), "PPWWPP", False, True),
                                  ((48, 32, 1), "PPWWPP", False, False)):
        h, md, cfg = _solver(B, N, bc, stretched, les)
        try:
            for call in range(2):                    # two calls in a row on fields that changed in between
                got, ref = _both(B, h, md, cfg, with_nut=les)
                assert ref > 0.0 and np.isfinite(ref)
                assert got == pytest.approx(ref, rel=TOL), (N, bc, stretched, les, call, got, ref)
                seen.append(ref)
                B.check(B.lib.cfdnn_solver_set_dt(h, min(ref, 0.002)))
                B.step(h)
        finally:
            B.lib.cfdnn_solver_destroy(h)
    assert len(set(seen)) > 4                        # the calls really computed different values


def test_noalloc_rejects_sizes_beyond_the_static_scratch(gpu):
    B = gpu
    md = B.mesh(8, MAX_NY + 1, 8)
    cfg = B.config(nu=0.01, dt=0.01)
    h, md_ok, cfg_ok = _solver(B, (16, 16, 16), "PPPPPP", False)
    try:
        out = C.c_double(-1.0)
        rc = B.lib.cfdnn_adaptive_dt_noalloc(C.byref(md), C.byref(cfg), _dev(B, h, "u"), _dev(B, h, "v"), _dev(B, h, "w"),
                                             None, None, C.byref(out))
        assert rc == ERR_INVALID
        msg = B.lib.cfdnn_last_error().decode()
        assert "CFDNN_NOALLOC_MAX_NY" in msg and str(MAX_NY + 1) in msg, msg
        assert out.value == -1.0
        # the library is still usable afterwards
        got, ref = _both(B, h, md_ok, cfg_ok, with_nut=False)
        assert got == pytest.approx(ref, rel=TOL)
    finally:
        B.lib.cfdnn_solver_destroy(h)
