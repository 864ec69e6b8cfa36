"""`-m "not gpu"`: host logic of the backend (V-cycle recursion, convergence checks, RK buffer
choreography, BC pass order) run on the CPU test double (tests/devsim) against the oracle
(the unmodified reference CPU build).  The GPU parity tests proper are in test_gpu_parity.py."""
import tempfile

import numpy as np
import pytest

import cases


@pytest.fixture(autouse=True)
def _plain_smoother_on_the_test_double(monkeypatch, request):
    """The temporally blocked smoother runs as ~500 fibers per CTA on the CPU test double (slow); the
    host-logic tests here use the plain one-sweep kernels (CFDNN_MG_TB=0, same arithmetic) except the
    tests marked `tiled`, which keep the blocked kernel covered on CPU.  On the GPU every test runs it."""
    if "tiled" not in request.keywords:
        monkeypatch.setenv("CFDNN_MG_TB", "0")


@pytest.mark.tiled
def test_blocked_smoother_on_the_test_double(sim, oracle):
    case = cases.MG_CASES[1]                                   # per32_conv1e8: 32^3, tiled levels 32 and 16
    with tempfile.TemporaryDirectory() as td:
        meta, pref = cases.run_mg_ref(case, td)
    st, p = cases.run_mg_backend(sim, case)
    assert st["cycles"] == meta["cycles"] and st["converged"] == meta["converged"]
    assert oracle.rel_l2(p, pref) <= 1e-12
    step = ("tgv16_rk3_tiled", (16, 16, 16), "PPPPPP", "rk3", "skew", 4, 1, "tgv", (0, 0, 0), {})
    with tempfile.TemporaryDirectory() as td:
        meta, ref = cases.run_step_ref(step, td)
    got_meta, got = cases.run_step_backend(sim, step)
    assert got_meta["cycles"] == meta["cycles"]
    cases.compare_states(got, ref, 1e-10, step[0])


@pytest.mark.tiled
@pytest.mark.parametrize("bc", ["PPNNPP", "PPDNPP"])
def test_blocked_smoother_with_y_walls(sim, oracle, monkeypatch, bc):
    """channel levels (x periodic, z periodic, Neumann / Dirichlet walls in y) take the 2 x 2 patch kernel with wall
    rows (k_sweep_tbq<.., WALLY>): exact divisions -> bit-identical to the general blocked kernel and to the one-sweep
    kernel, and the default path within 1e-12 of the oracle"""
    case = ("chan_walls_" + bc, (32, 16, 32), bc, 3, {})
    out = {}
    for tag, env in (("plain", dict(CFDNN_MG_FASTDIV="0", CFDNN_MG_TB="0")),
                     ("general", dict(CFDNN_MG_FASTDIV="0", CFDNN_MG_TB="2", CFDNN_MG_TBO="0")),
                     ("patch", dict(CFDNN_MG_FASTDIV="0", CFDNN_MG_TB="2", CFDNN_MG_TBO="1", CFDNN_MG_TBO_MODE="3")),
                     ("default", dict(CFDNN_MG_TB="2"))):
        for k in ("CFDNN_MG_FASTDIV", "CFDNN_MG_TB", "CFDNN_MG_TBO", "CFDNN_MG_TBO_MODE"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        _, out[tag] = cases.run_mg_backend(sim, case)
    assert np.array_equal(out["plain"], out["general"])
    assert np.array_equal(out["plain"], out["patch"])
    with tempfile.TemporaryDirectory() as td:
        meta, pref = cases.run_mg_ref(case, td)
    assert oracle.rel_l2(out["default"], pref) <= 1e-12


@pytest.mark.tiled
def test_blocked_smoother_variants_agree(sim, monkeypatch):
    cases.check_smoother_variants_bitwise(sim, monkeypatch, N=(16, 16, 16), cycles=1)


@pytest.mark.tiled
def test_tiled_convergence_norms_agree(sim, monkeypatch):
    cases.check_tiled_norms(sim, monkeypatch, cases.MG_CASES[5])        # mixed BCs, convergence mode, hits max cycles


@pytest.mark.parametrize("case", cases.MG_CASES[:7], ids=[c[0] for c in cases.MG_CASES[:7]])
def test_fused_residual_restrict_agrees(sim, monkeypatch, case):
    cases.check_fused_residual_restrict(sim, monkeypatch, case, bitwise="D" in case[2])


@pytest.mark.parametrize("case", cases.MG_CASES, ids=[c[0] for c in cases.MG_CASES])
def test_mg_solve_matches_reference(sim, oracle, case):
    with tempfile.TemporaryDirectory() as td:
        meta, pref = cases.run_mg_ref(case, td)
    st, p = cases.run_mg_backend(sim, case)
    assert st["cycles"] == meta["cycles"]                      # BASELINE gate: same MG iteration count
    assert st["converged"] == meta["converged"] and st["fixed_cycle_mode"] == meta["fixed_mode"]
    assert oracle.rel_l2(p, pref) <= 1e-12                     # all cells incl. ghosts (solve :3769-3783)
    for k in ("b_inf", "b_l2", "r0", "r0_l2"):
        assert st[k] == pytest.approx(meta[k], rel=1e-9, abs=1e-300)
    for k in ("residual", "residual_l2"):      # converged residuals are cancellation noise: looser
        assert st[k] == pytest.approx(meta[k], rel=1e-4, abs=1e-300)


@pytest.mark.parametrize("case", cases.MG_2D_CASES, ids=[c[0] for c in cases.MG_2D_CASES])
def test_mg_2d_matches_reference(sim, oracle, case):
    """2-D meshes (Nz == 1): 5-point Chebyshev / Jacobi V-cycles, x-only semi-coarsening when y is stretched"""
    cases.check_mg_2d(sim, oracle, case)


@pytest.mark.parametrize("case", cases.MG_SEMI_CASES, ids=[c[0] for c in cases.MG_SEMI_CASES])
def test_mg_semi_coarsening_matches_reference(sim, oracle, case):
    """y-stretched mesh: y-line Thomas (lexicographic line Gauss-Seidel reproduced by wavefronts),
    x-z RBGS, PCG coarse solve, K-cycle-lite — against MultigridPoissonSolver::solve of the oracle"""
    cases.check_mg_semi(sim, oracle, case)


@pytest.mark.parametrize("case", cases.STEP_CASES, ids=[c[0] for c in cases.STEP_CASES])
def test_step_matches_reference(sim, oracle, case):
    with tempfile.TemporaryDirectory() as td:
        meta, ref = cases.run_step_ref(case, td)
    got_meta, got = cases.run_step_backend(sim, case)
    if meta["cycles"]:                                         # POISSON_STATS lines exist on the RK path only
        assert got_meta["cycles"] == meta["cycles"]
    assert got_meta["residuals"] == pytest.approx(meta["residuals"], rel=1e-9)
    cases.compare_states(got, ref, 1e-10, case[0])


@pytest.mark.parametrize("case", cases.STEP_STRETCHED_CASES, ids=[c[0] for c in cases.STEP_STRETCHED_CASES])
def test_stretched_channel_step_matches_reference(sim, oracle, case):
    cases.check_step_stretched(sim, oracle, case)


@pytest.mark.parametrize("case", cases.STEP_2D_CASES, ids=[c[0] for c in cases.STEP_2D_CASES])
def test_2d_step_matches_reference(sim, oracle, case):
    """2-D meshes (BASELINE config 1 and variants): the solver's periodic-z-plane embedding + the 2-D MG"""
    cases.check_step_2d(sim, oracle, case)


def test_2d_step_with_tbnn_closure(sim, oracle):
    cases.check_step_2d_with_tbnn(sim, oracle)


def test_invalid_periodic_pair_is_invalid_argument(sim, R):
    import ctypes as C
    cfg = sim.config()
    h = C.c_void_p()
    m = sim.mesh(16, 16, 16)
    sim.check(sim.lib.cfdnn_solver_create(C.byref(m), C.byref(cfg), C.byref(h)))
    arr = (C.c_int * 6)(1, 0, 0, 0, 1, 1)                    # x_lo periodic, x_hi wall
    rc = sim.lib.cfdnn_solver_set_velocity_bc(h, arr)
    assert rc == 1 and b"Periodic BC in x" in sim.lib.cfdnn_last_error()   # std::invalid_argument (solver.cpp:653-661)
    sim.lib.cfdnn_solver_destroy(h)


def test_nan_guard_raises_runtime_error(sim):
    cases.check_nan_guard(sim)


@pytest.mark.parametrize("stretched", [False, True], ids=["uniform", "stretched_dyv_quirk"])
def test_adaptive_dt_matches_reference(sim, oracle, stretched):
    cases.check_adaptive_dt(sim, oracle, stretched=stretched)


def test_adaptive_dt_with_implicit_y_diffusion(sim, oracle):
    cases.check_adaptive_dt(sim, oracle, implicit=True)


def test_diffusion_cfl_recheck_with_closure(sim, oracle):
    cases.check_diffusion_cfl_recheck(sim, oracle)


def test_stretched_metrics_match_reference_mesh(oracle):
    """backend.stretched_y_metrics (the package's restatement of Mesh::init_stretched_y, used by bench.py's
    config-3 workload) against the oracle's own mesh arrays, bit for bit"""
    import os
    with tempfile.TemporaryDirectory() as td:
        oracle.run_ref("mesh", Nx=16, Ny=96, Nz=16, stretch_beta=2.0, x1=4 * np.pi, y0=-1.0, y1=1.0, z1=2 * np.pi,
                       out=os.path.join(td, "m"))
        ref = oracle.load_mesh_arrays(os.path.join(td, "m_mesh"))
    got = oracle.backend.stretched_y_metrics(96, -1.0, 1.0, 2.0)
    for k in ref:
        assert np.array_equal(ref[k], got[k]), k


# ---- implicit y-diffusion
@pytest.mark.parametrize("case", cases.STEP_IMPLICIT_CASES, ids=[c[0] for c in cases.STEP_IMPLICIT_CASES])
def test_implicit_y_diffusion_step_matches_reference(sim, oracle, case):
    with tempfile.TemporaryDirectory() as td:
        meta, ref = cases.run_step_ref(case, td)
    got_meta, got = cases.run_step_backend(sim, case)
    if meta["cycles"]:
        assert got_meta["cycles"] == meta["cycles"]
    assert got_meta["residuals"] == pytest.approx(meta["residuals"], rel=1e-9)
    cases.compare_states(got, ref, 1e-10, case[0])


@pytest.mark.parametrize("case", cases.STEP_IMPLICIT_STRETCHED_CASES, ids=[c[0] for c in cases.STEP_IMPLICIT_STRETCHED_CASES])
def test_implicit_y_diffusion_stretched_matches_reference(sim, oracle, case):
    cases.check_step_stretched(sim, oracle, case)


@pytest.mark.parametrize("case", cases.STEP_IMPLICIT_2D_CASES, ids=[c[0] for c in cases.STEP_IMPLICIT_2D_CASES])
def test_implicit_y_diffusion_2d_matches_reference(sim, oracle, case):
    cases.check_step_2d(sim, oracle, case)


# ---- static LES SGS closures
@pytest.mark.parametrize("model", cases.LES_MODELS)
def test_les_sgs_update_matches_reference(oracle, sim, model):
    cases.check_les_update(sim, oracle, model, N=(24, 16, 12))


@pytest.mark.parametrize("model", ["smagorinsky", "sigma"])
def test_les_sgs_update_stretched_and_2d(oracle, sim, model):
    cases.check_les_update(sim, oracle, model, N=(16, 24, 8), stretched=True)
    cases.check_les_update(sim, oracle, model, N=(32, 24, 1))


def test_step_with_smagorinsky_matches_reference(sim, oracle):
    cases.check_step_with_les(sim, oracle, "smagorinsky")


# ---- FFT Poisson solver: host logic + kernel arithmetic on the test double
@pytest.mark.parametrize("case", cases.FFT_CASES[:4], ids=[c[0] for c in cases.FFT_CASES[:4]])
def test_fft_poisson_matches_restatement(sim, case):
    cases.check_fft_solve(sim, case)


def test_fft_poisson_rejections(sim):
    cases.check_fft_rejections(sim)


def test_step_with_fft_poisson_matches_converged_reference(oracle, sim):
    cases.check_step_with_fft(sim, oracle)


# ---- trip forcing + velocity filter
@pytest.mark.parametrize("kind", ["uniform", "stretched", "no_w", "2d"])
def test_trip_forcing_adds_the_reference_term(sim, kind):
    cases.check_trip_forcing(sim, N=(16, 16, 1) if kind == "2d" else (16, 12, 8), stretched=kind == "stretched",
                             force_w=kind != "no_w")


def test_velocity_filter_matches_restatement(sim):
    cases.check_velocity_filter(sim)


# ---- SST k-omega transport + closure on 2-D meshes
@pytest.mark.parametrize("kind,ti", [("uniform", "rk3"), ("uniform", "euler"), ("stretched", "rk3")])
def test_step_with_sst_k_omega_matches_reference(sim, oracle, kind, ti):
    cases.check_step_with_sst(sim, oracle, stretched=kind == "stretched", ti=ti)


@pytest.mark.parametrize("kind", ["uniform", "stretched"])
def test_step_with_standard_k_omega_matches_reference(sim, oracle, kind):
    cases.check_step_with_sst(sim, oracle, stretched=kind == "stretched", ti="rk3", model="komega")


@pytest.mark.parametrize("model,kind", [("baseline", "uniform"), ("baseline", "stretched"), ("algebraic_komega", "uniform"),
                                        ("algebraic_komega", "stretched")])
def test_step_with_algebraic_rans_baselines(sim, oracle, model, kind):
    cases.check_step_with_algebraic_rans(sim, oracle, model=model, stretched=kind == "stretched")


@pytest.mark.parametrize("variant,kind", [(0, "uniform"), (1, "stretched"), (2, "uniform")])
def test_step_with_gep_closure(sim, oracle, variant, kind):
    cases.check_step_with_algebraic_rans(sim, oracle, model="gep", stretched=kind == "stretched", gep_variant=variant)


def test_sst_rejects_3d_meshes(sim):
    cases.check_sst_rejects_3d(sim)


@pytest.mark.parametrize("model,kind", [("earsm_wj", "uniform"), ("earsm_gs", "stretched"), ("earsm_pope", "uniform"),
                                        ("earsm_wj", "stretched")])
def test_step_with_sst_earsm_closures(sim, oracle, model, kind):
    cases.check_step_with_sst(sim, oracle, stretched=kind == "stretched", ti="rk3", model=model)
