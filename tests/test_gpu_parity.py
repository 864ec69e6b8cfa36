"""`-m gpu`: parity of the CUDA path (through the C ABI) against the oracle = the unmodified
reference CPU build (oracle/_ref/ref_harness, built in the container, shipped as a built file).
Gates are BASELINE.json's: per-field rel L2 <= 1e-10 after one step, <= 1e-8 after 100 steps,
identical V-cycle counts, bitwise-identical reruns."""
import tempfile

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", cases.MG_CASES, ids=[c[0] for c in cases.MG_CASES])
def test_mg_solve_matches_reference(gpu, oracle, case):
    with tempfile.TemporaryDirectory() as td:
        meta, pref = cases.run_mg_ref(case, td)
    st, p = cases.run_mg_backend(gpu, case)
    assert st["cycles"] == meta["cycles"]
    assert st["converged"] == meta["converged"] and st["fixed_cycle_mode"] == meta["fixed_mode"]
    assert oracle.rel_l2(p, pref) <= 1e-12
    for k in ("b_inf", "b_l2", "r0", "r0_l2"):
        assert st[k] == pytest.approx(meta[k], rel=1e-9, abs=1e-300)


@pytest.mark.parametrize("case", cases.STEP_CASES, ids=[c[0] for c in cases.STEP_CASES])
def test_step_matches_reference(gpu, oracle, case):
    with tempfile.TemporaryDirectory() as td:
        meta, ref = cases.run_step_ref(case, td)
    got_meta, got = cases.run_step_backend(gpu, case)
    if meta["cycles"]:
        assert got_meta["cycles"] == meta["cycles"]
    assert got_meta["residuals"] == pytest.approx(meta["residuals"], rel=1e-9)
    cases.compare_states(got, ref, 1e-10, case[0])


def test_tgv64_one_step_1e10(gpu, oracle):
    """BASELINE config 2 (3D TGV 64^3, RK3 + MG, convergence mode to 1e-8): one step."""
    case = ("tgv64", (64, 64, 64), "PPPPPP", "rk3", "skew", 0, 1, "tgv", (0, 0, 0), dict(tol_rhs=1e-8, maxv=50))
    with tempfile.TemporaryDirectory() as td:
        meta, ref = cases.run_step_ref(case, td)
    got_meta, got = cases.run_step_backend(gpu, case)
    assert got_meta["cycles"] == meta["cycles"]
    cases.compare_states(got, ref, 1e-10, "tgv64 1 step")


def test_tgv128_one_step_big_tiles_vs_oracle(gpu, oracle):
    """The kernels that make the headline number against the ORACLE, not against other variants of this
    library: at 128^3 level 0 runs k_sweep_tbo<64,26> (the 512^3 level-0 smoother) and the fused
    k_residual_restrict<32,8>, levels 1-2 the 32x12 tiles.  One RK3 step, MG to 1e-8 (default path:
    reciprocal multiplies + FMA, i.e. not bit-identical by design) within BASELINE's 1e-10."""
    case = ("tgv128", (128, 128, 128), "PPPPPP", "rk3", "skew", 0, 1, "tgv", (0, 0, 0), dict(tol_rhs=1e-8, maxv=50))
    with tempfile.TemporaryDirectory() as td:
        meta, ref = cases.run_step_ref(case, td)
    got_meta, got = cases.run_step_backend(gpu, case)
    assert got_meta["cycles"] == meta["cycles"]
    assert got_meta["residuals"] == pytest.approx(meta["residuals"], rel=1e-9)
    worst = cases.compare_states(got, ref, 1e-10, "tgv128 1 step")
    print("tgv128 per-field errors vs oracle", worst)


def test_channel_128x64x64_one_step_general_big_tiles_vs_oracle(gpu, oracle):
    """same for the general blocked kernel k_sweep_tb<2,64,26> (walls in y: the lean kernel does not apply)
    and the non-periodic variant of the fused residual + restriction"""
    case = ("chan128", (128, 64, 64), "PPWWPP", "rk3", "skew", 0, 1, "shear", (1.0, 0, 0), dict(tol_rhs=1e-8, maxv=50))
    with tempfile.TemporaryDirectory() as td:
        meta, ref = cases.run_step_ref(case, td)
    got_meta, got = cases.run_step_backend(gpu, case)
    assert got_meta["cycles"] == meta["cycles"]
    cases.compare_states(got, ref, 1e-10, "chan128 1 step")


def test_mg_256x128x128_solve_vs_oracle(gpu, oracle):
    """MultigridPoissonSolver::solve on 256 x 128 x 128 (big tiles on levels 0 and 1, fused residual +
    restriction on both), convergence mode to 1e-8, against the oracle: same cycle count, same solution."""
    case = ("mg256", (256, 128, 128), "PPPPPP", 0, dict(smooth=1, tol_rhs=1e-8, maxv=30, tol_rel=0, tol=0))
    with tempfile.TemporaryDirectory() as td:
        meta, pref = cases.run_mg_ref(case, td)
    st, p = cases.run_mg_backend(gpu, case)
    assert st["cycles"] == meta["cycles"] and st["converged"] == meta["converged"]
    err = oracle.rel_l2(p, pref)
    print("mg256 rel L2 vs oracle", err, "cycles", st["cycles"])
    assert err <= 1e-10
    for k in ("b_inf", "b_l2", "r0", "r0_l2"):
        assert st[k] == pytest.approx(meta[k], rel=1e-9, abs=1e-300)


def test_tgv32_100_steps_1e8(gpu, oracle):
    case = ("tgv32x100", (32, 32, 32), "PPPPPP", "rk3", "skew", 0, 100, "tgv", (0, 0, 0), dict(tol_rhs=1e-8, maxv=50))
    with tempfile.TemporaryDirectory() as td:
        meta, ref = cases.run_step_ref(case, td)
    got_meta, got = cases.run_step_backend(gpu, case)
    assert got_meta["cycles"] == meta["cycles"]          # 300 Poisson solves, same count each
    cases.compare_states(got, ref, 1e-8, "tgv32 100 steps")


def test_channel32_100_steps_walls(gpu, oracle):
    case = ("chan32x100", (32, 32, 32), "PPWWPP", "rk3", "skew", 8, 100, "shear", (1.0, 0, 0), {})
    with tempfile.TemporaryDirectory() as td:
        meta, ref = cases.run_step_ref(case, td)
    got_meta, got = cases.run_step_backend(gpu, case)
    cases.compare_states(got, ref, 1e-8, "channel32 100 steps")


def test_rerun_is_bitwise_identical(gpu):
    case = ("tgv64r", (64, 64, 64), "PPPPPP", "rk3", "skew", 8, 3, "tgv", (0, 0, 0), {})
    _, a = cases.run_step_backend(gpu, case)
    _, b = cases.run_step_backend(gpu, case)
    for n in cases.FIELDS:
        assert np.array_equal(a[n], b[n]), n


def test_run_equals_repeated_step(gpu):
    case = ("tgv32", (32, 32, 32), "PPPPPP", "rk3", "skew", 8, 4, "tgv", (0, 0, 0), {})
    _, a = cases.run_step_backend(gpu, case)
    h = cases.make_solver(gpu, case)
    gpu.run(h, 4)
    b = gpu.state(h, 32, 32, 32)
    gpu.lib.cfdnn_solver_destroy(h)
    for n in cases.FIELDS:
        assert np.array_equal(a[n], b[n]), n


def test_tgv256_properties(gpu, R):
    """Full-size, oracle-free properties (TGV invariants of reference tests/test_tgv_3d_invariants.cpp):
    divergence after projection at solver tolerance, kinetic energy decays, zero mean velocity."""
    N = 256
    case = ("tgv256", (N, N, N), "PPPPPP", "rk3", "skew", 0, 3, "tgv", (0, 0, 0), dict(tol_rhs=1e-8, maxv=50))
    h = cases.make_solver(gpu, case)

    def ke(st):
        return 0.5 * sum(float(np.mean(R.interior(n, st[n])[..., :N, :N, :N] ** 2)) if False else
                         float(np.mean(R.interior(n, st[n]) ** 2)) for n in ("u", "v", "w"))

    st0 = gpu.state(h, N, N, N, names=("u", "v", "w"))
    k0 = ke(st0)
    gpu.run(h, 3)
    st = gpu.state(h, N, N, N, names=("u", "v", "w", "div", "rhs"))
    gpu.lib.cfdnn_solver_destroy(h)
    k1 = ke(st)
    assert k1 < k0 and k1 > 0.9 * k0
    # zero mean velocity (the reference's own one-sided prolongation + 1e-3 relative MG exit leave
    # an O(1e-5) symmetry error, so this is a physical bound, not a round-off one)
    for n in ("u", "v", "w"):
        assert abs(float(np.mean(R.interior(n, st[n])))) < 1e-5
    # div field holds the pre-projection divergence of the last stage; recompute post-projection
    u, v, w = st["u"], st["v"], st["w"]
    d = 2 * np.pi / N
    div = ((u[1:-1, 1:-1, 2:N + 2] - u[1:-1, 1:-1, 1:N + 1]) + (v[1:-1, 2:N + 2, 1:-1] - v[1:-1, 1:N + 1, 1:-1]) +
           (w[2:N + 2, 1:-1, 1:-1] - w[1:N + 1, 1:-1, 1:-1])) / d
    assert np.abs(div).max() < 1e-6


# ---- y-stretched meshes: semi-coarsening MG and BASELINE config 3 in miniature
@pytest.mark.parametrize("case", cases.MG_SEMI_CASES, ids=[c[0] for c in cases.MG_SEMI_CASES])
def test_mg_semi_coarsening_matches_reference(gpu, oracle, case):
    cases.check_mg_semi(gpu, oracle, case)


def test_mg_semi_64x32x64_converges_like_reference(gpu, oracle):
    """channel box 64x32x64, tanh beta=2, convergence mode to 1e-8: same cycle count, same solution"""
    case = ("semi64", (64, 32, 64), "PPNNPP", 0, dict(smooth=1, tol_rhs=1e-8, maxv=20, tol_rel=0, tol=0, **cases.CHANNEL_BOX))
    cases.check_mg_semi(gpu, oracle, case)


@pytest.mark.parametrize("case", cases.STEP_STRETCHED_CASES, ids=[c[0] for c in cases.STEP_STRETCHED_CASES])
def test_stretched_channel_step_matches_reference(gpu, oracle, case):
    cases.check_step_stretched(gpu, oracle, case)


def test_stretched_channel_48x32x48_rk3_16_cycles(gpu, oracle):
    """the production recipe of config 3 (poisson_fixed_cycles = 16, RK3, skew) on a 48x32x48 channel"""
    case = ("chan48", (48, 32, 48), "PPWWPP", "rk3", "skew", 16, 2, "pchan", (1.0, 0, 0), dict(**cases.CHANNEL_BOX))
    err, sens = cases.check_step_within_reference_sensitivity(gpu, oracle, case)
    print("chan48 errors", err, "reference sensitivity to 1e-15 input noise", sens)


def test_stretched_channel_96x48x96_rk3_16_cycles(gpu, oracle):
    """BASELINE config 3 at half resolution (192x96x192 costs the single-threaded oracle ~2 min per step): tanh
    beta = 2 channel, RK3, skew-symmetric, 16 fixed V-cycles of the semi-coarsening solver.  Gate: the reference's
    own sensitivity to 1e-15 input noise (BASELINE.md "config 3 gate"), floor 1e-10."""
    case = ("chan96", (96, 48, 96), "PPWWPP", "rk3", "skew", 16, 1, "pchan", (1.0, 0, 0), dict(**cases.CHANNEL_BOX))
    err, sens = cases.check_step_within_reference_sensitivity(gpu, oracle, case)
    print("chan96 errors", err, "reference sensitivity to 1e-15 input noise", sens)


# ---- 2-D meshes (Nz == 1; BASELINE config 1): 2-D multigrid + the solver's periodic-z-plane embedding
@pytest.mark.parametrize("case", cases.MG_2D_CASES, ids=[c[0] for c in cases.MG_2D_CASES])
def test_mg_2d_matches_reference(gpu, oracle, case):
    cases.check_mg_2d(gpu, oracle, case)


@pytest.mark.parametrize("case", cases.STEP_2D_CASES, ids=[c[0] for c in cases.STEP_2D_CASES])
def test_2d_step_matches_reference(gpu, oracle, case):
    cases.check_step_2d(gpu, oracle, case)


def test_2d_step_with_tbnn_closure(gpu, oracle):
    """RANSSolver::step with the NN-TBNN closure attached (DMMA kernel inside the step), 2-D channel"""
    cases.check_step_2d_with_tbnn(gpu, oracle)


def test_poiseuille_2d_config1_100_steps(gpu, oracle):
    """BASELINE config 1: 2-D laminar Poiseuille channel Nx=32 Ny=64 nu=0.01, Euler; 100 steps <= 1e-8"""
    case = ("poiseuille2d_100", (32, 64, 1), "PPWWPP", "euler", "central", 8, 100, "shear", (1.0, 0, 0),
            dict(x1=2 * np.pi, y0=-1.0, y1=1.0))
    cases.check_step_2d(gpu, oracle, case, tol=1e-8)


def test_blocked_smoother_variants_agree(gpu, monkeypatch):
    """128^3 puts level 0 on the 64x26 tiles and levels 1-2 on the 32x12 tiles of both blocked kernels"""
    cases.check_smoother_variants_bitwise(gpu, monkeypatch, N=(128, 128, 128), cycles=2)


@pytest.mark.parametrize("N,bc", [((256, 128, 128), "PPPPPP"), ((160, 96, 64), "PPNNPP"), ((128, 64, 48), "DNNDND")],
                         ids=["periodic", "channel_ragged_tiles", "mixed_dirichlet"])
def test_fused_residual_restrict_agrees(gpu, monkeypatch, N, bc):
    """k_residual_restrict vs k_sweep_tb<RES> + k_restrict on levels that take it by default (Nx >= 128)
    and, forced, on every level down to 8 cells"""
    case = ("rr_" + bc, N, bc, 2, {})
    cases.check_fused_residual_restrict(gpu, monkeypatch, case, bitwise="D" in bc)


# ---- reductions of the step on the GPU
@pytest.mark.parametrize("stretched", [False, True], ids=["uniform", "stretched_dyv_quirk"])
def test_adaptive_dt_matches_reference(gpu, oracle, stretched):
    cases.check_adaptive_dt(gpu, oracle, stretched=stretched)


def test_adaptive_dt_with_implicit_y_diffusion(gpu, oracle):
    cases.check_adaptive_dt(gpu, oracle, implicit=True)


def test_nan_guard_raises_runtime_error(gpu):
    cases.check_nan_guard(gpu)


def test_diffusion_cfl_recheck_with_closure(gpu, oracle):
    dts, used = cases.check_diffusion_cfl_recheck(gpu, oracle)
    print("adaptive dts", dts, "dt used by step()", used)


# ---- implicit y-diffusion: Thomas solve per y-line + xz-only explicit diffusion
@pytest.mark.parametrize("case", cases.STEP_IMPLICIT_CASES, ids=[c[0] for c in cases.STEP_IMPLICIT_CASES])
def test_implicit_y_diffusion_step_matches_reference(gpu, oracle, case):
    with tempfile.TemporaryDirectory() as td:
        meta, ref = cases.run_step_ref(case, td)
    got_meta, got = cases.run_step_backend(gpu, case)
    if meta["cycles"]:
        assert got_meta["cycles"] == meta["cycles"]
    assert got_meta["residuals"] == pytest.approx(meta["residuals"], rel=1e-9)
    cases.compare_states(got, ref, 1e-10, case[0])


@pytest.mark.parametrize("case", cases.STEP_IMPLICIT_STRETCHED_CASES, ids=[c[0] for c in cases.STEP_IMPLICIT_STRETCHED_CASES])
def test_implicit_y_diffusion_stretched_matches_reference(gpu, oracle, case):
    cases.check_step_stretched(gpu, oracle, case)


@pytest.mark.parametrize("case", cases.STEP_IMPLICIT_2D_CASES, ids=[c[0] for c in cases.STEP_IMPLICIT_2D_CASES])
def test_implicit_y_diffusion_2d_matches_reference(gpu, oracle, case):
    cases.check_step_2d(gpu, oracle, case)


def test_implicit_y_diffusion_channel_128x64x64_vs_oracle(gpu, oracle):
    """the Thomas lines and the xz-only explicit term at a size where the momentum tiles and the big smoother tiles run"""
    case = ("chan128_implicit", (128, 64, 64), "PPWWPP", "rk3", "skew", 0, 1, "pchan", (1.0, 0, 0),
            dict(tol_rhs=1e-8, maxv=50, implicit_y=1, **cases.CHANNEL_UNIFORM))
    with tempfile.TemporaryDirectory() as td:
        meta, ref = cases.run_step_ref(case, td)
    got_meta, got = cases.run_step_backend(gpu, case)
    assert got_meta["cycles"] == meta["cycles"]
    cases.compare_states(got, ref, 1e-10, "chan128 implicit-y 1 step")


# ---- static LES SGS closures
@pytest.mark.parametrize("model", cases.LES_MODELS)
def test_les_sgs_update_matches_reference(oracle, gpu, model):
    cases.check_les_update(gpu, oracle, model, N=(24, 16, 12))


@pytest.mark.parametrize("model", ["smagorinsky", "sigma"])
def test_les_sgs_update_stretched_and_2d(oracle, gpu, model):
    cases.check_les_update(gpu, oracle, model, N=(16, 24, 8), stretched=True)
    cases.check_les_update(gpu, oracle, model, N=(32, 24, 1))


def test_step_with_smagorinsky_matches_reference(gpu, oracle):
    cases.check_step_with_les(gpu, oracle, "smagorinsky")


# ---- FFT Poisson solver
@pytest.mark.parametrize("case", cases.FFT_CASES, ids=[c[0] for c in cases.FFT_CASES])
def test_fft_poisson_matches_restatement(gpu, case):
    cases.check_fft_solve(gpu, case)


def test_fft_poisson_128x64x128(gpu):
    cases.check_fft_solve(gpu, ("fft_128x64x128", (128, 64, 128), False))


def test_fft_poisson_rejections(gpu):
    cases.check_fft_rejections(gpu)


@pytest.mark.parametrize("stretched", [False, True], ids=["uniform", "stretched"])
def test_step_with_fft_poisson_matches_converged_reference(gpu, oracle, stretched):
    cases.check_step_with_fft(gpu, oracle, stretched=stretched)


# ---- trip forcing + velocity filter
@pytest.mark.parametrize("kind", ["uniform", "stretched", "2d"])
def test_trip_forcing_adds_the_reference_term(gpu, kind):
    cases.check_trip_forcing(gpu, N=(32, 16, 1) if kind == "2d" else (32, 24, 16), stretched=kind == "stretched")


def test_velocity_filter_matches_restatement(gpu):
    cases.check_velocity_filter(gpu, N=(40, 24, 32))


# ---- SST k-omega transport + closure on 2-D meshes
@pytest.mark.parametrize("kind,ti", [("uniform", "rk3"), ("uniform", "euler"), ("stretched", "rk3")])
def test_step_with_sst_k_omega_matches_reference(gpu, oracle, kind, ti):
    cases.check_step_with_sst(gpu, oracle, stretched=kind == "stretched", ti=ti)


def test_step_with_sst_k_omega_192x96(gpu, oracle):
    """the x-y plane of BASELINE config 3 (192 x 96, tanh beta = 2), 5 steps"""
    cases.check_step_with_sst(gpu, oracle, stretched=True, ti="rk3", N=(192, 96, 1), nsteps=5, tol=1e-9, turb_tol=1e-9)


@pytest.mark.parametrize("kind", ["uniform", "stretched"])
def test_step_with_standard_k_omega_matches_reference(gpu, oracle, kind):
    cases.check_step_with_sst(gpu, oracle, stretched=kind == "stretched", ti="rk3", model="komega")


@pytest.mark.parametrize("model,kind", [("baseline", "uniform"), ("baseline", "stretched"), ("algebraic_komega", "uniform"),
                                        ("algebraic_komega", "stretched")])
def test_step_with_algebraic_rans_baselines(gpu, oracle, model, kind):
    cases.check_step_with_algebraic_rans(gpu, oracle, model=model, stretched=kind == "stretched")


@pytest.mark.parametrize("variant,kind", [(0, "uniform"), (1, "stretched"), (2, "uniform")])
def test_step_with_gep_closure(gpu, oracle, variant, kind):
    cases.check_step_with_algebraic_rans(gpu, oracle, model="gep", stretched=kind == "stretched", gep_variant=variant)


@pytest.mark.parametrize("model,kind", [("earsm_wj", "uniform"), ("earsm_gs", "stretched"), ("earsm_pope", "uniform"),
                                        ("earsm_wj", "stretched")])
def test_step_with_sst_earsm_closures(gpu, oracle, model, kind):
    cases.check_step_with_sst(gpu, oracle, stretched=kind == "stretched", ti="rk3", model=model)


def test_sst_rejects_3d_meshes(gpu):
    cases.check_sst_rejects_3d(gpu)
