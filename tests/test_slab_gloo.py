"""world_size-2 `gloo` run of the z-slab path on the CPU test double: the distributed V-cycle,
the velocity / pressure halo exchanges and the fused scalar all-reduces must reproduce the
SINGLE-DOMAIN result (oracle = the unmodified reference CPU build)."""
import os
import socket
import subprocess
import sys
import tempfile

import numpy as np
import pytest

import cases

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "slab_worker.py")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run_slabs(world, N, bc, ti, nsteps, fixed, td, gpu=False):
    port = _free_port()
    procs = [subprocess.Popen([sys.executable, WORKER, str(r), str(world), str(port), td, str(N[0]), str(N[1]), str(N[2]),
                               bc, ti, str(nsteps), str(fixed)] + (["gpu"] if gpu else []),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(world)]
    outs = [p.communicate(timeout=1500)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o[-3000:]
    return [np.load(os.path.join(td, f"slab{r}.npz")) for r in range(world)]


def _assemble(slabs, name, Nz):
    """Global ghost-inclusive array from the slabs' owned planes (ghost z planes from the end ranks)."""
    parts = []
    for r, s in enumerate(slabs):
        a = s[name]
        lo = 0 if r == 0 else 1
        hi = a.shape[0] if r == len(slabs) - 1 else 1 + Nz      # owned planes: 1..Nz (w: faces 1..Nz)
        parts.append(a[lo:hi])
    return np.concatenate(parts, axis=0)


@pytest.mark.parametrize("bc,ti,fixed,nsteps,N",
                         [("PPPPPP", "rk3", 4, 1, (16, 16, 32)), ("PPWWPP", "rk2", 0, 1, (16, 16, 32)),
                          ("PPWWWW", "euler", 3, 2, (16, 16, 32)), ("PPPPPP", "euler", 2, 1, (16, 16, 64))],
                         ids=["tgv_periodic_rk3", "channel_rk2_converge", "duct_z_walls_euler", "tgv_overlapped_halo"])
def test_two_slabs_match_single_domain(R, oracle, sim, bc, ti, fixed, nsteps, N, monkeypatch):
    _check_mg_slabs(R, 2, bc, ti, fixed, nsteps, N, monkeypatch)


def test_four_slabs_match_single_domain(R, oracle, sim, monkeypatch):
    """four ranks: interior ranks have two slab interfaces, the end ranks wrap around periodically"""
    _check_mg_slabs(R, 4, "PPPPPP", "rk3", 3, 1, (16, 16, 64), monkeypatch)


def _check_mg_slabs(R, world, bc, ti, fixed, nsteps, N, monkeypatch):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    # small boxes: keep the coarse levels distributed (the default replicates levels of <= 128^3 cells) and
    # put the fused residual + restriction on them, so the slab paths of both stay covered on CPU
    monkeypatch.setenv("CFDNN_MG_REPLICATE_CELLS", "600")
    if ti == "rk3":
        monkeypatch.setenv("CFDNN_MG_RR_MIN", "8")
    import slab_worker
    # (16, 16, 64): 32 planes per slab = 4 z chunks of the blocked smoother, so the halo exchange is overlapped
    # with the inner chunks (split launch, Multigrid::sweeps_tb)
    with tempfile.TemporaryDirectory() as td:
        slabs = _run_slabs(world, N, bc, ti, nsteps, fixed, td)
        u, v, w = slab_worker.global_init(*N, bc)
        pre = os.path.join(td, "in")
        u.tofile(pre + "_u.f64"); v.tofile(pre + "_v.f64"); w.tofile(pre + "_w.f64")
        force = (0.5, 0, 0) if "W" in bc else (0, 0, 0)
        extra = dict(tol_rhs=1e-8, maxv=50) if fixed == 0 else {}
        meta = R.run_ref("step", Nx=N[0], Ny=N[1], Nz=N[2], bc=bc, ti=ti, scheme="skew", fixed=fixed, nsteps=nsteps,
                         dt=0.01, nu=0.01, init="file:" + pre, call_initialize=1, fx=force[0], fy=force[1], fz=force[2],
                         out=os.path.join(td, "ref"), **extra)
        ref = R.load_state(os.path.join(td, "ref_final"), *N)
    Nz = N[2] // world
    got = {n: _assemble(slabs, n, Nz) for n in ("u", "v", "w", "p", "pcorr", "div", "rhs")}
    for n in got:
        assert got[n].shape == ref[n].shape, (n, got[n].shape, ref[n].shape)
    cases.compare_states(got, ref, 1e-10, f"{world} slabs {bc} {ti}")
    assert all(list(s_["cycles"]) == list(slabs[0]["cycles"]) for s_ in slabs)
    if meta["cycles"]:
        assert list(slabs[0]["cycles"]) == meta["cycles"]
    assert slabs[0]["residuals"] == pytest.approx(meta["residuals"], rel=1e-9)
    assert float(slabs[0]["adaptive_dt"]) == float(slabs[1]["adaptive_dt"])
    assert slabs[0]["stats"][0] > 0 and slabs[0]["stats"][1] > 0     # halos and all-reduces really ran


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_gpu_slabs_match_single_domain(R, oracle, gpu, world):
    """P B200s over NCCL vs the single-domain oracle (skipped when the box has fewer GPUs)."""
    if gpu.lib.cfdnn_device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import slab_worker
    N = (32, 32, 64)
    bc, ti, fixed, nsteps = "PPWWPP", "rk3", 0, 2
    with tempfile.TemporaryDirectory() as td:
        slabs = _run_slabs(world, N, bc, ti, nsteps, fixed, td, gpu=True)
        u, v, w = slab_worker.global_init(*N, bc)
        pre = os.path.join(td, "in")
        u.tofile(pre + "_u.f64"); v.tofile(pre + "_v.f64"); w.tofile(pre + "_w.f64")
        meta = R.run_ref("step", Nx=N[0], Ny=N[1], Nz=N[2], bc=bc, ti=ti, scheme="skew", fixed=fixed, nsteps=nsteps,
                         dt=0.01, nu=0.01, init="file:" + pre, call_initialize=1, fx=0.5, fy=0, fz=0,
                         out=os.path.join(td, "ref"), tol_rhs=1e-8, maxv=50)
        ref = R.load_state(os.path.join(td, "ref_final"), *N)
    Nz = N[2] // world
    got = {n: _assemble(slabs, n, Nz) for n in ("u", "v", "w", "p", "pcorr", "div", "rhs")}
    cases.compare_states(got, ref, 1e-10, f"{world} GPU slabs")
    assert list(slabs[0]["cycles"]) == meta["cycles"]


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_gpu_slabs_big_tiles_distributed_levels(R, oracle, gpu, world, monkeypatch):
    """The multi-GPU paths behind the scaling numbers, on real hardware over NCCL: a periodic 128 x 64 x (64 P)
    box keeps THREE multigrid levels distributed (CFDNN_MG_REPLICATE_CELLS small), so level 0 runs the 64x26-tile
    lean smoother with halo z sides, the split launch on the communication stream (overlap on) or the serial
    exchange (overlap off), slab coarse levels with 2-deep halos, the slab-aware fused residual + restriction
    (CFDNN_MG_RR_MIN=8) and the agglomerating all-gather.  Checked three ways: P slabs vs the single-GPU run
    of the same box (<= 1e-12), overlap on vs off (bitwise), and vs the single-domain oracle (<= 1e-10)."""
    if gpu.lib.cfdnn_device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import slab_worker
    monkeypatch.setenv("CFDNN_MG_REPLICATE_CELLS", "10000")
    monkeypatch.setenv("CFDNN_MG_RR_MIN", "8")
    N = (128, 64, 64 * world)
    bc, ti, fixed, nsteps = "PPPPPP", "rk3", 0, 1
    names = ("u", "v", "w", "p", "pcorr", "div", "rhs")
    got = {}
    for ov in ("1", "0"):
        monkeypatch.setenv("CFDNN_MG_OVERLAP", ov)
        with tempfile.TemporaryDirectory() as td:
            slabs = _run_slabs(world, N, bc, ti, nsteps, fixed, td, gpu=True)
            got[ov] = {n: _assemble(slabs, n, N[2] // world) for n in names}
            got[ov]["cycles"] = list(slabs[0]["cycles"])
    for n in names:
        assert np.array_equal(got["1"][n], got["0"][n]), f"overlap on/off differ in {n}"
    # single-GPU run of the same global box (same library, same environment, no communicator)
    u, v, w = slab_worker.global_init(*N, bc)
    cfg = gpu.config(nu=0.01, dt=0.01, time_integrator=R.TI[ti], convective_scheme=R.SCHEME["skew"], adaptive_dt=0,
                     poisson_fixed_cycles=fixed, poisson_tol_rhs=1e-8, poisson_max_vcycles=50)
    h = gpu.solver(gpu.mesh(*N), cfg, bc)
    gpu.upload(h, "u", u); gpu.upload(h, "v", v); gpu.upload(h, "w", w)
    gpu.check(gpu.lib.cfdnn_solver_initialize(h))
    gpu.step(h)
    single = gpu.state(h, *N, names=names)
    cyc_single = gpu.cycles(h)
    gpu.lib.cfdnn_solver_destroy(h)
    assert got["1"]["cycles"] == cyc_single
    for n in names:
        a, b = R.interior(n, got["1"][n]), R.interior(n, single[n])
        scale = max(np.abs(b).max(), 1e-300)
        assert np.abs(a - b).max() / scale <= (1e-12 if n not in ("div", "rhs") else 1e-9), n
    with tempfile.TemporaryDirectory() as td:
        pre = os.path.join(td, "in")
        u.tofile(pre + "_u.f64"); v.tofile(pre + "_v.f64"); w.tofile(pre + "_w.f64")
        meta = R.run_ref("step", Nx=N[0], Ny=N[1], Nz=N[2], bc=bc, ti=ti, scheme="skew", fixed=fixed, nsteps=nsteps,
                         dt=0.01, nu=0.01, init="file:" + pre, call_initialize=1, out=os.path.join(td, "ref"),
                         tol_rhs=1e-8, maxv=50)
        ref = R.load_state(os.path.join(td, "ref_final"), *N)
    assert got["1"]["cycles"] == meta["cycles"]
    cases.compare_states({n: got["1"][n] for n in names}, ref, 1e-10, f"{world} GPU slabs, big tiles")


def test_inconsistent_decomposition_is_rejected(sim):
    """comm_init validates the slab layout: equal slabs only (the reference's remainder rule,
    include/decomposition.hpp:81-92, is not supported), consistent Nz_global / k_offset — CFDNN_ERR_INVALID
    (std::invalid_argument through the shim) instead of silently wrong coarse right-hand sides."""
    import ctypes as C

    class Callbacks(C.Structure):
        _fields_ = [("ctx", C.c_void_p), ("sendrecv", C.c_void_p), ("allreduce", C.c_void_p), ("allgather", C.c_void_p)]

    cbs = Callbacks()
    fn = sim.lib.cfdnn_devsim_solver_comm_init
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(Callbacks)]
    cfg = sim.config()
    for kw, rank, what in ((dict(Nz_global=16, k_offset=0), 0, b"Nz_global"),          # global count not given
                           (dict(Nz_global=40, k_offset=0), 0, b"unequal z-slabs"),     # 40 planes on 2 ranks of 16
                           (dict(Nz_global=32, k_offset=0), 1, b"k_offset")):            # rank 1 must start at plane 16
        h = C.c_void_p()
        m = sim.mesh(16, 16, 16, **kw)
        sim.check(sim.lib.cfdnn_solver_create(C.byref(m), C.byref(cfg), C.byref(h)))
        rc = fn(h, rank, 2, C.byref(cbs))
        assert rc == 1 and what in sim.lib.cfdnn_last_error(), (rc, sim.lib.cfdnn_last_error())
        sim.lib.cfdnn_solver_destroy(h)


# ---- FFT Poisson on z-slabs (the reference's FFT_MPI role): local x transform, all-to-all, z transform + Thomas on
# this rank's kx columns, all-to-all back
def _single_domain_fft(B, R, slab_worker, N, bc, ti, nsteps, stretched=False):
    u, v, w = slab_worker.global_init(*N, bc)
    cfg = B.config(nu=0.01, dt=0.01, time_integrator=R.TI[ti], convective_scheme=R.SCHEME["skew"], adaptive_dt=0,
                   poisson_fixed_cycles=0, poisson_tol_rhs=1e-8, poisson_max_vcycles=50, poisson_solver=1)
    if stretched:
        L = 2 * np.pi
        md = B.stretched_mesh(*N, R.backend.stretched_y_metrics(N[1], -1.0, 1.0, 2.0), Lx=L, Lz=L, y_lo=-1.0, y_hi=1.0)
    else:
        md = B.mesh(*N)
    h = B.solver(md, cfg, bc, (0.5, 0.0, 0.0))
    B.upload(h, "u", u); B.upload(h, "v", v); B.upload(h, "w", w)
    B.check(B.lib.cfdnn_solver_initialize(h))
    for _ in range(nsteps):
        B.step(h)
    st = B.state(h, *N, names=("u", "v", "w", "p", "pcorr", "div", "rhs"))
    cyc = B.cycles(h)
    B.lib.cfdnn_solver_destroy(h)
    return st, cyc


def _check_fft_slabs(B, R, world, N, ti, nsteps, monkeypatch, on_gpu, p2p=None, stretched=False):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import slab_worker
    bc = "PPWWPP"
    monkeypatch.setenv("SLAB_POISSON", "fft")
    if stretched:
        monkeypatch.setenv("SLAB_STRETCH", "1")
    if p2p is not None:
        monkeypatch.setenv("CFDNN_FFT_P2P", "1" if p2p else "0")
    with tempfile.TemporaryDirectory() as td:
        slabs = _run_slabs(world, N, bc, ti, nsteps, 0, td, gpu=on_gpu)
    if p2p is not None:                             # the transposes really took the requested route on every rank
        assert [int(s_["fft_p2p"]) for s_ in slabs] == [1 if p2p else 0] * world
    names = ("u", "v", "w", "p", "pcorr", "div", "rhs")
    got = {n: _assemble(slabs, n, N[2] // world) for n in names}
    single, cyc = _single_domain_fft(B, R, slab_worker, N, bc, ti, nsteps, stretched)
    assert list(slabs[0]["cycles"]) == cyc and all(c == 1 for c in cyc)
    for n in names:
        a, b = R.interior(n, got[n]), R.interior(n, single[n])
        scale = max(np.abs(b).max(), 1e-300) if n not in ("div", "rhs", "pcorr") else max(np.abs(R.interior("p", single["p"])).max(), np.abs(b).max())
        assert np.abs(a - b).max() / scale <= 1e-12, (n, np.abs(a - b).max(), scale)


def test_two_slabs_fft_poisson_match_single_domain(R, sim, monkeypatch):
    _check_fft_slabs(sim, R, 2, (24, 12, 16), "rk3", 2, monkeypatch, on_gpu=False)


@pytest.mark.parametrize("world,N", [(4, (12, 6, 16)), (3, (10, 5, 12))], ids=["4_ranks", "3_ranks"])
def test_more_slabs_fft_poisson_ragged_columns(R, sim, monkeypatch, world, N):
    """more ranks than kx column blocks fill evenly: Nxh = 7 (6) columns over 4 (3) ranks of 8 padded columns each, so the
    last ranks own no mode at all; z-slabs of 4 planes"""
    _check_fft_slabs(sim, R, world, N, "rk3", 1, monkeypatch, on_gpu=False)


def test_two_slabs_fft_poisson_stretched_y(R, sim, monkeypatch):
    """a y-stretched channel on z-slabs: possible with the FFT solver only (the semi-coarsening multigrid has no slab variant)"""
    _check_fft_slabs(sim, R, 2, (16, 16, 16), "rk3", 1, monkeypatch, on_gpu=False, stretched=True)


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_gpu_slabs_fft_poisson(R, gpu, world, monkeypatch):
    if gpu.lib.cfdnn_device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    # peer-to-peer transposes (the transform kernels store into the other GPUs' arrays through CUDA IPC) and NCCL all-to-alls
    _check_fft_slabs(gpu, R, world, (128, 48, 32 * world), "rk3", 1, monkeypatch, on_gpu=True, p2p=True)
    _check_fft_slabs(gpu, R, world, (128, 48, 32 * world), "rk3", 2, monkeypatch, on_gpu=True, p2p=False)
    monkeypatch.delenv("CFDNN_FFT_P2P")
    _check_fft_slabs(gpu, R, world, (64, 48, 16 * world), "rk3", 1, monkeypatch, on_gpu=True, stretched=True)
