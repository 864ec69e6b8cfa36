#!/usr/bin/env python
"""bench.py — Mcell-steps/s of the fractional-step hot path on N B200s; one JSON line on stdout (rank 0).

  python bench.py --gpus 1 --steps K --warmup W                 # this backend (CUDA, C ABI)
  python bench.py --impl reference --steps K --warmup W          # reference CPU build (oracle/_ref)

Default workload = the north star: 3-D Taylor-Green vortex 512^3 per GPU, SSP-RK3, skew-symmetric
convection, multigrid Poisson to 1e-8 relative residual, NN-TBNN closure (tbnn_channel_caseholdout,
per-cell 5-64-64-64-4 MLP on the FP64 tensor cores) evaluated on every cell every step.

value      : device-resident throughput (fields live in HBM; CUDA events around K steps)
e2e        : the same steps driven through the C ABI with HOST buffers — every step uploads u,v,w,p
             from pinned host memory and downloads u,v,w,p
roofline   : dominant kernel (level-0 temporally blocked smoother): compulsory bytes of one launch
             (u, f in, u out once = 24 B per cell) over its CUDA-event time vs MEASURED_PEAKS.json hbm_gbs;
             .nn = the closure kernel inside the step (FP64 tensor pipe); .vcycle / .step = whole V-cycle /
             whole step in the per-sweep traffic model of SURVEY §8(d), the fused-kernel model and, where a
             committed ncu capture exists, measured DRAM bytes
cpu_baseline: the unmodified reference CPU build (single-threaded by construction) on a bounded
             sample (64^3 sub-grid, a few steps) of the same workload, timed on this box's host cores
parity     : a small slab-decomposed box stepped on the same ranks and compared with the single-domain
             run of the same box on rank 0 (N > 1), or rerun bitwise (N = 1)
"""
import argparse
import ctypes as C
import importlib.util
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")          # cpu_baseline / --impl reference only
PRESETS_NPZ = os.path.join(ROOT, "tests", "golden", "models.npz")      # data fixture (weights of the reference presets)
PRESET = "tbnn_channel_caseholdout"

# workloads: tgvN = periodic Taylor-Green box, N^3 cells PER GPU (weak scaling, BASELINE config 2 scaled up);
# chan* = channel (periodic x/z, no-slip walls in y), ONE global box split into z-slabs (strong scaling):
# chan1024x512x1024 = BASELINE config 5, chan192x96x192 = BASELINE config 3 (tanh-stretched y, 16 fixed cycles)
WORKLOADS = {
    "tgv64": dict(kind="tgv", N=64), "tgv128": dict(kind="tgv", N=128), "tgv256": dict(kind="tgv", N=256),
    "tgv384": dict(kind="tgv", N=384), "tgv512": dict(kind="tgv", N=512),
    "chan1024x512x1024": dict(kind="chan", N=(1024, 512, 1024), nu=0.005556, fixed=0),
    "chan512x256x512": dict(kind="chan", N=(512, 256, 512), nu=0.005556, fixed=0),
    "chan256x128x256": dict(kind="chan", N=(256, 128, 256), nu=0.005556, fixed=0),
    "chan192x96x192": dict(kind="chan", N=(192, 96, 192), nu=0.005556, fixed=16, beta=2.0),
}
VCYCLE_BYTES_PER_CELL = 500.0      # 438 B at level 0, x 8/7 for the coarse levels (one sweep per pass)
VCYCLE_FUSED_BYTES_PER_CELL = 281.0  # same V-cycle with two sweeps per pass and the fused residual+restriction
STEP_BYTES_FIXED, STEP_BYTES_PER_VC = 1770.0, 508.0   # RK3 step = 1770 + 508 n_vc B per cell
# compulsory bytes of THIS backend's fused kernels: copy2 72 + nu_eff 8 + momentum 56+80+80 + blends 144
# + 3 projections x (div 32 + rhs 16 + copy_in 32 + norms 24 + exit mean 8 + copy_out 16 + copy_correct 96) + max_change 48
# + NaN guard 40/5; per V-cycle 281 + 16/3 for the convergence check every third cycle; closure +80 (u, v, k, omega,
# k-provided test in; nu_t, 3 tau out; nu_t into nu_eff)
STEP_FUSED_FIXED, STEP_FUSED_PER_VC, STEP_FUSED_CLOSURE = 1168.0, 286.0, 80.0
NN_FLOP_PER_CELL = 17536.0         # 2 (5*64 + 64*64 + 64*64 + 64*4)
MG_TOL_RHS, MG_MAXV = 1e-8, 50     # north_star: "MG Poisson to 1e-8 relative residual"


def load_package():
    """cfd-nn_b200/ (hyphenated directory) imported from its path as module `cfdnn_b200`"""
    if "cfdnn_b200" in sys.modules:
        return sys.modules["cfdnn_b200"]
    spec = importlib.util.spec_from_file_location(
        "cfdnn_b200", os.path.join(ROOT, "cfd-nn_b200", "__init__.py"),
        submodule_search_locations=[os.path.join(ROOT, "cfd-nn_b200")])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["cfdnn_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def fp64_tensor_peak():
    """FP64 tensor (DMMA) peak measured by scripts/fp64_peaks.cu on this pool (profiles/fp64_peaks.json)"""
    try:
        j = json.load(open(os.path.join(ROOT, "profiles", "fp64_peaks.json")))
        return j["dmma_TF"], "measured (profiles/fp64_peaks.json, scripts/fp64_peaks.cu)"
    except Exception:
        return None, "unmeasured"


def traffic_entry(name):
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(name)
    except Exception:
        return None


class ClockSampler(threading.Thread):
    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                o = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i",
                                    str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                f = [x.strip() for x in o.split(",")]
                self.samples.append(float(f[0])); self.max_mhz = float(f[1])
                for n, v in zip(names, f[2:]):
                    if v.lower().startswith("active"):
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.2)

    def result(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def pin_to_gpu_numa(local):
    """Bind this rank's host threads (and so the first-touch of its pinned staging buffers) to the NUMA node
    of its GPU: with 8 ranks on one socket the device->host copies of the e2e leg fought over one memory
    controller (r1: 76 -> 465 ms per step per rank from 1 to 8 GPUs)."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(local)       # honours CUDA_VISIBLE_DEVICES, unlike NVML indices
        bus = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        allowed = os.sched_getaffinity(0) & cpus
        if allowed:
            os.sched_setaffinity(0, allowed)
            return {"numa_node": node, "cpus": len(allowed)}
    except Exception:
        pass
    return None


def run_ref(mode, **kw):
    o = subprocess.run([HARNESS, f"mode={mode}"] + [f"{k}={v}" for k, v in kw.items()], capture_output=True, text=True)
    if o.returncode != 0:
        raise RuntimeError(f"ref_harness failed ({o.returncode}): {o.stderr[-2000:]}")
    return json.loads(o.stdout.strip().splitlines()[-1])


def cpu_reference(workload, closure, steps, warmup):
    """The UNMODIFIED reference CPU build on a bounded sample of the workload: a 64^3 sub-grid (the reference
    is single-threaded; 512^3 is ~9 min per step).  With the closure the reference's own 3-D step evaluates
    TurbulenceModel::update on ONE z-plane only (it is a 2-D model; SURVEY §8 N6 defines the 3-D extension as
    the 2-D update on every interior plane) — the remaining Nz-1 plane updates are timed with the
    reference's 2-D update() on a plane of the same size and added."""
    import numpy as np
    w = WORKLOADS[workload]
    N = 64
    kw = dict(Nx=N, Ny=N, Nz=N, ti="rk3", scheme="skew", nsteps=steps, warmup=warmup)
    if w["kind"] == "tgv":
        kw.update(bc="PPPPPP", init="tgv", fixed=0, tol_rhs=MG_TOL_RHS, maxv=MG_MAXV,
                  dt=0.01 if closure == "none" else 1e-4 * 8, nu=0.01 if closure == "none" else 1.0 / 180.0)
        what = f"{N}^3 TGV sub-grid"
    else:
        kw.update(bc="PPWWPP", init="shear", fx=1.0, fixed=w["fixed"], tol_rhs=MG_TOL_RHS, maxv=MG_MAXV, dt=2e-3,
                  nu=w["nu"], x1=4 * np.pi, y0=-1.0, y1=1.0, z1=2 * np.pi)
        if "beta" in w:
            kw.update(stretch_beta=w["beta"])
        what = f"{N}^3 channel sub-grid" + (" (tanh-stretched y, 16 fixed cycles)" if "beta" in w else "")
    plane_ms = 0.0
    with tempfile.TemporaryDirectory() as td:
        if closure not in ("none", "tbnn"):          # static LES SGS model: the reference's own 3-D LESModel::update
            kw.update(model=closure)
        if closure == "tbnn":
            pkg = load_package()
            W, b, means, stds = pkg.backend.load_npz_preset(PRESETS_NPZ, PRESET)
            d = os.path.join(td, "m")
            os.makedirs(d)
            for l, (wm, bb) in enumerate(zip(W, b)):
                np.savetxt(os.path.join(d, f"layer{l}_W.txt"), wm, fmt="%.17g")
                np.savetxt(os.path.join(d, f"layer{l}_b.txt"), bb, fmt="%.17g")
            np.savetxt(os.path.join(d, "input_means.txt"), means, fmt="%.17g")
            np.savetxt(os.path.join(d, "input_stds.txt"), stds, fmt="%.17g")
            kw.update(model="nn_tbnn", nn_weights=d)
            # 2-D update() on one N x N plane, `rep` times
            sh = pkg.backend.shapes(N, N, 1)
            u2, v2, _ = pkg.backend.tgv_init(N, N, 1)
            pre = os.path.join(td, "p")
            np.ascontiguousarray(u2.reshape(sh["u"])).tofile(pre + "_u.f64")
            np.ascontiguousarray(v2.reshape(sh["v"])).tofile(pre + "_v.f64")
            rep = 8
            mn = run_ref("nn", Nx=N, Ny=N, Nz=1, model="nn_tbnn", nn_weights=d, nu=1.0 / 180.0, repeat=rep, **{"in": pre})
            plane_ms = mn["seconds"] / rep * 1e3
        m = run_ref("bench", **kw)
    ms = m["ms_per_step"] + plane_ms * (N - 1 if closure == "tbnn" else 0)
    v = N ** 3 / (ms * 1e-3) / 1e6
    sample = (f"{what} of the workload, {steps} steps after {warmup} warm-up, unmodified reference CPU build "
              "(single-threaded by construction)")
    if closure == "tbnn":
        sample += (f"; closure: the reference step evaluates 1 z-plane ({m['ms_per_step']:.0f} ms/step incl.), the other "
                   f"{N - 1} planes timed with its 2-D update() ({plane_ms:.1f} ms per {N}x{N} plane) and added")
    return v, ms, sample


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    v, ms, sample = cpu_reference(args.workload, args.closure, args.steps, args.warmup)
    line = {"impl": "reference", "metric": "Mcell-steps/s (3D RK3+MG step)", "value": v, "unit": "Mcell-steps/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": scaling_of(args.workload), "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": {"workload": args.workload, "closure": args.closure, "sample": sample},
            "cpu_baseline": {"value": v, "unit": "Mcell-steps/s", "cores": 1, "kind": "reference", "sample": sample},
            "e2e": {"value": v, "unit": "Mcell-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def scaling_of(workload):
    return "weak" if WORKLOADS[workload]["kind"] == "tgv" else "strong"


def split_planes(nz, nranks, rank):
    """block distribution with remainder (reference include/decomposition.hpp:81-92)"""
    base, rem = divmod(nz, nranks)
    n = base + (1 if rank < rem else 0)
    off = rank * base + min(rank, rem)
    return n, off


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="tgv512", choices=sorted(WORKLOADS))
    ap.add_argument("--closure", default=None, choices=["none", "tbnn", "smagorinsky", "wale", "vreman", "sigma"],
                    help="tbnn (default for tgv*): the NN-TBNN closure is attached to the step — the north star's "
                         "'RK3 + MG + NN-TBNN' workload; none: closure-free step")
    ap.add_argument("--adaptive-dt", action="store_true",
                    help="call compute_adaptive_dt + set_dt before every step (the CFL reductions in the timed region)")
    ap.add_argument("--implicit-y", action="store_true",
                    help="implicit y-diffusion (Thomas solve per y-line after every RK stage; channel workloads)")
    ap.add_argument("--poisson", default="mg", choices=["mg", "fft"],
                    help="Poisson back end of the projection: mg (default, the north-star path) or fft (FFTPoissonSolver: "
                         "periodic x/z + walls y, i.e. the chan* workloads, single GPU)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-second", action="store_true", help="skip the closure-free second measurement")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if args.poisson == "fft" and wl["kind"] != "chan":
        raise SystemExit("--poisson fft needs a channel workload (periodic x/z, walls y): the solver falls back to the multigrid "
                         "on a fully periodic box, which is not what this flag is for")
    if args.closure is None:
        args.closure = "tbnn" if wl["kind"] == "tgv" else "none"
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    pkg = load_package()
    K = pkg.backend
    capi = pkg._capi

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    numa = pin_to_gpu_numa(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B = K.Backend()
    B.check(B.lib.cfdnn_set_device(local))

    def comm_init(h):
        idbuf = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            raw = (C.c_char * 128)()
            B.check(B.lib.cfdnn_comm_unique_id(raw))
            idbuf = torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8).clone()
        idbuf = idbuf.cuda()
        dist.broadcast(idbuf, 0)
        raw = (C.c_char * 128).from_buffer_copy(bytes(idbuf.cpu().numpy().tobytes()))
        B.check(B.lib.cfdnn_solver_comm_init(h, rank, world, raw))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------------------------------------------------------- the workload
    L = 2 * np.pi
    stretched = None
    if wl["kind"] == "tgv":
        N = wl["N"]
        Nx = Ny = Nzl = N
        Nzg, koff = N * world, rank * N                 # weak scaling: every GPU owns an N^3 slab
        Lx, Ly, Lz, y_lo = L, L, L * world, 0.0
        bc = "PPPPPP"
        # dt keeps the CFL number of BASELINE config 2 (64^3, dt = 0.01): dt = 0.01 * 64 / N (dt = 0.01 at 512^3
        # is CFL ~ 0.8 and diverges).  With the closure nu_t reaches 10 nu: the explicit diffusive limit
        # dx^2 / (6 nu_eff) forces the smaller step.
        nu_, dt_ = (0.01, 0.01 * 64.0 / N) if args.closure == "none" else (1.0 / 180.0, 1e-4 * 512.0 / N)
        fixed, force = 0, (0.0, 0.0, 0.0)
    else:
        Nx, Ny, Nzg = wl["N"]
        Nzl, koff = split_planes(Nzg, world, rank)      # strong scaling: one global box, z-slabs
        Lx, Ly, Lz, y_lo = 4 * np.pi, 2.0, 2 * np.pi, -1.0
        bc = "PPWWPP"
        nu_, fixed, force = wl["nu"], wl["fixed"], (1.0, 0.0, 0.0)
        dy_min = Ly / Ny
        if "beta" in wl:
            if world > 1:
                raise SystemExit("the y-stretched (semi-coarsening) workload is single-GPU")
            stretched = K.stretched_y_metrics(Ny, y_lo, y_lo + Ly, wl["beta"])
            dy_min = float(stretched["dyv"].min())
        dt_ = min(0.3 * (Lx / Nx) / 1.1, 0.2 * dy_min * dy_min / nu_)
    if stretched is not None:
        mesh = B.stretched_mesh(Nx, Ny, Nzl, stretched, Lx=Lx, Lz=Lz, y_lo=y_lo, y_hi=y_lo + Ly)
    else:
        mesh = B.mesh(Nx, Ny, Nzl, Lx=Lx, Ly=Ly, Lz=Lz, Nz_global=Nzg, k_offset=koff)
        mesh.y_lo, mesh.y_hi = y_lo, y_lo + Ly
    cfg = B.config(nu=nu_, dt=dt_, time_integrator=K.TI["rk3"], convective_scheme=K.SCHEME["skew"],
                   adaptive_dt=1 if args.adaptive_dt else 0, poisson_fixed_cycles=fixed, poisson_tol_rhs=MG_TOL_RHS,
                   poisson_max_vcycles=MG_MAXV, implicit_y_diffusion=1 if args.implicit_y else 0,
                   poisson_solver=1 if args.poisson == "fft" else 0)
    h = B.solver(mesh, cfg, bc, force)
    nn = None
    if args.closure == "tbnn":
        W, b, means, stds = K.load_npz_preset(PRESETS_NPZ, PRESET)
        nn = K.nn_create(B, "tbnn", W, b, means, stds)
        B.check(B.lib.cfdnn_nn_set_params(nn, nu_, 1.0, 1.0, 1e-10, 1.0))
        B.check(B.lib.cfdnn_solver_set_turbulence_model(h, nn))
    elif args.closure in K.LES_MODEL:       # static LES SGS closure, default constant (reference turbulence_les.hpp)
        B.check(B.lib.cfdnn_solver_set_les_model(h, K.LES_MODEL[args.closure], -1.0))
    if world > 1:
        comm_init(h)
    fft_p2p = bool(B.lib.cfdnn_solver_fft_peer_to_peer(h))

    # synthetic input, this rank's slab of the global box (pinned host arrays)
    sh = K.shapes(Nx, Ny, Nzl)
    host = {n: torch.zeros(sh[k], dtype=torch.float64).pin_memory() for n, k in
            (("u", "u"), ("v", "v"), ("w", "w"), ("p", "c"))}
    dx, dy, dz = Lx / Nx, Ly / Ny, Lz / Nzg
    xc = (np.arange(Nx + 2) - 0.5) * dx; xf = (np.arange(Nx + 3) - 1) * dx
    yc = (np.arange(Ny + 2) - 0.5) * dy; yf = (np.arange(Ny + 3) - 1) * dy
    zc = (np.arange(Nzl + 2) - 0.5 + koff) * dz; zf = (np.arange(Nzl + 3) - 1 + koff) * dz
    if wl["kind"] == "tgv":        # Taylor-Green vortex on faces (reference app/main_taylor_green_3d.cpp:253-276)
        cz = np.cos(zc)[1:-1, None, None]
        host["u"].numpy()[1:-1, 1:-1, 1:Nx + 2] = cz * np.cos(yc[1:-1])[None, :, None] * np.sin(xf[1:Nx + 2])[None, None, :]
        host["v"].numpy()[1:-1, 1:Ny + 2, 1:-1] = -cz * np.sin(yf[1:Ny + 2])[None, :, None] * np.cos(xc[1:-1])[None, None, :]
    else:
        # perturbed channel (reference app/main_channel.cpp create_perturbed_channel_field: parabolic profile plus a
        # divergence-free sin/cos perturbation, amplitude 0.05, deterministic), written on normalised coordinates
        if stretched is not None:
            Yc = (stretched["yc"] - y_lo) / Ly
            Yf = (stretched["yf"] - y_lo) / Ly
        else:
            Yc, Yf = yc / Ly, yf / Ly
        X_f, X_c = xf / Lx, xc / Lx
        Z_c, Z_f = zc / Lz, zf / Lz
        a = 0.05
        u = host["u"].numpy(); v = host["v"].numpy(); w = host["w"].numpy()
        u[:] = ((4 * Yc * (1 - Yc))[None, :, None] + a * np.sin(2 * np.pi * Z_c)[:, None, None] *
                np.cos(np.pi * Yc)[None, :, None] * np.sin(4 * np.pi * X_f)[None, None, :])
        v[:] = a * np.cos(2 * np.pi * Z_c)[:, None, None] * np.sin(np.pi * Yf)[None, :, None] * np.cos(4 * np.pi * X_c)[None, None, :]
        w[:] = a * np.cos(2 * np.pi * Z_f)[:, None, None] * np.sin(np.pi * Yc)[None, :, None] * np.sin(2 * np.pi * X_c)[None, None, :]
    fid = K.FIELD_ID

    def upload_all():
        for n, t in host.items():
            B.check(B.lib.cfdnn_solver_upload(h, fid[n], C.cast(t.data_ptr(), capi.c_double_p), t.numel()))

    def download_all():
        for n, t in host.items():
            B.check(B.lib.cfdnn_solver_download(h, fid[n], C.cast(t.data_ptr(), capi.c_double_p), t.numel()))

    def timing():
        a_, b_, c_ = C.c_double(), C.c_double(), C.c_int()
        B.lib.cfdnn_solver_last_timing(h, C.byref(a_), C.byref(b_), C.byref(c_))
        ph = (C.c_double * 4)()
        B.lib.cfdnn_solver_phase_timing(h, ph)
        return a_.value, b_.value, c_.value, list(ph)

    r = C.c_double()
    dts = []

    def run_steps(n):
        if not args.adaptive_dt:
            B.check(B.lib.cfdnn_solver_run(h, n, C.byref(r)))
            return
        for _ in range(n):                      # app time loop: dt = compute_adaptive_dt(); set_dt(dt); step()
            d = C.c_double()
            B.check(B.lib.cfdnn_solver_compute_adaptive_dt(h, C.byref(d)))
            B.check(B.lib.cfdnn_solver_set_dt(h, d.value))
            dts.append(d.value)
            B.check(B.lib.cfdnn_solver_run(h, 1, C.byref(r)))

    def timed(n):
        """exactly n steps between barriers; device ms (CUDA events inside the library, summed per run call)"""
        barrier()
        l0 = B.lib.cfdnn_kernel_launches()
        t0 = time.perf_counter()
        if not args.adaptive_dt:
            run_steps(n)
            step_ms, vc_ms, n_vc, ph = timing()
        else:
            step_ms = vc_ms = 0.0; n_vc = 0; ph = [0.0] * 4
            for _ in range(n):
                run_steps(1)
                a_, b_, c_, p_ = timing()
                step_ms += a_; vc_ms += b_; n_vc += c_; ph = [x + y for x, y in zip(ph, p_)]
        barrier()
        wall_ms = (time.perf_counter() - t0) * 1e3
        return dict(step_ms=max_over_ranks(step_ms), vc_ms=vc_ms, n_vc=n_vc, phases=ph, wall_ms=wall_ms,
                    launches=B.lib.cfdnn_kernel_launches() - l0)

    upload_all()
    B.check(B.lib.cfdnn_solver_initialize(h))
    run_steps(args.warmup)
    # ---- device-resident timed region: exactly K steps
    sampler = ClockSampler(local)
    sampler.start()
    T = timed(args.steps)
    sampler.stop_flag = True
    sampler.join(timeout=2)
    cycles = B.cycles(h)
    cells_local = Nx * Ny * Nzl
    cells = Nx * Ny * Nzg
    value = cells * args.steps / (T["step_ms"] * 1e-3) / 1e6

    # ---- the closure-free variant of the same workload (second key, not the headline)
    second = None
    if nn and not args.no_second:
        B.check(B.lib.cfdnn_solver_set_turbulence_model(h, None))
        run_steps(1)
        T2 = timed(args.steps)
        second = {"value": cells * args.steps / (T2["step_ms"] * 1e-3) / 1e6, "unit": "Mcell-steps/s",
                  "ms_per_step": T2["step_ms"] / args.steps, "vcycles_per_solve": B.cycles(h)[-3:],
                  "note": "same run, closure detached (nu_t keeps its last value)"}
        B.check(B.lib.cfdnn_solver_set_turbulence_model(h, nn))

    # ---- end to end through the C ABI with host buffers
    e2e = None
    if not args.no_e2e:
        nbytes = sum(t_.numel() * 8 for t_ in host.values())
        download_all()
        upload_all(); run_steps(1); download_all()
        barrier()
        t0 = time.perf_counter()
        ph = [0.0, 0.0, 0.0]
        for _ in range(args.steps):
            ta = time.perf_counter()
            upload_all()
            tb = time.perf_counter()
            run_steps(1)
            tc = time.perf_counter()
            download_all()
            td = time.perf_counter()
            ph[0] += tb - ta; ph[1] += tc - tb; ph[2] += td - tc
        barrier()
        e2e_s = max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": cells * args.steps / e2e_s / 1e6, "unit": "Mcell-steps/s",
               "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": nbytes,
               "ms_per_step": {"h2d": ph[0] / args.steps * 1e3, "step": ph[1] / args.steps * 1e3,
                               "d2h": ph[2] / args.steps * 1e3}, "host_numa": numa}

    # dominant kernel, timed live with CUDA events on the solver's stream (after the timed steps, same
    # resident level-0 arrays, >> L2): the level-0 temporally blocked smoother
    sm_ms = C.c_double(0.0)
    B.check(B.lib.cfdnn_solver_time_smoother(h, 20, C.byref(sm_ms)))
    B.lib.cfdnn_solver_destroy(h)
    if nn:
        B.lib.cfdnn_nn_destroy(nn)

    # ---- parity evidence carried by the line itself
    parity = None
    if not args.no_parity:
        parity = parity_check(B, K, np, torch, dist, rank, world, comm_init)

    if rank == 0:
        peak, which = peaks()
        n_vc = T["n_vc"]
        vc_avg_ms = T["vc_ms"] / max(n_vc, 1)
        per_step_vc = n_vc / args.steps
        step_s = T["step_ms"] / args.steps * 1e-3

        def gbs(bytes_, ms):
            return bytes_ / (ms * 1e-3) / 1e9 if ms and ms > 0 else None

        def frac(x):
            return x / peak if x else None

        # level-0 smoother: one launch = two sweeps; compulsory traffic of the fused launch = u, f in, u out once
        comp = 24.0 * cells_local
        tj = traffic_entry("k_sweep_tbo<64,26>" if bc == "PPPPPP" else "k_sweep_tb<2,64,26>")
        traffic = tj["dram_bytes_per_launch"] if tj and tj.get("cells") == cells_local else None
        ach = gbs(comp, sm_ms.value)
        roof = {"bound": "hbm",
                "kernel": ("k_sweep_tbo<64,26>" if bc == "PPPPPP" else "k_sweep_tb<2,64,26>") +
                          ": level-0 smoother, 2 Chebyshev sweeps per launch",
                "achieved": ach, "peak": peak, "unit": "GB/s", "frac": frac(ach), "traffic": traffic,
                "frac_of_traffic": frac(gbs(traffic, sm_ms.value)) if traffic else None,
                "peak_source": which, "algorithmic_bytes_per_launch": comp,
                "algorithmic_bytes_note": "u, f read once, u written once per launch (24 B per cell per 2 sweeps)",
                "launch_ms": sm_ms.value, "launches_timed": 20,
                "model_achieved": gbs(2 * comp, sm_ms.value), "model_frac": frac(gbs(2 * comp, sm_ms.value)),
                "model_note": "SURVEY 8(d) per-sweep model (24 B per cell per SWEEP): what an unblocked smoother would "
                              "have to sustain; > 1 is the temporal-blocking gain, not a hardware fraction",
                "vcycle": {"ms": vc_avg_ms, "timed": n_vc, "share_of_step": T["vc_ms"] / T["step_ms"],
                           "fused_bytes": VCYCLE_FUSED_BYTES_PER_CELL * cells_local,
                           "achieved": gbs(VCYCLE_FUSED_BYTES_PER_CELL * cells_local, vc_avg_ms),
                           "frac": frac(gbs(VCYCLE_FUSED_BYTES_PER_CELL * cells_local, vc_avg_ms)),
                           "model_bytes": VCYCLE_BYTES_PER_CELL * cells_local,
                           "model_achieved": gbs(VCYCLE_BYTES_PER_CELL * cells_local, vc_avg_ms),
                           "model_frac": frac(gbs(VCYCLE_BYTES_PER_CELL * cells_local, vc_avg_ms))} if n_vc else None}
        sj = traffic_entry("step:" + args.workload + ":" + args.closure)
        fused_b = (STEP_FUSED_FIXED + STEP_FUSED_PER_VC * per_step_vc + (STEP_FUSED_CLOSURE if nn else 0.0)) * cells_local
        model_b = (STEP_BYTES_FIXED + STEP_BYTES_PER_VC * per_step_vc) * cells_local
        roof["step"] = {"vcycles_per_step": per_step_vc,
                        "fused_frac": frac(fused_b / step_s / 1e9), "model_frac": frac(model_b / step_s / 1e9),
                        "measured_dram_frac": frac(sj["dram_bytes_per_step"] / step_s / 1e9) if sj else None,
                        "note": "fused = compulsory bytes of this backend's fused kernels; model = SURVEY 8(d) per-sweep "
                                "traffic model (exceeds 1 through temporal blocking); measured = ncu DRAM bytes of a "
                                "committed capture of the same workload (profiles/traffic.json)"}
        if nn:
            nn_ms = T["phases"][0] / args.steps
            tf = cells_local * NN_FLOP_PER_CELL / (nn_ms * 1e-3) / 1e12 if nn_ms > 0 else None
            tpeak, tsrc = fp64_tensor_peak()
            nj = traffic_entry("k_nn_update_mmah")
            roof["nn"] = {"bound": "tensor", "kernel": "k_nn_update_mmah<TBNN,8> (mma.sync.m8n8k4.f64 = DMMA.8x8x4)",
                          "ms_per_step": nn_ms, "share_of_step": T["phases"][0] / T["step_ms"],
                          "mcells_per_s": cells_local / (nn_ms * 1e-3) / 1e6 if nn_ms > 0 else None,
                          "achieved": tf, "unit": "TFLOP/s", "peak": tpeak, "peak_source": tsrc,
                          "frac": tf / tpeak if tf and tpeak else None,
                          "dmma_pipe_pct": nj.get("dmma_pipe_pct") if nj else None}
        roof["momentum"] = {"ms_per_step": T["phases"][1] / args.steps, "launches_per_step": 3,
                            "achieved": gbs(3 * 200.0 * cells_local, T["phases"][1] / args.steps),
                            "frac": frac(gbs(3 * 200.0 * cells_local, T["phases"][1] / args.steps)),
                            "note": "fused conv + diff + advance, 200 B per cell per stage (SURVEY 8(d))"}
        if args.closure in K.LES_MODEL:
            les_ms = T["phases"][0] / args.steps
            roof["les"] = {"bound": "hbm", "kernel": f"k_les_sgs<{args.closure}> (velocity-gradient tensor + nu_sgs, one pass)",
                           "ms_per_step": les_ms, "share_of_step": T["phases"][0] / T["step_ms"],
                           "achieved": gbs(32.0 * cells_local, les_ms), "frac": frac(gbs(32.0 * cells_local, les_ms)),
                           "note": "compulsory 32 B per cell: u, v, w read once, nu_t written"}
        if len(T["phases"]) > 3:
            cells_b = 100.0 * cells_local
            pms = T["phases"][3] / args.steps
            roof["poisson"] = {"solver": args.poisson, "ms_per_step": pms, "solves_per_step": 3,
                               "share_of_step": T["phases"][3] / T["step_ms"]}
            if args.poisson == "fft":
                if world > 1:
                    roof["poisson"]["transposes"] = ("peer-to-peer stores over NVLink (CUDA IPC), fused into the transform kernels"
                                                     if fft_p2p else "NCCL all-to-all")
                roof["poisson"].update({"kernels": "k_fft_x_fwd, k_fft_z, k_fft_thomas, k_fft_z<inv>, k_fft_x_inv",
                                        "achieved": gbs(3 * cells_b, pms), "frac": frac(gbs(3 * cells_b, pms)),
                                        "note": "direct solve: 2-D FFT over x/z (shared-memory Stockham) + Thomas in y; "
                                                "100 B per cell per solve (DESIGN 3.5)"})
        if args.implicit_y:
            roof["implicit_y"] = {"ms_per_step": T["phases"][2] / args.steps, "launches_per_step": 9,
                                  "note": "Thomas solves per y-line of u, v, w after each of the 3 RK stages "
                                          "(src/solver_time_kernels_implicit.cpp); ~64 B per cell and component"}
        cpu = None
        if not args.no_cpu_baseline and os.path.exists(HARNESS):
            v, ms, sample = cpu_reference(args.workload, args.closure, 3, 1)
            cpu = {"value": v, "unit": "Mcell-steps/s", "cores": 1, "kind": "reference", "sample": sample}
        desc = (f"3D Taylor-Green vortex {Nx}x{Ny}x{Nzg}" if wl["kind"] == "tgv" else
                f"3D channel {Nx}x{Ny}x{Nzg} (periodic x/z, walls y" + (", tanh-stretched y beta=2" if stretched else "") + ")")
        line = {"metric": "Mcell-steps/s (3D RK3+MG step)", "value": value, "unit": "Mcell-steps/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": T["step_ms"] / args.steps,
                "higher_is_better": True, "scaling": scaling_of(args.workload), "vs_baseline": None, "dtype": "f64",
                "data": "synthetic",
                "config": {"workload": f"{args.workload}: {desc}, SSP-RK3, skew-symmetric convection, nu={nu_:.6f} "
                                       f"dt={dt_:g}, " +
                                       ("FFT Poisson solver (direct: FFT in x/z + tridiagonal systems in y)" if args.poisson == "fft" else
                                        "multigrid Poisson " + (f"{fixed} fixed V-cycles" if fixed else
                                                                f"to tol_rhs={MG_TOL_RHS} (max {MG_MAXV} V-cycles)")) +
                                       (f", NN-TBNN closure ({PRESET}) on every cell every step (FP64 DMMA)" if nn else
                                        f", static LES SGS closure ({args.closure})" if args.closure in K.LES_MODEL else ", no closure") +
                                       (", adaptive dt (compute_adaptive_dt every step)" if args.adaptive_dt else "") +
                                       (", implicit y-diffusion" if args.implicit_y else ""),
                           "closure": args.closure,
                           "l2": "inputs larger than L2" if cells_local >= 256 ** 3 else "small grid: fits L2",
                           "vcycles_per_solve": cycles[-3:], "decomposition": f"z-slabs x{world}",
                           "planes_per_rank": Nzl},
                "wall_ms_per_step": T["wall_ms"] / args.steps, "gpu_launches": int(T["launches"]), "clocks": sampler.result(),
                "e2e": e2e, "roofline": roof, "cpu_baseline": cpu, "no_closure": second, "parity": parity}
        if args.adaptive_dt:
            line["config"]["dt_last"] = dts[-1] if dts else None
        if world > 1:
            # NVLink-halo roofline term of SURVEY 8(d): 2 (Nx+2)(Ny+2) 8 B per exchanged field per exchange per GPU
            plane_b = 2.0 * (Nx + 2) * (Ny + 2) * 8
            line["roofline"]["halo"] = {"bytes_per_exchange_per_field": plane_b, "link_GBs": 770.0,
                                        "us_per_exchange_at_link_rate": plane_b / 770e9 * 1e6,
                                        "note": "measured peer-copy bandwidth 770 GB/s per direction (B200_PROFILING.md); "
                                                "one exchange moves 4 MB at 1024x512: ~5 us of wire time vs ~25 us of NCCL "
                                                "point-to-point latency, i.e. the halo term is latency-, not bandwidth-bound"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def parity_check(B, K, np, torch, dist, rank, world, comm_init):
    """Evidence inside the bench line: a 128x64x(64 P) periodic box (level 0 on the 64x26 tiles, three levels
    distributed, fused residual+restriction on slabs) is stepped twice (RK3, MG to 1e-8) on the P ranks and
    compared with the single-domain run of the same box on rank 0's GPU.  P = 1: the box is run twice and must
    be bitwise identical."""
    Nx, Ny, Nzl = 128, 64, 64
    Nzg = Nzl * world
    L = 2 * np.pi
    env_keep = {k: os.environ.get(k) for k in ("CFDNN_MG_REPLICATE_CELLS", "CFDNN_MG_RR_MIN")}
    os.environ["CFDNN_MG_REPLICATE_CELLS"] = "10000"
    os.environ["CFDNN_MG_RR_MIN"] = "8"

    def run(nz_local, koff, nzg, multi):
        mesh = B.mesh(Nx, Ny, nz_local, Lx=L, Ly=L, Lz=L * nzg / Nzl, Nz_global=nzg, k_offset=koff)
        cfg = B.config(nu=0.01, dt=0.005, time_integrator=K.TI["rk3"], convective_scheme=K.SCHEME["skew"], adaptive_dt=0,
                       poisson_fixed_cycles=0, poisson_tol_rhs=MG_TOL_RHS, poisson_max_vcycles=MG_MAXV)
        h = B.solver(mesh, cfg, "PPPPPP")
        if multi:
            comm_init(h)
        d = L / Nx
        zc = (np.arange(nz_local + 2) - 0.5 + koff) * (L / Nzl)
        yc = (np.arange(Ny + 2) - 0.5) * (L / Ny); yf = (np.arange(Ny + 3) - 1) * (L / Ny)
        xc = (np.arange(Nx + 2) - 0.5) * d; xf = (np.arange(Nx + 3) - 1) * d
        sh = K.shapes(Nx, Ny, nz_local)
        u = np.zeros(sh["u"]); v = np.zeros(sh["v"]); w = np.zeros(sh["w"])
        cz = np.cos(zc)[1:-1, None, None]
        u[1:-1, 1:-1, 1:Nx + 2] = cz * np.cos(yc[1:-1])[None, :, None] * np.sin(xf[1:Nx + 2])[None, None, :]
        v[1:-1, 1:Ny + 2, 1:-1] = -cz * np.sin(yf[1:Ny + 2])[None, :, None] * np.cos(xc[1:-1])[None, None, :]
        B.upload(h, "u", u); B.upload(h, "v", v); B.upload(h, "w", w)
        B.check(B.lib.cfdnn_solver_initialize(h))
        B.run(h, 2)
        st = B.state(h, Nx, Ny, nz_local, names=("u", "v", "w", "p"))
        cyc = B.cycles(h)
        B.lib.cfdnn_solver_destroy(h)
        return st, cyc

    try:
        if world == 1:
            a, ca = run(Nzl, 0, Nzl, False)
            b, cb = run(Nzl, 0, Nzl, False)
            ok = all(np.array_equal(a[n], b[n]) for n in a) and ca == cb
            return {"ok": bool(ok), "kind": "1 GPU: 128x64x64 box run twice, bitwise comparison of u, v, w, p",
                    "vcycles": ca, "checksum_p": float(np.abs(a["p"]).sum()), "checksum_u": float(np.abs(a["u"]).sum())}
        mine, cyc = run(Nzl, rank * Nzl, Nzg, True)
        full = {}
        shapes = K.shapes(Nx, Ny, Nzg)
        if rank == 0:
            ref, cref = run(Nzg, 0, Nzg, False)
        worst = 0.0
        for n, key in (("u", "u"), ("v", "v"), ("w", "w"), ("p", "c")):
            t = torch.empty(shapes[key], dtype=torch.float64, device="cuda")
            if rank == 0:
                t.copy_(torch.from_numpy(ref[n]))
            dist.broadcast(t, 0)
            full = t.cpu().numpy()
            k0 = rank * Nzl
            sl = full[k0 + 1:k0 + 1 + Nzl]                      # this rank's interior planes of the global array
            got = mine[n][1:1 + Nzl]
            scale = max(float(np.abs(full).max()), 1e-300)
            worst = max(worst, float(np.abs(got - sl).max()) / scale)
        cyc_t = torch.tensor(cyc[:6] + [0] * (6 - len(cyc[:6])), dtype=torch.int64, device="cuda")
        dist.broadcast(cyc_t, 0)          # rank 0's slab-run cycle counts (identical on all ranks by construction)
        w_all = torch.tensor([worst], dtype=torch.float64, device="cuda")
        dist.all_reduce(w_all, op=dist.ReduceOp.MAX)
        if rank == 0:
            return {"ok": bool(w_all.item() <= 1e-12 and cyc == cref), "max_rel_diff_vs_single_domain": float(w_all.item()),
                    "gate": 1e-12, "vcycles_slab": cyc, "vcycles_single": cref,
                    "kind": f"{Nx}x{Ny}x{Nzg} periodic box, 2 RK3 steps, {world} z-slabs (3 distributed MG levels, fused "
                            "residual+restriction on slabs) vs the single-domain run on rank 0"}
        return None
    finally:
        for k, v in env_keep.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


if __name__ == "__main__":
    sys.exit(main())
