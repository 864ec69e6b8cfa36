#!/usr/bin/env python
"""bench.py — Mcell-steps/s of the fractional-step hot path (RK3 + multigrid Poisson step of a
3-D Taylor-Green vortex) on N B200s; one JSON line on stdout (rank 0).

  python bench.py --gpus 1 --steps K --warmup W                 # this backend (CUDA, C ABI)
  python bench.py --impl reference --steps K --warmup W          # reference CPU build (oracle/_ref)

value      : device-resident throughput (fields live in HBM; CUDA events around K steps)
e2e        : same steps driven through the C ABI with HOST buffers — every step uploads u,v,w,p
             from pinned host memory and downloads u,v,w,p
roofline   : multigrid V-cycle, algorithmic bytes (~500 B per fine cell per V-cycle)
             over its CUDA-event time, vs MEASURED_PEAKS.json hbm_gbs
cpu_baseline: the unmodified reference CPU build (single-threaded by construction) on a bounded
             sample (64^3 grid, a few steps) of the same workload, timed on this box's host cores
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {"tgv64": 64, "tgv128": 128, "tgv256": 256, "tgv384": 384, "tgv512": 512}
VCYCLE_BYTES_PER_CELL = 500.0      # 438 B at level 0, x 8/7 for the coarse levels
MG_TOL_RHS, MG_MAXV = 1e-8, 50     # north_star: "MG Poisson to 1e-8 relative residual"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


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


def run_reference(args):
    import refharness as R
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    N = 64
    meta = R.run_ref("bench", Nx=N, Ny=N, Nz=N, bc="PPPPPP", ti="rk3", scheme="skew", fixed=0, tol_rhs=MG_TOL_RHS,
                     maxv=MG_MAXV, nsteps=args.steps, warmup=args.warmup, dt=0.01, nu=0.01, init="tgv")
    v = meta["mcell_steps_per_s"]
    sample = f"{N}^3 TGV sub-grid of the workload, {args.steps} steps after {args.warmup} warm-up, reference CPU build"
    line = {"impl": "reference", "metric": "Mcell-steps/s (3D RK3+MG step)", "value": v, "unit": "Mcell-steps/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": meta["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": args.workload, "sample": sample},
            "cpu_baseline": {"value": v, "unit": "Mcell-steps/s", "cores": 1, "kind": "reference", "sample": sample},
            "e2e": {"value": v, "unit": "Mcell-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="tgv512", choices=sorted(WORKLOADS))
    ap.add_argument("--closure", default="none", choices=["none", "tbnn"],
                    help="tbnn: attach the NN-TBNN closure (tbnn_channel_caseholdout, per-cell 5-64-64-64-4 MLP on the "
                         "FP64 tensor cores) to the step — the north_star's 'RK3 + MG + NN-TBNN' variant of the workload")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import refharness as R

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B = R.Backend()
    B.check(B.lib.cfdnn_set_device(local))

    N = WORKLOADS[args.workload]
    Nzg = N * world                      # weak scaling: every GPU owns an N x N x N slab
    L = 2 * np.pi
    mesh = B.mesh(N, N, N, Lx=L, Ly=L, Lz=L * world, Nz_global=Nzg, k_offset=rank * N)
    # with the closure nu_t reaches 10 nu: the explicit diffusive limit dx^2 / (6 nu_eff) forces a smaller step
    # dt keeps the CFL number of BASELINE config 2 (64^3, dt = 0.01): dt = 0.01 * 64 / N.  (dt = 0.01 at
    # 512^3 is CFL ~ 0.8 and the run diverges within 10 steps.)
    nu_, dt_ = (0.01, 0.01 * 64.0 / N) if args.closure == "none" else (1.0 / 180.0, 1e-4)
    cfg = B.config(nu=nu_, dt=dt_, time_integrator=R.TI["rk3"], convective_scheme=R.SCHEME["skew"],
                   adaptive_dt=0, poisson_fixed_cycles=0, poisson_tol_rhs=MG_TOL_RHS, poisson_max_vcycles=MG_MAXV)
    h = B.solver(mesh, cfg, "PPPPPP")
    nn = None
    if args.closure == "tbnn":
        import nnmodels as M
        nn = M.create_preset(B, "tbnn_channel_caseholdout")
        B.check(B.lib.cfdnn_nn_set_params(nn, nu_, 1.0, 1.0, 1e-10, 1.0))
        B.check(B.lib.cfdnn_solver_set_turbulence_model(h, nn))
    if world > 1:
        idbuf = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            raw = (C.c_char * 128)()
            B.check(B.lib.cfdnn_comm_unique_id(raw))
            idbuf = torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8).clone()
        idbuf = idbuf.cuda()
        dist.broadcast(idbuf, 0)
        raw = (C.c_char * 128).from_buffer_copy(bytes(idbuf.cpu().numpy().tobytes()))
        B.check(B.lib.cfdnn_solver_comm_init(h, rank, world, raw))

    # synthetic input: Taylor-Green vortex of the global box, this rank's slab (pinned host arrays)
    sh = R.shapes(N, N, N)
    host = {n: torch.zeros(sh[k], dtype=torch.float64).pin_memory() for n, k in
            (("u", "u"), ("v", "v"), ("w", "w"), ("p", "c"))}
    d = L / N
    zc = (np.arange(N + 2) - 1 + 0.5 + rank * N) * d
    yc = (np.arange(N + 2) - 1 + 0.5) * d
    xc = (np.arange(N + 2) - 1 + 0.5) * d
    xf = (np.arange(N + 3) - 1) * d
    yf = (np.arange(N + 3) - 1) * d
    cz = np.cos(zc)[1:-1, None, None]
    host["u"].numpy()[1:-1, 1:-1, 1:N + 2] = cz * np.cos(yc[1:-1])[None, :, None] * np.sin(xf[1:N + 2])[None, None, :]
    host["v"].numpy()[1:-1, 1:N + 2, 1:-1] = -cz * np.sin(yf[1:N + 2])[None, :, None] * np.cos(xc[1:-1])[None, None, :]
    fid = R.FIELD_ID

    def upload_all():
        for n, t in host.items():
            B.check(B.lib.cfdnn_solver_upload(h, fid[n], C.cast(t.data_ptr(), R.capi.c_double_p), t.numel()))

    def download_all():
        for n, t in host.items():
            B.check(B.lib.cfdnn_solver_download(h, fid[n], C.cast(t.data_ptr(), R.capi.c_double_p), t.numel()))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timing():
        a, b, c = C.c_double(), C.c_double(), C.c_int()
        B.lib.cfdnn_solver_last_timing(h, C.byref(a), C.byref(b), C.byref(c))
        return a.value, b.value, c.value

    upload_all()
    B.check(B.lib.cfdnn_solver_initialize(h))
    r = C.c_double()
    for _ in range(args.warmup):
        B.check(B.lib.cfdnn_solver_run(h, 1, C.byref(r)))
    # ---- device-resident timed region: exactly K steps, CUDA events inside the library
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    l0 = B.lib.cfdnn_kernel_launches()
    t0 = time.perf_counter()
    B.check(B.lib.cfdnn_solver_run(h, args.steps, C.byref(r)))
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    launches = B.lib.cfdnn_kernel_launches() - l0
    step_ms, vc_ms, n_vc = timing()
    sampler.stop_flag = True
    sampler.join(timeout=2)
    cycles = B.cycles(h)
    t = torch.tensor([step_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    step_ms = float(t.item())
    cells = N * N * Nzg
    value = cells * args.steps / (step_ms * 1e-3) / 1e6

    # ---- end to end through the C ABI with host buffers
    e2e = None
    if not args.no_e2e:
        nbytes = sum(t_.numel() * 8 for t_ in host.values())
        download_all()
        for _ in range(1):
            upload_all(); B.check(B.lib.cfdnn_solver_run(h, 1, C.byref(r))); download_all()
        barrier()
        t0 = time.perf_counter()
        ph = [0.0, 0.0, 0.0]
        for _ in range(args.steps):
            ta = time.perf_counter()
            upload_all()
            tb = time.perf_counter()
            B.check(B.lib.cfdnn_solver_run(h, 1, C.byref(r)))
            tc = time.perf_counter()
            download_all()
            td = time.perf_counter()
            ph[0] += tb - ta; ph[1] += tc - tb; ph[2] += td - tc
        barrier()
        e2e_s = time.perf_counter() - t0
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e = {"value": cells * args.steps / float(t.item()) / 1e6, "unit": "Mcell-steps/s",
               "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": nbytes,
               "ms_per_step": {"h2d": ph[0] / args.steps * 1e3, "step": ph[1] / args.steps * 1e3,
                               "d2h": ph[2] / args.steps * 1e3}}

    # dominant kernel, timed live with CUDA events on the solver's stream (after the timed steps, same
    # resident 512^3 level-0 arrays, >> L2): the level-0 temporally blocked smoother
    sm_ms = C.c_double(0.0)
    B.check(B.lib.cfdnn_solver_time_smoother(h, 20, C.byref(sm_ms)))

    if rank == 0:
        peak, which = peaks()
        vc_avg_ms = vc_ms / max(n_vc, 1)
        cells_local = N * N * N
        vc_achieved = VCYCLE_BYTES_PER_CELL * cells_local / (vc_avg_ms * 1e-3) / 1e9 if n_vc else None
        # 24 B per cell per sweep (u, f in; u out); one launch = 2 sweeps.  With the two
        # sweeps fused the compulsory traffic of a launch is 24 B per cell (what ncu's DRAM bytes show).
        alg = 2 * 24.0 * cells_local
        achieved = alg / (sm_ms.value * 1e-3) / 1e9 if sm_ms.value > 0 else None
        traffic = None
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["k_sweep_tbo<64,26>"]
            if tj["cells"] == cells_local:
                traffic = tj["dram_bytes_per_launch"]
        except Exception:
            pass
        roof = {"bound": "hbm", "kernel": "k_sweep_tbo<64,26>: level-0 smoother, 2 Chebyshev sweeps per launch",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak if achieved else None,
                "traffic": traffic, "peak_source": which, "algorithmic_bytes_per_launch": alg,
                "fused_compulsory_bytes_per_launch": 24.0 * cells_local, "launch_ms": sm_ms.value, "launches_timed": 20,
                "dram_GBs_at_fused_traffic": (24.0 * cells_local / (sm_ms.value * 1e-3) / 1e9) if sm_ms.value > 0 else None,
                "vcycle": {"algorithmic_bytes": VCYCLE_BYTES_PER_CELL * cells_local, "ms": vc_avg_ms, "timed": n_vc,
                           "achieved": vc_achieved, "frac": vc_achieved / peak if vc_achieved else None,
                           "share_of_step": vc_ms / step_ms}}
        cpu = None
        if not args.no_cpu_baseline and R.have_oracle():
            ns = 4
            m = R.run_ref("bench", Nx=64, Ny=64, Nz=64, bc="PPPPPP", ti="rk3", scheme="skew", fixed=0,
                          tol_rhs=MG_TOL_RHS, maxv=MG_MAXV, nsteps=ns, warmup=1, dt=0.01, nu=0.01, init="tgv")
            cpu = {"value": m["mcell_steps_per_s"], "unit": "Mcell-steps/s", "cores": 1, "kind": "reference",
                   "sample": f"64^3 TGV sub-grid of the workload, {ns} steps, unmodified reference CPU build "
                             "(single-threaded by construction)"}
        line = {"metric": "Mcell-steps/s (3D RK3+MG step)", "value": value, "unit": "Mcell-steps/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_ms / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic",
                "config": {"workload": f"{args.workload}: 3D Taylor-Green vortex {N}x{N}x{Nzg}, SSP-RK3, skew-symmetric "
                                       f"convection, nu={nu_:.5f} dt={dt_:g}, multigrid Poisson to tol_rhs={MG_TOL_RHS} "
                                       f"(max {MG_MAXV} V-cycles)"
                                       + (", NN-TBNN closure on every cell (FP64 DMMA)" if nn else ""),
                           "l2": "inputs larger than L2" if N >= 256 else "small grid: fits L2",
                           "vcycles_per_solve": cycles[-3:], "decomposition": f"z-slabs x{world}"},
                "wall_ms_per_step": wall_ms / args.steps, "gpu_launches": int(launches), "clocks": sampler.result(),
                "e2e": e2e, "roofline": roof, "cpu_baseline": cpu}
        print(json.dumps(line), flush=True)
    B.lib.cfdnn_solver_destroy(h)
    if nn:
        B.lib.cfdnn_nn_destroy(nn)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
