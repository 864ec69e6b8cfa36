"""One rank of a z-slab run on the CPU test double (tests/devsim) with a torch.distributed `gloo`
transport behind the backend's host-callback communicator.  TEST INFRASTRUCTURE.
usage: slab_worker.py rank world port outdir Nx Ny NzGlobal bc integrator nsteps fixed_cycles [gpu]
With the trailing `gpu` argument the rank runs the PRODUCT library on cuda:<rank> over NCCL instead
(the gloo group then only carries the ncclUniqueId)."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import refharness as R  # noqa: E402


def global_init(Nx, Ny, Nz, bc):
    """Smooth field with all three components non-zero (not divergence free: the projection
    takes care of that); walls get zero normal velocity via initialize()."""
    u, v, w = R.tgv_init(Nx, Ny, Nz)
    L = 2 * np.pi
    sh = R.shapes(Nx, Ny, Nz)
    zf = (np.arange(sh["w"][0]) - 1) * (L / Nz)
    yc = (np.arange(sh["w"][1]) - 1 + 0.5) * (L / Ny)
    xc = (np.arange(sh["w"][2]) - 1 + 0.5) * (L / Nx)
    w[1:Nz + 2, 1:-1, 1:-1] = 0.3 * (np.sin(zf[1:Nz + 2])[:, None, None] * np.cos(yc[1:-1])[None, :, None] *
                                     np.sin(2 * xc[1:-1])[None, None, :])
    zc = (np.arange(sh["u"][0]) - 1 + 0.5) * (L / Nz)
    u[1:-1, 1:-1, 1:Nx + 2] += 0.2 * np.sin(2 * zc[1:-1])[:, None, None]
    return u, v, w


def main():
    rank, world, port = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    outdir = sys.argv[4]
    Nx, Ny, NzG = int(sys.argv[5]), int(sys.argv[6]), int(sys.argv[7])
    bc, ti, nsteps, fixed = sys.argv[8], sys.argv[9], int(sys.argv[10]), int(sys.argv[11])
    on_gpu = len(sys.argv) > 12 and sys.argv[12] == "gpu"
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)

    def view(ptr, n):
        return torch.from_numpy(np.ctypeslib.as_array(ptr, shape=(n,)))

    SENDRECV = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.POINTER(C.c_double), C.c_size_t, C.POINTER(C.c_double),
                           C.c_size_t, C.c_int, C.POINTER(C.c_double), C.c_size_t, C.POINTER(C.c_double), C.c_size_t)
    ALLREDUCE = C.CFUNCTYPE(None, C.c_void_p, C.POINTER(C.c_double), C.c_int, C.c_int)
    ALLGATHER = C.CFUNCTYPE(None, C.c_void_p, C.POINTER(C.c_double), C.c_size_t)
    stats = {"sendrecv": 0, "allreduce": 0, "allgather": 0}

    def sendrecv(ctx, lo_peer, s_lo, ns_lo, r_lo, nr_lo, hi_peer, s_hi, ns_hi, r_hi, nr_hi):
        stats["sendrecv"] += 1
        reqs = []
        # tag 1: messages travelling upward (to the hi peer), tag 2: downward
        if hi_peer >= 0 and ns_hi:
            reqs.append(dist.isend(view(s_hi, ns_hi).clone(), hi_peer, tag=1))
        if lo_peer >= 0 and ns_lo:
            reqs.append(dist.isend(view(s_lo, ns_lo).clone(), lo_peer, tag=2))
        if lo_peer >= 0 and nr_lo:
            reqs.append(dist.irecv(view(r_lo, nr_lo), lo_peer, tag=1))
        if hi_peer >= 0 and nr_hi:
            reqs.append(dist.irecv(view(r_hi, nr_hi), hi_peer, tag=2))
        for q in reqs:
            q.wait()

    def allreduce(ctx, d, n, op):
        stats["allreduce"] += 1
        t = view(d, n)
        dist.all_reduce(t, op={0: dist.ReduceOp.SUM, 1: dist.ReduceOp.MAX, 2: dist.ReduceOp.MIN}[op])

    def allgather(ctx, base, count):
        stats["allgather"] += 1
        whole = view(base, count * world)
        mine = whole[rank * count:(rank + 1) * count].clone()
        parts = [torch.empty(count, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(parts, mine)
        for r_, p in enumerate(parts):
            whole[r_ * count:(r_ + 1) * count] = p

    class Callbacks(C.Structure):
        _fields_ = [("ctx", C.c_void_p), ("sendrecv", SENDRECV), ("allreduce", ALLREDUCE), ("allgather", ALLGATHER)]

    cbs = Callbacks(None, SENDRECV(sendrecv), ALLREDUCE(allreduce), ALLGATHER(allgather))
    B = R.Backend() if on_gpu else R.Backend(R.DEVSIM_LIB)
    if on_gpu:
        B.check(B.lib.cfdnn_set_device(rank))
    Nz = NzG // world
    k0 = rank * Nz
    L = 2 * np.pi
    if os.environ.get("SLAB_STRETCH"):          # tanh-stretched y (y in [-1, 1], beta = 2): FFT Poisson only on slabs
        arrs = R.backend.stretched_y_metrics(Ny, -1.0, 1.0, 2.0)
        mesh = B.stretched_mesh(Nx, Ny, Nz, arrs, Lx=L, Lz=L, y_lo=-1.0, y_hi=1.0)
        mesh.dz, mesh.Nz_global, mesh.k_offset = L / NzG, NzG, k0
    else:
        mesh = B.mesh(Nx, Ny, Nz, Lz=L, Nz_global=NzG, k_offset=k0)
    cfg = B.config(nu=0.01, dt=0.01, time_integrator=R.TI[ti], convective_scheme=R.SCHEME["skew"], adaptive_dt=0,
                   poisson_fixed_cycles=fixed, poisson_tol_rhs=1e-8, poisson_max_vcycles=50,
                   poisson_solver=1 if os.environ.get("SLAB_POISSON") == "fft" else 0)
    h = B.solver(mesh, cfg, bc, (0.5, 0.0, 0.0) if "W" in bc else (0.0, 0.0, 0.0))
    if on_gpu:
        raw = (C.c_char * 128)()
        if rank == 0:
            B.check(B.lib.cfdnn_comm_unique_id(raw))
        idt = torch.tensor(list(raw.raw), dtype=torch.uint8)
        dist.broadcast(idt, 0)
        raw = (C.c_char * 128).from_buffer_copy(bytes(idt.numpy().tobytes()))
        B.check(B.lib.cfdnn_solver_comm_init(h, rank, world, raw))
    else:
        B.lib.cfdnn_devsim_solver_comm_init.restype = C.c_int
        B.lib.cfdnn_devsim_solver_comm_init.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(Callbacks)]
        B.check(B.lib.cfdnn_devsim_solver_comm_init(h, rank, world, C.byref(cbs)))
    u, v, w = global_init(Nx, Ny, NzG, bc)
    B.upload(h, "u", u[k0:k0 + Nz + 2]); B.upload(h, "v", v[k0:k0 + Nz + 2]); B.upload(h, "w", w[k0:k0 + Nz + 3])
    B.check(B.lib.cfdnn_solver_initialize(h))
    res = [B.step(h) for _ in range(nsteps)]
    dts = []
    dt = C.c_double()
    B.check(B.lib.cfdnn_solver_compute_adaptive_dt(h, C.byref(dt)))
    fft_p2p = int(B.lib.cfdnn_solver_fft_peer_to_peer(h))
    st = B.state(h, Nx, Ny, Nz, names=("u", "v", "w", "p", "pcorr", "div", "rhs"))
    np.savez(os.path.join(outdir, f"slab{rank}.npz"), residuals=np.array(res), cycles=np.array(B.cycles(h)),
             adaptive_dt=dt.value, fft_p2p=fft_p2p, stats=np.array([stats["sendrecv"], stats["allreduce"], stats["allgather"]]), **st)
    B.lib.cfdnn_solver_destroy(h)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
