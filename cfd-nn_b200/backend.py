"""Host-side Python mirror of the reference's RANSSolver / MultigridPoissonSolver call sequence over the
C ABI of include/cfdnn_b200.h (what include/cfdnn_b200_shim.hpp does in C++).  Used by bench.py,
__graft_entry__.smoke() and the tests.  Product code: it never touches oracle/."""
import ctypes as C

import numpy as np

from . import _capi as capi

VBC = {"W": 0, "P": 1, "I": 2, "O": 3}
PBC = {"D": 0, "N": 1, "P": 2}
TI = {"euler": 0, "rk2": 1, "rk3": 2}
SCHEME = {"central": 0, "skew": 2}



def shapes(Nx, Ny, Nz, Ng=1):
    return {
        "u": (Nz + 2 * Ng, Ny + 2 * Ng, Nx + 1 + 2 * Ng),
        "v": (Nz + 2 * Ng, Ny + 1 + 2 * Ng, Nx + 2 * Ng),
        "w": (Nz + 1 + 2 * Ng, Ny + 2 * Ng, Nx + 2 * Ng),
        "c": (Nz + 2 * Ng, Ny + 2 * Ng, Nx + 2 * Ng),
    }



def interior(name, a, Ng=1):
    """Interior faces/cells of a ghosted array, the ranges the reference's loops own."""
    s = slice(Ng, -Ng)
    if name == "u":
        return a[s, s, Ng:a.shape[2] - Ng]
    if name == "v":
        return a[s, Ng:a.shape[1] - Ng, s]
    if name == "w":
        return a[Ng:a.shape[0] - Ng, s, s]
    return a[s, s, s]


def rel_l2(a, b):
    d = np.linalg.norm((a - b).ravel())
    n = np.linalg.norm(b.ravel())
    return d / n if n > 0 else d


FIELD_ID = {"u": 0, "v": 1, "w": 2, "p": 3, "pcorr": 4, "div": 5, "rhs": 6, "nut": 7, "k": 8, "omega": 9,
            "nueff": 10, "tau_xx": 11, "tau_xy": 12, "tau_yy": 13, "u_star": 14, "v_star": 15, "w_star": 16}



class Backend:
    """Thin Python mirror of the reference's RANSSolver / MultigridPoissonSolver call sequence
    over the C ABI (what INTEGRATION.md's C++ shim does)."""

    def __init__(self, lib_path=None):
        self.lib = capi.load(lib_path)

    def check(self, rc):
        capi.check(self.lib, rc)

    def mesh(self, Nx, Ny, Nz, Lx=2 * np.pi, Ly=2 * np.pi, Lz=2 * np.pi, Ng=1, Nz_global=0, k_offset=0):
        return capi.MeshDesc(Nx=Nx, Ny=Ny, Nz=Nz, Ng=Ng, dx=Lx / Nx, dy=Ly / Ny, dz=Lz / max(Nz_global or Nz, 1),
                             y_stretched=0, Nz_global=Nz_global or Nz, k_offset=k_offset)

    def stretched_mesh(self, Nx, Ny, Nz, pre, Lx=2 * np.pi, Lz=2 * np.pi, y_lo=-1.0, y_hi=1.0):
        """MeshDesc of a tanh-stretched-y mesh from the metric arrays the oracle dumped
        (`<pre>_dyv.f64` ... written by ref_harness dump_mesh = the reference's Mesh::init_stretched_y)."""
        arrs = pre if isinstance(pre, dict) else _load_mesh_arrays(pre)
        md = capi.MeshDesc(Nx=Nx, Ny=Ny, Nz=Nz, Ng=1, dx=Lx / Nx, dy=(y_hi - y_lo) / Ny, dz=Lz / max(Nz, 1),
                           y_stretched=1, Nz_global=Nz, k_offset=0, y_lo=y_lo, y_hi=y_hi)
        md.dyv, md.dyc = capi.dptr(arrs["dyv"]), capi.dptr(arrs["dyc"])
        md.yLap_aS, md.yLap_aN, md.yLap_aP = capi.dptr(arrs["aS"]), capi.dptr(arrs["aN"]), capi.dptr(arrs["aP"])
        md.yc = capi.dptr(arrs["yc"])
        if "yf" in arrs:
            md.yf = capi.dptr(arrs["yf"])
        md._keep = arrs                      # keep the arrays alive as long as the descriptor
        return md

    def config(self, **kw):
        cfg = capi.Config()
        self.lib.cfdnn_config_default(C.byref(cfg))
        for k, v in kw.items():
            if not hasattr(cfg, k):
                raise KeyError(k)
            setattr(cfg, k, v)
        return cfg

    def solver(self, mesh, cfg, bc="PPPPPP", force=(0.0, 0.0, 0.0)):
        h = C.c_void_p()
        self.check(self.lib.cfdnn_solver_create(C.byref(mesh), C.byref(cfg), C.byref(h)))
        arr = (C.c_int * 6)(*[VBC[c] for c in bc])
        self.check(self.lib.cfdnn_solver_set_velocity_bc(h, arr))
        self.check(self.lib.cfdnn_solver_set_body_force(h, *force))
        return h

    def upload(self, h, name, a):
        a = np.ascontiguousarray(a, dtype=np.float64)
        self.check(self.lib.cfdnn_solver_upload(h, FIELD_ID[name], capi.dptr(a), a.size))

    def download(self, h, name, shape):
        a = np.empty(shape, dtype=np.float64)
        self.check(self.lib.cfdnn_solver_download(h, FIELD_ID[name], capi.dptr(a), a.size))
        return a

    def step(self, h):
        r = C.c_double()
        self.check(self.lib.cfdnn_solver_step(h, C.byref(r)))
        return r.value

    def run(self, h, n):
        r = C.c_double()
        self.check(self.lib.cfdnn_solver_run(h, n, C.byref(r)))
        return r.value

    def cycles(self, h):
        buf = (C.c_int * 4096)()
        n = C.c_int()
        self.lib.cfdnn_solver_cycle_log(h, buf, 4096, C.byref(n))
        return list(buf[:min(n.value, 4096)])

    def state(self, h, Nx, Ny, Nz, names=("u", "v", "w", "p", "pcorr", "div", "rhs", "nut")):
        sh = shapes(Nx, Ny, Nz)
        key = {"u": "u", "v": "v", "w": "w"}
        return {n: self.download(h, n, sh[key.get(n, "c")]) for n in names}



def _load_mesh_arrays(pre):
    """y-metric arrays of a stretched mesh (dyv, dyc, aS, aN, aP, yc, yf) from `<pre>_<name>.f64` files"""
    import os
    return {k: np.ascontiguousarray(np.fromfile(f"{pre}_{k}.f64")) for k in ("dyv", "dyc", "aS", "aN", "aP", "yc", "yf")
            if os.path.exists(f"{pre}_{k}.f64")}


def tgv_init(Nx, Ny, Nz, L=2 * np.pi, V0=1.0, Ng=1):
    """Taylor-Green vortex on faces (reference app/main_taylor_green_3d.cpp:253-276), interior only."""
    sh = shapes(Nx, Ny, Nz, Ng)
    dx, dy, dz = L / Nx, L / Ny, L / Nz
    xf = (np.arange(Nx + 1 + 2 * Ng) - Ng) * dx
    xc = (np.arange(Nx + 2 * Ng) - Ng + 0.5) * dx
    yf = (np.arange(Ny + 1 + 2 * Ng) - Ng) * dy
    yc = (np.arange(Ny + 2 * Ng) - Ng + 0.5) * dy
    zc = (np.arange(Nz + 2 * Ng) - Ng + 0.5) * dz
    u = np.zeros(sh["u"]); v = np.zeros(sh["v"]); w = np.zeros(sh["w"])
    s = slice(Ng, -Ng)
    u[s, s, Ng:Ng + Nx + 1] = V0 * (np.cos(zc[s])[:, None, None] * np.cos(yc[s])[None, :, None] *
                                    np.sin(xf[Ng:Ng + Nx + 1])[None, None, :])
    v[s, Ng:Ng + Ny + 1, s] = -V0 * (np.cos(zc[s])[:, None, None] * np.sin(yf[Ng:Ng + Ny + 1])[None, :, None] *
                                     np.cos(xc[s])[None, None, :])
    return u, v, w


# ---------------------------------------------------------------- y-stretched meshes
def stretched_y_metrics(Ny, y_lo, y_hi, beta, Ng=1):
    """The y-metric arrays of a tanh-stretched mesh as Mesh::init_stretched_y builds them
    (reference src/mesh.cpp:119-238, Mesh::tanh_stretching :243-249): faces yf (ghost faces extrapolated
    with the first / last interior spacing), centres yc, cell heights dyv, centre spacings dyc and the
    y-Laplacian coefficients aS / aN / aP.  Returned in the dict layout Backend.stretched_mesh takes."""
    import math
    tot = Ny + 2 * Ng
    Ly = y_hi - y_lo
    tb = math.tanh(beta)

    def stretch(eta):
        return 0.5 * (1.0 + math.tanh(beta * (2.0 * eta - 1.0)) / tb)

    yf = np.empty(tot + 1)
    for j in range(tot + 1):
        eta = float(j - Ng) / Ny
        if eta < 0:
            y0 = y_lo + stretch(0.0) * Ly
            y1 = y_lo + stretch(1.0 / Ny) * Ly
            yf[j] = y0 + eta * Ny * (y1 - y0)
        elif eta > 1:
            y0 = y_lo + stretch(1.0 - 1.0 / Ny) * Ly
            y1 = y_lo + stretch(1.0) * Ly
            yf[j] = y1 + (eta - 1.0) * Ny * (y1 - y0)
        else:
            yf[j] = y_lo + stretch(eta) * Ly
    yc = 0.5 * (yf[:-1] + yf[1:])
    dyv = yf[1:] - yf[:-1]
    dyc = np.empty(tot + 1)
    for j in range(1, tot + 1):
        dyc[j] = yc[j] - yc[j - 1] if j < tot else yf[tot] - yc[tot - 1]
    dyc[0] = dyc[1]
    aS = np.empty(tot); aN = np.empty(tot)
    for j in range(tot):
        south = yc[j] - yc[j - 1] if j > 0 else dyv[j]
        north = yc[j + 1] - yc[j] if j < tot - 1 else dyv[j]
        aS[j] = 1.0 / (dyv[j] * south)
        aN[j] = 1.0 / (dyv[j] * north)
    aP = -(aS + aN)
    return {k: np.ascontiguousarray(v) for k, v in
            dict(dyv=dyv, dyc=dyc, aS=aS, aN=aN, aP=aP, yc=yc, yf=yf).items()}


# ---------------------------------------------------------------- NN closures (cfdnn_nn handles)
NN_KIND = {"mlp": 1, "tbnn": 2}
LES_MODEL = {"smagorinsky": 1, "wale": 2, "vreman": 3, "sigma": 4}


def load_npz_preset(path, name):
    """(W, b, means, stds) of model `name` from an .npz holding `<name>/n_layers`, `<name>/W<l>`, `<name>/b<l>`,
    `<name>/means`, `<name>/stds` (row-major (out, in) weights, the reference's layerN_W.txt convention)."""
    z = np.load(path)
    n = int(z[f"{name}/n_layers"])
    W = [np.ascontiguousarray(z[f"{name}/W{l}"], dtype=np.float64) for l in range(n)]
    b = [np.ascontiguousarray(z[f"{name}/b{l}"], dtype=np.float64) for l in range(n)]
    return W, b, np.ascontiguousarray(z[f"{name}/means"]), np.ascontiguousarray(z[f"{name}/stds"])


def nn_create(B, kind, W, b, means=None, stds=None, act=None):
    """cfdnn_nn_create from numpy arrays; kind = "tbnn" | "mlp" """
    n = len(W)
    dims = [W[0].shape[1]] + [w.shape[0] for w in W]
    dims_c = (C.c_int * (n + 1))(*dims)
    act_c = (C.c_int * n)(*act) if act is not None else None
    Wp = (capi.c_double_p * n)(*[capi.dptr(w) for w in W])
    bp = (capi.c_double_p * n)(*[capi.dptr(x) for x in b])
    h = C.c_void_p()
    B.check(B.lib.cfdnn_nn_create(NN_KIND[kind], n, dims_c, act_c, Wp, bp, capi.dptr(means) if means is not None else None,
                                  capi.dptr(stds) if stds is not None else None, C.byref(h)))
    return h


def trip_tables(Nx, Ny, Nz, x_lo, Lx, yf, z_lo, Lz, fx, y_lo, y_hi, amplitude=3.0, n_modes=8, x_start=-1.0, x_end=-1.0,
                Ng=1, k_offset=0, Nz_global=None):
    """The tables of the reference's trip-region forcing, RANSSolver::initialize_trip_forcing (src/solver.cpp:818-923),
    for cfdnn_solver_set_trip_forcing: x envelope on cell centres, g(y) = y (1 - y^2) on the y faces `yf` (Ny+2Ng+1
    values), spanwise mode sums on cell centres / z faces with the phases of std::srand(12345); std::rand() (the C
    library's generator, called through ctypes so the sequence is the reference's).  Returns (amplitude * u_tau^2, dict)."""
    import ctypes
    import ctypes.util
    import math
    libc = ctypes.CDLL(ctypes.util.find_library("c"))
    libc.srand(12345)
    RAND_MAX = 2147483647
    phases = [2.0 * math.pi * float(libc.rand()) / RAND_MAX for _ in range(n_modes)]
    delta = 0.5 * (y_hi - y_lo)
    u_tau = math.sqrt(delta * abs(fx))
    x0 = x_lo + 0.1 * Lx if x_start < 0 else x_start
    x1 = x_lo + 0.2 * Lx if x_end < 0 else x_end
    width = x1 - x0
    dx = Lx / Nx
    env = np.zeros(Nx + 2 * Ng)
    for i in range(Nx + 2 * Ng):
        x = x_lo + (i - Ng + 0.5) * dx
        if x0 <= x <= x1:
            xi = (x - x0) / width
            env[i] = 0.5 * (1.0 - math.cos(2.0 * math.pi * xi))
    yf = np.asarray(yf, dtype=np.float64)
    g_y = yf * (1.0 - yf * yf)
    out = {"env_x": env, "g_y": np.ascontiguousarray(g_y)}
    if Nz > 1:
        NzG = Nz_global or Nz
        dz = Lz / NzG
        wsum = 0.0
        for m in range(n_modes):
            wsum += 1.0 / (m + 1)
        fzv = np.zeros(Nz + 2 * Ng)
        fzw = np.zeros(Nz + 2 * Ng + 1)
        for k in range(Nz + 2 * Ng):
            z = z_lo + (k + k_offset - Ng + 0.5) * dz
            s = 0.0
            for m in range(n_modes):
                kz = 2.0 * math.pi * (m + 1) / Lz
                s += ((1.0 / (m + 1)) / wsum) * math.sin(kz * z + phases[m])
            fzv[k] = s
        for k in range(Nz + 2 * Ng + 1):
            z = z_lo + (k + k_offset - Ng) * dz
            s = 0.0
            for m in range(n_modes):
                kz = 2.0 * math.pi * (m + 1) / Lz
                s += ((1.0 / (m + 1)) / wsum) * math.cos(kz * z + phases[m])
            fzw[k] = s
        out["Fz_v"], out["Fz_w"] = fzv, fzw
    return amplitude * u_tau * u_tau, out


def set_trip(B, h, A, tabs, duration=2.0, ramp_off_start=1.5, w_scale=1.0, force_w=True):
    d = capi.TripDesc(amplitude=A, duration=duration, ramp_off_start=ramp_off_start, w_scale=w_scale)
    d.env_x, d.g_y = capi.dptr(tabs["env_x"]), capi.dptr(tabs["g_y"])
    if "Fz_v" in tabs:
        d.Fz_v = capi.dptr(tabs["Fz_v"])
        if force_w:
            d.Fz_w = capi.dptr(tabs["Fz_w"])
    B.check(B.lib.cfdnn_solver_set_trip_forcing(h, C.byref(d)))


def velocity_filter_reference(u, v, w, strength):
    """apply_velocity_filter (src/solver.cpp:4065-4162) with every second difference taken from the unfiltered field
    (ghost-inclusive 3-D arrays, Ng = 1); returns filtered copies."""
    alpha = strength * 0.25
    alpha_y = alpha * 0.5
    out = []
    for q, comp in ((u, 0), (v, 1), (w, 2)):
        r = q.copy()
        c = q[1:-1, 1:-1, 1:-1]
        Lx = q[1:-1, 1:-1, :-2] - 2.0 * c + q[1:-1, 1:-1, 2:]
        Lz = q[:-2, 1:-1, 1:-1] - 2.0 * c + q[2:, 1:-1, 1:-1]
        Ly = q[1:-1, :-2, 1:-1] - 2.0 * c + q[1:-1, 2:, 1:-1]
        Ly = Ly.copy()
        Ly[:, 0, :] = 0.0            # j = Ng: first cell row (u, w) / lower wall face (v)
        Ly[:, -1, :] = 0.0           # last cell row (u, w) / upper wall face (v)
        r[1:-1, 1:-1, 1:-1] = c + (alpha * (Lx + Lz) + alpha_y * Ly)
        out.append(r)
    return out
