"""NN-closure parity cases shared by the devsim (not gpu) and gpu tests."""
import ctypes as C
import os

import numpy as np

import nnmodels as M
import refharness as R

NU = 1.0 / 180.0


def channel_fields(Nx, Ny, seed=1, amp=0.05, k_val=None, om_val=None, planes=1, beta=None, yc=None):
    """perturbed parabolic channel profile on a 2-D mesh (Nz=1 -> 3 planes, plane 0 is the 2-D plane)"""
    rng = np.random.default_rng(seed)
    sh = R.shapes(Nx, Ny, 1)
    if yc is None:
        yc = -1 + (np.arange(Ny + 2) - 0.5) * 2.0 / Ny
    u = np.zeros(sh["u"]); v = np.zeros(sh["v"])
    u[0] = (1 - yc ** 2)[:, None] + amp * rng.standard_normal(sh["u"][1:])
    v[0] = amp * rng.standard_normal(sh["v"][1:])
    k = om = None
    if k_val is not None:
        k = np.zeros(sh["c"]); om = np.zeros(sh["c"])
        k[0] = k_val * (1 + 0.3 * rng.random(sh["c"][1:]))
        om[0] = om_val * (1 + 0.3 * rng.random(sh["c"][1:]))
    return u, v, k, om


def ref_update(td, model_dir, kind, Nx, Ny, u, v, k, om, Lx=4.0, extra=None):
    pre = os.path.join(td, "nnin")
    u.tofile(pre + "_u.f64"); v.tofile(pre + "_v.f64")
    for n in ("_k.f64", "_omega.f64"):
        if os.path.exists(pre + n):
            os.remove(pre + n)
    if k is not None:
        k.tofile(pre + "_k.f64"); om.tofile(pre + "_omega.f64")
    out = os.path.join(td, "nnout")
    kw = dict(Nx=Nx, Ny=Ny, Nz=1, x0=0, x1=Lx, y0=-1, y1=1, model=kind, nn_weights=model_dir, out=out, nu=NU, delta=1.0)
    kw["in"] = pre
    kw.update(extra or {})
    meta = R.run_ref("nn", **kw)
    sh = R.shapes(Nx, Ny, 1)["c"]
    res = {"nut": np.fromfile(out + "_nut.f64").reshape(sh)}
    for n in ("xx", "xy", "yy"):
        res["tau_" + n] = np.fromfile(out + f"_tau_{n}.f64").reshape(sh)
    return res, meta


def backend_update(B, h, Nx, Ny, u, v, k, om, Lx=4.0, Nz=1, want_tau=True, mesh_kw=None):
    sh = R.shapes(Nx, Ny, Nz)["c"]
    md = R.capi.MeshDesc(Nx=Nx, Ny=Ny, Nz=Nz, Ng=1, dx=Lx / Nx, dy=2.0 / Ny, dz=1.0, y_lo=-1.0, y_hi=1.0, Nz_global=Nz)
    keep = []
    for kk, vv in (mesh_kw or {}).items():
        if isinstance(vv, np.ndarray):
            keep.append(vv)
            vv = R.capi.dptr(vv)
        setattr(md, kk, vv)
    nut = np.zeros(sh)
    tau = [np.zeros(sh) for _ in range(3)] if want_tau else [None] * 3
    B.check(B.lib.cfdnn_nn_update(h, C.byref(md), R.capi.dptr(u), R.capi.dptr(v), R.capi.dptr(k) if k is not None else None,
                                  R.capi.dptr(om) if om is not None else None, R.capi.dptr(nut),
                                  *[R.capi.dptr(t) if t is not None else None for t in tau]))
    return {"nut": nut, "tau_xx": tau[0], "tau_xy": tau[1], "tau_yy": tau[2]}


def random_mlp(seed=3, dims=(6, 32, 32, 1), out_bias=0.02):
    """6-32-32-1 tanh net (the reference's NN-MLP architecture, data/models/mlp_*) with weights chosen
    so that outputs fall inside (0, nu_t_max): the shipped MLP presets clip to 0 on these inputs."""
    rng = np.random.default_rng(seed)
    W = [rng.standard_normal((dims[l + 1], dims[l])) / np.sqrt(dims[l]) for l in range(len(dims) - 1)]
    b = [0.1 * rng.standard_normal(dims[l + 1]) for l in range(len(dims) - 1)]
    W[-1] *= 0.01
    b[-1][:] = out_bias
    means = 0.1 * rng.standard_normal(dims[0])
    stds = 1.0 + rng.random(dims[0])
    return W, b, means, stds


def write_model(d, W, b, means, stds):
    os.makedirs(d, exist_ok=True)
    for l, (w, bb) in enumerate(zip(W, b)):
        np.savetxt(os.path.join(d, f"layer{l}_W.txt"), np.atleast_2d(w), fmt="%.17g")
        np.savetxt(os.path.join(d, f"layer{l}_b.txt"), bb, fmt="%.17g")
    np.savetxt(os.path.join(d, "input_means.txt"), means, fmt="%.17g")
    np.savetxt(os.path.join(d, "input_stds.txt"), stds, fmt="%.17g")
    return d


def numpy_mlp(W, b, means, stds, x):
    a = (x - means) / stds if means is not None else x
    for l, (w, bb) in enumerate(zip(W, b)):
        a = a @ w.T + bb
        if l + 1 < len(W):
            a = np.tanh(a)
    return a
