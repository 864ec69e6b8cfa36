"""Model fixtures (tests/golden/models.npz) -> cfdnn_nn handles and text dirs for the oracle."""
import ctypes as C
import os

import numpy as np

import refharness as R

NPZ = os.path.join(R.ROOT, "tests", "golden", "models.npz")
KIND = {"mlp": 1, "tbnn": 2}


def load_preset(name):
    z = np.load(NPZ)
    n = int(z[f"{name}/n_layers"])
    W = [np.ascontiguousarray(z[f"{name}/W{l}"], dtype=np.float64) for l in range(n)]
    b = [np.ascontiguousarray(z[f"{name}/b{l}"], dtype=np.float64) for l in range(n)]
    return W, b, np.ascontiguousarray(z[f"{name}/means"]), np.ascontiguousarray(z[f"{name}/stds"])


def write_text_dir(name, d):
    """the reference's own on-disk format (layerN_W.txt ...), %.17g round-trips exactly"""
    W, b, means, stds = load_preset(name)
    os.makedirs(d, exist_ok=True)
    for l, (w, bb) in enumerate(zip(W, b)):
        np.savetxt(os.path.join(d, f"layer{l}_W.txt"), w, fmt="%.17g")
        np.savetxt(os.path.join(d, f"layer{l}_b.txt"), bb, fmt="%.17g")
    np.savetxt(os.path.join(d, "input_means.txt"), means, fmt="%.17g")
    np.savetxt(os.path.join(d, "input_stds.txt"), stds, fmt="%.17g")
    return d


def create(B, kind, W, b, means=None, stds=None, act=None):
    n = len(W)
    dims = [W[0].shape[1]] + [w.shape[0] for w in W]
    dims_c = (C.c_int * (n + 1))(*dims)
    act_c = (C.c_int * n)(*act) if act is not None else None
    Wp = (R.capi.c_double_p * n)(*[R.capi.dptr(w) for w in W])
    bp = (R.capi.c_double_p * n)(*[R.capi.dptr(x) for x in b])
    h = C.c_void_p()
    B.check(B.lib.cfdnn_nn_create(KIND[kind], n, dims_c, act_c, Wp, bp, R.capi.dptr(means) if means is not None else None,
                                  R.capi.dptr(stds) if stds is not None else None, C.byref(h)))
    return h


def create_preset(B, name):
    W, b, means, stds = load_preset(name)
    return create(B, "tbnn" if name.startswith("tbnn") else "mlp", W, b, means, stds)
