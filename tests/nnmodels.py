"""Model fixtures (tests/golden/models.npz) -> cfdnn_nn handles and text dirs for the oracle."""
import ctypes as C
import os

import numpy as np

import refharness as R

NPZ = os.path.join(R.ROOT, "tests", "golden", "models.npz")
KIND = {"mlp": 1, "tbnn": 2}


def load_preset(name):
    return R.backend.load_npz_preset(NPZ, name)


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
    return R.backend.nn_create(B, kind, W, b, means, stds, act)


def create_preset(B, name):
    W, b, means, stds = load_preset(name)
    return create(B, "tbnn" if name.startswith("tbnn") else "mlp", W, b, means, stds)
