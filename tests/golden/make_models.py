"""Packs the reference's model assets (data/models/<preset>/layerN_{W,b}.txt, input_means/stds.txt —
the text files the reference's MLP::load_weights reads, src/nn_core.cpp:501-611) into one small
.npz fixture so the GPU box (no /root/reference) can feed identical weights to both the oracle
binary and the CUDA path.  Run in the build container:  python tests/golden/make_models.py"""
import os

import numpy as np

REF = "/root/reference/data/models"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "models.npz")
PRESETS = ["tbnn_channel_caseholdout", "mlp_channel_caseholdout", "tbnn_phll_caseholdout", "mlp_phll_caseholdout"]

out = {}
for p in PRESETS:
    d = os.path.join(REF, p)
    l = 0
    while os.path.exists(os.path.join(d, f"layer{l}_W.txt")):
        out[f"{p}/W{l}"] = np.atleast_2d(np.loadtxt(os.path.join(d, f"layer{l}_W.txt")))
        out[f"{p}/b{l}"] = np.atleast_1d(np.loadtxt(os.path.join(d, f"layer{l}_b.txt")))
        l += 1
    out[f"{p}/means"] = np.atleast_1d(np.loadtxt(os.path.join(d, "input_means.txt")))
    out[f"{p}/stds"] = np.atleast_1d(np.loadtxt(os.path.join(d, "input_stds.txt")))
    out[f"{p}/n_layers"] = np.array(l)
np.savez_compressed(OUT, **out)
print("wrote", OUT, os.path.getsize(OUT), "bytes")
