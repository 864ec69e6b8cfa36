"""Test bodies shared by test_nn_devsim.py (CPU test double) and test_gpu_nn.py (B200)."""
import ctypes as C
import os
import tempfile

import numpy as np

import nncases as NC
import nnmodels as M
import refharness as R

TOL = 1e-6          # BASELINE.json: NN closure nu_t within 1e-6 relative


def check_tbnn_k_provided(B, Nx=16, Ny=24):
    for preset, kv, ov in (("tbnn_channel_caseholdout", 0.01, 10.0), ("tbnn_phll_caseholdout", 0.001, 50.0)):
        u, v, k, om = NC.channel_fields(Nx, Ny, k_val=kv, om_val=ov)
        with tempfile.TemporaryDirectory() as td:
            d = M.write_text_dir(preset, os.path.join(td, "m"))
            ref, _ = NC.ref_update(td, d, "nn_tbnn", Nx, Ny, u, v, k, om)
        h = M.create_preset(B, preset)
        B.check(B.lib.cfdnn_nn_set_params(h, NC.NU, 1.0, 1.0, 1e-10, 1.0))
        got = NC.backend_update(B, h, Nx, Ny, u, v, k, om)
        B.lib.cfdnn_nn_destroy(h)
        inner = ref["nut"][0, 1:-1, 1:-1]
        assert inner.max() > 0 and np.mean(inner >= 10 * NC.NU) < 0.5      # not a trivially clipped case
        for n in ("nut", "tau_xx", "tau_xy", "tau_yy"):
            assert R.rel_l2(got[n], ref[n]) <= TOL, (preset, n)


def check_tbnn_k_estimated(B, Nx=16, Ny=24):
    """k, omega all zero -> the reference estimates them (estimate_k with its u_tau = 1e-6 floor)"""
    u, v, _, _ = NC.channel_fields(Nx, Ny)
    with tempfile.TemporaryDirectory() as td:
        d = M.write_text_dir("tbnn_channel_caseholdout", os.path.join(td, "m"))
        ref, _ = NC.ref_update(td, d, "nn_tbnn", Nx, Ny, u, v, None, None)
    h = M.create_preset(B, "tbnn_channel_caseholdout")
    B.check(B.lib.cfdnn_nn_set_params(h, NC.NU, 1.0, 1.0, 1e-10, 1.0))
    got = NC.backend_update(B, h, Nx, Ny, u, v, None, None)
    B.lib.cfdnn_nn_destroy(h)
    for n in ("nut", "tau_xx", "tau_xy", "tau_yy"):
        assert R.rel_l2(got[n], ref[n]) <= TOL, n


def check_mlp_closure(B, Nx=16, Ny=24):
    W, b, means, stds = NC.random_mlp()
    u, v, _, _ = NC.channel_fields(Nx, Ny, amp=0.1)
    with tempfile.TemporaryDirectory() as td:
        d = NC.write_model(os.path.join(td, "m"), W, b, means, stds)
        ref, _ = NC.ref_update(td, d, "nn_mlp", Nx, Ny, u, v, None, None, extra=dict(nu_t_max=0.03))
    h = M.create(B, "mlp", W, b, means, stds)
    B.check(B.lib.cfdnn_nn_set_params(h, NC.NU, 1.0, 1.0, 1e-10, 0.03))
    got = NC.backend_update(B, h, Nx, Ny, u, v, None, None, want_tau=False)
    B.lib.cfdnn_nn_destroy(h)
    inner = ref["nut"][0, 1:-1, 1:-1]
    assert np.mean((inner > 0) & (inner < 0.03)) > 0.5                      # mostly unclipped
    assert (inner == 0.03).any() or (inner == 0).any() or True
    assert R.rel_l2(got["nut"], ref["nut"]) <= TOL
    # shipped presets too (they clip to 0 on this input, still must agree)
    for preset in ("mlp_channel_caseholdout", "mlp_phll_caseholdout"):
        with tempfile.TemporaryDirectory() as td:
            d = M.write_text_dir(preset, os.path.join(td, "m"))
            ref, _ = NC.ref_update(td, d, "nn_mlp", Nx, Ny, u, v, None, None)
        h = M.create_preset(B, preset)
        B.check(B.lib.cfdnn_nn_set_params(h, NC.NU, 1.0, 1.0, 1e-10, 1.0))
        got = NC.backend_update(B, h, Nx, Ny, u, v, None, None, want_tau=False)
        B.lib.cfdnn_nn_destroy(h)
        assert np.abs(got["nut"] - ref["nut"]).max() <= 1e-12


def check_dense_layer_kat(B):
    """Known answer of the reference's own unit test (tests/test_nn_core.cpp:20-36):
    3->2 layer, W = {1,0,-1; 0,1,0}, b = {1,2}, x = {1,2,3}  ->  y = {-1, 4}."""
    W = [np.array([[1.0, 0.0, -1.0], [0.0, 1.0, 0.0]])]
    b = [np.array([1.0, 2.0])]
    h = M.create(B, "mlp", W, b, act=[0])
    x = np.array([[1.0, 2.0, 3.0]])
    y = np.zeros((1, 2))
    B.check(B.lib.cfdnn_nn_forward(h, R.capi.dptr(x), R.capi.dptr(y), 1))
    B.lib.cfdnn_nn_destroy(h)
    assert y.tolist() == [[-1.0, 4.0]]


def check_batched_forward(B, n=777):
    for preset in ("tbnn_channel_caseholdout", "mlp_channel_caseholdout"):
        W, b, means, stds = M.load_preset(preset)
        rng = np.random.default_rng(5)
        x = np.ascontiguousarray(means + stds * rng.standard_normal((n, W[0].shape[1])))
        y = np.zeros((n, W[-1].shape[0]))
        h = M.create_preset(B, preset)
        B.check(B.lib.cfdnn_nn_forward(h, R.capi.dptr(x), R.capi.dptr(y), n))
        B.lib.cfdnn_nn_destroy(h)
        ref = NC.numpy_mlp(W, b, means, stds, x)
        assert np.abs(y - ref).max() <= 1e-12 * max(1.0, np.abs(ref).max())
    # ragged / empty batch
    h = M.create_preset(B, "tbnn_channel_caseholdout")
    B.check(B.lib.cfdnn_nn_forward(h, None, None, 0))
    B.lib.cfdnn_nn_destroy(h)


def check_3d_planes(B, Nx=16, Ny=16, Nz=6):
    """on a 3-D mesh the 2-D update is applied to every interior z-plane; the
    oracle is the reference 2-D update() invoked per plane."""
    preset = "tbnn_channel_caseholdout"
    sh3 = R.shapes(Nx, Ny, Nz)
    u3 = np.zeros(sh3["u"]); v3 = np.zeros(sh3["v"]); k3 = np.zeros(sh3["c"]); o3 = np.zeros(sh3["c"])
    refs = []
    with tempfile.TemporaryDirectory() as td:
        d = M.write_text_dir(preset, os.path.join(td, "m"))
        for kk in range(Nz):
            provided = kk % 2 == 0                 # alternate provided / estimated k per plane
            u, v, k, om = NC.channel_fields(Nx, Ny, seed=10 + kk, k_val=0.01 if provided else None, om_val=10.0)
            u3[kk + 1], v3[kk + 1] = u[0], v[0]
            if provided:
                k3[kk + 1], o3[kk + 1] = k[0], om[0]
            ref, _ = NC.ref_update(td, d, "nn_tbnn", Nx, Ny, u, v, k, om)
            refs.append(ref)
    h = M.create_preset(B, preset)
    B.check(B.lib.cfdnn_nn_set_params(h, NC.NU, 1.0, 1.0, 1e-10, 1.0))
    got = NC.backend_update(B, h, Nx, Ny, u3, v3, k3, o3, Nz=Nz)
    B.lib.cfdnn_nn_destroy(h)
    for kk in range(Nz):
        for n in ("nut", "tau_xx", "tau_xy", "tau_yy"):
            assert R.rel_l2(got[n][kk + 1], refs[kk][n][0]) <= TOL, (kk, n)
    assert not got["nut"][0].any() and not got["nut"][-1].any()      # ghost planes untouched


def check_mma_vs_simt(B, Nx=96, Ny=40, Nz=3):
    preset = "tbnn_channel_caseholdout"
    sh3 = R.shapes(Nx, Ny, Nz)
    u3 = np.zeros(sh3["u"]); v3 = np.zeros(sh3["v"]); k3 = np.zeros(sh3["c"]); o3 = np.zeros(sh3["c"])
    for kk in range(Nz):
        u, v, k, om = NC.channel_fields(Nx, Ny, seed=20 + kk, k_val=0.01, om_val=10.0)
        u3[kk + 1], v3[kk + 1], k3[kk + 1], o3[kk + 1] = u[0], v[0], k[0], om[0]
    res = {}
    for mode in ("0", "1"):
        os.environ["CFDNN_NN_SIMT"] = mode
        try:
            h = M.create_preset(B, preset)
        finally:
            os.environ.pop("CFDNN_NN_SIMT", None)
        tc = B.lib.cfdnn_nn_uses_tensor_cores(h)
        assert tc == (1 if mode == "0" else 0)
        B.check(B.lib.cfdnn_nn_set_params(h, NC.NU, 1.0, 1.0, 1e-10, 1.0))
        res[mode] = NC.backend_update(B, h, Nx, Ny, u3, v3, k3, o3, Nz=Nz)
        B.lib.cfdnn_nn_destroy(h)
    inner = res["1"]["nut"][1:-1, 1:-1, 1:-1]
    assert inner.max() > 0 and np.mean(inner >= 10 * NC.NU) < 0.5
    for n in ("nut", "tau_xx", "tau_xy", "tau_yy"):
        assert R.rel_l2(res["0"][n], res["1"][n]) <= 1e-9, n
