"""`-m gpu`: NN closures on the B200 through the C ABI vs the oracle (reference CPU update())."""
import pytest

import nn_parity_common as P

pytestmark = pytest.mark.gpu


def test_dense_layer_known_answer(gpu):
    P.check_dense_layer_kat(gpu)


def test_batched_forward_vs_numpy(gpu):
    P.check_batched_forward(gpu, n=100003)


def test_tbnn_update_k_provided(gpu, oracle):
    P.check_tbnn_k_provided(gpu, Nx=64, Ny=48)


def test_tbnn_update_k_estimated(gpu, oracle):
    P.check_tbnn_k_estimated(gpu, Nx=64, Ny=48)


def test_mlp_update(gpu, oracle):
    P.check_mlp_closure(gpu, Nx=64, Ny=48)


def test_tbnn_3d_per_plane(gpu, oracle):
    P.check_3d_planes(gpu, Nx=32, Ny=32, Nz=6)


def test_tbnn_update_ragged_tiles(gpu, oracle):
    """Nx not a multiple of the 32-cell warp tile, Ny not a multiple of the 4-row CTA"""
    P.check_tbnn_k_provided(gpu, Nx=50, Ny=37)
    P.check_mlp_closure(gpu, Nx=50, Ny=37)


def test_tensor_core_vs_scalar_kernel(gpu):
    """The FP64 tensor-core (DMMA) kernel against the scalar per-thread kernel of the same library
    (CFDNN_NN_SIMT=1 selects it at closure creation): same cells, nu_t/tau within 1e-9."""
    P.check_mma_vs_simt(gpu, Nx=96, Ny=40, Nz=3)
