"""`-m "not gpu"`: NN-closure host logic + plain-kernel arithmetic on the CPU test double vs the oracle."""
import nn_parity_common as P


def test_dense_layer_known_answer(sim):
    P.check_dense_layer_kat(sim)


def test_batched_forward_vs_numpy(sim):
    P.check_batched_forward(sim, n=65)


def test_tbnn_update_k_provided(sim, oracle):
    P.check_tbnn_k_provided(sim)


def test_tbnn_update_k_estimated(sim, oracle):
    P.check_tbnn_k_estimated(sim)


def test_mlp_update(sim, oracle):
    P.check_mlp_closure(sim)


def test_tbnn_3d_per_plane(sim, oracle):
    P.check_3d_planes(sim, Nz=3)
