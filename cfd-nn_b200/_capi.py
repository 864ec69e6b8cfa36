"""ctypes binding of the C ABI in include/cfdnn_b200.h (used by tests/ and bench.py).

The product library is cfd-nn_b200/libcfdnn_b200.so (nvcc, sm_100a).  There is no CPU path:
`load()` raises if the library is missing, and every compute entry returns
CFDNN_ERR_NO_DEVICE without a GPU.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_LIB = os.path.join(HERE, "libcfdnn_b200.so")

c_double_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int)


class MeshDesc(C.Structure):
    _fields_ = [("Nx", C.c_int), ("Ny", C.c_int), ("Nz", C.c_int), ("Ng", C.c_int),
                ("dx", C.c_double), ("dy", C.c_double), ("dz", C.c_double),
                ("y_stretched", C.c_int),
                ("dyv", c_double_p), ("dyc", c_double_p),
                ("yLap_aS", c_double_p), ("yLap_aN", c_double_p), ("yLap_aP", c_double_p),
                ("yc", c_double_p), ("y_lo", C.c_double), ("y_hi", C.c_double),
                ("Nz_global", C.c_int), ("k_offset", C.c_int), ("yf", c_double_p)]


class PoissonCfg(C.Structure):
    _fields_ = [("tol", C.c_double), ("tol_abs", C.c_double), ("tol_rhs", C.c_double),
                ("tol_rel", C.c_double), ("max_vcycles", C.c_int), ("check_interval", C.c_int),
                ("use_l2_norm", C.c_int), ("linf_safety_factor", C.c_double),
                ("fixed_cycles", C.c_int), ("nu1", C.c_int), ("nu2", C.c_int),
                ("chebyshev_degree", C.c_int)]


class PoissonStats(C.Structure):
    _fields_ = [("cycles", C.c_int), ("converged", C.c_int), ("fixed_cycle_mode", C.c_int),
                ("status", C.c_int), ("residual", C.c_double), ("residual_l2", C.c_double),
                ("b_inf", C.c_double), ("b_l2", C.c_double), ("r0", C.c_double),
                ("r0_l2", C.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class Config(C.Structure):
    _fields_ = [("nu", C.c_double), ("dt", C.c_double),
                ("time_integrator", C.c_int), ("convective_scheme", C.c_int),
                ("adaptive_dt", C.c_int),
                ("CFL_max", C.c_double), ("CFL_xz", C.c_double), ("dt_safety", C.c_double),
                ("poisson_fixed_cycles", C.c_int), ("poisson_max_vcycles", C.c_int),
                ("poisson_tol_rhs", C.c_double), ("poisson_tol", C.c_double),
                ("poisson_abs_tol_floor", C.c_double), ("poisson_tol_abs", C.c_double),
                ("poisson_tol_rel", C.c_double),
                ("poisson_check_interval", C.c_int), ("poisson_use_l2_norm", C.c_int),
                ("poisson_linf_safety", C.c_double),
                ("poisson_nu1", C.c_int), ("poisson_nu2", C.c_int), ("poisson_chebyshev_degree", C.c_int),
                ("turb_guard_enabled", C.c_int), ("turb_guard_interval", C.c_int),
                ("turb_model", C.c_int), ("nu_t_max", C.c_double), ("implicit_y_diffusion", C.c_int),
                ("poisson_solver", C.c_int)]


class TripDesc(C.Structure):
    _fields_ = [("amplitude", C.c_double), ("duration", C.c_double), ("ramp_off_start", C.c_double), ("w_scale", C.c_double),
                ("env_x", c_double_p), ("g_y", c_double_p), ("Fz_v", c_double_p), ("Fz_w", c_double_p)]


class CfdnnError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"cfdnn error {code}: {msg}")
        self.code = code


_SIGS = {
    "cfdnn_last_error": (C.c_char_p, []),
    "cfdnn_version": (C.c_int, []),
    "cfdnn_kernel_launches": (C.c_ulonglong, []),
    "cfdnn_device_count": (C.c_int, []),
    "cfdnn_set_device": (C.c_int, [C.c_int]),
    "cfdnn_device_synchronize": (C.c_int, []),
    "cfdnn_poisson_cfg_default": (None, [C.POINTER(PoissonCfg)]),
    "cfdnn_config_default": (None, [C.POINTER(Config)]),
    "cfdnn_mg_create": (C.c_int, [C.POINTER(MeshDesc), C.POINTER(C.c_void_p)]),
    "cfdnn_mg_destroy": (None, [C.c_void_p]),
    "cfdnn_mg_set_bc": (C.c_int, [C.c_void_p, c_int_p]),
    "cfdnn_mg_num_levels": (C.c_int, [C.c_void_p]),
    "cfdnn_mg_solve": (C.c_int, [C.c_void_p, c_double_p, c_double_p, C.POINTER(PoissonCfg), C.POINTER(PoissonStats)]),
    "cfdnn_mg_solve_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(PoissonCfg), C.POINTER(PoissonStats)]),
    "cfdnn_mg_last_timing": (C.c_int, [C.c_void_p, c_double_p, c_int_p]),
    "cfdnn_solver_create": (C.c_int, [C.POINTER(MeshDesc), C.POINTER(Config), C.POINTER(C.c_void_p)]),
    "cfdnn_solver_destroy": (None, [C.c_void_p]),
    "cfdnn_solver_set_velocity_bc": (C.c_int, [C.c_void_p, c_int_p]),
    "cfdnn_solver_set_body_force": (C.c_int, [C.c_void_p, C.c_double, C.c_double, C.c_double]),
    "cfdnn_solver_set_dt": (C.c_int, [C.c_void_p, C.c_double]),
    "cfdnn_solver_get_dt": (C.c_int, [C.c_void_p, c_double_p]),
    "cfdnn_solver_upload": (C.c_int, [C.c_void_p, C.c_int, c_double_p, C.c_size_t]),
    "cfdnn_solver_download": (C.c_int, [C.c_void_p, C.c_int, c_double_p, C.c_size_t]),
    "cfdnn_solver_field_size": (C.c_size_t, [C.c_void_p, C.c_int]),
    "cfdnn_solver_initialize": (C.c_int, [C.c_void_p]),
    "cfdnn_solver_step": (C.c_int, [C.c_void_p, c_double_p]),
    "cfdnn_solver_run": (C.c_int, [C.c_void_p, C.c_int, c_double_p]),
    "cfdnn_solver_compute_adaptive_dt": (C.c_int, [C.c_void_p, c_double_p]),
    "cfdnn_solver_poisson_stats": (C.c_int, [C.c_void_p, C.POINTER(PoissonStats)]),
    "cfdnn_solver_cycle_log": (C.c_int, [C.c_void_p, c_int_p, C.c_int, c_int_p]),
    "cfdnn_solver_set_turbulence_model": (C.c_int, [C.c_void_p, C.c_void_p]),
    "cfdnn_solver_last_timing": (C.c_int, [C.c_void_p, c_double_p, c_double_p, c_int_p]),
    "cfdnn_solver_device_ptr": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    "cfdnn_solver_set_external_nu_t": (C.c_int, [C.c_void_p, C.c_int]),
    "cfdnn_solver_phase_timing": (C.c_int, [C.c_void_p, c_double_p]),
    "cfdnn_solver_time_smoother": (C.c_int, [C.c_void_p, C.c_int, c_double_p]),
    "cfdnn_comm_unique_id": (C.c_int, [C.c_void_p]),
    "cfdnn_solver_comm_init": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "cfdnn_nn_create": (C.c_int, [C.c_int, C.c_int, c_int_p, c_int_p, C.POINTER(c_double_p), C.POINTER(c_double_p),
                                  c_double_p, c_double_p, C.POINTER(C.c_void_p)]),
    "cfdnn_nn_destroy": (None, [C.c_void_p]),
    "cfdnn_nn_set_params": (C.c_int, [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double]),
    "cfdnn_nn_update": (C.c_int, [C.c_void_p, C.POINTER(MeshDesc), c_double_p, c_double_p, c_double_p, c_double_p,
                                  c_double_p, c_double_p, c_double_p, c_double_p]),
    "cfdnn_nn_update_dev": (C.c_int, [C.c_void_p, C.POINTER(MeshDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "cfdnn_adaptive_dt_noalloc": (C.c_int, [C.POINTER(MeshDesc), C.POINTER(Config), C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]),
    "cfdnn_fft_create": (C.c_int, [C.POINTER(MeshDesc), C.POINTER(C.c_void_p)]),
    "cfdnn_fft_destroy": (None, [C.c_void_p]),
    "cfdnn_fft_set_bc": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "cfdnn_fft_is_suitable": (C.c_int, [C.POINTER(C.c_int), C.c_int, C.c_int]),
    "cfdnn_fft_solve_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int)]),
    "cfdnn_fft_solve": (C.c_int, [C.c_void_p, c_double_p, c_double_p, C.POINTER(C.c_int)]),
    "cfdnn_solver_fft_peer_to_peer": (C.c_int, [C.c_void_p]),
    "cfdnn_solver_set_sst_model": (C.c_int, [C.c_void_p, C.c_int]),
    "cfdnn_solver_set_rans_model": (C.c_int, [C.c_void_p, C.c_int]),
    "cfdnn_solver_set_mixing_length_params": (C.c_int, [C.c_void_p, C.c_double, C.c_double, C.c_double]),
    "cfdnn_solver_set_gep_params": (C.c_int, [C.c_void_p, C.c_int, C.c_double]),
    "cfdnn_solver_set_pope_constants": (C.c_int, [C.c_void_p, C.c_double, C.c_double]),
    "cfdnn_solver_set_trip_forcing": (C.c_int, [C.c_void_p, C.POINTER(TripDesc)]),
    "cfdnn_solver_get_time": (C.c_int, [C.c_void_p, C.POINTER(C.c_double)]),
    "cfdnn_solver_set_time": (C.c_int, [C.c_void_p, C.c_double]),
    "cfdnn_solver_apply_velocity_filter": (C.c_int, [C.c_void_p, C.c_double]),
    "cfdnn_les_update": (C.c_int, [C.POINTER(MeshDesc), C.c_int, C.c_double, c_double_p, c_double_p, c_double_p, c_double_p]),
    "cfdnn_solver_set_les_model": (C.c_int, [C.c_void_p, C.c_int, C.c_double]),
    "cfdnn_nn_forward": (C.c_int, [C.c_void_p, c_double_p, c_double_p, C.c_size_t]),
    "cfdnn_nn_last_timing": (C.c_int, [C.c_void_p, c_double_p]),
    "cfdnn_nn_uses_tensor_cores": (C.c_int, [C.c_void_p]),
}

EXPORTS = tuple(_SIGS)


def load(path=None):
    """dlopen the backend and declare every prototype of include/cfdnn_b200.h."""
    path = path or DEFAULT_LIB
    if not os.path.exists(path):
        raise ImportError(
            f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). This backend has no CPU fallback.")
    lib = C.CDLL(path, mode=C.RTLD_GLOBAL)
    for name, (res, args) in _SIGS.items():
        fn = getattr(lib, name)       # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    return lib


def dptr(a):
    if a is None:
        return None
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(c_double_p)


def check(lib, rc):
    if rc != 0:
        raise CfdnnError(rc, lib.cfdnn_last_error().decode(errors="replace"))
