// capi.cu — the C ABI of libcfdnn_b200.so (declared in include/cfdnn_b200.h).
#include "mg.cuh"
#include "fft.cuh"
#include "comm.cuh"
#ifndef CFDNN_MG_ONLY
#include "solver.cuh"
#include "nn.cuh"
#include "les.cuh"
#endif
#include <mutex>
#include <cstring>
#include <memory>
#include <algorithm>

namespace cfdnn {
std::atomic<unsigned long long> g_launches{0};
static thread_local std::string t_last_error;
void set_last_error(const std::string& m) { t_last_error = m; }

void require_device() {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0)
        throw Error(CFDNN_ERR_NO_DEVICE,
                    "libcfdnn_b200: no usable CUDA device (this backend has no CPU path)");
}
}  // namespace cfdnn

using namespace cfdnn;

struct cfdnn_mg { Multigrid impl; explicit cfdnn_mg(const cfdnn_mesh_desc& m) : impl(m) {} };
struct cfdnn_fft { FftPoisson impl; explicit cfdnn_fft(const cfdnn_mesh_desc& m) : impl(m) {} };
#ifndef CFDNN_MG_ONLY
struct cfdnn_nn {
    NNClosure impl;
    cfdnn_nn(int kind, int nl, const int* dims, const int* act, const double* const* W, const double* const* b,
             const double* means, const double* stds) : impl(kind, nl, dims, act, W, b, means, stds) {}
};
struct cfdnn_solver {
    Solver impl;
    std::unique_ptr<Comm> comm;
    cfdnn_solver(const cfdnn_mesh_desc& m, const cfdnn_config& c) : impl(m, c) {}
};
#endif

extern "C" {

const char* cfdnn_last_error(void) { return t_last_error.c_str(); }
int cfdnn_version(void) { return 100; }
unsigned long long cfdnn_kernel_launches(void) { return g_launches.load(); }
int cfdnn_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}
int cfdnn_set_device(int d) { return guarded([&] { CFDNN_CUDA(cudaSetDevice(d)); }); }
int cfdnn_device_synchronize(void) { return guarded([&] { CFDNN_CUDA(cudaDeviceSynchronize()); check_launch("sync"); }); }

void cfdnn_poisson_cfg_default(cfdnn_poisson_cfg* c) {   // include/poisson_solver.hpp:17-52
    c->tol = 1e-6; c->tol_abs = 0.0; c->tol_rhs = 1e-3; c->tol_rel = 1e-3;
    c->max_vcycles = 10000; c->check_interval = 3; c->use_l2_norm = 1; c->linf_safety_factor = 10.0;
    c->fixed_cycles = 0; c->nu1 = 0; c->nu2 = 0; c->chebyshev_degree = 4;
}
void cfdnn_config_default(cfdnn_config* c) {             // include/config.hpp:65-234
    std::memset(c, 0, sizeof(*c));
    c->nu = 0.001; c->dt = 0.001;
    c->time_integrator = CFDNN_TI_EULER; c->convective_scheme = CFDNN_CONV_CENTRAL;
    c->adaptive_dt = 1; c->CFL_max = 0.5; c->CFL_xz = -1.0; c->dt_safety = 1.0;
    c->poisson_fixed_cycles = 8; c->poisson_max_vcycles = 20; c->poisson_tol_rhs = 1e-6;
    c->poisson_tol = 1e-6; c->poisson_abs_tol_floor = 1e-8; c->poisson_tol_abs = 0.0; c->poisson_tol_rel = 1e-3;
    c->poisson_check_interval = 3; c->poisson_use_l2_norm = 1; c->poisson_linf_safety = 10.0;
    c->poisson_nu1 = 0; c->poisson_nu2 = 0; c->poisson_chebyshev_degree = 4;
    c->turb_guard_enabled = 1; c->turb_guard_interval = 5;
    c->turb_model = CFDNN_TURB_NONE; c->nu_t_max = 1.0;
    c->implicit_y_diffusion = 0;
    c->poisson_solver = CFDNN_POISSON_MG;
}

// ---------------------------------------------------------------- multigrid
int cfdnn_mg_create(const cfdnn_mesh_desc* mesh, cfdnn_mg** out) {
    return guarded([&] {
        if (!mesh || !out) throw std::invalid_argument("cfdnn_mg_create: null argument");
        *out = new cfdnn_mg(*mesh);
    });
}
void cfdnn_mg_destroy(cfdnn_mg* mg) { delete mg; }

int cfdnn_fft_create(const cfdnn_mesh_desc* mesh, cfdnn_fft** out) {
    return guarded([&] {
        if (!mesh || !out) throw std::invalid_argument("cfdnn_fft_create: null argument");
        *out = new cfdnn_fft(*mesh);
    });
}
void cfdnn_fft_destroy(cfdnn_fft* f) { delete f; }
int cfdnn_fft_set_bc(cfdnn_fft* f, const int bc[6]) { return guarded([&] { f->impl.set_bc(bc); }); }
int cfdnn_fft_is_suitable(const int bc[6], int uniform_x, int uniform_z) {       // poisson_solver_fft.cpp:284-292
    return bc[0] == CFDNN_PBC_PERIODIC && bc[1] == CFDNN_PBC_PERIODIC && bc[4] == CFDNN_PBC_PERIODIC &&
           bc[5] == CFDNN_PBC_PERIODIC && uniform_x && uniform_z;
}
int cfdnn_fft_solve_dev(cfdnn_fft* f, const double* rhs_dev, double* p_dev, void* stream, int* iterations) {
    return guarded([&] {
        const int it = f->impl.solve_dev(rhs_dev, p_dev, static_cast<cudaStream_t>(stream));
        if (iterations) *iterations = it;
    });
}
int cfdnn_fft_solve(cfdnn_fft* f, const double* rhs, double* p, int* iterations) {
    return guarded([&] {
        const int it = f->impl.solve_host(rhs, p);
        if (iterations) *iterations = it;
    });
}
int cfdnn_mg_set_bc(cfdnn_mg* mg, const int bc[6]) { return guarded([&] { mg->impl.set_bc(bc); }); }
int cfdnn_mg_num_levels(const cfdnn_mg* mg) { return mg->impl.num_levels(); }
int cfdnn_mg_solve(cfdnn_mg* mg, const double* rhs, double* p, const cfdnn_poisson_cfg* cfg,
                   cfdnn_poisson_stats* stats) {
    return guarded([&] {
        mg->impl.solve_host(rhs, p, *cfg);
        if (stats) *stats = mg->impl.stats;
    });
}
int cfdnn_mg_solve_dev(cfdnn_mg* mg, const double* rhs_dev, double* p_dev, const cfdnn_poisson_cfg* cfg,
                       cfdnn_poisson_stats* stats) {
    return guarded([&] {
        cudaStream_t st = mg->impl.own_stream();
        mg->impl.solve_dev(rhs_dev, p_dev, *cfg, st);
        CFDNN_CUDA(cudaStreamSynchronize(st));
        check_launch("mg solve_dev");
        mg->impl.collect_timing();
        if (stats) *stats = mg->impl.stats;
    });
}
int cfdnn_mg_last_timing(const cfdnn_mg* mg, double* ms, int* n) {
    if (ms) *ms = mg->impl.vcycle_ms_total;
    if (n) *n = mg->impl.vcycle_count;
    return CFDNN_OK;
}


#ifndef CFDNN_MG_ONLY
// ---------------------------------------------------------------- solver
int cfdnn_solver_create(const cfdnn_mesh_desc* mesh, const cfdnn_config* cfg, cfdnn_solver** out) {
    return guarded([&] {
        if (!mesh || !cfg || !out) throw std::invalid_argument("cfdnn_solver_create: null argument");
        *out = new cfdnn_solver(*mesh, *cfg);
    });
}
void cfdnn_solver_destroy(cfdnn_solver* s) { delete s; }
int cfdnn_solver_set_velocity_bc(cfdnn_solver* s, const int bc[6]) { return guarded([&] { s->impl.set_velocity_bc(bc); }); }
int cfdnn_solver_set_body_force(cfdnn_solver* s, double fx, double fy, double fz) {
    return guarded([&] { s->impl.set_body_force(fx, fy, fz); });
}
int cfdnn_solver_set_dt(cfdnn_solver* s, double dt) { return guarded([&] { s->impl.set_dt(dt); }); }
int cfdnn_solver_get_dt(const cfdnn_solver* s, double* dt) { *dt = s->impl.dt(); return CFDNN_OK; }
int cfdnn_solver_upload(cfdnn_solver* s, int field, const double* host, size_t n) {
    return guarded([&] { s->impl.upload(field, host, n); });
}
int cfdnn_solver_download(cfdnn_solver* s, int field, double* host, size_t n) {
    return guarded([&] { s->impl.download(field, host, n); });
}
size_t cfdnn_solver_field_size(const cfdnn_solver* s, int field) {
    size_t n = 0;
    guarded([&] { n = s->impl.field_size(field); });
    return n;
}
int cfdnn_solver_initialize(cfdnn_solver* s) { return guarded([&] { s->impl.initialize(); }); }
int cfdnn_solver_step(cfdnn_solver* s, double* residual_out) {
    return guarded([&] { double r = s->impl.step(); if (residual_out) *residual_out = r; });
}
int cfdnn_solver_run(cfdnn_solver* s, int nsteps, double* last_residual_out) {
    return guarded([&] { s->impl.run(nsteps, last_residual_out); });
}
int cfdnn_solver_compute_adaptive_dt(cfdnn_solver* s, double* dt_out) {
    return guarded([&] { *dt_out = s->impl.compute_adaptive_dt(); });
}
int cfdnn_solver_poisson_stats(const cfdnn_solver* s, cfdnn_poisson_stats* out) {
    *out = s->impl.poisson_stats();
    return CFDNN_OK;
}
int cfdnn_solver_cycle_log(cfdnn_solver* s, int* cycles, int cap, int* n_out) {
    auto& log = s->impl.cycle_log;
    const int n = (int)std::min<size_t>(log.size(), (size_t)std::max(cap, 0));
    for (int i = 0; i < n; ++i) cycles[i] = log[log.size() - n + i];
    if (n_out) *n_out = (int)log.size();
    log.clear();
    return CFDNN_OK;
}
int cfdnn_solver_set_turbulence_model(cfdnn_solver* s, cfdnn_nn* nn) {
    return guarded([&] { s->impl.set_turbulence_model(nn ? &nn->impl : nullptr); });
}
int cfdnn_solver_last_timing(const cfdnn_solver* s, double* step_ms, double* vcycle_ms, int* n_vcycles) {
    if (step_ms) *step_ms = s->impl.last_step_ms;
    if (vcycle_ms) *vcycle_ms = s->impl.last_vcycle_ms;
    if (n_vcycles) *n_vcycles = s->impl.last_vcycles;
    return CFDNN_OK;
}
int cfdnn_solver_device_ptr(cfdnn_solver* s, int field, double** dev_out, size_t* n_out) {
    return guarded([&] {
        size_t n = 0;
        double* p = s->impl.view_ptr(field, &n);
        if (dev_out) *dev_out = p;
        if (n_out) *n_out = n;
    });
}
int cfdnn_solver_set_external_nu_t(cfdnn_solver* s, int on) {
    return guarded([&] { s->impl.set_external_nu_t(on != 0, on == 2); });
}
int cfdnn_solver_phase_timing(const cfdnn_solver* s, double ms[CFDNN_PHASE_COUNT]) {
    for (int q = 0; q < CFDNN_PHASE_COUNT; ++q) ms[q] = s->impl.last_phase_ms[q];
    return CFDNN_OK;
}
int cfdnn_solver_time_smoother(cfdnn_solver* s, int reps, double* ms_per_launch) {
    return guarded([&] {
        const double ms = s->impl.mg().time_level0_smoother(reps, s->impl.stream());
        if (ms_per_launch) *ms_per_launch = ms;
    });
}
int cfdnn_comm_unique_id(void* id128) { return guarded([&] { nccl_unique_id(id128); }); }
int cfdnn_solver_comm_init(cfdnn_solver* s, int rank, int nranks, const void* id128) {
    return guarded([&] {
        s->comm.reset(make_nccl_comm(rank, nranks, id128));
        s->impl.comm_init(s->comm.get());
    });
}
#ifdef CFDNN_DEVSIM
// test double only: slab communicator over host callbacks (gloo in tests/)
int cfdnn_devsim_solver_comm_init(cfdnn_solver* s, int rank, int nranks, const cfdnn_comm_callbacks_t* cb) {
    return guarded([&] {
        s->comm.reset(make_callback_comm(rank, nranks, *cb));
        s->impl.comm_init(s->comm.get());
    });
}
#endif

// ---------------------------------------------------------------- NN closures
int cfdnn_nn_create(int kind, int n_layers, const int* dims, const int* act, const double* const* W,
                    const double* const* b, const double* means, const double* stds, cfdnn_nn** out) {
    return guarded([&] {
        if (!dims || !W || !b || !out) throw std::invalid_argument("cfdnn_nn_create: null argument");
        *out = new cfdnn_nn(kind, n_layers, dims, act, W, b, means, stds);
    });
}
void cfdnn_nn_destroy(cfdnn_nn* nn) { delete nn; }
int cfdnn_nn_set_params(cfdnn_nn* nn, double nu, double delta, double u_ref, double k_min, double nu_t_max) {
    return guarded([&] { nn->impl.set_params(nu, delta, u_ref, k_min, nu_t_max); });
}
int cfdnn_nn_update(cfdnn_nn* nn, const cfdnn_mesh_desc* mesh, const double* u, const double* v, const double* k,
                    const double* omega, double* nu_t, double* tau_xx, double* tau_xy, double* tau_yy) {
    return guarded([&] { nn->impl.update_host(*mesh, u, v, k, omega, nu_t, tau_xx, tau_xy, tau_yy); });
}
int cfdnn_nn_update_dev(cfdnn_nn* nn, const cfdnn_mesh_desc* mesh, const double* u_dev, const double* v_dev,
                        const double* k_dev, const double* omega_dev, double* nu_t_dev, double* tau_xx_dev,
                        double* tau_xy_dev, double* tau_yy_dev) {
    return guarded([&] { nn->impl.update_dev_ptrs(*mesh, u_dev, v_dev, k_dev, omega_dev, nu_t_dev, tau_xx_dev, tau_xy_dev, tau_yy_dev); });
}
int cfdnn_les_update(const cfdnn_mesh_desc* mesh, int model, double constant, const double* u, const double* v,
                     const double* w, double* nu_t) {
    return guarded([&] {
        if (!mesh) throw std::invalid_argument("cfdnn_les_update: null mesh");
        les_update_host(*mesh, model, constant, u, v, w, nu_t);
    });
}
int cfdnn_solver_set_les_model(cfdnn_solver* s, int model, double constant) {
    return guarded([&] { s->impl.set_les_model(model, constant); });
}
int cfdnn_solver_fft_peer_to_peer(const cfdnn_solver* s) { return s->impl.fft_peer_to_peer() ? 1 : 0; }
int cfdnn_solver_set_sst_model(cfdnn_solver* s, int kind) { return guarded([&] { s->impl.set_sst_model(kind); }); }
int cfdnn_solver_set_rans_model(cfdnn_solver* s, int kind) { return cfdnn_solver_set_sst_model(s, kind); }
int cfdnn_solver_set_mixing_length_params(cfdnn_solver* s, double kappa, double A_plus, double delta) {
    return guarded([&] { s->impl.set_mixing_length_params(kappa, A_plus, delta); });
}
int cfdnn_solver_set_gep_params(cfdnn_solver* s, int variant, double model_nu) {
    return guarded([&] { s->impl.set_gep_params(variant, model_nu); });
}
int cfdnn_solver_set_pope_constants(cfdnn_solver* s, double C1, double C2) {
    return guarded([&] { s->impl.set_pope_constants(C1, C2); });
}
int cfdnn_solver_set_trip_forcing(cfdnn_solver* s, const cfdnn_trip_desc* desc) {
    return guarded([&] { s->impl.set_trip_forcing(desc); });
}
int cfdnn_solver_get_time(const cfdnn_solver* s, double* t) { *t = s->impl.time(); return CFDNN_OK; }
int cfdnn_solver_set_time(cfdnn_solver* s, double t) { s->impl.set_time(t); return CFDNN_OK; }
int cfdnn_solver_apply_velocity_filter(cfdnn_solver* s, double strength) {
    return guarded([&] { s->impl.apply_velocity_filter(strength); });
}
int cfdnn_nn_forward(cfdnn_nn* nn, const double* x, double* y, size_t n) {
    return guarded([&] { nn->impl.forward_host(x, y, n); });
}
int cfdnn_nn_last_timing(const cfdnn_nn* nn, double* ms) { if (ms) *ms = nn->impl.last_ms; return CFDNN_OK; }
int cfdnn_nn_uses_tensor_cores(const cfdnn_nn* nn) { return nn && nn->impl.uses_tensor_cores() ? 1 : 0; }
#endif  // CFDNN_MG_ONLY

}  // extern "C"
