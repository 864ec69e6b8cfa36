// capi.cu — the C ABI of libcfdnn_b200.so (declared in include/cfdnn_b200.h).
#include "mg.cuh"
#include "comm.cuh"
#ifndef CFDNN_MG_ONLY
#include "solver.cuh"
#include "nn.cuh"
#endif
#include <mutex>
#include <cstring>

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
#ifndef CFDNN_MG_ONLY
struct cfdnn_nn { NNClosure impl; };
struct cfdnn_solver { Solver impl; cfdnn_solver(const cfdnn_mesh_desc& m, const cfdnn_config& c) : impl(m, c) {} };
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
    c->poisson_fixed_cycles = 8; c->poisson_max_vcycles = 20; c->poisson_tol_rhs = 1e-3;
    c->turb_guard_enabled = 1; c->turb_guard_interval = 5;
    c->turb_model = CFDNN_TURB_NONE; c->nu_t_max = 1.0;
}

// ---------------------------------------------------------------- multigrid
int cfdnn_mg_create(const cfdnn_mesh_desc* mesh, cfdnn_mg** out) {
    return guarded([&] {
        if (!mesh || !out) throw std::invalid_argument("cfdnn_mg_create: null argument");
        *out = new cfdnn_mg(*mesh);
    });
}
void cfdnn_mg_destroy(cfdnn_mg* mg) { delete mg; }
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

}  // extern "C"
