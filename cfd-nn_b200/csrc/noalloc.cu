// noalloc.cu — cfdnn_adaptive_dt_noalloc: compute_adaptive_dt on caller-owned device fields WITHOUT any allocation.
//
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  The arithmetic is the one of
// Solver::compute_adaptive_dt (reference src/solver.cpp:3369-3729) and runs the same two kernels
// (k_cfl_maxima + k_final_max_rows); what differs is where their scratch lives: the per-block partial maxima, the five
// final maxima and (y-stretched meshes) the device copy of dyv sit in ONE file-scope `static __device__` array of fixed
// size instead of the solver's cudaMalloc'ed buffers.
//
//   * every GPU that loads the library holds the array (CFDNN_NOALLOC_SCRATCH_BYTES = 80 224 B) until exit;
//   * the array's name is a host-side stand-in in host code: its device address is taken once per device, with that
//     device current, by cudaGetSymbolAddress and that pointer is what the kernels get;
//   * every call shares the array and it keeps the last call's contents: each call rewrites everything it reads
//     (partials [5][nb], the 5 maxima, dyv) before reading it, nothing has to start at zero;
//   * calls that overlap on the GPU (two streams, two host threads) RACE on the array — serialise them;
//   * no cudaMalloc / cudaFree / allocating library call on this path (cudaMemcpyAsync + launches only).
#include "solver.cuh"
#include "solver_kernels.cuh"
#include <algorithm>
#include <mutex>

namespace cfdnn {
namespace {

constexpr int NA_MAX_BLOCKS = 148 * 8;                          // CTAs of k_cfl_maxima = partials per quantity
constexpr int NA_PART = 5 * NA_MAX_BLOCKS;                      // [5][nb] partial maxima
constexpr int NA_OUT = NA_PART;                                 // 5 final maxima (8 slots)
constexpr int NA_DYV = NA_PART + 8;                             // dyv copy, CFDNN_NOALLOC_MAX_NY + 2 doubles
constexpr int NA_DOUBLES = NA_DYV + CFDNN_NOALLOC_MAX_NY + 2 + 2;   // padded to a multiple of 4 doubles
static_assert(NA_DOUBLES * sizeof(double) == CFDNN_NOALLOC_SCRATCH_BYTES, "header and scratch array disagree");
static_assert(NA_DYV % 4 == 0, "dyv slice must stay 32-byte aligned");

static __device__ __align__(256) double g_noalloc_scratch[NA_DOUBLES];

constexpr int NA_MAX_DEVICES = 64;
double* g_scratch_ptr[NA_MAX_DEVICES] = {};
std::mutex g_scratch_mu;

// device address of the scratch array on the CURRENT device (looked up once per device)
double* scratch_on_current_device() {
    int dev = 0;
    CFDNN_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NA_MAX_DEVICES) throw std::invalid_argument("adaptive_dt_noalloc: device ordinal out of range");
    std::lock_guard<std::mutex> lk(g_scratch_mu);
    if (!g_scratch_ptr[dev]) {
        double* p = nullptr;
        CFDNN_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        if (!p) throw Error(CFDNN_ERR_CUDA, "adaptive_dt_noalloc: cudaGetSymbolAddress returned a null address");
        g_scratch_ptr[dev] = p;
    }
    return g_scratch_ptr[dev];
}

double adaptive_dt_noalloc(const cfdnn_mesh_desc& m, const cfdnn_config& c, const double* u, const double* v, const double* w,
                           const double* nu_t, cudaStream_t st) {
    require_device();
    if (!u || !v || !w) throw std::invalid_argument("adaptive_dt_noalloc: null velocity pointer");
    if (m.Ng != 1) throw std::invalid_argument("adaptive_dt_noalloc: only Nghost == 1 is supported");
    if (m.Nx < 1 || m.Ny < 1 || m.Nz < 1) throw std::invalid_argument("adaptive_dt_noalloc: bad mesh size");
    if (m.Ny > CFDNN_NOALLOC_MAX_NY)
        throw std::invalid_argument("adaptive_dt_noalloc: Ny = " + std::to_string(m.Ny) + " exceeds CFDNN_NOALLOC_MAX_NY = " +
                                    std::to_string(CFDNN_NOALLOC_MAX_NY) + " (fixed static scratch; use "
                                    "cfdnn_solver_compute_adaptive_dt for larger meshes)");
    if (m.Nz_global > 0 && m.Nz_global != m.Nz)
        throw std::invalid_argument("adaptive_dt_noalloc: single-domain meshes only (no z-slab reduction on this path)");
    if (m.y_stretched && !m.dyv) throw std::invalid_argument("adaptive_dt_noalloc: stretched mesh needs dyv");
    const bool is2d = m.Nz == 1;
    cfdnn_mesh_desc mm = m;
    if (is2d && !(mm.dz > 0.0)) mm.dz = 1.0;
    double* scr = scratch_on_current_device();
    Geom g = make_geom(mm);
    double dy_min = mm.dy;
    if (m.y_stretched) {
        const int nv = m.Ny + 2 * m.Ng;
        CFDNN_CUDA(cudaMemcpyAsync(scr + NA_DYV, m.dyv, nv * sizeof(double), cudaMemcpyHostToDevice, st));
        g.dyv = scr + NA_DYV;
        dy_min = *std::min_element(m.dyv, m.dyv + nv);                                  // src/solver.cpp:3672-3675
    }
    Vel vel;
    vel.u = const_cast<double*>(u); vel.v = const_cast<double*>(v); vel.w = const_cast<double*>(w);
    const int nb = std::min(NA_MAX_BLOCKS, m.Ny * m.Nz);
    CFDNN_LAUNCH_COOP(k_cfl_maxima, nb, 256, 0, st, g, vel, nu_t, c.nu, nu_t ? 1 : 0, scr, nb);
    CFDNN_LAUNCH_COOP(k_final_max_rows, 1, 1024, 0, st, scr, nb, 5, scr + NA_OUT);
    double h[5];
    CFDNN_CUDA(cudaMemcpyAsync(h, scr + NA_OUT, sizeof(h), cudaMemcpyDeviceToHost, st));
    CFDNN_CUDA(cudaStreamSynchronize(st));
    check_launch("adaptive_dt_noalloc");
    // Solver::compute_adaptive_dt, same expressions in the same order (src/solver.cpp:3660-3729)
    const double u_abs_max = std::max(1e-10, h[0]);
    const double w_abs_max = std::max(1e-10, h[2]);
    const double v_dy_ratio_max = std::max(1e-10, h[3]);
    const double nu_eff_max = std::max(c.nu, h[4]);
    const double cfl_xz = c.CFL_xz > 0 ? c.CFL_xz : c.CFL_max;
    const double dt_cfl_x = cfl_xz * mm.dx / u_abs_max;
    const double dt_cfl_y = c.CFL_max / v_dy_ratio_max;
    const double dt_cfl_z = is2d ? 1e10 : cfl_xz * mm.dz / w_abs_max;
    const double dx_min_full = is2d ? std::min(mm.dx, dy_min) : std::min({mm.dx, dy_min, mm.dz});
    const double dx_min = c.implicit_y_diffusion ? (is2d ? mm.dx : std::min(mm.dx, mm.dz)) : dx_min_full;
    const double dt_diff = 0.25 * dx_min * dx_min / nu_eff_max;
    return c.dt_safety * std::min({dt_cfl_x, dt_cfl_y, dt_cfl_z, dt_diff});
}

}  // namespace
}  // namespace cfdnn

extern "C" int cfdnn_adaptive_dt_noalloc(const cfdnn_mesh_desc* mesh, const cfdnn_config* cfg, const double* u_dev,
                                         const double* v_dev, const double* w_dev, const double* nu_t_dev, void* stream,
                                         double* dt_out) {
    return cfdnn::guarded([&] {
        if (!mesh || !cfg || !dt_out) throw std::invalid_argument("cfdnn_adaptive_dt_noalloc: null argument");
        *dt_out = cfdnn::adaptive_dt_noalloc(*mesh, *cfg, u_dev, v_dev, w_dev, nu_t_dev, static_cast<cudaStream_t>(stream));
    });
}
