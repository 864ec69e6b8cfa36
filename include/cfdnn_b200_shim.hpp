// cfdnn_b200_shim.hpp — reference-side C++ binding of libcfdnn_b200.so.
//
// Header-only.  Compiled INSIDE the reference tree (it includes the reference's own headers) so
// that reference host code — apps, tests, bench_mg — can switch to the B200 backend by changing a
// type name: the classes below present the public signatures of
//   nncfd::MultigridPoissonSolver   (include/poisson_solver_multigrid.hpp:26-102)
//   nncfd::RANSSolver               (include/solver.hpp:86-311; hot-path subset)
//   nncfd::TurbulenceNNTBNN / NNMLP (include/turbulence_nn_tbnn.hpp:35-56, turbulence_nn_mlp.hpp) — derived from
//                                   nncfd::TurbulenceModel (include/turbulence_model.hpp:81-165)
// and forward to the C ABI of include/cfdnn_b200.h.  Errors come back as the exception types
// the reference throws (std::invalid_argument / std::runtime_error).
//
// Nothing here computes: there is no CPU path.  See INTEGRATION.md for the build line.
#pragma once
#include <algorithm>
#include <memory>
#include <cstdlib>
#include <cmath>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "cfdnn_b200.h"
#include "config.hpp"
#include "numerics.hpp"
#include "fields.hpp"
#include "mesh.hpp"
#include "nn_core.hpp"
#include "poisson_solver.hpp"
#include "solver.hpp"   // VelocityBC, RANSSolver::PoissonStats
#include "decomposition.hpp"
#include "turbulence_model.hpp"

namespace nncfd {
namespace b200 {

inline void check(int rc) {
    if (rc == CFDNN_OK) return;
    const std::string msg = cfdnn_last_error();
    if (rc == CFDNN_ERR_INVALID) throw std::invalid_argument(msg);
    throw std::runtime_error(msg);
}

inline cfdnn_mesh_desc mesh_desc(const Mesh& m) {
    cfdnn_mesh_desc d{};
    d.Nx = m.Nx; d.Ny = m.Ny; d.Nz = m.Nz; d.Ng = m.Nghost;
    d.dx = m.dx; d.dy = m.dy; d.dz = m.dz;
    d.y_stretched = m.is_y_stretched() ? 1 : 0;
    if (d.y_stretched) {
        d.dyv = m.dyv.data(); d.dyc = m.dyc.data();
        d.yLap_aS = m.yLap_aS.data(); d.yLap_aN = m.yLap_aN.data(); d.yLap_aP = m.yLap_aP.data();
    }
    d.yc = m.yc.data();
    d.yf = m.yf.data();
    d.y_lo = m.y_min; d.y_hi = m.y_max;
    d.Nz_global = m.Nz; d.k_offset = 0;
    return d;
}

inline cfdnn_poisson_cfg poisson_cfg(const PoissonConfig& c) {
    cfdnn_poisson_cfg o;
    cfdnn_poisson_cfg_default(&o);
    o.tol = c.tol; o.tol_abs = c.tol_abs; o.tol_rhs = c.tol_rhs; o.tol_rel = c.tol_rel;
    o.max_vcycles = c.max_vcycles; o.check_interval = c.check_interval;
    o.use_l2_norm = c.use_l2_norm ? 1 : 0; o.linf_safety_factor = c.linf_safety_factor;
    o.fixed_cycles = c.fixed_cycles; o.nu1 = c.nu1; o.nu2 = c.nu2; o.chebyshev_degree = c.chebyshev_degree;
    return o;
}

// ------------------------------------------------------------------ MultigridPoissonSolver
class MultigridPoissonSolverB200 {
public:
    explicit MultigridPoissonSolverB200(const Mesh& mesh) {
        const cfdnn_mesh_desc d = mesh_desc(mesh);
        check(cfdnn_mg_create(&d, &h_));
    }
    ~MultigridPoissonSolverB200() { cfdnn_mg_destroy(h_); }
    MultigridPoissonSolverB200(const MultigridPoissonSolverB200&) = delete;
    MultigridPoissonSolverB200& operator=(const MultigridPoissonSolverB200&) = delete;

    void set_bc(PoissonBC x_lo, PoissonBC x_hi, PoissonBC y_lo, PoissonBC y_hi) {
        set_bc(x_lo, x_hi, y_lo, y_hi, PoissonBC::Periodic, PoissonBC::Periodic);
    }
    void set_bc(PoissonBC x_lo, PoissonBC x_hi, PoissonBC y_lo, PoissonBC y_hi, PoissonBC z_lo, PoissonBC z_hi) {
        const int bc[6] = {(int)x_lo, (int)x_hi, (int)y_lo, (int)y_hi, (int)z_lo, (int)z_hi};
        check(cfdnn_mg_set_bc(h_, bc));
    }
    /// nabla^2 p = rhs; returns the number of V-cycles (reference :46)
    int solve(const ScalarField& rhs, ScalarField& p, const PoissonConfig& cfg) {
        const cfdnn_poisson_cfg c = poisson_cfg(cfg);
        check(cfdnn_mg_solve(h_, rhs.data().data(), p.data().data(), &c, &st_));
        return st_.cycles;
    }
    /// device twin (reference :59): here the arguments are DEVICE pointers in the ghosted cell layout
    int solve_device(double* rhs_dev, double* p_dev, const PoissonConfig& cfg) {
        const cfdnn_poisson_cfg c = poisson_cfg(cfg);
        check(cfdnn_mg_solve_dev(h_, rhs_dev, p_dev, &c, &st_));
        return st_.cycles;
    }
    double residual() const { return st_.residual; }
    double residual_l2() const { return st_.residual_l2; }
    double rhs_norm() const { return st_.b_inf; }
    double rhs_norm_l2() const { return st_.b_l2; }
    double initial_residual() const { return st_.r0; }
    double initial_residual_l2() const { return st_.r0_l2; }
    bool converged() const { return st_.converged != 0; }
    bool used_fixed_cycles() const { return st_.fixed_cycle_mode != 0; }
    void disable_vcycle_graph() {}
    cfdnn_mg* handle() const { return h_; }

private:
    cfdnn_mg* h_ = nullptr;
    cfdnn_poisson_stats st_{};
};

// ------------------------------------------------------------------ NN closures
/// Stands in for TurbulenceNNTBNN / TurbulenceNNMLP (include/turbulence_nn_tbnn.hpp:35-56,
/// include/turbulence_nn_mlp.hpp) and IS a nncfd::TurbulenceModel (include/turbulence_model.hpp:81-165): it can be
/// handed to the reference's own RANSSolver::set_turbulence_model(std::unique_ptr<TurbulenceModel>) or to
/// RANSSolverB200 below.  Weights come from the reference's own nncfd::MLP loader (src/nn_core.cpp:501-611) so
/// the on-disk format stays the reference's.
///   update(..., device_view == nullptr): host fields in, host fields out (copies inside the call);
///   update(..., device_view != nullptr): the view's pointers are DEVICE twins (RANSSolverB200::get_device_view),
///                                        nothing crosses PCIe; the host ScalarField arguments are not touched.
class TurbulenceNNB200 : public TurbulenceModel {
public:
    explicit TurbulenceNNB200(int kind /* CFDNN_TURB_NN_MLP | CFDNN_TURB_NN_TBNN */) : kind_(kind) {}
    ~TurbulenceNNB200() override { cfdnn_nn_destroy(h_); }
    TurbulenceNNB200(const TurbulenceNNB200&) = delete;
    TurbulenceNNB200& operator=(const TurbulenceNNB200&) = delete;

    void load(const std::string& weights_dir, const std::string& scaling_dir) {
        MLP mlp;
        mlp.load_weights(weights_dir);
        try {
            mlp.load_scaling(scaling_dir + "/input_means.txt", scaling_dir + "/input_stds.txt");
        } catch (...) { /* no scaling files: raw features, as the reference */ }
        const int nl = mlp.num_layers();
        std::vector<int> dims(nl + 1), act(nl);
        std::vector<const double*> W(nl), b(nl);
        for (int l = 0; l < nl; ++l) {
            const DenseLayer& L = mlp.layer(l);
            dims[l] = L.in_dim; dims[l + 1] = L.out_dim;
            W[l] = L.W.data(); b[l] = L.b.data();
            const Activation a = mlp.activations()[l];
            if (a == Activation::Linear) act[l] = 0;
            else if (a == Activation::Tanh) act[l] = 1;
            else if (a == Activation::ReLU) act[l] = 2;
            else throw std::invalid_argument("TurbulenceNNB200: activation not supported by the B200 backend");
        }
        cfdnn_nn_destroy(h_); h_ = nullptr;
        check(cfdnn_nn_create(kind_, nl, dims.data(), act.data(), W.data(), b.data(),
                              mlp.has_scaling() ? mlp.input_means_ptr() : nullptr,
                              mlp.has_scaling() ? mlp.input_stds_ptr() : nullptr, &h_));
        push_params();
    }
    // set_nu / set_delta are non-virtual members of the base (turbulence_model.hpp:153-158): the values are read
    // from the base at every update()
    void set_u_ref(double u) { u_ref_ = u; }
    void set_k_min(double k) { k_min_ = k; }
    void set_nu_t_max(double v) { nu_t_max_ = v; }
    std::string name() const override { return kind_ == CFDNN_TURB_NN_TBNN ? "NNTBNN" : "NNMLP"; }
    bool provides_reynolds_stresses() const override { return kind_ == CFDNN_TURB_NN_TBNN; }
    bool uses_transport_equations() const override { return false; }
    void initialize_gpu_buffers(const Mesh&) override {}       // weights are resident from load() on
    void cleanup_gpu_buffers() override {}
    bool is_gpu_ready() const override { return h_ != nullptr; }

    /// TurbulenceModel::update (include/turbulence_model.hpp:100-108)
    void update(const Mesh& mesh, const VectorField& velocity, const ScalarField& k, const ScalarField& omega,
                ScalarField& nu_t, TensorField* tau_ij = nullptr,
                const TurbulenceDeviceView* device_view = nullptr) override {
        if (!h_) throw std::runtime_error("TurbulenceNNB200::update: load() has not been called");
        push_params();
        const cfdnn_mesh_desc d = mesh_desc(mesh);
        if (device_view) {
            if (!device_view->u_face || !device_view->v_face || !device_view->nu_t)
                throw std::runtime_error("TurbulenceNNB200::update: device view without u/v/nu_t pointers");
            check(cfdnn_nn_update_dev(h_, &d, device_view->u_face, device_view->v_face, device_view->k,
                                      device_view->omega, device_view->nu_t, device_view->tau_xx,
                                      device_view->tau_xy, device_view->tau_yy));
            return;
        }
        check(cfdnn_nn_update(h_, &d, velocity.u_data().data(), velocity.v_data().data(), k.data().data(),
                              omega.data().data(), nu_t.data().data(),
                              tau_ij ? tau_ij->xx_data().data() : nullptr,
                              tau_ij ? tau_ij->xy_data().data() : nullptr,
                              tau_ij ? tau_ij->yy_data().data() : nullptr));
    }
    cfdnn_nn* handle() const { return h_; }
    /// parameters pushed to the device closure (called by update(); RANSSolverB200 calls it before every step)
    void push_params() {
        if (h_) check(cfdnn_nn_set_params(h_, nu_, delta_, u_ref_, k_min_, nu_t_max_));
    }

private:
    int kind_;
    cfdnn_nn* h_ = nullptr;
    double u_ref_ = 1.0, k_min_ = 1e-10, nu_t_max_ = 1.0;
};

// ------------------------------------------------------------------ RANSSolver (hot-path subset)
inline cfdnn_config solver_cfg(const Config& c) {
    cfdnn_config o;
    cfdnn_config_default(&o);
    o.nu = c.nu; o.dt = c.dt;
    o.time_integrator = c.time_integrator == TimeIntegrator::RK3 ? CFDNN_TI_RK3
                      : c.time_integrator == TimeIntegrator::RK2 ? CFDNN_TI_RK2 : CFDNN_TI_EULER;
    switch (c.convective_scheme) {
        case ConvectiveScheme::Central: o.convective_scheme = CFDNN_CONV_CENTRAL; break;
        case ConvectiveScheme::Skew: o.convective_scheme = CFDNN_CONV_SKEW; break;
        default: throw std::invalid_argument("B200 backend: convective_scheme must be central or skew");
    }
    o.adaptive_dt = c.adaptive_dt ? 1 : 0;
    o.CFL_max = c.CFL_max; o.CFL_xz = c.CFL_xz; o.dt_safety = c.dt_safety;
    o.poisson_fixed_cycles = c.poisson_fixed_cycles; o.poisson_max_vcycles = c.poisson_max_vcycles;
    o.poisson_tol_rhs = c.poisson_tol_rhs; o.poisson_tol = c.poisson_tol;
    o.poisson_abs_tol_floor = c.poisson_abs_tol_floor; o.poisson_tol_abs = c.poisson_tol_abs;
    o.poisson_tol_rel = c.poisson_tol_rel; o.poisson_check_interval = c.poisson_check_interval;
    o.poisson_use_l2_norm = c.poisson_use_l2_norm ? 1 : 0; o.poisson_linf_safety = c.poisson_linf_safety;
    o.poisson_nu1 = c.poisson_nu1; o.poisson_nu2 = c.poisson_nu2;
    o.poisson_chebyshev_degree = c.poisson_chebyshev_degree;
    o.turb_guard_enabled = c.turb_guard_enabled ? 1 : 0; o.turb_guard_interval = c.turb_guard_interval;
    o.turb_model = c.turb_model == TurbulenceModelType::NNTBNN ? CFDNN_TURB_NN_TBNN
                 : c.turb_model == TurbulenceModelType::NNMLP ? CFDNN_TURB_NN_MLP : CFDNN_TURB_NONE;
    o.nu_t_max = c.nu_t_max;
    o.implicit_y_diffusion = c.implicit_y_diffusion ? 1 : 0;
    // Config::poisson_solver (include/config.hpp:151): an explicit FFT request maps to this backend's FFT solver; Auto
    // stays on the multigrid (what the reference's CPU build, the parity oracle, selects)
    o.poisson_solver = c.poisson_solver == PoissonSolverType::FFT ? CFDNN_POISSON_FFT : CFDNN_POISSON_MG;
    return o;
}

inline int vbc_code(VelocityBC::Type t) {
    switch (t) {
        case VelocityBC::NoSlip: return CFDNN_VBC_NOSLIP;
        case VelocityBC::Periodic: return CFDNN_VBC_PERIODIC;
        case VelocityBC::Inflow: return CFDNN_VBC_INFLOW;
        case VelocityBC::Outflow: return CFDNN_VBC_OUTFLOW;
    }
    return CFDNN_VBC_NOSLIP;
}

/// Same construction / stepping / accessor surface as nncfd::RANSSolver for the fractional-step
/// hot path.  Host fields (std::vector<double>, reference layouts) are authoritative at
/// initialize()/sync_to_gpu() and refreshed by sync_from_gpu()/sync_solution_from_gpu(); between
/// those calls the fields live in HBM (the reference's "Model 1" contract, turbulence_model.hpp:24-26).
class RANSSolverB200 {
public:
    using PoissonStats = RANSSolver::PoissonStats;
    using PoissonSolveStatus = RANSSolver::PoissonSolveStatus;

    RANSSolverB200(const Mesh& mesh, const Config& config)
        : mesh_(&mesh), config_(config), velocity_(mesh), pressure_(mesh), pressure_correction_(mesh),
          div_velocity_(mesh), rhs_poisson_(mesh), nu_t_(mesh), k_(mesh), omega_(mesh), nu_eff_(mesh), tau_ij_(mesh) {
        const cfdnn_mesh_desc d = mesh_desc(mesh);
        const cfdnn_config c = solver_cfg(config);
        check(cfdnn_solver_create(&d, &c, &h_));
        current_dt_ = config.dt;
        nu_ = config.nu;
        nu_t_max_ = config.nu_t_max;
    }
    ~RANSSolverB200() { cfdnn_solver_destroy(h_); }
    RANSSolverB200(const RANSSolverB200&) = delete;
    RANSSolverB200& operator=(const RANSSolverB200&) = delete;

    void set_velocity_bc(const VelocityBC& bc) {
        const int v[6] = {vbc_code(bc.x_lo), vbc_code(bc.x_hi), vbc_code(bc.y_lo), vbc_code(bc.y_hi),
                          vbc_code(bc.z_lo), vbc_code(bc.z_hi)};
        check(cfdnn_solver_set_velocity_bc(h_, v));
        bc_ = bc; have_bc_ = true;
    }
    void set_body_force(double fx, double fy, double fz = 0.0) {
        fx_ = fx; fy_ = fy; fz_ = fz;
        check(cfdnn_solver_set_body_force(h_, fx, fy, fz));
    }
    /// set_turbulence_model(std::unique_ptr<TurbulenceModel>) (include/solver.hpp:107).  A TurbulenceNNB200 runs
    /// device-resident inside step(); so do the reference's own static LES models (SmagorinskyModel, WALEModel,
    /// VremanModel, SigmaModel with their default constants: recognised by name(), evaluated by k_les_sgs).  Any other
    /// reference TurbulenceModel (e.g. SST k-omega, which has no B200 kernels yet) is BRIDGED: every step() downloads the velocity, calls the model's own host update() /
    /// advance_turbulence() exactly where the reference does (src/solver.cpp:1284-1324), and uploads nu_t — correct,
    /// but PCIe-bound; it exists so that such a run works at all, not as a fast path.
    void set_turbulence_model(std::unique_ptr<TurbulenceModel> model) {
        turb_ = std::move(model);
        native_ = dynamic_cast<TurbulenceNNB200*>(turb_.get());
        les_ = CFDNN_LES_NONE;
        if (turb_ && !native_) {       // the reference's own static LES models run as device kernels (default constants)
            const std::string n = turb_->name();
            // DynamicSmagorinsky: the reference's CPU path (the parity target) evaluates it as static Smagorinsky with
            // Cs = 0.17 — the Germano procedure exists in its GPU build only (src/turbulence_les_dynamic.cpp:56-67)
            les_ = (n == "Smagorinsky" || n == "DynamicSmagorinsky") ? CFDNN_LES_SMAGORINSKY : n == "WALE" ? CFDNN_LES_WALE
                 : n == "Vreman" ? CFDNN_LES_VREMAN : n == "Sigma" ? CFDNN_LES_SIGMA : CFDNN_LES_NONE;
        }
        // the reference's SSTKOmegaTransport (default SST closure) on a 2-D mesh: native wavefront sweep + closure
        // ... its standard KOmegaTransport (default Boussinesq closure) and the two algebraic baselines
        sst_ = CFDNN_RANS_NONE;
        if (turb_ && !native_ && mesh_->is2D())
            sst_ = turb_->name() == "SSTKOmega" ? CFDNN_RANS_SST : turb_->name() == "KOmega" ? CFDNN_RANS_KOMEGA
                 : turb_->name() == "MixingLength" ? CFDNN_RANS_MIXING_LENGTH       // kappa, A+ at their defaults (no getters)
                 : turb_->name() == "AlgebraicKOmega" ? CFDNN_RANS_ALGEBRAIC_KOMEGA
                 : turb_->name() == "SST+WJ-EARSM" ? CFDNN_RANS_EARSM_WJ : turb_->name() == "SST+GS-EARSM" ? CFDNN_RANS_EARSM_GS
                 : turb_->name() == "SST+Pope-Quadratic" ? CFDNN_RANS_EARSM_POPE : CFDNN_RANS_NONE;
        // TurbulenceGEP keeps its variant and its own nu private (no getters): it stays a bridged host model; the device
        // kernel is available through cfdnn_solver_set_sst_model(CFDNN_RANS_GEP) + cfdnn_solver_set_gep_params
        if (turb_) turb_->set_nu(nu_);
        if (native_) native_->set_nu_t_max(nu_t_max_);
        attach_model();
    }
    /// set_decomposition (include/solver.hpp:93): this rank's mesh is the z-slab [k_global_start, +nz_local) of
    /// nz_global planes; the NCCL communicator replaces halo_exchange.cpp.  `nccl_id` is the 128-byte
    /// ncclUniqueId made by rank 0 with cfdnn_comm_unique_id() and broadcast by the caller (MPI_Bcast in an MPI
    /// build).  Call before initialize(), like the reference.
    void set_decomposition(Decomposition* decomp, const void* nccl_id = nullptr) {
        decomp_ = decomp;
        if (!decomp || !decomp->is_parallel()) return;
        if (!nccl_id) throw std::invalid_argument("RANSSolverB200::set_decomposition: parallel run needs the ncclUniqueId");
        cfdnn_mesh_desc d = mesh_desc(*mesh_);
        d.Nz_global = decomp->nz_global();
        d.k_offset = decomp->k_global_start();
        const cfdnn_config c = solver_cfg(config_);
        cfdnn_solver_destroy(h_); h_ = nullptr;
        check(cfdnn_solver_create(&d, &c, &h_));
        if (have_bc_) set_velocity_bc(bc_);
        check(cfdnn_solver_set_body_force(h_, fx_, fy_, fz_));
        if (turb_) attach_model();
        check(cfdnn_solver_comm_init(h_, decomp->rank(), decomp->nprocs(), nccl_id));
    }
    Decomposition* decomposition() const { return decomp_; }
    void initialize(const VectorField& initial_velocity) {
        velocity_ = initial_velocity;
        sync_to_gpu();
        check(cfdnn_solver_initialize(h_));
        if (turb_ && turb_->uses_transport_equations()) {        // k / omega seeding, src/solver.cpp:1064-1103
            down(CFDNN_F_U, velocity_.u_data()); down(CFDNN_F_V, velocity_.v_data());    // with the BCs applied
            double u_max = 0.0;
            for (int j = mesh_->j_begin(); j < mesh_->j_end(); ++j)
                for (int i = mesh_->i_begin(); i < mesh_->i_end(); ++i) {
                    const double uc = 0.5 * (velocity_.u(i, j) + velocity_.u(i + 1, j));
                    const double vc = 0.5 * (velocity_.v(i, j) + velocity_.v(i, j + 1));
                    u_max = std::max(u_max, std::sqrt(uc * uc + vc * vc));
                }
            seed_k_omega(std::max(u_max, 0.01), /*guard_wall=*/true);
        }
        if (turb_) turb_->initialize(*mesh_, velocity_);
        initialize_trip_forcing();
    }
    /// the uniform k, omega of RANSSolver::initialize / initialize_uniform with the near-wall omega rows
    void seed_k_omega(double u_ref, bool guard_wall) {
        const double Ti = 0.05;
        double k_init = 1.5 * (u_ref * Ti) * (u_ref * Ti);
        k_init = std::max(k_init, 1e-7);
        double omega_init = k_init / (numerics::C_MU * config_.nu * 100.0);
        omega_init = std::max(omega_init, 1e-6);
        k_.fill(k_init);
        omega_.fill(omega_init);
        for (int i = mesh_->i_begin(); i < mesh_->i_end(); ++i) {
            const int jb = mesh_->j_begin(), jt = mesh_->j_end() - 1;
            const double yb = mesh_->wall_distance(i, jb), yt = mesh_->wall_distance(i, jt);
            if (!guard_wall || yb > 1e-10) omega_(i, jb) = 10.0 * 6.0 * config_.nu / (0.075 * yb * yb);
            if (!guard_wall || yt > 1e-10) omega_(i, jt) = 10.0 * 6.0 * config_.nu / (0.075 * yt * yt);
        }
        up(CFDNN_F_K, k_.data()); up(CFDNN_F_OMEGA, omega_.data());
    }
    /// initialize_trip_forcing (src/solver.cpp:818-923): the envelope / profile / spanwise-mode tables, built here with the
    /// reference's expressions (std::srand(12345) phases) from its Mesh and Config and handed to the backend, which adds
    /// the force to v* and w* in the Euler predictor while current_time() < trip_duration (:1724-1810).
    void initialize_trip_forcing() {
        if (!config_.trip_enabled) { check(cfdnn_solver_set_trip_forcing(h_, nullptr)); return; }
        const Mesh& m = *mesh_;
        const double delta = 0.5 * (m.y_max - m.y_min);
        const double u_tau = std::sqrt(delta * std::abs(fx_));
        const double Lx = m.x_max - m.x_min, Lz = m.z_max - m.z_min;
        const double x0 = config_.trip_x_start < 0 ? m.x_min + 0.1 * Lx : config_.trip_x_start;
        const double x1 = config_.trip_x_end < 0 ? m.x_min + 0.2 * Lx : config_.trip_x_end;
        const double width = x1 - x0;
        const int nm = config_.trip_n_modes_z;
        std::srand(12345);
        std::vector<double> ph(nm);
        for (int q = 0; q < nm; ++q) ph[q] = 2.0 * M_PI * static_cast<double>(std::rand()) / RAND_MAX;
        std::vector<double> env(m.total_Nx(), 0.0), gy(m.total_Ny() + 1, 0.0), fzv, fzw;
        for (int i = 0; i < m.total_Nx(); ++i) {
            const double x = m.xc[i];
            if (x >= x0 && x <= x1) env[i] = 0.5 * (1.0 - std::cos(2.0 * M_PI * ((x - x0) / width)));
        }
        for (int j = 0; j < static_cast<int>(gy.size()) && j < static_cast<int>(m.yf.size()); ++j)
            gy[j] = m.yf[j] * (1.0 - m.yf[j] * m.yf[j]);
        const bool is3d = !m.is2D();
        if (is3d) {
            double wsum = 0.0;
            for (int q = 0; q < nm; ++q) wsum += 1.0 / (q + 1);
            fzv.assign(m.total_Nz(), 0.0);
            fzw.assign(m.total_Nz() + 1, 0.0);
            for (int k = 0; k < m.total_Nz(); ++k) {
                double sum = 0.0;
                for (int q = 0; q < nm; ++q)
                    sum += ((1.0 / (q + 1)) / wsum) * std::sin(2.0 * M_PI * (q + 1) / Lz * m.zc[k] + ph[q]);
                fzv[k] = sum;
            }
            for (int k = 0; k < static_cast<int>(fzw.size()) && k < static_cast<int>(m.zf.size()); ++k) {
                double sum = 0.0;
                for (int q = 0; q < nm; ++q)
                    sum += ((1.0 / (q + 1)) / wsum) * std::cos(2.0 * M_PI * (q + 1) / Lz * m.zf[k] + ph[q]);
                fzw[k] = sum;
            }
        }
        cfdnn_trip_desc d{};
        d.amplitude = config_.trip_amplitude * u_tau * u_tau;
        d.duration = config_.trip_duration;
        d.ramp_off_start = config_.trip_ramp_off_start;
        d.w_scale = config_.trip_w_scale;
        d.env_x = env.data(); d.g_y = gy.data();
        d.Fz_v = is3d ? fzv.data() : nullptr;
        d.Fz_w = (is3d && config_.trip_force_w) ? fzw.data() : nullptr;
        check(cfdnn_solver_set_trip_forcing(h_, &d));
    }
    /// apply_velocity_filter (include/solver.hpp:260, src/solver.cpp:4065-4162): explicit Laplacian filter between steps
    void apply_velocity_filter(double strength) { check(cfdnn_solver_apply_velocity_filter(h_, strength)); }
    double current_time() const { double t = 0.0; cfdnn_solver_get_time(h_, &t); return t; }
    /// initialize_uniform (src/solver.cpp:1132-1167; the k / omega seeding of transport models is the model's own
    /// initialize() here)
    void initialize_uniform(double u0, double v0) {
        velocity_.fill(u0, v0);
        sync_to_gpu();
        check(cfdnn_solver_initialize(h_));
        if (turb_ && turb_->uses_transport_equations())          // src/solver.cpp:1136-1160
            seed_k_omega(std::max(std::abs(u0), 0.01), /*guard_wall=*/false);
        if (turb_) turb_->initialize(*mesh_, velocity_);
    }
    void sync_to_gpu() {
        up(CFDNN_F_U, velocity_.u_data()); up(CFDNN_F_V, velocity_.v_data());
        if (!mesh_->is2D()) up(CFDNN_F_W, velocity_.w_data());
        up(CFDNN_F_P, pressure_.data()); up(CFDNN_F_NUT, nu_t_.data());
        up(CFDNN_F_K, k_.data()); up(CFDNN_F_OMEGA, omega_.data());
    }
    void sync_solution_from_gpu() {
        down(CFDNN_F_U, velocity_.u_data()); down(CFDNN_F_V, velocity_.v_data());
        if (!mesh_->is2D()) down(CFDNN_F_W, velocity_.w_data());
        down(CFDNN_F_P, pressure_.data()); down(CFDNN_F_NUT, nu_t_.data());
    }
    void sync_from_gpu() {
        sync_solution_from_gpu();
        down(CFDNN_F_PCORR, pressure_correction_.data()); down(CFDNN_F_DIV, div_velocity_.data());
        down(CFDNN_F_RHS, rhs_poisson_.data()); down(CFDNN_F_NUEFF, nu_eff_.data());
        down(CFDNN_F_K, k_.data()); down(CFDNN_F_OMEGA, omega_.data());
    }
    /// one time step; returns max|u^{n+1} - u^n| like the reference (:208)
    double step() {
        double r = 0.0;
        if (native_) native_->push_params();
        if (bridged()) bridge_host_model();
        check(cfdnn_solver_step(h_, &r));
        refresh_stats();
        return r;
    }
    /// n steps without host round trips (app time loops)
    double advance(int nsteps) {
        double r = 0.0;
        if (bridged()) {
            for (int s = 0; s < nsteps; ++s) r = step();
            return r;
        }
        if (native_) native_->push_params();
        check(cfdnn_solver_run(h_, nsteps, &r));
        refresh_stats();
        return r;
    }
    /// get_solver_view / get_device_view (include/solver.hpp:1118-1122): DEVICE twins of the solver's arrays in the
    /// reference layouts.  Valid until the next step() (the RK stages rotate the velocity buffers).  On 2-D meshes the
    /// pointers address the data plane, as the reference's k = 0 plane does.
    SolverDeviceView get_solver_view() const {
        SolverDeviceView v;
        v.u_face = dev(CFDNN_F_U); v.v_face = dev(CFDNN_F_V);
        v.u_star_face = dev(CFDNN_F_U_STAR); v.v_star_face = dev(CFDNN_F_V_STAR);
        v.u_old_face = dev(CFDNN_F_U_OLD); v.v_old_face = dev(CFDNN_F_V_OLD);
        v.u_stride = velocity_.u_stride(); v.v_stride = velocity_.v_stride();
        if (!mesh_->is2D()) {
            v.w_face = dev(CFDNN_F_W); v.w_star_face = dev(CFDNN_F_W_STAR); v.w_old_face = dev(CFDNN_F_W_OLD);
            v.w_stride = velocity_.w_stride();
            v.u_plane_stride = velocity_.u_plane_stride(); v.v_plane_stride = velocity_.v_plane_stride();
            v.w_plane_stride = velocity_.w_plane_stride();
            v.conv_w = dev(CFDNN_F_CONV_W); v.diff_w = dev(CFDNN_F_DIFF_W);
        }
        v.p = dev(CFDNN_F_P); v.p_corr = dev(CFDNN_F_PCORR); v.nu_t = dev(CFDNN_F_NUT); v.nu_eff = dev(CFDNN_F_NUEFF);
        v.rhs = dev(CFDNN_F_RHS); v.div = dev(CFDNN_F_DIV);
        v.cell_stride = mesh_->total_Nx();
        v.cell_plane_stride = mesh_->total_Nx() * mesh_->total_Ny();
        v.conv_u = dev(CFDNN_F_CONV_U); v.conv_v = dev(CFDNN_F_CONV_V);
        v.diff_u = dev(CFDNN_F_DIFF_U); v.diff_v = dev(CFDNN_F_DIFF_V);
        v.Nx = mesh_->Nx; v.Ny = mesh_->Ny; v.Nz = mesh_->Nz; v.Ng = mesh_->Nghost;
        v.dx = mesh_->dx; v.dy = mesh_->dy; v.dz = mesh_->dz; v.dt = current_dt_;
        v.y_stretched = mesh_->is_y_stretched();
        return v;
    }
    TurbulenceDeviceView get_device_view() const {
        TurbulenceDeviceView v;
        v.u_face = dev(CFDNN_F_U); v.v_face = dev(CFDNN_F_V);
        v.u_stride = velocity_.u_stride(); v.v_stride = velocity_.v_stride();
        if (!mesh_->is2D()) {
            v.w_face = dev(CFDNN_F_W); v.w_stride = velocity_.w_stride();
            v.u_plane_stride = velocity_.u_plane_stride(); v.v_plane_stride = velocity_.v_plane_stride();
            v.w_plane_stride = velocity_.w_plane_stride();
        }
        v.k = dev(CFDNN_F_K); v.omega = dev(CFDNN_F_OMEGA); v.nu_t = dev(CFDNN_F_NUT);
        v.cell_stride = mesh_->total_Nx();
        v.cell_plane_stride = mesh_->total_Nx() * mesh_->total_Ny();
        v.tau_xx = dev(CFDNN_F_TAU_XX); v.tau_xy = dev(CFDNN_F_TAU_XY); v.tau_yy = dev(CFDNN_F_TAU_YY);
        v.is_y_stretched = mesh_->is_y_stretched();
        v.Nx = mesh_->Nx; v.Ny = mesh_->Ny; v.Nz = mesh_->Nz; v.Ng = mesh_->Nghost;
        v.dx = mesh_->dx; v.dy = mesh_->dy; v.dz = mesh_->dz;
        return v;
    }
    double compute_adaptive_dt() const {
        double dt = 0.0;
        check(cfdnn_solver_compute_adaptive_dt(h_, &dt));
        return dt;
    }
    double current_dt() const { return current_dt_; }
    void set_dt(double dt) { current_dt_ = dt; check(cfdnn_solver_set_dt(h_, dt)); }

    // accessors: host mirrors (call sync_*_from_gpu() first, as with the reference's GPU build)
    const VectorField& velocity() const { return velocity_; }
    VectorField& velocity() { return velocity_; }
    const ScalarField& pressure() const { return pressure_; }
    ScalarField& pressure() { return pressure_; }
    const ScalarField& nu_t() const { return nu_t_; }
    const ScalarField& k() const { return k_; }
    const ScalarField& omega() const { return omega_; }
    const ScalarField& nu_eff() const { return nu_eff_; }
    const ScalarField& div_velocity() const { return div_velocity_; }
    const ScalarField& rhs_poisson() const { return rhs_poisson_; }
    const ScalarField& pressure_correction() const { return pressure_correction_; }
    const PoissonStats& poisson_stats() const { return poisson_stats_; }
    cfdnn_solver* handle() const { return h_; }

private:
    void up(int f, const std::vector<double>& v) { check(cfdnn_solver_upload(h_, f, v.data(), v.size())); }
    void down(int f, std::vector<double>& v) { check(cfdnn_solver_download(h_, f, v.data(), v.size())); }
    bool bridged() const { return turb_ && !native_ && les_ == CFDNN_LES_NONE && !sst_; }
    void attach_model() {
        check(cfdnn_solver_set_turbulence_model(h_, native_ ? native_->handle() : nullptr));
        check(cfdnn_solver_set_les_model(h_, les_, 0.0));
        check(cfdnn_solver_set_sst_model(h_, sst_));
        if (sst_ == CFDNN_RANS_MIXING_LENGTH) check(cfdnn_solver_set_mixing_length_params(h_, 0.41, 26.0, turb_->delta()));
        // the Pope constants as the reference's factory sets them (create_turbulence_model, src/turbulence_baseline.cpp:542)
        if (sst_ == CFDNN_RANS_EARSM_POPE) check(cfdnn_solver_set_pope_constants(h_, config_.pope_C1, config_.pope_C2));
        check(cfdnn_solver_set_external_nu_t(h_, bridged() ? (turb_->uses_transport_equations() ? 2 : 1) : 0));
    }
    double* dev(int f) const {
        double* p = nullptr;
        size_t n = 0;
        check(cfdnn_solver_device_ptr(h_, f, &p, &n));
        if (p && mesh_->is2D()) {   // 2-D: the device keeps the data in plane Ng of its 3-plane arrays
            const size_t planes = (f == CFDNN_F_W || f == CFDNN_F_W_STAR || f == CFDNN_F_W_OLD || f == CFDNN_F_CONV_W ||
                                   f == CFDNN_F_DIFF_W) ? 4 : 3;
            p += (n / planes) * (size_t)mesh_->Nghost;
        }
        return p;
    }
    /// a host-side reference TurbulenceModel inside step(): the reference's call site src/solver.cpp:1284-1324
    void bridge_host_model() {
        down(CFDNN_F_U, velocity_.u_data()); down(CFDNN_F_V, velocity_.v_data());
        if (!mesh_->is2D()) down(CFDNN_F_W, velocity_.w_data());
        if (turb_->uses_transport_equations())
            turb_->advance_turbulence(*mesh_, velocity_, current_dt_, k_, omega_, nu_t_, nullptr);
        turb_->update(*mesh_, velocity_, k_, omega_, nu_t_, turb_->provides_reynolds_stresses() ? &tau_ij_ : nullptr, nullptr);
        up(CFDNN_F_NUT, nu_t_.data());
    }
    void refresh_stats() {
        cfdnn_poisson_stats s{};
        check(cfdnn_solver_poisson_stats(h_, &s));
        poisson_stats_.cycles = s.cycles;
        poisson_stats_.status = s.status == CFDNN_PSTAT_FIXED ? PoissonSolveStatus::FixedCycles
                              : s.status == CFDNN_PSTAT_HIT_MAX ? PoissonSolveStatus::HitMaxCycles
                                                                : PoissonSolveStatus::ConvergedToTol;
        poisson_stats_.rhs_norm_l2 = s.b_l2; poisson_stats_.rhs_norm_inf = s.b_inf;
        poisson_stats_.res_norm_l2 = s.residual_l2; poisson_stats_.res_norm_inf = s.residual;
        poisson_stats_.res_over_rhs = s.b_l2 > 0.0 ? s.residual_l2 / s.b_l2 : 0.0;
        cfdnn_solver_get_dt(h_, &current_dt_);
    }
    const Mesh* mesh_;
    Config config_;
    cfdnn_solver* h_ = nullptr;
    std::unique_ptr<TurbulenceModel> turb_;
    TurbulenceNNB200* native_ = nullptr;        // turb_ when it is one of this backend's device closures
    int les_ = CFDNN_LES_NONE;                  // turb_ is one of the reference's static LES models: device kernel
    int sst_ = CFDNN_RANS_NONE;                 // turb_ is the reference's SST / standard k-omega on a 2-D mesh: device kernels
    Decomposition* decomp_ = nullptr;
    VelocityBC bc_;
    bool have_bc_ = false;
    double fx_ = 0.0, fy_ = 0.0, fz_ = 0.0;
    VectorField velocity_;
    ScalarField pressure_, pressure_correction_, div_velocity_, rhs_poisson_, nu_t_, k_, omega_, nu_eff_;
    TensorField tau_ij_;
    PoissonStats poisson_stats_;
    double current_dt_ = 0.0, nu_ = 0.0, nu_t_max_ = 1.0;
};

}  // namespace b200
}  // namespace nncfd
