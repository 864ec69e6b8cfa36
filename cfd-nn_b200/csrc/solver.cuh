// solver.cuh — device-resident fractional-step solver on the staggered MAC grid.
//
// Stands in for RANSSolver (reference include/solver.hpp:86-311): step() with the Euler,
// SSP-RK2 and SSP-RK3 integrators (src/solver.cpp:1200-2786, src/solver_time.cpp:269-1513),
// velocity BCs, convection (central / skew), variable-viscosity diffusion, divergence,
// Poisson RHS, multigrid projection, velocity correction, CFL / residual / NaN reductions.
//
// HBM layout: every solver field keeps the reference's ghost-inclusive layout
// (include/fields.hpp:118-136), i fastest, so upload/download are flat copies and the
// stale-ghost semantics of the RK stages are reproduced by construction:
//   cell : (Nx+2Ng)(Ny+2Ng)(Nz+2Ng)         u : row length Nx+1+2Ng
//   v    : Ny+1+2Ng rows                      w : Nz+1+2Ng planes
// The multigrid levels use their own ghost-free layout (mg.cuh).
#pragma once
#include "common.cuh"
#include "mg.cuh"
#include "fft.cuh"
#include <vector>
#include <memory>

namespace cfdnn {

struct Comm;
class NNClosure;
struct LesMetrics;

struct Geom {                 // by-value kernel argument
    int Nx, Ny, Nz, Ng;
    int us, vs, ws, cs;       // row strides of u, v, w, cell arrays
    long long up, vp, wp, cp; // plane strides
    Div dx, dy, dz;           // spacings
    Div dx_2, dy_2, dz_2;     // 2*dx ...   (central differences)
    Div dx2, dy2, dz2;        // dx*dx ...  (diffusion)
    int y_stretched;
    const double* dyv;        // Ny+2Ng   (divergence, stretched y)
    const double* dyc;        // Ny+2Ng+1 (pressure gradient, stretched y)
    const double* rdyv;       // reciprocals of the above (for the corrected division)
    const double* rdyc;
};

struct Vel { double *u = nullptr, *v = nullptr, *w = nullptr; };

// strides / spacings of a mesh in the reference layout (y-metric pointers are left null)
Geom make_geom(const cfdnn_mesh_desc& m);

class Solver {
public:
    Solver(const cfdnn_mesh_desc& m, const cfdnn_config& c);
    ~Solver();

    void set_velocity_bc(const int bc[6]);
    void set_body_force(double fx, double fy, double fz) { fx_ = fx; fy_ = fy; fz_ = fz; }
    void set_dt(double dt) { dt_ = dt; }
    double dt() const { return dt_; }
    void set_turbulence_model(NNClosure* nn) { nn_ = nn; }
    void set_external_nu_t(bool on, bool transport = false) { ext_nut_ = on; ext_transport_ = on && transport; }   // nu_t uploaded by the caller (bridged host model)
    bool fft_peer_to_peer() const { return fft_ && fft_active_ && fft_->peer_to_peer(); }
    void set_mixing_length_params(double kappa, double A_plus, double delta);
    void set_gep_params(int variant, double model_nu);
    void set_pope_constants(double C1, double C2);
    void set_sst_model(int kind);                            // native SST / standard k-omega transport + closure (2-D meshes)
    void set_trip_forcing(const cfdnn_trip_desc* d);         // trip-region forcing in the Euler predictor
    void apply_velocity_filter(double strength);             // explicit Laplacian filter between steps
    double time() const { return time_; }
    void set_time(double t) { time_ = t; }
    void set_les_model(int model, double constant);          // static LES SGS closure evaluated at the top of step()
    bool has_model() const { return nn_ != nullptr || ext_nut_ || les_model_ != 0 || sst_ != nullptr; }
    bool transport_model() const {               // uses_transport_equations(): SST / standard k-omega, not the algebraic models
        return ext_transport_ || (sst_ != nullptr && (sst_kind_ == CFDNN_RANS_SST || sst_kind_ == CFDNN_RANS_KOMEGA ||
                                                      sst_kind_ >= CFDNN_RANS_EARSM_WJ));
    }
    double* view_ptr(int field, size_t* n) const;           // device twin of a field (cfdnn_solver_device_ptr)
    void comm_init(Comm* c);

    size_t field_size(int field) const;
    size_t plane_elems(int field) const;
    void upload(int field, const double* host, size_t n);
    void download(int field, double* host, size_t n);
    void initialize();                    // apply_velocity_bc on the uploaded velocity (:1060-1062)

    double step();                        // double RANSSolver::step()
    void run(int nsteps, double* last_residual);
    double compute_adaptive_dt();

    cfdnn_poisson_stats poisson_stats() const { return mg_->stats; }
    std::vector<int> cycle_log;           // V-cycles of every Poisson solve since last read
    double last_step_ms = 0.0, last_vcycle_ms = 0.0;
    int last_vcycles = 0;
    double last_phase_ms[CFDNN_PHASE_COUNT] = {};   // device ms per phase of the last run() (cfdnn_solver_phase_timing)
    Multigrid& mg() { return *mg_; }
    cudaStream_t stream() const { return st_; }

private:
    cfdnn_mesh_desc mesh_;
    cfdnn_config cfg_;
    Geom g_{};
    size_t su_ = 0, sv_ = 0, sw_ = 0, sc_ = 0;
    int vbc_[6] = {1, 1, 0, 0, 1, 1};     // VelocityBC defaults (include/solver.hpp:63-70)
    double fx_ = 0, fy_ = 0, fz_ = 0, dt_ = 0, time_ = 0;
    int step_count_ = 0;
    bool is2d_ = false;

    std::vector<DevBuf<double>> store_;
    Vel vel_, star_, old_, rk_, conv_, diff_;
    double *p_ = nullptr, *pcorr_ = nullptr, *div_ = nullptr, *rhs_ = nullptr, *nut_ = nullptr,
           *nueff_ = nullptr, *k_ = nullptr, *omega_ = nullptr, *tau_[3] = {nullptr, nullptr, nullptr};
    DevBuf<double> dyv_, dyc_, rdyv_, rdyc_, yc_;
    double dy_min_ = 0.0;
    std::vector<double> h_dyv_, h_dyc_;
    DevBuf<double> partial_, scalars_;
    double* h_scalars_ = nullptr;          // pinned
    size_t npartial_ = 0;

    std::unique_ptr<Multigrid> mg_;
    std::unique_ptr<class SstModel> sst_;
    int sst_kind_ = CFDNN_RANS_NONE;
    // trip forcing (src/solver.cpp:818-950): device copies of the caller's tables
    bool trip_on_ = false, trip_w_ = false;
    double trip_A_ = 0, trip_duration_ = 0, trip_ramp_start_ = 0, trip_w_scale_ = 1;
    DevBuf<double> trip_env_x_, trip_g_y_, trip_fz_v_, trip_fz_w_;
    void trip_force(Vel& star, double dt);
    std::unique_ptr<FftPoisson> fft_;        // cfg.poisson_solver == CFDNN_POISSON_FFT (periodic x/z, walls y)
    bool fft_active_ = false;                // the current velocity BCs allow it (else the multigrid takes over)
    NNClosure* nn_ = nullptr;
    bool ext_nut_ = false, ext_transport_ = false;
    int les_model_ = 0;                    // CFDNN_LES_*
    double les_c_ = 0.0;
    std::unique_ptr<LesMetrics> les_met_;
    cfdnn_mesh_desc les_mesh_{};           // mesh descriptor with the y arrays the LES closure needs (host copies below)
    std::vector<double> h_yc_, h_yf_;
    bool mom_z_ = true;                    // z-marching momentum kernel (env CFDNN_MOMENTUM_Z=0: one thread per point)
    bool mom_exact_ = false;               // env CFDNN_MOMENTUM_EXACT=1: exact divisions on uniform 3-D meshes too
    Comm* comm_ = nullptr;
    cudaStream_t st_ = nullptr;
    cudaEvent_t ev0_ = nullptr, ev1_ = nullptr;

    Vel alloc_vel();
    double* alloc_cell();
    double* field_ptr(int field, size_t* n) const;

    // operators (all act on vel_ like the reference's, which ignore their arguments)
    void apply_velocity_bc();                                   // solver_operators.cpp:43-491
    void compute_convective_term();                             // :493-867
    void compute_diffusive_term();                              // :869-1070
    void compute_divergence(const Vel& v, bool uniform_dy = false);                      // :1072-1245
    void correct_velocity();                                    // :1259-1604
    void blend(const Vel& a, const Vel& b, const Vel& c, double a1, double a2);  // c = a1 a + a2 b
    void momentum(const Vel& in, Vel& out, double dt, bool seam_average);   // fused conv + diff + advance
    void seam(const Vel& out, bool seam_average);
    void copy_correct_velocity();
    void build_rhs(double dt);
    int poisson_solve(const cfdnn_poisson_cfg& pc);
    void euler_substep(const Vel& in, Vel& out, double dt);         // solver_time.cpp:269-661
    void implicit_y_diffusion_step(const Vel& v, double dt);        // solver_time.cpp:667-735 (Thomas solves per y-line)
    bool implicit_y() const;                                         // config flag and y not periodic
    void project_velocity(Vel* which, double dt);                    // :736-1211
    void ssprk3_step(double dt);                                     // :1303-1513
    void ssprk2_step(double dt);                                     // :1213-1301
    void euler_step();                                               // solver.cpp:1505-2507
    void update_turbulence();
    void build_nu_eff();
    void step_device(bool want_residual);
    // ---- z-slab links (multi-GPU): internal interfaces and the periodic wrap between the end ranks
    bool multi() const;
    bool lo_int_ = false, hi_int_ = false, lo_wrap_ = false, hi_wrap_ = false;
    DevBuf<double> seam_scratch_;
    void update_links();
    // bc_site = true : at apply_velocity_bc call sites (all links, incl. the periodic wrap: this is
    //                  the z pass of the reference's BC routine, ghosts go stale like the reference's)
    // bc_site = false: refresh of internal interfaces before a stencil reads across them (an
    //                  interior plane of the single-domain run is never stale)
    void halo_vel(const Vel& v, bool bc_site);
    void halo_cell(double* c, bool bc_site);
    void seam_z(const Vel& v, bool average);   // periodic seam of w across the end ranks
    void check_nan(int step);
    double max_change_host();
    // per-phase device timing of run(): CUDA event pairs recorded around the phase, read after the final sync
    struct PhaseEv { cudaEvent_t a = nullptr, b = nullptr; int phase = 0; };
    std::vector<PhaseEv> phpool_;
    size_t phused_ = 0;
    PhaseEv* ph_begin(int phase);
    void ph_end(PhaseEv* e);
};

}  // namespace cfdnn
