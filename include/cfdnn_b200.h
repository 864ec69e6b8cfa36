/* cfdnn_b200.h — C ABI of libcfdnn_b200.so
 *
 * B200-native (sm_100a) backend for the per-timestep fractional-step hot path of
 * comp-physics/cfd-nn.  Plain pointers and sizes only; no torch / C++ types.
 * Every entry point names the reference interface it stands in for
 * (paths relative to the reference tree, commit c73310a).
 *
 * All `const double*` / `double*` field arguments are HOST pointers in the
 * reference's own ghost-inclusive layouts (include/fields.hpp:118-136):
 *   cell fields : (Nx+2Ng)(Ny+2Ng)(Nz+2Ng), i fastest
 *   u faces     : row length Nx+1+2Ng ; v faces: Ny+1+2Ng rows ; w faces: Nz+1+2Ng planes
 * unless the function name ends in `_dev`, in which case they are device pointers.
 *
 * Return value: 0 on success, non-zero on error; cfdnn_last_error() gives the text.
 * The library has NO CPU fallback: every compute entry fails with CFDNN_ERR_NO_DEVICE
 * when no CUDA device is usable.
 */
#ifndef CFDNN_B200_H
#define CFDNN_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CFDNN_OK 0
#define CFDNN_ERR_INVALID 1    /* std::invalid_argument in the reference            */
#define CFDNN_ERR_RUNTIME 2    /* std::runtime_error (NaN guard, unsupported BC...) */
#define CFDNN_ERR_NO_DEVICE 3  /* no CUDA device / kernel image not loadable        */
#define CFDNN_ERR_CUDA 4

/* PoissonBC — include/poisson_solver.hpp:9-13 (same numeric values) */
enum { CFDNN_PBC_DIRICHLET = 0, CFDNN_PBC_NEUMANN = 1, CFDNN_PBC_PERIODIC = 2 };
/* VelocityBC::Type — include/solver.hpp (Periodic, NoSlip, Inflow, Outflow) */
enum { CFDNN_VBC_NOSLIP = 0, CFDNN_VBC_PERIODIC = 1, CFDNN_VBC_INFLOW = 2, CFDNN_VBC_OUTFLOW = 3 };
/* TimeIntegrator / ConvectiveScheme — include/config.hpp */
enum { CFDNN_TI_EULER = 0, CFDNN_TI_RK2 = 1, CFDNN_TI_RK3 = 2 };
enum { CFDNN_CONV_CENTRAL = 0, CFDNN_CONV_SKEW = 2 };  /* Upwind(1)/Upwind2(3): out of scope */
/* closure kinds — TurbulenceModelType::{None,NNMLP,NNTBNN} */
enum { CFDNN_TURB_NONE = 0, CFDNN_TURB_NN_MLP = 1, CFDNN_TURB_NN_TBNN = 2 };
/* static LES subgrid-scale models — TurbulenceModelType::{Smagorinsky, WALE, Vreman, Sigma}, include/turbulence_les.hpp */
enum { CFDNN_LES_NONE = 0, CFDNN_LES_SMAGORINSKY = 1, CFDNN_LES_WALE = 2, CFDNN_LES_VREMAN = 3, CFDNN_LES_SIGMA = 4 };
/* PoissonSolveStatus — include/solver.hpp:130-140 */
enum { CFDNN_PSTAT_CONVERGED = 0, CFDNN_PSTAT_HIT_MAX = 1, CFDNN_PSTAT_FIXED = 2 };

/* field ids for upload/download (RANSSolver accessors, include/solver.hpp:230-300) */
enum {
    CFDNN_F_U = 0, CFDNN_F_V = 1, CFDNN_F_W = 2,      /* velocity()            */
    CFDNN_F_P = 3,                                      /* pressure()            */
    CFDNN_F_PCORR = 4,                                  /* pressure_correction() */
    CFDNN_F_DIV = 5,                                    /* div_velocity()        */
    CFDNN_F_RHS = 6,                                    /* rhs_poisson()         */
    CFDNN_F_NUT = 7,                                    /* nu_t()                */
    CFDNN_F_K = 8, CFDNN_F_OMEGA = 9,                   /* k(), omega()          */
    CFDNN_F_NUEFF = 10,
    CFDNN_F_TAU_XX = 11, CFDNN_F_TAU_XY = 12, CFDNN_F_TAU_YY = 13,
    CFDNN_F_COUNT = 14,
    /* device-view only (cfdnn_solver_device_ptr): the work arrays of SolverDeviceView,
     * include/turbulence_model.hpp:27-78 — not transferable with upload/download */
    CFDNN_F_U_STAR = 14, CFDNN_F_V_STAR = 15, CFDNN_F_W_STAR = 16,
    CFDNN_F_U_OLD = 17, CFDNN_F_V_OLD = 18, CFDNN_F_W_OLD = 19,
    CFDNN_F_CONV_U = 20, CFDNN_F_CONV_V = 21, CFDNN_F_CONV_W = 22,
    CFDNN_F_DIFF_U = 23, CFDNN_F_DIFF_V = 24, CFDNN_F_DIFF_W = 25,
    CFDNN_F_VIEW_COUNT = 26
};

/* Mesh — include/mesh.hpp:16-182 (only what the hot path reads).
 * y-metric arrays have Ny+2Ng entries (dyv, yLap_*) / Ny+2Ng+1 (dyc, yf) and may be NULL
 * when y_stretched == 0.  Nz == 1 means a 2-D mesh (reference Mesh::is2D()). */
typedef struct cfdnn_mesh_desc {
    int Nx, Ny, Nz, Ng;
    double dx, dy, dz;
    int y_stretched;
    const double* dyv;
    const double* dyc;
    const double* yLap_aS;
    const double* yLap_aN;
    const double* yLap_aP;
    const double* yc;       /* cell-centre y, Ny+2Ng                                   */
    double y_lo, y_hi;      /* wall positions (Mesh::y_min/y_max) for wall distance      */
    /* z-slab decomposition (replaces include/decomposition.hpp): this rank owns global
     * planes [k_offset, k_offset+Nz) of Nz_global; single rank: k_offset=0, Nz_global=Nz */
    int Nz_global, k_offset;
    const double* yf;       /* face y, Ny+2Ng+1 (Mesh::yf); needed by the LES closures on y-stretched meshes, else NULL */
} cfdnn_mesh_desc;

/* PoissonConfig — include/poisson_solver.hpp:16-53 (same defaults via cfdnn_poisson_cfg_default) */
typedef struct cfdnn_poisson_cfg {
    double tol, tol_abs, tol_rhs, tol_rel;
    int max_vcycles, check_interval, use_l2_norm;
    double linf_safety_factor;
    int fixed_cycles, nu1, nu2, chebyshev_degree;
} cfdnn_poisson_cfg;

/* MultigridPoissonSolver accessors residual()/residual_l2()/rhs_norm()/... —
 * include/poisson_solver_multigrid.hpp:61-96 */
typedef struct cfdnn_poisson_stats {
    int cycles, converged, fixed_cycle_mode, status;
    double residual, residual_l2, b_inf, b_l2, r0, r0_l2;
} cfdnn_poisson_stats;

/* Hot-path keys of Config — include/config.hpp:65-234 (same names, same defaults via
 * cfdnn_config_default) */
typedef struct cfdnn_config {
    double nu, dt;
    int time_integrator, convective_scheme;
    int adaptive_dt;
    double CFL_max, CFL_xz, dt_safety;
    /* Poisson keys; the RK path uses only max_vcycles, tol_rhs, fixed_cycles
     * (src/solver_time.cpp:868-871), the Euler path all of them (src/solver.cpp:2181-2205) */
    int poisson_fixed_cycles, poisson_max_vcycles;
    double poisson_tol_rhs, poisson_tol, poisson_abs_tol_floor, poisson_tol_abs, poisson_tol_rel;
    int poisson_check_interval, poisson_use_l2_norm;
    double poisson_linf_safety;
    int poisson_nu1, poisson_nu2, poisson_chebyshev_degree;
    int turb_guard_enabled, turb_guard_interval;
    int turb_model;
    double nu_t_max;
    /* implicit y-diffusion (include/config.hpp:103): the d2/dy2 part of the momentum diffusion is taken out of the
     * explicit term (src/solver_operators.cpp:915) and solved per y-line with the Thomas algorithm after every
     * predictor / RK stage (src/solver_time.cpp:658-735, src/solver_time_kernels_implicit.cpp); no effect when y is
     * periodic; relaxes the diffusive dt limit to the x/z spacings (src/solver.cpp:1481, 3697-3712) */
    int implicit_y_diffusion;
    /* Poisson back end of the projection (Config::poisson_solver, include/config.hpp:47-56, 151): CFDNN_POISSON_MG
     * (default; also what Auto means here — the reference's CPU build, the parity oracle, always takes MG) or
     * CFDNN_POISSON_FFT = FFTPoissonSolver (periodic x/z, walls y, uniform or stretched y; single GPU) */
    int poisson_solver;
} cfdnn_config;
enum { CFDNN_POISSON_MG = 0, CFDNN_POISSON_FFT = 1 };

typedef struct cfdnn_solver cfdnn_solver;   /* stands in for RANSSolver              */
typedef struct cfdnn_mg cfdnn_mg;           /* stands in for MultigridPoissonSolver  */
typedef struct cfdnn_nn cfdnn_nn;           /* stands in for TurbulenceNNTBNN / NNMLP */
typedef struct cfdnn_fft cfdnn_fft;         /* stands in for FFTPoissonSolver        */

const char* cfdnn_last_error(void);
int cfdnn_version(void);
/* number of CUDA kernels this library has launched in this process (bench.py: gpu_launches) */
unsigned long long cfdnn_kernel_launches(void);
int cfdnn_device_count(void);
int cfdnn_set_device(int device);
int cfdnn_device_synchronize(void);

void cfdnn_poisson_cfg_default(cfdnn_poisson_cfg* cfg);
void cfdnn_config_default(cfdnn_config* cfg);

/* ---- MultigridPoissonSolver (include/poisson_solver_multigrid.hpp:26-102) ---- */
int cfdnn_mg_create(const cfdnn_mesh_desc* mesh, cfdnn_mg** out);      /* ctor :29            */
void cfdnn_mg_destroy(cfdnn_mg* mg);                                    /* dtor :32            */
int cfdnn_mg_set_bc(cfdnn_mg* mg, const int bc[6]);                     /* set_bc :35-42       */
int cfdnn_mg_num_levels(const cfdnn_mg* mg);
/* solve(const ScalarField& rhs, ScalarField& p, cfg) :46 — host arrays, copies inside */
int cfdnn_mg_solve(cfdnn_mg* mg, const double* rhs, double* p, const cfdnn_poisson_cfg* cfg,
                   cfdnn_poisson_stats* stats);
/* solve_device(double* rhs, double* p, cfg) :59 — device-resident twin */
int cfdnn_mg_solve_dev(cfdnn_mg* mg, const double* rhs_dev, double* p_dev,
                       const cfdnn_poisson_cfg* cfg, cfdnn_poisson_stats* stats);
/* device timing of the last solve's V-cycles (CUDA events), for bench.py's roofline object */
int cfdnn_mg_last_timing(const cfdnn_mg* mg, double* vcycle_ms_total, int* n_vcycles);

/* ---- FFTPoissonSolver (include/poisson_solver_fft.hpp:18-50; src/poisson_solver_fft.cpp, GPU builds only) ----
 * 2-D FFT over the periodic x/z directions + one tridiagonal system in y per mode; Nx, Nz = 2^a 3^b 5^c <= 1024 */
int cfdnn_fft_create(const cfdnn_mesh_desc* mesh, cfdnn_fft** out);    /* ctor :20            */
void cfdnn_fft_destroy(cfdnn_fft* f);                                   /* dtor :21            */
/* set_bc :27-30 — x, z periodic (else CFDNN_ERR_RUNTIME like the reference's runtime_error), y Neumann */
int cfdnn_fft_set_bc(cfdnn_fft* f, const int bc[6]);
/* is_suitable :38-41 */
int cfdnn_fft_is_suitable(const int bc[6], int uniform_x, int uniform_z);
/* solve_device(double* rhs, double* p) :46 — device arrays in the ghosted cell layout, on `stream` (a cudaStream_t or
 * NULL); *iterations = 1 (direct solve).  cfdnn_fft_solve: the same with host arrays, copies inside. */
int cfdnn_fft_solve_dev(cfdnn_fft* f, const double* rhs_dev, double* p_dev, void* stream, int* iterations);
int cfdnn_fft_solve(cfdnn_fft* f, const double* rhs, double* p, int* iterations);

/* ---- RANSSolver (include/solver.hpp:86-311) ---- */
int cfdnn_solver_create(const cfdnn_mesh_desc* mesh, const cfdnn_config* cfg, cfdnn_solver** out); /* ctor :88 */
void cfdnn_solver_destroy(cfdnn_solver* s);
int cfdnn_solver_set_velocity_bc(cfdnn_solver* s, const int bc[6]);     /* set_velocity_bc :101 */
int cfdnn_solver_set_body_force(cfdnn_solver* s, double fx, double fy, double fz); /* :104 */
int cfdnn_solver_set_dt(cfdnn_solver* s, double dt);                    /* set_dt             */
int cfdnn_solver_get_dt(const cfdnn_solver* s, double* dt);             /* current_dt         */
int cfdnn_solver_upload(cfdnn_solver* s, int field, const double* host, size_t n);   /* sync_to_gpu   */
int cfdnn_solver_download(cfdnn_solver* s, int field, double* host, size_t n);       /* sync_from_gpu; also the CFDNN_F_*_STAR / OLD / CONV / DIFF work fields */
size_t cfdnn_solver_field_size(const cfdnn_solver* s, int field);
/* initialize(const VectorField&) :1060-1062 — apply_velocity_bc on the uploaded velocity */
int cfdnn_solver_initialize(cfdnn_solver* s);
int cfdnn_solver_step(cfdnn_solver* s, double* residual_out);           /* double step() :208 */
/* run n steps without reading the residual back between them (app time loops) */
int cfdnn_solver_run(cfdnn_solver* s, int nsteps, double* last_residual_out);
int cfdnn_solver_compute_adaptive_dt(cfdnn_solver* s, double* dt_out);  /* compute_adaptive_dt :214 */
int cfdnn_solver_poisson_stats(const cfdnn_solver* s, cfdnn_poisson_stats* out); /* poisson_stats() */
/* NO COUNTERPART IN THE REFERENCE — added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  compute_adaptive_dt on caller-owned
 * DEVICE fields (the solver's ghosted u, v, w layouts, e.g. from cfdnn_solver_device_ptr; nu_t_dev = NULL when no
 * closure is attached) that allocates nothing: the kernels' scratch (per-block partial maxima, the five final maxima,
 * the device copy of dyv) lives in one file-scope static __device__ array of CFDNN_NOALLOC_SCRATCH_BYTES that every
 * GPU which loads the library holds until exit.  Limits: Ny <= CFDNN_NOALLOC_MAX_NY (else CFDNN_ERR_INVALID), Nghost 1,
 * single domain.  `stream` is a cudaStream_t or NULL; the call synchronises that stream before it returns.  All calls
 * share the array and it keeps the previous call's contents (each call rewrites what it reads): calls that overlap on
 * the GPU — other streams, other host threads — race and must be serialised by the caller.  Same value as
 * cfdnn_solver_compute_adaptive_dt on the same fields (bridged transport models excepted, see that entry). */
#define CFDNN_NOALLOC_MAX_NY 4096
#define CFDNN_NOALLOC_SCRATCH_BYTES 80224
int cfdnn_adaptive_dt_noalloc(const cfdnn_mesh_desc* mesh, const cfdnn_config* cfg, const double* u_dev,
                              const double* v_dev, const double* w_dev, const double* nu_t_dev, void* stream,
                              double* dt_out);
/* V-cycles of every Poisson solve since the last call (oracle: POISSON_STATS lines) */
int cfdnn_solver_cycle_log(cfdnn_solver* s, int* cycles, int cap, int* n_out);
/* 1 when the FFT Poisson solver of a multi-GPU run does its transposes by peer-to-peer stores (CUDA IPC over NVLink: the
 * transform kernels write straight into the owning GPU's array), 0 when it uses NCCL all-to-alls or is not in use */
int cfdnn_solver_fft_peer_to_peer(const cfdnn_solver* s);
/* set_turbulence_model(unique_ptr<TurbulenceModel>) :98 — solver borrows the closure */
int cfdnn_solver_set_turbulence_model(cfdnn_solver* s, cfdnn_nn* nn);
/* device timing of the last step()/run(): total ms and ms spent inside V-cycles */
int cfdnn_solver_last_timing(const cfdnn_solver* s, double* step_ms, double* vcycle_ms, int* n_vcycles);
/* get_solver_view() / get_device_view() (include/solver.hpp:1118-1122): the DEVICE twin of a solver field
 * (SolverDeviceView / TurbulenceDeviceView pointers, include/turbulence_model.hpp:27-78,
 * include/turbulence_device_view.hpp:12-55) and its element count.  On 2-D meshes the data plane is k = Ng.
 * The RK stages rotate the velocity buffers, so pointers are valid until the next step()/run(). */
int cfdnn_solver_device_ptr(cfdnn_solver* s, int field, double** dev_out, size_t* n_out);
/* nu_t supplied by the caller (on = 1; on = 2 when that model also solves transport equations, which keeps the
 * dy term in the diffusive dt limit under implicit y-diffusion, src/solver.cpp:1481, 3709): a reference
 * TurbulenceModel that is NOT one of this backend's closures
 * (e.g. SST k-omega running on the host) is bridged by download(velocity) -> model.update() ->
 * upload(CFDNN_F_NUT); with this flag on, step() builds nu_eff = nu + nu_t (src/solver.cpp:1326-1410) and
 * takes the closure branches of the diffusion-CFL re-check and the NaN guard as if a model were attached */
int cfdnn_solver_set_external_nu_t(cfdnn_solver* s, int on);
/* SST k-omega transport + SST closure (SSTKOmegaTransport, include/turbulence_transport.hpp:32-58,
 * src/turbulence_transport.cpp:56-190, 259-313, 547-603, 768-865) evaluated device-resident at the top of every step():
 * advance_turbulence (one in-place sweep in the reference CPU path's (j, i) order, as a wavefront) with the previous
 * nu_t, then the closure.  Default SSTConstants.  k and omega are the solver's CFDNN_F_K / CFDNN_F_OMEGA fields — seed
 * them with cfdnn_solver_upload the way RANSSolver::initialize does (src/solver.cpp:1064-1103).  2-D meshes only, like
 * the reference model (CFDNN_ERR_INVALID on a 3-D mesh); replaces the NN / LES closure slot.  kind = CFDNN_RANS_SST, or
 * CFDNN_RANS_KOMEGA for the standard Wilcox model KOmegaTransport with its Boussinesq closure (:202-228, 1012-1118;
 * no wall BC pass, face velocities advect), or CFDNN_RANS_NONE to detach. */
enum { CFDNN_RANS_NONE = 0, CFDNN_RANS_SST = 1, CFDNN_RANS_KOMEGA = 2, CFDNN_RANS_MIXING_LENGTH = 3,
       CFDNN_RANS_ALGEBRAIC_KOMEGA = 4, CFDNN_RANS_GEP = 5,
       /* SSTWithEARSM (src/turbulence_earsm.cpp:205-346, 349-690): SST transport + an explicit algebraic Reynolds-stress
        * closure on the TBNN tensor basis — Wallin-Johansson 2000, Gatski-Speziale 1993, Pope 1975 quadratic
        * (turb_model = earsm_wj / earsm_gs / earsm_pope); nu_t and tau_ij (CFDNN_F_TAU_*) are written */
       CFDNN_RANS_EARSM_WJ = 6, CFDNN_RANS_EARSM_GS = 7, CFDNN_RANS_EARSM_POPE = 8 };
int cfdnn_solver_set_sst_model(cfdnn_solver* s, int kind);
/* the same entry under the name of what it has become: every 2-D RANS closure of the reference (CFDNN_RANS_*) */
int cfdnn_solver_set_rans_model(cfdnn_solver* s, int kind);
/* The algebraic 2-D baselines of src/turbulence_baseline.cpp through the same slot: CFDNN_RANS_MIXING_LENGTH =
 * MixingLengthModel (van Driest mixing length, nu_t under-relaxed with its previous value; `turb_model = baseline`,
 * :36-76, 330-402) and CFDNN_RANS_ALGEBRAIC_KOMEGA = AlgebraicKOmegaModel (:408-483).  Mixing-length constants
 * (set_kappa / set_A_plus / set_delta, include/turbulence_baseline.hpp:38-45; defaults 0.41, 26, 1): */
int cfdnn_solver_set_mixing_length_params(cfdnn_solver* s, double kappa, double A_plus, double delta);
/* CFDNN_RANS_GEP = TurbulenceGEP (Weatheritt-Sandberg algebraic model, src/turbulence_gep.cpp:21-83); variant as
 * TurbulenceGEP::Variant (include/turbulence_gep.hpp:53-60: 0 WS2016_Channel, 1 WS2016_PeriodicHill, 2 Simple = default).
 * model_nu: the viscosity inside the model.  TurbulenceGEP keeps its OWN nu_ = 0.001 and a non-virtual set_nu
 * (include/turbulence_gep.hpp:48, 63), so the set_nu the solver / factory issue through a TurbulenceModel pointer never
 * reaches it: a reference run evaluates the model with nu = 0.001 whatever Config::nu is.  model_nu <= 0 selects that
 * effective value (parity with a reference run); pass the physical nu for a model whose concrete set_nu was called. */
int cfdnn_solver_set_gep_params(cfdnn_solver* s, int variant, double model_nu);
/* SSTWithEARSM::set_pope_constants (Config::pope_C1 / pope_C2, include/config.hpp:122-123; defaults 0.1, 0.1) */
int cfdnn_solver_set_pope_constants(cfdnn_solver* s, double C1, double C2);
/* Trip-region forcing (Config::trip_*, include/config.hpp:202-210): a localized body force on v* (and w*) in the Euler
 * predictor while time < duration (src/solver.cpp:1724-1810, get_trip_time_ramp :939-950).  The tables are the ones
 * initialize_trip_forcing (:818-923) builds — x envelope on cell centres (Nx+2Ng), g(y) = y (1 - y^2) on the y faces
 * (Ny+2Ng+1), the spanwise mode sums on cell centres (Nz+2Ng) and z faces (Nz+2Ng+1) with std::srand(12345) phases — and
 * are computed by the caller (the C++ shim does it with the reference's expressions; a z-slab passes its own planes).
 * amplitude = trip_amplitude * u_tau^2.  Fz_v = NULL on 2-D meshes; Fz_w = NULL: no w forcing.  desc = NULL: off.
 * As in the reference the RK integrators do not apply it. */
typedef struct cfdnn_trip_desc {
    double amplitude, duration, ramp_off_start, w_scale;
    const double* env_x;
    const double* g_y;
    const double* Fz_v;
    const double* Fz_w;
} cfdnn_trip_desc;
int cfdnn_solver_set_trip_forcing(cfdnn_solver* s, const cfdnn_trip_desc* desc);
int cfdnn_solver_get_time(const cfdnn_solver* s, double* t);             /* current_time()            */
int cfdnn_solver_set_time(cfdnn_solver* s, double t);                    /* restart / tests           */
/* apply_velocity_filter(strength) (include/solver.hpp:260, src/solver.cpp:4065-4162): explicit Laplacian filter of u, v, w
 * between steps (3-D meshes; strength <= 0: no-op).  The reference has this kernel in its GPU build only and updates
 * the field in place while neighbouring threads read it; here every second difference is taken from the unfiltered
 * field (the order-independent reading of the same formula). */
int cfdnn_solver_apply_velocity_filter(cfdnn_solver* s, double strength);
/* device ms of the last step()/run() spent in the closure update (TurbulenceModel::update inside step(),
 * src/solver.cpp:1284-1324) and in the fused momentum kernels (convective + diffusive + advance) */
enum { CFDNN_PHASE_CLOSURE = 0, CFDNN_PHASE_MOMENTUM = 1, CFDNN_PHASE_IMPLICIT_Y = 2, CFDNN_PHASE_POISSON = 3,
       CFDNN_PHASE_COUNT = 4 };   /* POISSON: the Poisson solves (multigrid or FFT), copies in and out included */
int cfdnn_solver_phase_timing(const cfdnn_solver* s, double ms[CFDNN_PHASE_COUNT]);
/* measurement aid for bench.py's roofline: ms per launch of the level-0 temporally blocked smoother
 * kernel (2 Chebyshev sweeps per launch), `reps` launches timed with CUDA events on the solver's
 * stream; 0 when level 0 does not use the tiled kernel.  Leaves the MG scratch iterate modified. */
int cfdnn_solver_time_smoother(cfdnn_solver* s, int reps, double* ms_per_launch);
/* z-slab multi-GPU (replaces Decomposition + HaloExchange, src/decomposition.cpp, src/halo_exchange.cpp):
 * nccl_unique_id is the 128-byte ncclUniqueId produced by rank 0 and broadcast by the caller */
int cfdnn_comm_unique_id(void* id128);
int cfdnn_solver_comm_init(cfdnn_solver* s, int rank, int nranks, const void* id128);

/* ---- NN closures (include/turbulence_model.hpp:81-165, include/turbulence_nn_tbnn.hpp:35-56) ---- */
/* weights: row-major (out,in) per layer as in layerN_W.txt (src/nn_core.cpp:501-611);
 * act[l]: 0 linear, 1 tanh, 2 relu ; means/stds may be NULL (no input scaling) */
int cfdnn_nn_create(int kind, int n_layers, const int* dims /* n_layers+1 */, const int* act,
                    const double* const* W, const double* const* b,
                    const double* means, const double* stds, cfdnn_nn** out);
void cfdnn_nn_destroy(cfdnn_nn* nn);
int cfdnn_nn_set_params(cfdnn_nn* nn, double nu, double delta, double u_ref, double k_min,
                        double nu_t_max);
/* TurbulenceModel::update(mesh, velocity, k, omega, nu_t, tau_ij, view) on host arrays of a
 * 2-D mesh (reference semantics) or on every interior z-plane of a 3-D mesh.
 * k/omega may be NULL (== all zero); tau_* may be NULL. */
int cfdnn_nn_update(cfdnn_nn* nn, const cfdnn_mesh_desc* mesh, const double* u, const double* v,
                    const double* k, const double* omega, double* nu_t,
                    double* tau_xx, double* tau_xy, double* tau_yy);
/* TurbulenceModel::update with a device_view (include/turbulence_model.hpp:100-108): the same update on the
 * DEVICE twins of the fields (pointers from cfdnn_solver_device_ptr or the caller's own allocations) */
int cfdnn_nn_update_dev(cfdnn_nn* nn, const cfdnn_mesh_desc* mesh, const double* u_dev, const double* v_dev,
                        const double* k_dev, const double* omega_dev, double* nu_t_dev,
                        double* tau_xx_dev, double* tau_xy_dev, double* tau_yy_dev);
/* ---- static LES SGS closures (include/turbulence_les.hpp:24-115, src/turbulence_les.cpp, src/velocity_gradient.cpp) ----
 * LESModel::update(mesh, velocity, k, omega, nu_t, ...) on host arrays: nu_t = nu_sgs on every interior cell
 * (2-D mesh: plane 0).  constant <= 0 selects the reference default (Cs 0.17, Cw 0.325, Cv 0.07, C_sigma 1.35). */
int cfdnn_les_update(const cfdnn_mesh_desc* mesh, int model, double constant, const double* u, const double* v,
                     const double* w, double* nu_t);
/* set_turbulence_model(std::make_unique<SmagorinskyModel|WALEModel|VremanModel|SigmaModel>(constant)): the model is
 * evaluated device-resident at the top of every step() (src/solver.cpp:1284-1324); CFDNN_LES_NONE detaches it */
int cfdnn_solver_set_les_model(cfdnn_solver* s, int model, double constant);
/* batched MLP forward (MLP::forward, src/nn_core.cpp:144-204): x[n][in] -> y[n][out] */
int cfdnn_nn_forward(cfdnn_nn* nn, const double* x, double* y, size_t n);
int cfdnn_nn_last_timing(const cfdnn_nn* nn, double* ms);
/* 1 when update() runs the hidden-layer GEMMs on the FP64 tensor cores (mma.sync.m8n8k4.f64), 0 when the
 * network only fits the scalar kernel (more than 8 inputs/outputs, or fragments larger than shared memory) */
int cfdnn_nn_uses_tensor_cores(const cfdnn_nn* nn);

#ifdef __cplusplus
}
#endif
#endif /* CFDNN_B200_H */
