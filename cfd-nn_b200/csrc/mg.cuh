// mg.cuh — geometric-multigrid pressure-Poisson solver, device-resident.
//
// Stands in for MultigridPoissonSolver (reference include/poisson_solver_multigrid.hpp:26-102,
// src/poisson_solver_multigrid.cpp) — CPU-path numerics, not the reference's
// mg_cuda_kernels path.
//
// HBM layout of one level array (u, f, r, tmp):
//   * NO x/y ghost cells: ghost values are a pure function of interior values
//     (apply_bc, :307-554), so every stencil resolves them on the fly from the BC type.
//   * Hz halo planes on each z side (k = -Hz..-1 and k = Nz..Nz+Hz-1) — used only when the z side
//     is a slab interface (BC code CFDNN_PBC_HALO; Hz = 2 there so that two blocked sweeps need one
//     exchange); physical z BCs are resolved on the fly too.
//   * row pitch Px = Nx rounded up to 16 doubles (128 B) so rows start on a line boundary.
//   index(i,j,k) = (k+Hz)*Px*Ny + j*Px + i
#pragma once
#include "common.cuh"
#include "mg_semi.cuh"
#include "mg2d.cuh"
#include <vector>
#include <memory>

namespace cfdnn {

enum { CFDNN_PBC_HALO = 3 };  // z side is a slab interface: read the halo plane as is

struct LevelGeom {            // by-value kernel argument
    int Nx, Ny, Nz;           // interior size of the local slab
    int Px;                   // row pitch
    int Hz;                   // halo planes on each z side (1; T of the blocked smoother on slab levels)
    long long plane;          // Px * Ny
    int bc[6];                // xlo,xhi,ylo,yhi,zlo,zhi
    Div dx2, dy2, dz2, coeff; // dx^2 ..., coeff = 2/dx2 + 2/dy2 + 2/dz2
    int nonuniform_y;         // 1: use aS/aN/aP[j + joff] (stretched y)
    int joff;
    const double *aS, *aN, *aP;
    double inv_dx2_2, inv_dz2_2;   // 2.0/dx2, 2.0/dz2 (nonuniform diag)
    int fastdiv;                   // blocked smoother: x * RN(1/d) instead of the correctly rounded x / d —
                                   // 13 instead of 22 FP64 instructions per cell-sweep, each quotient within
                                   // 1 ulp of the reference's.  Default on; CFDNN_MG_FASTDIV=0 restores the
                                   // correctly rounded divisions (last-bit agreement with the x86 build).
};

struct Level {
    LevelGeom g{};
    double dx, dy, dz;
    int NzGlobal;
    size_t size = 0;          // (Nz+2*Hz)*plane
    DevBuf<double> u, f, r, tmp;
    double lambda_max = 2.2;
    long long cells() const { return (long long)g.Nx * g.Ny * g.Nz; }
};

struct Comm;  // slab communicator (comm.cuh)

class Multigrid {
public:
    explicit Multigrid(const cfdnn_mesh_desc& m);
    ~Multigrid();
    void set_bc(const int bc[6]);
    int num_levels() const {
        return mg2d_impl_ ? mg2d_impl_->num_levels() : semi_impl_ ? semi_impl_->num_levels() : (int)levels_.size();
    }
    // 2-D meshes: which z-plane of the caller's arrays holds the data (0 = the reference's 2-D
    // convention, the default; 1 = interior plane of the solver's periodic-z embedding, which also
    // wants the result copied to the two ghost planes)
    void set_2d_plane(int plane, bool dup) { plane2d_ = plane; dup2d_ = dup; }

    // rhs/p: device pointers, reference ghosted cell layout. Returns V-cycles used.
    int solve_dev(const double* rhs, double* p, const cfdnn_poisson_cfg& cfg, cudaStream_t st);
    int solve_host(const double* rhs, double* p, const cfdnn_poisson_cfg& cfg);

    cfdnn_poisson_stats stats{};
    double vcycle_ms_total = 0.0;   // device time of the V-cycles of the last solve
    int vcycle_count = 0;
    // accumulated over solves since reset (solver step timing)
    double acc_vcycle_ms = 0.0;
    int acc_vcycles = 0;
    bool time_vcycles = true;

    void set_comm(Comm* c);
    // Measurement aid (bench.py roofline): `reps` launches of the level-0 temporally blocked smoother
    // (2 Chebyshev sweeps per launch) on stream st, timed with CUDA events; returns ms per launch
    // (0 when level 0 is not eligible for the tiled kernel).  The level-0 iterate is scratch afterwards.
    double time_level0_smoother(int reps, cudaStream_t st);
    void collect_timing();          // read the V-cycle events of the solves since the last call (syncs)
    cudaStream_t own_stream();
    const cfdnn_mesh_desc& mesh() const { return mesh_; }

private:
    cfdnn_mesh_desc mesh_;
    std::vector<std::unique_ptr<Level>> levels_;
    int bc_[6] = {2, 2, 1, 1, 2, 2};   // reference defaults (header :185-190)
    bool semi_ = false, ystretched_ = false;
    std::unique_ptr<Multigrid2D> mg2d_impl_;     // Nz == 1: 2-D solver (mg2d.cuh)
    int plane2d_ = 0;
    bool dup2d_ = false;
    std::unique_ptr<MultigridSemi> semi_impl_;   // y-stretched meshes: semi-coarsening solver (mg_semi.cuh)
    DevBuf<double> aS_, aN_, aP_;
    DevBuf<double> partial_;        // per-block partials for reductions
    DevBuf<double> scalars_;        // device scalars: [0]=sum/mean [1]=max [2]=sumsq ...
    double* h_scalars_ = nullptr;   // pinned
    struct EvPair { cudaEvent_t a = nullptr, b = nullptr; int cycles = 0; };
    std::vector<EvPair> evpool_;    // one pair per solve since the last collect_timing()
    size_t evused_ = 0;
    EvPair* cur_ev_ = nullptr;
    Comm* comm_ = nullptr;
    cudaStream_t st_ = nullptr;

    // CUDA graph of one V-cycle, keyed by (nu1, nu2, degree)
    cudaGraphExec_t graph_ = nullptr;
    int graph_key_[3] = {-1, -1, -1};
    bool use_graph_ = true;
    unsigned long long graph_launches_ = 0;
    DevBuf<double> h_rhs_, h_p_;    // device staging for solve_host
    cudaStream_t own_st_ = nullptr;

    void apply_bc_codes();
    bool has_nullspace() const;
    void vcycle(int level, int nu1, int nu2, int degree);
    void run_vcycle(int nu1, int nu2, int degree);
    void descend(int level, int nu1, int nu2, int degree);   // vcycle(level), replayed from a graph when communication-free
    cudaGraphExec_t sub_graph_ = nullptr;                    // sub-V-cycle below the first replicated level (z-slab runs)
    int sub_key_[4] = {-1, -1, -1, -1};
    unsigned long long sub_launches_ = 0;
    bool sub_ok_ = true;
    void smooth_chebyshev(int level, int degree);
    void smooth_jacobi(int level, int iters, double omega);
    void sweep(int level, const double* uin, double* uout, double omega);
    void tb_setup();
    bool sweeps_tb(int level, const double* in, double* out, const double* w, int n);
    int tb_max_ = 2;                // sweeps fused per pass (env CFDNN_MG_TB: 0 = plain k_sweep only)
    int tbo_mode_ = 3;              // lean smoother variant (env CFDNN_MG_TBO_MODE): 3 = k_sweep_tbq (2 x 2 patch per thread,
                                    // mg_tbq_kernels.cuh), 0-2 = the barrier variants of k_sweep_tbo (mg_tbo_kernels.cuh)
    bool tbo_ = true;               // lean kernel on levels without physical faces in reach (env CFDNN_MG_TBO=0: general kernel)
    void residual(int level);
    void residual_norms(int level, double* d_max, double* d_sumsq);
    void restrict_to(int fine, bool fused);
    bool rr_eligible(int fine) const;
    bool norms_tb_ = true;          // convergence-check residual norms with the tiled kernel (env CFDNN_MG_NORMS_TB=0: k_residual)
    bool overlap_ = true;           // slab levels: halo exchange overlapped with the inner z chunks (env CFDNN_MG_OVERLAP=0: serial)
    cudaStream_t comm_st_ = nullptr;
    cudaEvent_t ev_fork_ = nullptr, ev_join_ = nullptr;
    int rr_min_ = 128;              // smallest level width that takes the fused kernel (env CFDNN_MG_RR_MIN, tests)
    bool rr_ = true;                // fused residual + restriction on big levels (env CFDNN_MG_RR=0: two kernels)
    void prolong_from(int coarse);
    void mean_free(int level, double* field, bool subtract = true);
    void halo(int level, double* field, int depth = 1);   // slab interface exchange of `depth` planes (no-op single GPU)
    void invalidate_graph();
};

}  // namespace cfdnn
