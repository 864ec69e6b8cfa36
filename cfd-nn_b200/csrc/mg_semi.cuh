// mg_semi.cuh — semi-coarsening multigrid for y-stretched meshes.
//
// Stands in for the stretched-y branch of MultigridPoissonSolver (reference
// src/poisson_solver_multigrid.cpp): x/z-only coarsening (:181-207), y-line Thomas relaxation
// (:1420-1716), x-z red-black Gauss-Seidel (:1247-1418), nonuniform residual (:1797-1813), 9-point
// x-z restriction (:2069-2154), bilinear x-z prolongation (:2267-2362), PCG coarse solve with the
// y-line preconditioner (:2490-3119) and the K-cycle-lite V-cycle (:3121-3322).
//
// The CPU path these follow updates u IN PLACE: the y-line pass is Gauss-Seidel over lines in
// (k, i) lexicographic order and the "red-black" pass is Gauss-Seidel in j inside each colour, and
// both read ghost cells that were filled by the previous apply_bc (stale during the pass).  To
// match that bit for bit the levels here keep the reference's ghosted layout
// (index = k*plane + j*stride + i, Ng = 1), apply_bc runs as explicit kernels at exactly the
// reference's call sites, lines are relaxed along i + k = const wavefronts (the dependency order
// of the lexicographic sweep) and every colour column is marched in ascending j by one thread.
#pragma once
#include "common.cuh"
#include <vector>
#include <memory>

namespace cfdnn {

struct SemiGeom {                 // by-value kernel argument
    int Nx, Ny, Nz;
    int S;                        // row stride Nx + 2
    long long P;                  // plane stride S * (Ny + 2)
    long long total;              // P * (Nz + 2)
    int bc[6];
    Div dx2, dz2;
    double inv_dx2, inv_dz2;      // 1.0 / dx2, 1.0 / dz2 (y-line RHS, :1443-1444)
    double two_inv_dx2, two_inv_dz2;   // 2.0 / dx2, 2.0 / dz2 (RBGS diagonal, :1380)
    const double *aS, *aN, *aP;   // y-metric coefficients, Ny + 2 entries (Mesh::yLap_*)
    // Thomas factorisation of the y-line matrix (depends on the level only): off-diagonals,
    // c' and the (guarded) pivots with their reciprocals
    const double *ta, *tcp, *tden, *trden;
};

struct SemiLevel {
    SemiGeom g{};
    double dx, dz;
    DevBuf<double> u, f, r;
    DevBuf<double> ta, tcp, tden, trden;
    DevBuf<double> rbd, rbrd;       // RBGS diagonal 2/dx2 + (aS+aN) + 2/dz2 per row and its reciprocal
};

struct Comm;

class MultigridSemi {
public:
    explicit MultigridSemi(const cfdnn_mesh_desc& m);
    ~MultigridSemi();
    void set_bc(const int bc[6]);
    int num_levels() const { return (int)levels_.size(); }
    int solve_dev(const double* rhs, double* p, const cfdnn_poisson_cfg& cfg, cudaStream_t st);
    cfdnn_poisson_stats stats{};

private:
    cfdnn_mesh_desc mesh_;
    std::vector<std::unique_ptr<SemiLevel>> levels_;
    int bc_[6] = {2, 2, 1, 1, 2, 2};
    DevBuf<double> aS_, aN_, aP_;
    std::vector<double> h_aS_, h_aN_;
    DevBuf<double> partial_, scalars_;
    DevBuf<double> pcg_p_, pcg_z_, pcg_Ap_, pcg_x_, pcg_fsave_;
    double* h_scalars_ = nullptr;
    cudaStream_t st_ = nullptr;

    bool has_nullspace() const;
    void factor_lines();                       // per-level Thomas coefficients (needs the y BCs)
    void apply_bc(int level, double* a, bool residual_mode = false);
    void smooth_y_lines(int level, int iterations);
    void smooth_xz_rbgs(int level, int iterations);
    void do_smooth(int level, int nu);
    void residual(int level);
    void residual_norms(int level, double* d_max, double* d_sumsq);
    void restrict_xz(int fine);
    void prolong_xz(int coarse);
    void mean_free(int level, double* field);
    void solve_coarse_pcg(int max_iters, double tol);
    void vcycle(int level, int nu1, int nu2, int degree);
};

}  // namespace cfdnn
