// mg2d.cuh — geometric multigrid for 2-D meshes (Nz == 1; BASELINE config 1, 2-D Poiseuille).
//
// Stands in for the 2-D branches of MultigridPoissonSolver (reference
// src/poisson_solver_multigrid.cpp): 5-point Chebyshev / damped-Jacobi smoothing (:825-1001,
// :1003-1245), apply_bc with the periodic corner re-pass (:340-455), residual (:1759-1795), 9-point
// full weighting (:1995-2022), bilinear prolongation (:2183-2219), and for y-stretched 2-D meshes the
// x-only semi-coarsening with the 16-sweep Jacobi smoother (:189-196, :3164-3166, :2100-2122, :2296-2324).
//
// 2-D problems are small (the reference's own case is 32 x 64): every level fits one CTA's reach, so
// a whole sequence of V-cycles runs as ONE single-CTA launch (k_mg2d_vcycles) instead of ~600
// launch-latency-bound kernels per cycle.  Levels keep the reference's ghosted layout
// (index = j*(Nx+2) + i) and apply_bc runs where the reference runs it.  Interior means are summed
// in the reference's order (one running double).
#pragma once
#include "common.cuh"
#include <vector>
#include <memory>

namespace cfdnn {

constexpr int MG2D_MAX_LEVELS = 12;
constexpr int MG2D_MAX_DEGREE = 8;

struct Lvl2D {
    int Nx, Ny, S;                // interior size, row stride Nx + 2
    int total;                    // S * (Ny + 2)
    double dx2, dy2, coeff;       // dx^2, dy^2, 2/dx2 + 2/dy2
    double two_inv_dx2;           // 2.0 / dx2 (nonuniform diagonal)
    double *u, *f, *r, *tmp;
};

struct MG2DParams {               // by-value kernel argument
    int nl;
    Lvl2D L[MG2D_MAX_LEVELS];
    int bc[4];                    // xlo, xhi, ylo, yhi
    int semi;                     // 1: y-stretched (x-only coarsening, nonuniform y, Jacobi smoothing)
    int nullspace;
    const double *aS, *aN, *aP;   // y metrics (Ny + 2), semi only
    double omega[MG2D_MAX_DEGREE];   // Chebyshev weights (uniform levels all have lambda_max = 2.2)
};

class Multigrid2D {
public:
    explicit Multigrid2D(const cfdnn_mesh_desc& m);
    ~Multigrid2D();
    void set_bc(const int bc[6]);
    int num_levels() const { return P_.nl; }
    // rhs / p: device arrays in the reference's cell layout with 3 z-planes; `plane` = the plane that
    // holds the data (0: reference 2-D convention; 1: the solver's interior plane).  With
    // dup_planes the result is written to all three planes (periodic-z copies for the 3-D kernels).
    int solve_dev(const double* rhs, double* p, const cfdnn_poisson_cfg& cfg, cudaStream_t st, int plane, bool dup_planes);
    cfdnn_poisson_stats stats{};

private:
    cfdnn_mesh_desc mesh_;
    MG2DParams P_{};
    std::vector<DevBuf<double>> bufs_;
    DevBuf<double> aS_, aN_, aP_, scalars_;
    double* h_scalars_ = nullptr;
    int bc_[4] = {2, 2, 1, 1};
    int degree_cached_ = -1;
    void set_degree(int degree);
    bool has_nullspace() const;
};

}  // namespace cfdnn
