// fft.cuh — FFT-hybrid Poisson solver for channels (periodic x/z, Neumann walls in y): 2-D FFT over (x, z) with
// hand-written shared-memory Stockham kernels + one tridiagonal (Thomas) system in y per Fourier mode.
// Stands in for FFTPoissonSolver (reference include/poisson_solver_fft.hpp:18-50, src/poisson_solver_fft.cpp);
// see fft.cu.  Single GPU (no z-slab variant).
#pragma once
#include <vector>
#include "common.cuh"
#include "comm.cuh"

namespace cfdnn {

class FftPoisson {
public:
    explicit FftPoisson(const cfdnn_mesh_desc& m);
    ~FftPoisson() { release_peer_access(); }
    bool peer_to_peer() const { return p2p_; }   // transposes by direct stores into the other GPUs' arrays
    // x and z must be periodic, y Neumann on both walls (include/poisson_solver_fft.hpp:27-30 allows Dirichlet in y
    // as well; not supported here)
    void set_bc(const int bc[6]);
    // z-slab decomposition (the reference's FFT_MPI variant, src/poisson_solver_fft_mpi.cpp): c->nranks equal slabs;
    // the spectrum is redistributed by two all-to-alls per solve.  nullptr / one rank: single domain.
    void set_comm(Comm* c);
    // rhs, p: device arrays in the solver's ghosted cell layout; returns 1 like the reference's direct solve
    int solve_dev(const double* rhs, double* p, cudaStream_t st);
    // host arrays, same layout (test/FFI convenience: FFTPoissonSolver has only solve_device)
    int solve_host(const double* rhs, double* p);
    static bool size_supported(int n);          // n = 2^a 3^b 5^c, 2 <= n <= 1024
    size_t field_elems() const { return (size_t)cp_ * (Nz_ + 2 * Ng_); }

private:
    void setup_tables(const cfdnn_mesh_desc& m);
    void plan(Comm* c);
    void setup_peer_access();
    void release_peer_access();
    bool p2p_ = false;
    std::vector<double2*> peer_S_, peer_T_;     // S / T of every rank as mapped here (own entry: the local array)
    std::vector<void*> peer_opened_;            // IPC mappings to close
    DevBuf<double> sync_;                       // operand of the barrier all-reduce
    int Nx_, Ny_, Nz_, Ng_;         // Nz_: planes of this rank
    int Nzg_;                       // planes of the whole box
    int Nxh_, Pq_ = 0;              // kept x modes Nx/2+1; columns per rank (padded, complex elements)
    int P_ = 1, rank_ = 0;
    Comm* comm_ = nullptr;
    bool planned_ = false;
    int cs_, cp_;                   // row and plane pitch of the ghosted cell layout
    int Lx_, Lz_;                   // lines per CTA of the x and z transforms
    int nrx_, nrz_;
    int radx_[12], radz_[12];       // radix sequences
    double inv_Ly_;                 // 1 / sum dyv
    bool bc_set_ = false;
    DevBuf<double2> spec_, spec2_;  // S[q][k_local][j][c] and (P > 1) T[kz][j][c]
    DevBuf<double> cprime_, invden_;   // [Nz][Ny][Pxh] Thomas c' and 1 / den (depend on the mesh only)
    DevBuf<double2> twx_, twz_;     // exp(-2 pi i n / N)
    DevBuf<double> lamx_, lamz_;    // Laplacian symbols
    DevBuf<double> lo_, up_, base_, dyv_;   // tridiagonal rows, volume weights
};

}  // namespace cfdnn
