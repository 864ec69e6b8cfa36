// mg_tb_kernels.cuh — temporally blocked smoother of the multigrid V-cycle (FP64, sm_100a).
//
// T consecutive weighted-Jacobi sweeps (one Chebyshev pass is `degree` of them,
// smooth_chebyshev src/poisson_solver_multigrid.cpp:825-1001) are applied in ONE pass over HBM:
// u and f are read once, u is written once (24 B per cell per T sweeps instead of 24 B per sweep).
// Legal because the reference re-applies the BCs before every sweep (:909) and its ghost values
// are a pure function of interior values (:307-554) — so sweep s+1 at a cell only needs sweep s
// at the 7-point neighbourhood, with boundary neighbours resolved from the BC code on the fly.
//
// Scheme (2.5-D overlapped tiling, streaming in z):
//   * a CTA owns a BX x BY column of cells plus a halo of T cells (x halo rounded up to even so
//     16-byte pairs stay aligned) and marches through a chunk of z planes;
//   * a thread owns 4 consecutive x cells ("quad") of one tile row for every level 0..T-1
//     (level l = field after l sweeps).  Per level it keeps the planes z-1 and z of its quad in
//     registers; plane z+1 arrives from the level below in the same iteration (registers);
//   * only the x/y neighbours of the centre plane go through shared memory (N/S rows as
//     LDS.128) or warp shuffles (E/W), one double-buffered plane per level, one __syncthreads
//     per z iteration;
//   * tile cells whose dependency cone leaves the tile compute garbage that is never consumed
//     by a stored cell (the valid region shrinks by one cell per level, exactly the halo).
// Arithmetic of one sweep is expression-for-expression that of k_sweep (mg_kernels.cuh), so the
// result is bitwise identical to T launches of k_sweep.
#pragma once
#include "mg_kernels.cuh"

namespace cfdnn {

template <int T>
struct TbOmega {
    double w[T];      // omega of sweep s
};

template <int T, int BX, int BY>
struct TbShape {
    static_assert(BX % 4 == 0, "BX must be a multiple of 4");
    static constexpr int HX = (T + 1) / 2 * 2;
    static constexpr int TX = BX + 2 * HX;
    static constexpr int TY = BY + 2 * T;
    static constexpr int NQ = TX / 4;
    static constexpr int NACTIVE = NQ * TY;
    static constexpr int NTHREADS = (NACTIVE + 31) / 32 * 32;
    static constexpr int PLANE = TX * TY;
    static constexpr size_t SMEM = sizeof(double) * (size_t)PLANE * 2 * T;
};

__device__ __forceinline__ int wrap_idx(int i, int N) {
    i %= N;
    return i < 0 ? i + N : i;
}

struct Quad {
    double v[4];
};

__device__ __forceinline__ void ld_pair(const double* __restrict__ p, double& a, double& b) {
    const double2 t = *reinterpret_cast<const double2*>(p);
    a = t.x; b = t.y;
}
__device__ __forceinline__ void st_pair(double* p, double a, double b) {
    double2 t; t.x = a; t.y = b;
    *reinterpret_cast<double2*>(p) = t;
}

// One weighted-Jacobi update, same expression as k_sweep (smooth_chebyshev :970-983).
__device__ __forceinline__ double jacobi_point(const LevelGeom& g, double uc, double uW, double uE, double uS,
                                               double uN, double ub, double uf, double fv, double omega,
                                               double om1) {
    const double uj = ((uE + uW) / g.dx2 + (uN + uS) / g.dy2 + (uf + ub) / g.dz2 - fv) / g.coeff;
    return om1 * uc + omega * uj;
}

template <int T, int BX, int BY>
__global__ void __launch_bounds__(TbShape<T, BX, BY>::NTHREADS, 1)
k_sweep_tb(LevelGeom g, const double* __restrict__ u, const double* __restrict__ f,
           double* __restrict__ out, TbOmega<T> om, int ZC) {
    using S = TbShape<T, BX, BY>;
    CFDNN_DYN_SMEM(double, sm);
    const int t = threadIdx.x;
    const int lane = t & 31;
    const bool active = t < S::NACTIVE;
    const int row = active ? t / S::NQ : 0;
    const int q = active ? t % S::NQ : 0;
    const int lx = 4 * q;
    const int gx0 = (int)blockIdx.x * BX - S::HX + lx;
    const int gy = (int)blockIdx.y * BY - T + row;
    const int z0 = (int)blockIdx.z * ZC;
    const int z1 = min(z0 + ZC, g.Nz);
    const bool perx = g.bc[0] == CFDNN_PBC_PERIODIC, pery = g.bc[2] == CFDNN_PBC_PERIODIC,
               perz = g.bc[4] == CFDNN_PBC_PERIODIC;
    const double sxlo = g.bc[0] == CFDNN_PBC_DIRICHLET ? -1.0 : 1.0;
    const double sxhi = g.bc[1] == CFDNN_PBC_DIRICHLET ? -1.0 : 1.0;
    const double sylo = g.bc[2] == CFDNN_PBC_DIRICHLET ? -1.0 : 1.0;
    const double syhi = g.bc[3] == CFDNN_PBC_DIRICHLET ? -1.0 : 1.0;
    const double szlo = g.bc[4] == CFDNN_PBC_DIRICHLET ? -1.0 : 1.0;
    const double szhi = g.bc[5] == CFDNN_PBC_DIRICHLET ? -1.0 : 1.0;

    // ---- per-thread geometry: load offsets, store validity, boundary masks
    const bool rowok = active && (pery || (gy >= 0 && gy < g.Ny));
    const int gyw = pery ? wrap_idx(gy, g.Ny) : gy;
    long long off[2];
    bool ldok[2], stok[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int gxp = gx0 + 2 * h;
        const bool okx = perx || (gxp >= 0 && gxp + 1 < g.Nx);
        const int gxw = perx ? wrap_idx(gxp, g.Nx) : gxp;
        ldok[h] = rowok && okx;
        off[h] = ldok[h] ? (long long)gyw * g.Px + gxw : 0;
        const int tx = lx + 2 * h;
        stok[h] = active && tx >= S::HX && tx < S::HX + BX && row >= T && row < T + BY &&
                  gxp >= 0 && gxp + 1 < g.Nx && gy >= 0 && gy < g.Ny;
    }
    unsigned bw = 0, be = 0;            // point p sits on the x-lo / x-hi physical boundary
    if (!perx) {
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            if (gx0 + p == 0) bw |= 1u << p;
            if (gx0 + p == g.Nx - 1) be |= 1u << p;
        }
    }
    const bool selfS = !pery && gy == 0, selfN = !pery && gy == g.Ny - 1;
    const int si = row * S::TX + lx;                       // my quad in a smem plane
    const int sS = row > 0 ? si - S::TX : si;
    const int sN = row < S::TY - 1 ? si + S::TX : si;
    const bool wsm = (lane == 0) && q > 0;                 // W neighbour must come from smem
    const bool esm = (lane == 31) && q < S::NQ - 1;

    // ---- pipeline state
    double m[T][4], c[T][4], fr[T][4];
#pragma unroll
    for (int l = 0; l < T; ++l)
#pragma unroll
        for (int p = 0; p < 4; ++p) { m[l][p] = 0.0; c[l][p] = 0.0; fr[l][p] = 0.0; }

    // z range of input planes: virtual planes wrap when periodic, are clipped otherwise
    const int zi_first = perz ? z0 - T : max(z0 - T, 0);
    const int zi_last = z1 - 1 + T;                        // inclusive; loads stop at Nz-1 unless periodic
    double nu_[4], nf_[4];                                  // prefetched u[zi], f[zi-1]
    auto load_plane = [&](const double* __restrict__ a, int z, double* dst) {
        const bool zok = perz || (z >= 0 && z < g.Nz);
        const int zw = perz ? wrap_idx(z, g.Nz) : z;
        const double* base = a + (size_t)(zw + 1) * (size_t)g.plane;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (zok && ldok[h]) ld_pair(base + off[h], dst[2 * h], dst[2 * h + 1]);
            else { dst[2 * h] = 0.0; dst[2 * h + 1] = 0.0; }
        }
    };
    load_plane(u, zi_first, nu_);
    load_plane(f, zi_first - 1, nf_);

    int par = 0;
    for (int zi = zi_first; zi <= zi_last; ++zi, par ^= 1) {
        double p_in[4];
#pragma unroll
        for (int p = 0; p < 4; ++p) p_in[p] = nu_[p];
        // f ring: fr[s] must hold f[zi-1-s] while the stages run
#pragma unroll
        for (int s = T - 1; s > 0; --s)
#pragma unroll
            for (int p = 0; p < 4; ++p) fr[s][p] = fr[s - 1][p];
#pragma unroll
        for (int p = 0; p < 4; ++p) fr[0][p] = nf_[p];
        // prefetch the next iteration's planes
        if (zi < zi_last) {
            load_plane(u, zi + 1, nu_);
            load_plane(f, zi, nf_);
        }
#pragma unroll
        for (int s = 0; s < T; ++s) {
            // level-s plane (zi - s) arrives in p_in: publish it as next iteration's centre plane
            double* wbuf = sm + (size_t)(2 * s + par) * S::PLANE;
            const double* rbuf = sm + (size_t)(2 * s + (par ^ 1)) * S::PLANE;
            if (active) {
                st_pair(wbuf + si, p_in[0], p_in[1]);
                st_pair(wbuf + si + 2, p_in[2], p_in[3]);
            }
            const int zo = zi - 1 - s;                     // plane this stage produces (level s+1)
            const bool zvalid = perz || (zo >= 0 && zo < g.Nz);
            double o[4] = {0.0, 0.0, 0.0, 0.0};
            // E/W neighbours of the centre plane: registers of the adjacent lanes
            double uWq = __shfl_up_sync(0xffffffffu, c[s][3], 1);
            double uEq = __shfl_down_sync(0xffffffffu, c[s][0], 1);
            if (zvalid) {
                if (wsm) uWq = rbuf[si - 1];
                if (esm) uEq = rbuf[si + 4];
                double n_[4], s_[4];
                ld_pair(rbuf + sN, n_[0], n_[1]); ld_pair(rbuf + sN + 2, n_[2], n_[3]);
                ld_pair(rbuf + sS, s_[0], s_[1]); ld_pair(rbuf + sS + 2, s_[2], s_[3]);
                const bool zlo = !perz && zo == 0, zhi = !perz && zo == g.Nz - 1;
                const double omega = om.w[s], om1 = 1.0 - omega;
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    const double uc = c[s][p];
                    double uW = p == 0 ? uWq : c[s][p - 1];
                    double uE = p == 3 ? uEq : c[s][p + 1];
                    if (bw >> p & 1u) uW = sxlo * uc;
                    if (be >> p & 1u) uE = sxhi * uc;
                    const double uS = selfS ? sylo * uc : s_[p];
                    const double uN = selfN ? syhi * uc : n_[p];
                    const double ub = zlo ? szlo * uc : m[s][p];
                    const double uf = zhi ? szhi * uc : p_in[p];
                    o[p] = jacobi_point(g, uc, uW, uE, uS, uN, ub, uf, fr[s][p], omega, om1);
                }
            }
#pragma unroll
            for (int p = 0; p < 4; ++p) { m[s][p] = c[s][p]; c[s][p] = p_in[p]; p_in[p] = o[p]; }
        }
        // p_in now holds level-T plane zi - T
        const int zf = zi - T;
        if (zf >= z0 && zf < z1) {
            double* base = out + (size_t)(zf + 1) * (size_t)g.plane;
#pragma unroll
            for (int h = 0; h < 2; ++h)
                if (stok[h]) st_pair(base + off[h], p_in[2 * h], p_in[2 * h + 1]);
        }
        __syncthreads();
    }
}

}  // namespace cfdnn
