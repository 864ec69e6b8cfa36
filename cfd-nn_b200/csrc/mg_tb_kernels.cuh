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
//     conflict-free LDS.128: a plane is stored as [pair][thread] double2) or warp shuffles (E/W),
//     one double-buffered plane per level, one __syncthreads per z iteration;
//   * the z loop is unrolled by 3 so the three live planes of a level rotate through fixed
//     registers (no moves); tiles that touch a non-periodic face take a separate code path so
//     the interior path carries no boundary selects;
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
    static constexpr size_t SMEM = sizeof(double) * (size_t)(4 * NTHREADS) * 2 * T;   // [2T planes][2 pairs][NTHREADS] double2
};

__device__ __forceinline__ int wrap_idx(int i, int N) {
    i %= N;
    return i < 0 ? i + N : i;
}

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
    if (g.fastdiv) {   // reciprocal multiplies: 1 ulp per quotient off the reference's divisions (mg.cuh)
        const double uj = ((uE + uW) * g.dx2.r + (uN + uS) * g.dy2.r + (uf + ub) * g.dz2.r - fv) * g.coeff.r;
        return om1 * uc + omega * uj;
    }
    const double uj = ((uE + uW) / g.dx2 + (uN + uS) / g.dy2 + (uf + ub) / g.dz2 - fv) / g.coeff;
    return om1 * uc + omega * uj;
}

// r = f - L u, same expression as k_residual (compute_residual :1807-1816)
__device__ __forceinline__ double residual_point(const LevelGeom& g, double uc, double uW, double uE, double uS,
                                                 double uN, double ub, double uf, double fv) {
    const double two_c = 2.0 * uc;
    const double lap = (uE - two_c + uW) / g.dx2 + (uN - two_c + uS) / g.dy2 + (uf - two_c + ub) / g.dz2;
    return fv - lap;
}

// RES = true: the LAST stage evaluates the residual of its input level instead of a sweep
// (T = 1, RES: a tiled residual kernel; `out` is then the r array).
// NORMS (with RES): nothing is stored; per-CTA max|r| and sum r^2 go to part_max / part_sq
// (compute_residual_and_norms :1833-1970 — the convergence check, whose r nobody reads).
template <int T, int BX, int BY, bool RES = false, bool NORMS = false>
__global__ void __launch_bounds__(TbShape<T, BX, BY>::NTHREADS, 1)
k_sweep_tb(LevelGeom g, const double* __restrict__ u, const double* __restrict__ f,
           double* __restrict__ out, TbOmega<T> om, int ZC, double* __restrict__ part_max = nullptr,
           double* __restrict__ part_sq = nullptr) {
    using S = TbShape<T, BX, BY>;
    static_assert(T >= 1 && T <= 2, "f ring has 3 slots");
    CFDNN_DYN_SMEM(double, sm);
    constexpr int NT = S::NTHREADS;
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
    int off[2];                         // within-plane element offset (< 2^31)
    bool ldok[2], stok[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int gxp = gx0 + 2 * h;
        const bool okx = perx || (gxp >= 0 && gxp + 1 < g.Nx);
        const int gxw = perx ? wrap_idx(gxp, g.Nx) : gxp;
        ldok[h] = rowok && okx;
        off[h] = ldok[h] ? gyw * g.Px + gxw : 0;
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
    const bool edge = (bw | be) != 0u || selfS || selfN;
    // smem plane = [2 pairs][NT threads] double2: every LDS/STS.128 of a warp is contiguous
    const int tS = row > 0 ? t - S::NQ : t;               // thread holding my S / N neighbours
    const int tN = (active && row < S::TY - 1) ? t + S::NQ : t;
    const bool wsm = (lane == 0) && q > 0;                 // W neighbour must come from smem
    const bool esm = (lane == 31) && q < S::NQ - 1;
    constexpr int PL = 4 * NT;                             // doubles per smem plane

    // ---- pipeline state: lv[l][slot] = plane of level l that arrived in phase `slot` (mod 3)
    double lv[T][3][4], fr[3][4], nu_[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int l = 0; l < T; ++l)
#pragma unroll
        for (int k = 0; k < 3; ++k)
#pragma unroll
            for (int p = 0; p < 4; ++p) lv[l][k][p] = 0.0;
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int p = 0; p < 4; ++p) fr[k][p] = 0.0;

    // z range of input planes: virtual planes wrap when periodic, are clipped otherwise
    // a z side is "open" when planes beyond it exist: periodic wrap, or the halo planes of a slab
    // interface (exchanged T deep before the launch); a closed side is a physical face.
    const bool lo_open = perz || g.bc[4] == CFDNN_PBC_HALO, hi_open = perz || g.bc[5] == CFDNN_PBC_HALO;
    const int zi_first = lo_open ? z0 - T : max(z0 - T, 0);
    const int zi_last = z1 - 1 + T;
    const unsigned uplane = (unsigned)g.plane;
    auto load_plane = [&](const double* __restrict__ a, int z, double* dst) {
        int zw = z;                                        // z in [-T, Nz+T): one conditional wrap
        if (perz) zw = zw < 0 ? zw + g.Nz : (zw >= g.Nz ? zw - g.Nz : zw);
        const bool zok = zw >= (lo_open ? -g.Hz : 0) && zw < g.Nz + (hi_open ? g.Hz : 0);   // halo planes are stored planes
        const double* base = a + (size_t)(unsigned)(zw + g.Hz) * (size_t)uplane;   // one IMAD.WIDE.U32
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (zok && ldok[h]) ld_pair(base + off[h], dst[2 * h], dst[2 * h + 1]);   // else: keep the stale (finite) value
        }
    };
    load_plane(u, zi_first, nu_);
    load_plane(f, zi_first - 1, fr[0]);

    double nmx = 0.0, nsq = 0.0;
    int par = 0;
    for (int zb = zi_first; zb <= zi_last; zb += 3) {
#pragma unroll
        for (int ph = 0; ph < 3; ++ph) {
            const int zi = zb + ph;
            if (zi > zi_last) break;
            const int A = ph, C = (ph + 2) % 3, M = (ph + 1) % 3;   // arriving / centre / z-1 slots
#pragma unroll
            for (int p = 0; p < 4; ++p) lv[0][A][p] = nu_[p];
            if (zi < zi_last) {                            // prefetch next iteration's planes
                load_plane(u, zi + 1, nu_);
                load_plane(f, zi, fr[(ph + 1) % 3]);
            }
#pragma unroll
            for (int s = 0; s < T; ++s) {
                // level-s plane (zi - s) has arrived: publish it as next iteration's centre plane
                double* wbuf = sm + (size_t)(2 * s + par) * PL;
                const double* rbuf = sm + (size_t)(2 * s + (par ^ 1)) * PL;
                st_pair(wbuf + 2 * t, lv[s][A][0], lv[s][A][1]);
                st_pair(wbuf + 2 * NT + 2 * t, lv[s][A][2], lv[s][A][3]);
                const int zo = zi - 1 - s;                 // plane this stage produces (level s+1)
                const bool zvalid = (zo >= 0 || lo_open) && (zo < g.Nz || hi_open);
                double o[4];
                double uWq = __shfl_up_sync(0xffffffffu, lv[s][C][3], 1);
                double uEq = __shfl_down_sync(0xffffffffu, lv[s][C][0], 1);
                if (zvalid) {
                    if (wsm) uWq = rbuf[2 * NT + 2 * (t - 1) + 1];
                    if (esm) uEq = rbuf[2 * (t + 1)];
                    double n_[4], s_[4];
                    ld_pair(rbuf + 2 * tN, n_[0], n_[1]); ld_pair(rbuf + 2 * NT + 2 * tN, n_[2], n_[3]);
                    ld_pair(rbuf + 2 * tS, s_[0], s_[1]); ld_pair(rbuf + 2 * NT + 2 * tS, s_[2], s_[3]);
                    const double omega = om.w[s], om1 = 1.0 - omega;
                    const double* fv = fr[(ph + 3 - s) % 3];
                    const bool zlo = !lo_open && zo == 0, zhi = !hi_open && zo == g.Nz - 1;
                    if (!edge && !zlo && !zhi) {           // interior path: no boundary selects
#pragma unroll
                        for (int p = 0; p < 4; ++p) {
                            const double uW = p == 0 ? uWq : lv[s][C][p - 1];
                            const double uE = p == 3 ? uEq : lv[s][C][p + 1];
                            if (RES && s == T - 1)
                                o[p] = residual_point(g, lv[s][C][p], uW, uE, s_[p], n_[p], lv[s][M][p], lv[s][A][p], fv[p]);
                            else
                                o[p] = jacobi_point(g, lv[s][C][p], uW, uE, s_[p], n_[p], lv[s][M][p], lv[s][A][p],
                                                    fv[p], omega, om1);
                        }
                    } else {                               // faces: ghost = sign * adjacent interior value
#pragma unroll
                        for (int p = 0; p < 4; ++p) {
                            const double uc = lv[s][C][p];
                            double uW = p == 0 ? uWq : lv[s][C][p - 1];
                            double uE = p == 3 ? uEq : lv[s][C][p + 1];
                            if (bw >> p & 1u) uW = sxlo * uc;
                            if (be >> p & 1u) uE = sxhi * uc;
                            const double uS = selfS ? sylo * uc : s_[p];
                            const double uN = selfN ? syhi * uc : n_[p];
                            const double ub = zlo ? szlo * uc : lv[s][M][p];
                            const double uf = zhi ? szhi * uc : lv[s][A][p];
                            if (RES && s == T - 1) o[p] = residual_point(g, uc, uW, uE, uS, uN, ub, uf, fv[p]);
                            else o[p] = jacobi_point(g, uc, uW, uE, uS, uN, ub, uf, fv[p], omega, om1);
                        }
                    }
                } else {
#pragma unroll
                    for (int p = 0; p < 4; ++p) o[p] = lv[s][C][p];   // plane outside a closed z face: never consumed
                }
                if (s + 1 < T) {
#pragma unroll
                    for (int p = 0; p < 4; ++p) lv[(s + 1) % T][A][p] = o[p];
                } else {
                    const int zf = zi - T;                 // o = level-T plane zi - T
                    if (zf >= z0 && zf < z1) {
                        if (NORMS) {
#pragma unroll
                            for (int h = 0; h < 2; ++h)
                                if (stok[h]) {
                                    nmx = fmax(nmx, fmax(fabs(o[2 * h]), fabs(o[2 * h + 1])));
                                    nsq += o[2 * h] * o[2 * h];
                                    nsq += o[2 * h + 1] * o[2 * h + 1];
                                }
                        } else {
                            double* base = out + (size_t)(unsigned)(zf + g.Hz) * (size_t)uplane;
#pragma unroll
                            for (int h = 0; h < 2; ++h)
                                if (stok[h]) st_pair(base + off[h], o[2 * h], o[2 * h + 1]);
                        }
                    }
                }
            }
            __syncthreads();
            par ^= 1;
        }
    }
    if (NORMS) {
        const int bid = ((int)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        const double m = block_max(nmx, sm);
        if (t == 0) part_max[bid] = m;
        const double q2 = block_sum(nsq, sm);
        if (t == 0) part_sq[bid] = q2;
    }
}

}  // namespace cfdnn
