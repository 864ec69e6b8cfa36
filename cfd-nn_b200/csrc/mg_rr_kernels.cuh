// mg_rr_kernels.cuh — fused residual + full-weighting restriction of the multigrid V-cycle (FP64, sm_100a).
//
// The reference computes r = f - L u over the whole fine level (compute_residual
// src/poisson_solver_multigrid.cpp:1718-1831), fills the residual ghosts (apply_bc_to_residual
// :556-823: periodic wrap, else 0) and then gathers 27 residuals per coarse cell
// (restrict_residual :1972-2067).  Inside a V-cycle r has no other reader (the smoother's scratch
// aliases it), so this kernel never writes it to HBM: per fine cell it reads u and
// f once (16 B) and writes 1/8 coarse value (1 B) instead of 24 B + 9 B in two passes.
//
// Scheme: a CTA owns CX x CY coarse columns and marches through a chunk of z.  Per fine plane the
// u tile (with the ghosts resolved from the BC codes by the loader: +-1 halo of the residual tile,
// which itself reaches one fine cell below the first coarse cell) goes through a 3-plane ring in
// shared memory, the residual plane is evaluated into a second 3-plane ring (odd and even fine
// columns stored apart, so the stride-2 gather below is conflict-free), and after every second plane
// one thread per coarse cell sums its 27 residuals in the reference's order.  Global loads of plane
// z+2 are issued into registers before plane z is evaluated.
// Arithmetic: residual_point() (= k_residual) and the weighted sum of k_restrict, so the coarse
// right-hand side is bitwise that of the two-kernel path.
#pragma once
#include "mg_tb_kernels.cuh"

namespace cfdnn {

template <int CX, int CY>
struct RrShape {
    static constexpr int NT = CX * CY;
    static constexpr int UX = 2 * CX + 3, UY = 2 * CY + 3, UXP = UX + 1;
    static constexpr int RX = 2 * CX + 1, RY = 2 * CY + 1, RH = CX + 1, RXP = 2 * RH;
    static constexpr int UPL = UXP * UY, RPL = RXP * RY;
    static constexpr int NU = (UX * UY + NT - 1) / NT;
    static constexpr int NR = (RX * RY + NT - 1) / NT;
    static constexpr size_t SMEM = sizeof(double) * 3 * (size_t)(UPL + RPL);
};

// x / y index of a u-tile element -> stored index and ghost sign; i in [-2, N + tile overshoot]
__device__ __forceinline__ int rr_map(int i, int N, int lo, int hi, double& s) {
    if (i < 0) {
        if (lo == CFDNN_PBC_PERIODIC) return i + N;
        if (i == -1 && lo == CFDNN_PBC_DIRICHLET) s = -s;
        return 0;                                  // ghost -1 mirrors cell 0; -2 feeds no stored value
    }
    if (i >= N) {
        if (hi == CFDNN_PBC_PERIODIC) return min(i - N, N - 1);
        if (i == N && hi == CFDNN_PBC_DIRICHLET) s = -s;
        return N - 1;
    }
    return i;
}

template <int CX, int CY>
__global__ void __launch_bounds__(RrShape<CX, CY>::NT, 3)
k_residual_restrict(LevelGeom gf, LevelGeom gc, const double* __restrict__ u, const double* __restrict__ f,
                    double* __restrict__ fc, int kc_off, int KC, double* __restrict__ part_sum) {
    using S = RrShape<CX, CY>;
    CFDNN_DYN_SMEM(double, sm);
    double* const U = sm;
    double* const R = sm + 3 * S::UPL;
    const int t = threadIdx.x;
    const int c0x = (int)blockIdx.x * CX, c0y = (int)blockIdx.y * CY;
    const int nzc = gf.Nz >> 1;
    const int kc0 = (int)blockIdx.z * KC, kc1 = min(kc0 + KC, nzc);
    const bool perx = gf.bc[0] == CFDNN_PBC_PERIODIC, pery = gf.bc[2] == CFDNN_PBC_PERIODIC,
               perz = gf.bc[4] == CFDNN_PBC_PERIODIC;
    const bool lo_halo = gf.bc[4] == CFDNN_PBC_HALO, hi_halo = gf.bc[5] == CFDNN_PBC_HALO;
    const size_t plane = (size_t)gf.plane;

    // ---- loader elements of the u tile: offset inside a plane and ghost sign
    int uoff[S::NU], usm[S::NU];
    double usg[S::NU];
#pragma unroll
    for (int n = 0; n < S::NU; ++n) {
        const int e = t + n * S::NT;
        uoff[n] = -1; usm[n] = 0; usg[n] = 1.0;
        if (e < S::UX * S::UY) {
            const int tx = e % S::UX, ty = e / S::UX;
            double s = 1.0;
            const int ix = rr_map(2 * c0x - 2 + tx, gf.Nx, gf.bc[0], gf.bc[1], s);
            const int iy = rr_map(2 * c0y - 2 + ty, gf.Ny, gf.bc[2], gf.bc[3], s);
            uoff[n] = iy * gf.Px + ix;
            usm[n] = ty * S::UXP + tx;
            usg[n] = s;
        }
    }
    // ---- residual elements: centre in the u tile, slot in the r plane, offset of f
    int roff[S::NR], rcen[S::NR], rslot[S::NR];
    bool rok[S::NR];
#pragma unroll
    for (int n = 0; n < S::NR; ++n) {
        const int e = t + n * S::NT;
        roff[n] = 0; rcen[n] = 0; rslot[n] = -1; rok[n] = false;
        if (e < S::RX * S::RY) {
            const int rx = e % S::RX, ry = e / S::RX;
            const int fx = 2 * c0x - 1 + rx, fy = 2 * c0y - 1 + ry;
            rok[n] = (fx >= 0 || perx) && fx < gf.Nx && (fy >= 0 || pery) && fy < gf.Ny;   // else ghost = 0 (:556-823)
            const int wx = fx < 0 ? fx + gf.Nx : fx, wy = fy < 0 ? fy + gf.Ny : fy;
            roff[n] = rok[n] ? wy * gf.Px + wx : 0;
            rcen[n] = (ry + 1) * S::UXP + rx + 1;
            rslot[n] = ry * S::RXP + (rx & 1) * S::RH + (rx >> 1);   // even rx = odd fine column
        }
    }

    // u plane kz in [-2, Nz]: stored plane and ghost sign
    auto ldu = [&](int kz, double* dst) {
        double sz = 1.0;
        int pz = kz;
        if (kz < 0) {
            if (perz) pz = kz + gf.Nz;
            else if (lo_halo) pz = kz;                       // halo planes are stored planes
            else { pz = 0; if (kz == -1 && gf.bc[4] == CFDNN_PBC_DIRICHLET) sz = -1.0; }
        } else if (kz >= gf.Nz) {
            if (perz) pz = kz - gf.Nz;
            else if (hi_halo) pz = kz;
            else { pz = gf.Nz - 1; if (gf.bc[5] == CFDNN_PBC_DIRICHLET) sz = -1.0; }
        }
        const double* base = u + (size_t)(pz + gf.Hz) * plane;
#pragma unroll
        for (int n = 0; n < S::NU; ++n)
            if (uoff[n] >= 0) dst[n] = (usg[n] * sz) * base[uoff[n]];
    };
    auto stu = [&](int slot, const double* src) {
        double* d = U + slot * S::UPL;
#pragma unroll
        for (int n = 0; n < S::NU; ++n)
            if (uoff[n] >= 0) d[usm[n]] = src[n];
    };
    // f of residual plane kf in [-1, Nz)
    auto ldf = [&](int kf, double* dst) {
        int pz = kf;
        if (kf < 0) {
            if (perz) pz = kf + gf.Nz;
            else if (!lo_halo) return;                       // ghost plane of a physical face: residual 0
        }
        const double* base = f + (size_t)(pz + gf.Hz) * plane;
#pragma unroll
        for (int n = 0; n < S::NR; ++n)
            if (rok[n]) dst[n] = base[roff[n]];
    };

    const int kfirst = 2 * kc0 - 1, klast = 2 * kc1 - 1;
    double nu_[S::NU], nf_[S::NR];
#pragma unroll
    for (int n = 0; n < S::NU; ++n) nu_[n] = 0.0;
#pragma unroll
    for (int n = 0; n < S::NR; ++n) nf_[n] = 0.0;
    int ub = 0, uc = 1, uf = 2;          // ring slots of u planes kf-1, kf, kf+1
    int r0 = 0, r1 = 1, r2 = 2;          // ring slots of r planes kf-2, kf-1, kf
    ldu(kfirst - 1, nu_); stu(ub, nu_);
    ldu(kfirst, nu_);     stu(uc, nu_);
    ldu(kfirst + 1, nu_);
    ldf(kfirst, nf_);

    const int cx = t % CX, cy = t / CX;
    const int ic = c0x + cx, jc = c0y + cy;
    const bool cok = ic < gc.Nx && jc < gc.Ny;
    double acc = 0.0;

    for (int kf = kfirst; kf <= klast; ++kf) {
        stu(uf, nu_);
        double fcur[S::NR];
#pragma unroll
        for (int n = 0; n < S::NR; ++n) fcur[n] = nf_[n];
        if (kf < klast) {
            ldu(kf + 2, nu_);
            ldf(kf + 1, nf_);
        }
        __syncthreads();
        {
            const double* Ub = U + ub * S::UPL;
            const double* Uc = U + uc * S::UPL;
            const double* Uf = U + uf * S::UPL;
            double* Rw = R + r2 * S::RPL;
            const bool zok = kf >= 0 || perz || lo_halo;
#pragma unroll
            for (int n = 0; n < S::NR; ++n) {
                if (rslot[n] >= 0) {
                    double rv = 0.0;
                    if (zok && rok[n]) {
                        const int c = rcen[n];
                        rv = residual_point(gf, Uc[c], Uc[c - 1], Uc[c + 1], Uc[c - S::UXP], Uc[c + S::UXP], Ub[c],
                                            Uf[c], fcur[n]);
                    }
                    Rw[rslot[n]] = rv;
                }
            }
        }
        __syncthreads();
        if (kf > kfirst && ((kf - kfirst) & 1) == 0) {       // planes 2kc-1, 2kc, 2kc+1 are in the ring
            const int kc = (kf - 1) >> 1;
            if (cok) {
                const double* Rp[3] = {R + r0 * S::RPL, R + r1 * S::RPL, R + r2 * S::RPL};
                double v[3][3][3];   // [dk][dj][di]
#pragma unroll
                for (int c = 0; c < 3; ++c)
#pragma unroll
                    for (int b = 0; b < 3; ++b) {
                        const double* row = Rp[c] + (2 * cy + b) * S::RXP;
                        v[c][b][0] = row[cx];
                        v[c][b][1] = row[S::RH + cx];
                        v[c][b][2] = row[cx + 1];
                    }
#define RV(di, dj, dk) v[(dk) + 1][(dj) + 1][(di) + 1]
                double sum = 0.0;            // restrict_residual :2035-2062, same order as k_restrict
                sum += 0.125 * RV(0, 0, 0);
                sum += 0.0625 * (RV(-1, 0, 0) + RV(1, 0, 0) + RV(0, -1, 0) + RV(0, 1, 0) + RV(0, 0, -1) + RV(0, 0, 1));
                sum += 0.03125 * (RV(-1, -1, 0) + RV(1, -1, 0) + RV(-1, 1, 0) + RV(1, 1, 0)
                                + RV(-1, 0, -1) + RV(1, 0, -1) + RV(-1, 0, 1) + RV(1, 0, 1)
                                + RV(0, -1, -1) + RV(0, 1, -1) + RV(0, -1, 1) + RV(0, 1, 1));
                sum += 0.015625 * (RV(-1, -1, -1) + RV(1, -1, -1) + RV(-1, 1, -1) + RV(1, 1, -1)
                                 + RV(-1, -1, 1) + RV(1, -1, 1) + RV(-1, 1, 1) + RV(1, 1, 1));
#undef RV
                fc[lidx(gc, ic, jc, kc + kc_off)] = sum;
                acc += sum;
            }
        }
        { const int x = ub; ub = uc; uc = uf; uf = x; }
        { const int x = r0; r0 = r1; r1 = r2; r2 = x; }
    }
    if (part_sum) {                                          // make_rhs_mean_free partials (:3479-3548)
        __syncthreads();
        const double s = block_sum(acc, sm);
        if (t == 0) part_sum[((int)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = s;
    }
}

}  // namespace cfdnn
