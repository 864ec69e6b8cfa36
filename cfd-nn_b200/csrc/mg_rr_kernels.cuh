// mg_rr_kernels.cuh — fused residual + full-weighting restriction of the multigrid V-cycle (FP64, sm_100a).
//
// The reference computes r = f - L u over the whole fine level (compute_residual
// src/poisson_solver_multigrid.cpp:1718-1831), fills the residual ghosts (apply_bc_to_residual
// :556-823: periodic wrap, else 0) and then gathers 27 residuals per coarse cell
// (restrict_residual :1972-2067).  Inside a V-cycle r has no other reader (the smoother's scratch
// aliases it), so this kernel never writes it to HBM: per fine cell it reads u and
// f once (16 B) and writes 1/8 coarse value (1 B) instead of 24 B + 9 B in two passes.
//
// Scheme: a CTA owns CX x CY coarse columns plus one column / row of "virtual" coarse cells on the low
// side (the 27-point stencil of coarse cell ic covers fine cells 2ic-1 .. 2ic+1) and marches through a
// chunk of z.  A thread owns the 2 x 2 fine cells of one coarse column: their u values of the planes
// z-1, z, z+1 and their residuals of the last three planes live in registers.  Per fine plane the u
// tile (ghosts resolved from the BC codes by the loader; even and odd columns stored apart so every
// access is a conflict-free LDS.64) goes through a 3-slot ring in shared memory for the x / y
// neighbours, the residual planes go through a 4-slot ring, and after every second plane a thread gathers
// its 27 residuals (15 of them from the threads below / left) and sums them in the reference's order.  One __syncthreads per plane; global loads of plane z+2 are in flight while plane z is
// evaluated (ghost signs are applied when the registers are published, not when they are loaded).
// Arithmetic: residual_point() (= k_residual) and the weighted sum of k_restrict, so the coarse
// right-hand side is bitwise that of the two-kernel path.
#pragma once
#include "mg_tb_kernels.cuh"

namespace cfdnn {

template <int CX, int CY>
struct RrShape {
    static constexpr int VX = CX + 1, VY = CY + 1;          // virtual coarse tile
    static constexpr int NACTIVE = VX * VY;
    static constexpr int NT = (NACTIVE + 31) / 32 * 32;
    static constexpr int NP = CX + 3;                       // column pairs of the u tile: fine x in [2c0x-4, 2c0x+2CX+2)
    static constexpr int UH = NP + 1;                       // offset of the odd-column half of a row
    static constexpr int UP = 2 * UH;                       // row pitch
    static constexpr int UY = 2 * CY + 4;                   // fine y in [2c0y-3, 2c0y+2CY+1)
    static constexpr int UPL = UP * UY;
    static constexpr int NU = (NP * UY + NT - 1) / NT;      // loader pairs per thread
    static constexpr int RPL = 4 * NT;
    static constexpr size_t SMEM = sizeof(double) * (size_t)(3 * UPL + 4 * RPL);
};

// x / y index of a u-tile element -> stored index and ghost sign; i in [-4, N + tile overshoot]
__device__ __forceinline__ int rr_map(int i, int N, int lo, int hi, double& s) {
    if (i < 0) {
        if (lo == CFDNN_PBC_PERIODIC) return i + N;
        if (i == -1 && lo == CFDNN_PBC_DIRICHLET) s = -s;
        return 0;                                  // ghost -1 mirrors cell 0; further out feeds no stored value
    }
    if (i >= N) {
        if (hi == CFDNN_PBC_PERIODIC) return min(i - N, N - 1);
        if (i == N && hi == CFDNN_PBC_DIRICHLET) s = -s;
        return N - 1;
    }
    return i;
}

// OPEN = true: x and y are periodic (no ghost pairs, no sign flips, every tile cell is a stored cell) — the
// loader and the residual stage carry no boundary branches.
template <int CX, int CY, bool OPEN>
__global__ void __launch_bounds__(RrShape<CX, CY>::NT, 2)
k_residual_restrict(LevelGeom gf, LevelGeom gc, const double* __restrict__ u, const double* __restrict__ f,
                    double* __restrict__ fc, int kc_off, int KC, double* __restrict__ part_sum) {
    using S = RrShape<CX, CY>;
    CFDNN_DYN_SMEM(double, sm);
    double* const U = sm;                   // [3][UY][even half | odd half]
    double* const R = sm + 3 * S::UPL;      // [4][4][NT]
    const int t = threadIdx.x;
    const int c0x = (int)blockIdx.x * CX, c0y = (int)blockIdx.y * CY;
    const int nzc = gf.Nz >> 1;
    const int kc0 = (int)blockIdx.z * KC, kc1 = min(kc0 + KC, nzc);
    const bool perx = gf.bc[0] == CFDNN_PBC_PERIODIC, pery = gf.bc[2] == CFDNN_PBC_PERIODIC,
               perz = gf.bc[4] == CFDNN_PBC_PERIODIC;
    const bool lo_halo = gf.bc[4] == CFDNN_PBC_HALO, hi_halo = gf.bc[5] == CFDNN_PBC_HALO;
    const unsigned uplane = (unsigned)gf.plane;

    // ---- loader slots: one aligned column pair of one tile row each
    int urow[S::NU], ucol[S::NU], usm[S::NU];     // row offset, offset in the plane (or raw x of a ghost pair), smem index / -1
    bool ughost[S::NU], uneg[S::NU];              // pair touches a non-periodic x ghost; row ghost flips the sign
#pragma unroll
    for (int n = 0; n < S::NU; ++n) {
        const int e = t + n * S::NT;
        urow[n] = 0; ucol[n] = 0; usm[n] = -1; ughost[n] = false; uneg[n] = false;
        if (e < S::NP * S::UY) {
            const int m = e % S::NP, ty = e / S::NP;
            const int gx = 2 * c0x - 4 + 2 * m;
            double s = 1.0;
            const int iy = rr_map(2 * c0y - 3 + ty, gf.Ny, gf.bc[2], gf.bc[3], s);
            urow[n] = iy * gf.Px;
            uneg[n] = s < 0.0;
            usm[n] = ty * S::UP + m;
            if (OPEN || perx) {
                ucol[n] = urow[n] + (gx < 0 ? gx + gf.Nx : (gx >= gf.Nx ? min(gx - gf.Nx, gf.Nx - 2) : gx));
            } else if (gx >= 0 && gx + 1 < gf.Nx) {
                ucol[n] = urow[n] + gx;
            } else {
                ucol[n] = gx; ughost[n] = true;
            }
        }
    }
    // u plane kz in [-2, Nz]: stored plane and ghost sign
    auto zmap = [&](int kz, double& sz) {
        if (kz < 0) {
            if (perz) return kz + gf.Nz;
            if (lo_halo) return kz;                          // halo planes are stored planes
            if (kz == -1 && gf.bc[4] == CFDNN_PBC_DIRICHLET) sz = -1.0;
            return 0;
        }
        if (kz >= gf.Nz) {
            if (perz) return kz - gf.Nz;
            if (hi_halo) return kz;
            if (gf.bc[5] == CFDNN_PBC_DIRICHLET) sz = -1.0;
            return gf.Nz - 1;
        }
        return kz;
    };
    auto ldu = [&](int kz, double (*dst)[2]) {               // raw values (ghost pairs: x sign applied here)
        double sz = 1.0;
        const double* base = u + (size_t)(unsigned)(zmap(kz, sz) + gf.Hz) * (size_t)uplane;   // one IMAD.WIDE.U32
#pragma unroll
        for (int n = 0; n < S::NU; ++n) {
            if (usm[n] < 0) continue;
            if (OPEN || !ughost[n]) {
                ld_pair(base + (unsigned)ucol[n], dst[n][0], dst[n][1]);
            } else {
                double s0 = 1.0, s1 = 1.0;
                const int i0 = rr_map(ucol[n], gf.Nx, gf.bc[0], gf.bc[1], s0);
                const int i1 = rr_map(ucol[n] + 1, gf.Nx, gf.bc[0], gf.bc[1], s1);
                dst[n][0] = s0 * base[urow[n] + i0];
                dst[n][1] = s1 * base[urow[n] + i1];
            }
        }
    };
    auto stu = [&](int kz, int slot, const double (*src)[2]) {
        double sz = 1.0;
        (void)zmap(kz, sz);
        double* d = U + slot * S::UPL;
#pragma unroll
        for (int n = 0; n < S::NU; ++n) {
            if (usm[n] < 0) continue;
            const double sg = (!OPEN && uneg[n]) ? -sz : sz;
            d[usm[n]] = sg * src[n][0];
            d[usm[n] + S::UH] = sg * src[n][1];
        }
    };

    // ---- the thread's coarse column and its 2 x 2 fine cells
    const bool active = t < S::NACTIVE;
    const int vx = active ? t % S::VX : 0, vy = active ? t / S::VX : 0;
    const int ic = c0x - 1 + vx, jc = c0y - 1 + vy;
    const int x0 = 2 * ic, y0 = 2 * jc;
    const bool xok = active && (x0 >= 0 || perx) && x0 + 1 < gf.Nx;   // residual ghosts: periodic wrap, else 0 (:556-823)
    bool yok[2];
    int foff[2];
#pragma unroll
    for (int b = 0; b < 2; ++b) {
        const int y = y0 + b;
        yok[b] = xok && (y >= 0 || pery) && y < gf.Ny;
        const int wx = x0 < 0 ? x0 + gf.Nx : x0, wy = y < 0 ? y + gf.Ny : y;
        foff[b] = yok[b] ? wy * gf.Px + wx : 0;
    }
    auto ldf = [&](int kf, double (*dst)[2]) {               // f of residual plane kf in [-1, Nz)
        int pz = kf;
        if (kf < 0) {
            if (perz) pz = kf + gf.Nz;
            else if (!lo_halo) return;                       // ghost plane of a physical face: residual 0
        }
        const double* base = f + (size_t)(unsigned)(pz + gf.Hz) * (size_t)uplane;
#pragma unroll
        for (int b = 0; b < 2; ++b)
            if (yok[b]) ld_pair(base + (unsigned)foff[b], dst[b][0], dst[b][1]);
    };
    const int m0 = vx + 1, ty0 = 2 * vy + 1;                 // own pair column / first own row in the u tile
    const bool cok = active && vx >= 1 && vy >= 1 && ic < gc.Nx && jc < gc.Ny;

    const int kfirst = 2 * kc0 - 1, klast = 2 * kc1 - 1;     // residual planes of this chunk
    double nu_[S::NU][2];
#pragma unroll
    for (int n = 0; n < S::NU; ++n) nu_[n][0] = nu_[n][1] = 0.0;
    // rings in registers, rotated by the phase of the 3-way unrolled plane loop (no moves):
    //   X[ph] = own u of plane kf+1, X[ph+2] = plane kf, X[ph+1] = plane kf-1;  F[ph] = f of plane kf
    double X[3][2][2], F[3][2][2];
#pragma unroll
    for (int p = 0; p < 3; ++p)
#pragma unroll
        for (int b = 0; b < 2; ++b) X[p][b][0] = X[p][b][1] = F[p][b][0] = F[p][b][1] = 0.0;
    double acc = 0.0;
    // per-thread shared-memory offsets (doubles)
    const int oC = ty0 * S::UP + m0;                         // own even-column element of row y0
    const int tg = t - S::VX;                                // thread owning the coarse column below
    ldu(kfirst - 1, nu_);
    ldf(kfirst, F[2]);                                       // plane kfirst is evaluated in phase 2

    for (int kb = kfirst - 2; kb <= klast + 1; kb += 3) {
#pragma unroll
        for (int ph = 0; ph < 3; ++ph) {
            const int kf = kb + ph;
            if (kf > klast + 1) break;
            const int Pn = ph, Cn = (ph + 2) % 3, Mn = (ph + 1) % 3;
            if (kf <= klast) {
                stu(kf + 1, ph, nu_);                        // publish u plane kf+1 into ring slot ph
                if (kf < klast) ldu(kf + 2, nu_);
                if (kf >= kfirst && kf < klast) ldf(kf + 1, F[(ph + 1) % 3]);
            }
            __syncthreads();
            // ---- planes kf-3 .. kf-1 = 2kc-1 .. 2kc+1: full weighting (restrict_residual :2035-2062)
            if (kf >= kfirst + 3 && ((kf - kfirst) & 1) && cok) {
                const int kc = (kf - 2) >> 1;
                double v[3][3][3];   // [dk][dj][di]
#pragma unroll
                for (int p = 0; p < 3; ++p) {
                    const double* Rp = R + ((kf - 3 + p - kfirst) & 3) * S::RPL;
                    v[p][0][0] = Rp[3 * S::NT + tg - 1];             // (x0-1, y0-1): r11 of (vx-1, vy-1)
                    v[p][0][1] = Rp[2 * S::NT + tg];                 // (x0,   y0-1): r10 of (vx, vy-1)
                    v[p][0][2] = Rp[3 * S::NT + tg];                 // (x0+1, y0-1): r11 of (vx, vy-1)
                    v[p][1][0] = Rp[1 * S::NT + t - 1];              // (x0-1, y0  ): r01 of (vx-1, vy)
                    v[p][2][0] = Rp[3 * S::NT + t - 1];              // (x0-1, y0+1): r11 of (vx-1, vy)
                    v[p][1][1] = Rp[t];              v[p][1][2] = Rp[S::NT + t];        // own 2 x 2
                    v[p][2][1] = Rp[2 * S::NT + t];  v[p][2][2] = Rp[3 * S::NT + t];
                }
#define RV(di, dj, dk) v[(dk) + 1][(dj) + 1][(di) + 1]
                double sum = 0.0;                            // same order as k_restrict
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
            if (kf <= klast) {
                // ---- own u of plane kf+1, then the residual of plane kf
                const double* Uf = U + Pn * S::UPL + oC;
#pragma unroll
                for (int b = 0; b < 2; ++b) {
                    X[Pn][b][0] = Uf[b * S::UP];
                    X[Pn][b][1] = Uf[b * S::UP + S::UH];
                }
                if (kf >= kfirst) {
                    const double* Uc = U + Cn * S::UPL + oC;
                    double r[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
                    if ((OPEN || active) && (kf >= 0 || perz || lo_halo)) {
                        const double s0 = Uc[-S::UP], s1 = Uc[-S::UP + S::UH];
                        const double n0 = Uc[2 * S::UP], n1 = Uc[2 * S::UP + S::UH];
#pragma unroll
                        for (int b = 0; b < 2; ++b) {
                            if (!OPEN && !yok[b]) continue;
                            const double w = Uc[b * S::UP + S::UH - 1];
                            const double e = Uc[b * S::UP + 1];
                            const double uS0 = b == 0 ? s0 : X[Cn][0][0], uS1 = b == 0 ? s1 : X[Cn][0][1];
                            const double uN0 = b == 1 ? n0 : X[Cn][1][0], uN1 = b == 1 ? n1 : X[Cn][1][1];
                            r[b][0] = residual_point(gf, X[Cn][b][0], w, X[Cn][b][1], uS0, uN0, X[Mn][b][0], X[Pn][b][0],
                                                     F[ph][b][0]);
                            r[b][1] = residual_point(gf, X[Cn][b][1], X[Cn][b][0], e, uS1, uN1, X[Mn][b][1], X[Pn][b][1],
                                                     F[ph][b][1]);
                        }
                    }
                    double* Rw = R + ((kf - kfirst) & 3) * S::RPL + t;
                    Rw[0] = r[0][0];                         // [row][col] = r00, r01, r10, r11
                    Rw[S::NT] = r[0][1];
                    Rw[2 * S::NT] = r[1][0];
                    Rw[3 * S::NT] = r[1][1];
                }
            }
        }
    }
    if (part_sum) {                                          // make_rhs_mean_free partials (:3479-3548)
        __syncthreads();
        const double s = block_sum(acc, sm);
        if (t == 0) part_sum[((int)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = s;
    }
}

}  // namespace cfdnn
