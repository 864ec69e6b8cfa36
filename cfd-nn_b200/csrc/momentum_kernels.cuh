// momentum_kernels.cuh — fused momentum update, z-marching tile version (FP64, sm_100a).
//
//   out = in + dt * ( -conv(vel) + diff(vel, nu_eff) + f )   on interior faces of u, v, w
// = compute_convective_term + compute_diffusive_term + the advance loop of euler_substep
// (reference src/solver_operators.cpp:493-1070, src/solver_time.cpp:269-516; per-face formulas
// include/solver_kernels.hpp:845-1022 central, :1377-1569 skew-symmetric, :2199-2337 diffusion).
//
// The one-thread-per-point kernel it replaces (k_momentum, solver_kernels.cuh) gathered 93 values per
// cell through the LSU and rebuilt every address with 64-bit multiplies: instruction-bound at 0.26 of
// the HBM roofline.  Here
//   * a CTA owns a BX x BY column of cells and marches through a chunk of z planes; a thread owns one
//     (i, j) column and keeps the planes k-1, k, k+1 of u, v, w, nu_eff of that column in registers
//     (each value is loaded from HBM once per plane, 4 loads per cell instead of 93);
//   * the 31 in-plane neighbours the three face stencils need come from one shared-memory tile per
//     field and plane (ring of 4 z slots, one __syncthreads per plane), addressed with constant offsets
//     from three per-iteration slot registers; the global loads of plane k+2 (and of `in` on plane k+1) are
//     issued before plane k is evaluated and published into the ring afterwards, so HBM latency hides
//     behind the ~230 FP64 instructions of a cell;
//   * the three face updates of a cell share every loaded value (43 distinct inputs in all);
//   * EXACT = true keeps the reference's expression order and true divisions (correctly rounded fdiv):
//     bit-identical to the x86 build — required on y-stretched meshes, whose Poisson solver amplifies a
//     last-bit change.  EXACT = false (uniform meshes): divisions by the constant spacings
//     become multiplications by their reciprocals and products feed fused multiply-adds — ~225 instead of
//     ~345 FP64 instructions per cell, every face value within a few ulp of the reference's.
//
// Index space: raw array indices (I, J, K) over the ghost-inclusive arrays, like k_momentum: entries that
// are not interior faces are copied from vel when copy_ghosts is set (RK stages 2-3 write into a scratch
// buffer that is then swapped with vel, see Solver::momentum).
#pragma once
#include "solver.cuh"
#include <type_traits>

#ifndef CIDX
#define CIDX(i, j, k) ((size_t)(k) * g.cp + (size_t)(j) * g.cs + (i))
#endif

namespace cfdnn {

struct MomCoef {                  // spacings as divisor/reciprocal pairs and folded constants
    Div dx, dy, dz, dx_2, dy_2, dz_2, dx2, dy2, dz2;
    double h_rdx, h_rdy, h_rdz;           // 0.5 / dx ...      (EXACT = false)
    double h_r2dx, h_r2dy, h_r2dz;        // 0.5 / (2 dx) ...
    double dt, fx, fy, fz;
};

template <bool EXACT>
__device__ __forceinline__ double mom_fma(double a, double b, double c) {
    if (EXACT) return a * b + c;          // two roundings, as the x86 build (-fmad=false)
    return fma(a, b, c);
}

// One direction of the skew-symmetric form 0.5 * (advective + conservative) for face value q:
//   adv = a * (q_p - q_m) / (2h),  cons = (a_hi * 0.5 (q + q_p) - a_lo * 0.5 (q + q_m)) / h
// a: advecting velocity at the face, a_lo / a_hi: advecting velocity at the two neighbouring flux points.
template <bool EXACT>
__device__ __forceinline__ double skew_dir(double a, double a_lo, double a_hi, double q, double q_m, double q_p,
                                           const Div& h, const Div& h2, double h_rh, double h_r2h) {
    if (EXACT) {
        const double adv = a * (q_p - q_m) / h2;
        const double lo = 0.5 * (q + q_m), hi = 0.5 * (q + q_p);
        const double cons = (a_hi * hi - a_lo * lo) / h;
        return 0.5 * (adv + cons);
    } else {
        const double lo = 0.5 * (q + q_m), hi = 0.5 * (q + q_p);
        const double fl = fma(a_hi, hi, -(a_lo * lo));
        return fma(fl, h_rh, (a * (q_p - q_m)) * h_r2h);
    }
}
// own direction of the face: the advecting velocity is q itself, flux = (0.5 (q + q_p))^2
template <bool EXACT>
__device__ __forceinline__ double skew_own(double q, double q_m, double q_p, const Div& h, const Div& h2,
                                           double h_rh, double h_r2h) {
    if (EXACT) {
        const double adv = q * (q_p - q_m) / h2;
        const double lo = 0.5 * (q_m + q), hi = 0.5 * (q + q_p);
        const double cons = (hi * hi - lo * lo) / h;
        return 0.5 * (adv + cons);
    } else {
        const double lo = 0.5 * (q_m + q), hi = 0.5 * (q + q_p);
        const double fl = fma(hi, hi, -(lo * lo));
        return fma(fl, h_rh, (q * (q_p - q_m)) * h_r2h);
    }
}
template <bool EXACT>
__device__ __forceinline__ double central_dir(double a, double q_m, double q_p, const Div& h2) {
    if (EXACT) return a * ((q_p - q_m) / h2);
    return a * ((q_p - q_m) * h2.r);
}
// (nu_hi (q_p - q) - nu_lo (q - q_m)) / h^2
template <bool EXACT>
__device__ __forceinline__ double diff_dir(double nu_lo, double nu_hi, double q, double q_m, double q_p, const Div& hh) {
    if (EXACT) return (nu_hi * (q_p - q) - nu_lo * (q - q_m)) / hh;
    return fma(nu_hi, q_p - q, -(nu_lo * (q - q_m))) * hh.r;
}

template <int BX, int BY>
struct MomShape {
    static constexpr int TX = BX + 2, TY = BY + 2;
    static constexpr int PL = TX * TY;                    // doubles per tile plane
    static constexpr int NT = BX * BY;
    static constexpr int NHALO = PL - NT;
    static constexpr size_t SMEM = sizeof(double) * 4 /*fields*/ * 4 /*slots*/ * PL;
    static_assert(NHALO <= NT, "one halo element per thread at most");
};

// SKEW: skew-symmetric (true) or central (false) convection.  Launch: grid (ceil((ex+1)/BX), ceil((ey+1)/BY), z chunks),
// block BX*BY threads.  zc = planes per chunk.
// XZ: implicit y-diffusion — the explicit term carries d2/dx2 + d2/dz2 only (src/solver_operators.cpp:915)
template <bool SKEW, bool EXACT, bool XZ, int BX, int BY>
__global__ void __launch_bounds__(BX * BY, 2)
k_momentum_z(Geom g, MomCoef c, Vel vel, const double* __restrict__ nu, Vel in, Vel out, int copy_ghosts, int zc) {
    using S = MomShape<BX, BY>;
    CFDNN_DYN_SMEM(double, sm);
    constexpr int PL = S::PL, TX = S::TX;
    constexpr int FU = 0, FV = 4 * PL, FW = 8 * PL, FN = 12 * PL;    // field bases (4 slots each)
    const int Ng = g.Ng;
    const int ex = g.Nx + 2 * Ng, ey = g.Ny + 2 * Ng, ez = g.Nz + 2 * Ng;   // cell-array extents
    const int tx = threadIdx.x % BX, ty = threadIdx.x / BX;
    const int i0 = (int)blockIdx.x * BX, j0 = (int)blockIdx.y * BY;
    const int i = i0 + tx, j = j0 + ty;
    const int k0 = (int)blockIdx.z * zc, k1 = min(k0 + zc, ez + 1);   // raw planes [k0, k1), K in [0, ez]

    // ---- per-thread load geometry: own column and (for the first NHALO threads) one halo column
    // clamped coordinates per field (u has ex+1 columns, v ey+1 rows, w ez+1 planes)
    const int iu = min(i, ex), ic = min(i, ex - 1);
    const int jv = min(j, ey), jc = min(j, ey - 1);
    const unsigned own_u = (unsigned)(jc * g.us + iu), own_v = (unsigned)(jv * g.vs + ic);
    const unsigned own_w = (unsigned)(jc * g.ws + ic), own_n = (unsigned)(jc * g.cs + ic);
    const int t_own = (ty + 1) * TX + tx + 1;
    int t_halo = -1;
    unsigned hal_u = 0, hal_v = 0, hal_w = 0, hal_n = 0;
    if ((int)threadIdx.x < S::NHALO) {
        int h = threadIdx.x, hx, hy;
        if (h < TX) { hx = h; hy = 0; }
        else if (h < 2 * TX) { hx = h - TX; hy = S::TY - 1; }
        else if (h < 2 * TX + BY) { hx = 0; hy = h - 2 * TX + 1; }
        else { hx = TX - 1; hy = h - 2 * TX - BY + 1; }
        t_halo = hy * TX + hx;
        const int hi_ = i0 + hx - 1, hj_ = j0 + hy - 1;
        const int hiu = min(max(hi_, 0), ex), hic = min(max(hi_, 0), ex - 1);
        const int hjv = min(max(hj_, 0), ey), hjc = min(max(hj_, 0), ey - 1);
        hal_u = (unsigned)(hjc * g.us + hiu); hal_v = (unsigned)(hjv * g.vs + hic);
        hal_w = (unsigned)(hjc * g.ws + hic); hal_n = (unsigned)(hjc * g.cs + hic);
    }
    // which entries this column owns, and which of them are interior faces
    const bool col_u = i <= ex && j < ey, col_v = i < ex && j <= ey, col_w = i < ex && j < ey;
    const bool ix_c = i >= Ng && i < Ng + g.Nx, iy_c = j >= Ng && j < Ng + g.Ny;
    const bool ix_f = i >= Ng && i <= Ng + g.Nx, iy_f = j >= Ng && j <= Ng + g.Ny;
    const bool int_u = ix_f && iy_c, int_v = ix_c && iy_f, int_w = ix_c && iy_c;     // and the k condition, per plane

    // Own-column planes live in register rings of 4 slots (slot = (plane - k0 + 1) & 3, the same numbering as the
    // shared-memory ring); the z loop is unrolled by 4 so that every slot index is a compile-time constant: no
    // register moves between iterations (a move of a just-prefetched value would stall on its load) and constant
    // shared-memory offsets.  fetch() issues the global loads of a plane (clamped per field) into a ring slot and the
    // halo registers; publish() stores them into the tile ring one evaluation later.
    double RU[4], RV[4], RW[4], RN[4];
    double hu = 0.0, hv = 0.0, hw = 0.0, hn = 0.0;          // halo column (first NHALO threads)
    double INU[2] = {0.0, 0.0}, INV[2] = {0.0, 0.0}, INW[2] = {0.0, 0.0};   // `in` of the own faces, planes K / K+1
    auto fetch = [&](int K, double& pu, double& pv, double& pw, double& pn) {
        const int ku = min(max(K, 0), ez - 1), kw = min(max(K, 0), ez);
        const double* bu = vel.u + (size_t)ku * (size_t)g.up;
        const double* bv = vel.v + (size_t)ku * (size_t)g.vp;
        const double* bw = vel.w + (size_t)kw * (size_t)g.wp;
        const double* bn = nu + (size_t)ku * (size_t)g.cp;
        pu = bu[own_u]; pv = bv[own_v]; pw = bw[own_w]; pn = bn[own_n];
        if (t_halo >= 0) { hu = bu[hal_u]; hv = bv[hal_v]; hw = bw[hal_w]; hn = bn[hal_n]; }
    };
    auto publish = [&](int slot, double pu, double pv, double pw, double pn) {
        double* t = sm + slot * PL;
        t[FU + t_own] = pu; t[FV + t_own] = pv; t[FW + t_own] = pw; t[FN + t_own] = pn;
        if (t_halo >= 0) { t[FU + t_halo] = hu; t[FV + t_halo] = hv; t[FW + t_halo] = hw; t[FN + t_halo] = hn; }
    };
    auto fetch_in = [&](int K, double& a, double& b, double& cc) {       // clamped; unused where the face is not interior
        const int ku = min(K, ez - 1), kw = min(K, ez);
        if (int_u) a = in.u[(size_t)ku * (size_t)g.up + own_u];
        if (int_v) b = in.v[(size_t)ku * (size_t)g.vp + own_v];
        if (int_w) cc = in.w[(size_t)kw * (size_t)g.wp + own_w];
    };

    // ---- prologue: planes k0-1, k0, k0+1 into slots 0, 1, 2; `in` of plane k0
    fetch(k0 - 1, RU[0], RV[0], RW[0], RN[0]); publish(0, RU[0], RV[0], RW[0], RN[0]);
    fetch(k0, RU[1], RV[1], RW[1], RN[1]); publish(1, RU[1], RV[1], RW[1], RN[1]);
    fetch(k0 + 1, RU[2], RV[2], RW[2], RN[2]); publish(2, RU[2], RV[2], RW[2], RN[2]);
    fetch_in(k0, INU[0], INV[0], INW[0]);
    __syncthreads();

    auto step = [&](auto AA, const int K) {
        constexpr int A = decltype(AA)::value;
        constexpr int SM_ = A, SC = (A + 1) & 3, SP = (A + 2) & 3, SN = (A + 3) & 3;   // planes K-1, K, K+1, K+2
        constexpr int IC = A & 1, IN_ = (A + 1) & 1;
        fetch(K + 2, RU[SN], RV[SN], RW[SN], RN[SN]);       // in flight during the evaluation of plane K
        fetch_in(K + 1, INU[IN_], INV[IN_], INW[IN_]);
        const double um = RU[SM_], vm = RV[SM_], wm = RW[SM_], nm = RN[SM_];
        const double uc = RU[SC], vc = RV[SC], wc = RW[SC], nc = RN[SC];
        const double up_ = RU[SP], vp_ = RV[SP], wp_ = RW[SP], np_ = RN[SP];
        const double in_u = INU[IC], in_v = INV[IC], in_w = INW[IC];
        const double* tm = sm + SM_ * PL + t_own;           // tiles of planes K-1, K, K+1 at the own column
        const double* tc = sm + SC * PL + t_own;
        const double* tp = sm + SP * PL + t_own;
        const bool kz_c = K >= Ng && K < Ng + g.Nz, kz_f = K >= Ng && K <= Ng + g.Nz;
        const bool kin = K < ez;                            // u, v entries exist on planes < ez; w on planes <= ez

        // in-plane neighbours shared by the three faces (plane K)
        const double u_w = tc[FU - 1], u_e = tc[FU + 1], u_s = tc[FU - TX], u_n = tc[FU + TX], u_se = tc[FU - TX + 1];
        const double v_w = tc[FV - 1], v_e = tc[FV + 1], v_s = tc[FV - TX], v_n = tc[FV + TX], v_nw = tc[FV + TX - 1];
        const double w_w = tc[FW - 1], w_e = tc[FW + 1], w_s = tc[FW - TX], w_n = tc[FW + TX];
        const double n_w = tc[FN - 1], n_e = tc[FN + 1], n_s = tc[FN - TX], n_n = tc[FN + TX];
        const double n_sw = tc[FN - TX - 1], n_nw = tc[FN + TX - 1], n_se = tc[FN - TX + 1];

        if (col_u && kin) {
            const size_t o = (size_t)K * (size_t)g.up + own_u;
            if (int_u && kz_c) {
                // u face (i, j, K): x is the own direction
                const double wp_w = tp[FW - 1];
                const double vv = 0.25 * (v_w + vc + v_nw + v_n);
                const double ww = 0.25 * (w_w + wc + wp_w + wp_);
                double conv;
                if (SKEW) {
                    const double sx = skew_own<EXACT>(uc, u_w, u_e, c.dx, c.dx_2, c.h_rdx, c.h_r2dx);
                    const double sy = skew_dir<EXACT>(vv, 0.5 * (v_w + vc), 0.5 * (v_nw + v_n), uc, u_s, u_n, c.dy, c.dy_2, c.h_rdy, c.h_r2dy);
                    const double sz = skew_dir<EXACT>(ww, 0.5 * (w_w + wc), 0.5 * (wp_w + wp_), uc, um, up_, c.dz, c.dz_2, c.h_rdz, c.h_r2dz);
                    conv = sx + sy + sz;
                } else {
                    conv = central_dir<EXACT>(uc, u_w, u_e, c.dx_2) + central_dir<EXACT>(vv, u_s, u_n, c.dy_2) +
                           central_dir<EXACT>(ww, um, up_, c.dz_2);
                }
                const double nl = n_w, nr = nc;             // nu of the two cells adjacent in x
                const double d2x = diff_dir<EXACT>(nl, nr, uc, u_w, u_e, c.dx2);
                const double nu_n = 0.25 * (nl + nr + n_nw + n_n);
                const double nu_s = 0.25 * (nl + nr + n_sw + n_s);
                const double d2y = diff_dir<EXACT>(nu_s, nu_n, uc, u_s, u_n, c.dy2);
                const double nu_f = 0.25 * (nl + nr + tp[FN - 1] + np_);
                const double nu_b = 0.25 * (nl + nr + tm[FN - 1] + nm);
                const double d2z = diff_dir<EXACT>(nu_b, nu_f, uc, um, up_, c.dz2);
                const double diff = XZ ? d2x + d2z : d2x + d2y + d2z;
                out.u[o] = mom_fma<EXACT>(c.dt, -conv + diff + c.fx, in_u);
            } else if (copy_ghosts) {
                out.u[o] = uc;
            }
        }
        if (col_v && kin) {
            const size_t o = (size_t)K * (size_t)g.vp + own_v;
            if (int_v && kz_c) {
                // v face (i, j, K): y is the own direction
                const double wp_s = tp[FW - TX];
                const double uu = 0.25 * (u_s + u_se + uc + u_e);
                const double ww = 0.25 * (w_s + wc + wp_s + wp_);
                double conv;
                if (SKEW) {
                    const double sx = skew_dir<EXACT>(uu, 0.5 * (u_s + uc), 0.5 * (u_se + u_e), vc, v_w, v_e, c.dx, c.dx_2, c.h_rdx, c.h_r2dx);
                    const double sy = skew_own<EXACT>(vc, v_s, v_n, c.dy, c.dy_2, c.h_rdy, c.h_r2dy);
                    const double sz = skew_dir<EXACT>(ww, 0.5 * (w_s + wc), 0.5 * (wp_s + wp_), vc, vm, vp_, c.dz, c.dz_2, c.h_rdz, c.h_r2dz);
                    conv = sx + sy + sz;
                } else {
                    conv = central_dir<EXACT>(uu, v_w, v_e, c.dx_2) + central_dir<EXACT>(vc, v_s, v_n, c.dy_2) +
                           central_dir<EXACT>(ww, vm, vp_, c.dz_2);
                }
                const double nb = n_s, nt = nc;             // nu of the two cells adjacent in y
                const double nu_e = 0.25 * (nb + nt + n_se + n_e);
                const double nu_w = 0.25 * (nb + nt + n_sw + n_w);
                const double d2x = diff_dir<EXACT>(nu_w, nu_e, vc, v_w, v_e, c.dx2);
                const double d2y = diff_dir<EXACT>(nb, nt, vc, v_s, v_n, c.dy2);
                const double nu_f = 0.25 * (nb + nt + tp[FN - TX] + np_);
                const double nu_b = 0.25 * (nb + nt + tm[FN - TX] + nm);
                const double d2z = diff_dir<EXACT>(nu_b, nu_f, vc, vm, vp_, c.dz2);
                const double diff = XZ ? d2x + d2z : d2x + d2y + d2z;
                out.v[o] = mom_fma<EXACT>(c.dt, -conv + diff + c.fy, in_v);
            } else if (copy_ghosts) {
                out.v[o] = vc;
            }
        }
        if (col_w) {
            const size_t o = (size_t)K * (size_t)g.wp + own_w;
            if (int_w && kz_f) {
                // w face (i, j, K): z is the own direction; the transverse velocities sit on planes K-1 and K
                const double um_e = tm[FU + 1], vm_n = tm[FV + TX];
                const double uu = 0.25 * (um + um_e + uc + u_e);
                const double vv = 0.25 * (vm + vm_n + vc + v_n);
                double conv;
                if (SKEW) {
                    const double sx = skew_dir<EXACT>(uu, 0.5 * (um + uc), 0.5 * (um_e + u_e), wc, w_w, w_e, c.dx, c.dx_2, c.h_rdx, c.h_r2dx);
                    const double sy = skew_dir<EXACT>(vv, 0.5 * (vm + vc), 0.5 * (vm_n + v_n), wc, w_s, w_n, c.dy, c.dy_2, c.h_rdy, c.h_r2dy);
                    const double sz = skew_own<EXACT>(wc, wm, wp_, c.dz, c.dz_2, c.h_rdz, c.h_r2dz);
                    conv = sx + sy + sz;
                } else {
                    conv = central_dir<EXACT>(uu, w_w, w_e, c.dx_2) + central_dir<EXACT>(vv, w_s, w_n, c.dy_2) +
                           central_dir<EXACT>(wc, wm, wp_, c.dz_2);
                }
                const double nbk = nm, nfr = nc;            // nu of the two cells adjacent in z
                const double nu_e = 0.25 * (nbk + nfr + tm[FN + 1] + n_e);
                const double nu_w = 0.25 * (nbk + nfr + tm[FN - 1] + n_w);
                const double d2x = diff_dir<EXACT>(nu_w, nu_e, wc, w_w, w_e, c.dx2);
                const double nu_n = 0.25 * (nbk + nfr + tm[FN + TX] + n_n);
                const double nu_s = 0.25 * (nbk + nfr + tm[FN - TX] + n_s);
                const double d2y = diff_dir<EXACT>(nu_s, nu_n, wc, w_s, w_n, c.dy2);
                const double d2z = diff_dir<EXACT>(nbk, nfr, wc, wm, wp_, c.dz2);
                const double diff = XZ ? d2x + d2z : d2x + d2y + d2z;
                out.w[o] = mom_fma<EXACT>(c.dt, -conv + diff + c.fz, in_w);
            } else if (copy_ghosts) {
                out.w[o] = wc;
            }
        }
        publish(SN, RU[SN], RV[SN], RW[SN], RN[SN]);       // slot last read as plane K-2: free since the previous barrier
        __syncthreads();
    };
    for (int K = k0; K < k1; K += 4) {
        step(std::integral_constant<int, 0>{}, K);
        if (K + 1 >= k1) break;
        step(std::integral_constant<int, 1>{}, K + 1);
        if (K + 2 >= k1) break;
        step(std::integral_constant<int, 2>{}, K + 2);
        if (K + 3 >= k1) break;
        step(std::integral_constant<int, 3>{}, K + 3);
    }
}

// ------------------------------------------------------------------ implicit y-diffusion
// (I - dt d/dy(nu_eff d/dy)) q_new = q per y-line, Thomas algorithm: thomas_y_diffusion_{2d,3d}[_stretched]
// (reference src/solver_time_kernels_implicit.cpp:16-454).  One thread per (i, k) line, consecutive threads on
// consecutive i so every step of the march is one coalesced access per array; the forward sweep leaves d' in the
// field itself (q[j] is read once, at step j) and c' in a scratch array of the field's shape, the backward sweep
// reads them back.  Same operations in the same order as the reference (true divisions, no FMA): bit-identical.
//   COMP 0: u lines  i in [Ng, Ng+Nx], k cells, Ny unknowns, nu between cells i-1 and i, no-slip ghost = -interior
//   COMP 1: v lines  i cells, k cells, Ny-1 interior faces (wall faces stay 0), nu between cells j-1 and j
//   COMP 2: w lines  i cells, k in [Ng, Ng+Nz], Ny unknowns, nu between cells k-1 and k
template <int COMP, bool STRETCHED>
__global__ void __launch_bounds__(128)
k_thomas_y(Geom g, double* __restrict__ q, double* __restrict__ cp, const double* __restrict__ nu, double dt,
           double dt_inv_dy2, int k_first, int nk) {
    const int Ng = g.Ng;
    const int ni = COMP == 0 ? g.Nx + 1 : g.Nx;
    const int il = blockIdx.x * blockDim.x + threadIdx.x;
    const int kl = blockIdx.y;
    if (il >= ni || kl >= nk) return;
    const int i = il + Ng, k = k_first + kl;
    const int n = COMP == 1 ? g.Ny - 1 : g.Ny;                 // unknowns per line
    const int j0 = COMP == 1 ? Ng + 1 : Ng;
    const size_t qs = COMP == 0 ? g.us : (COMP == 1 ? g.vs : g.ws);
    const size_t qp = COMP == 0 ? (size_t)g.up : (COMP == 1 ? (size_t)g.vp : (size_t)g.wp);
    double* ql = q + (size_t)k * qp + i;
    double* cl = cp + (size_t)k * qp + i;
    double cprev = 0.0, dprev = 0.0;
    for (int jj = 0; jj < n; ++jj) {
        const int j = jj + j0;
        double nu_avg;
        if (COMP == 0) nu_avg = 0.5 * (nu[CIDX(i - 1, j, k)] + nu[CIDX(i, j, k)]);
        else if (COMP == 1) nu_avg = 0.5 * (nu[CIDX(i, j - 1, k)] + nu[CIDX(i, j, k)]);
        else nu_avg = 0.5 * (nu[CIDX(i, j, k - 1)] + nu[CIDX(i, j, k)]);
        double a_j, c_j, b_j;
        if (STRETCHED) {
            double alpha_lo, alpha_hi;
            if (COMP == 1) {
                alpha_lo = dt * nu_avg / (g.dyc[j] * g.dyv[j - 1]);
                alpha_hi = dt * nu_avg / (g.dyc[j] * g.dyv[j]);
            } else {
                alpha_lo = dt * nu_avg / (g.dyv[j] * g.dyc[j]);
                alpha_hi = dt * nu_avg / (g.dyv[j] * g.dyc[j + 1]);
            }
            a_j = jj > 0 ? -alpha_lo : 0.0;
            c_j = jj < n - 1 ? -alpha_hi : 0.0;
            b_j = 1.0 + alpha_lo + alpha_hi;
            if (COMP != 1) {
                if (jj == 0) b_j += alpha_lo;
                if (jj == n - 1) b_j += alpha_hi;
            }
        } else {
            const double alpha = dt_inv_dy2 * nu_avg;
            a_j = jj > 0 ? -alpha : 0.0;
            c_j = jj < n - 1 ? -alpha : 0.0;
            b_j = 1.0 + 2.0 * alpha;
            if (COMP != 1) {
                if (jj == 0) b_j += alpha;
                if (jj == n - 1) b_j += alpha;
            }
        }
        const double d_j = ql[(size_t)j * qs];
        double cj, dj;
        if (jj == 0) {
            cj = c_j / b_j;
            dj = d_j / b_j;
        } else {
            const double denom = b_j - a_j * cprev;
            cj = c_j / denom;
            dj = (d_j - a_j * dprev) / denom;
        }
        cl[(size_t)j * qs] = cj;
        ql[(size_t)j * qs] = dj;
        cprev = cj; dprev = dj;
    }
    double x = dprev;                                          // q[n-1] = d'[n-1], already in place
    for (int jj = n - 2; jj >= 0; --jj) {
        const int j = jj + j0;
        x = ql[(size_t)j * qs] - cl[(size_t)j * qs] * x;
        ql[(size_t)j * qs] = x;
    }
}

}  // namespace cfdnn
