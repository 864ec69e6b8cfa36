// mg_kernels.cuh — device kernels of the multigrid V-cycle (FP64, HBM-bound).
// Arithmetic follows the reference CPU path expression by expression so results track the
// x86 oracle to rounding; file:line citations are to src/poisson_solver_multigrid.cpp.
#pragma once
#include "mg.cuh"

namespace cfdnn {

// Map an index in [-1, N] to a stored location + sign according to the BC of that side.
//   Periodic : u[g] = u[N+g]        Neumann : u[g] = u[first]     Dirichlet(0): u[g] = -u[first]
//   HALO     : the halo plane itself (slab interface)                          (:307-554)
__device__ __forceinline__ int bc_map(int i, int N, int lo, int hi, double& s) {
    if (i < 0) {
        if (lo == CFDNN_PBC_PERIODIC) return N - 1;
        if (lo == CFDNN_PBC_HALO) return -1;
        if (lo == CFDNN_PBC_DIRICHLET) s = -s;
        return 0;
    }
    if (i >= N) {
        if (hi == CFDNN_PBC_PERIODIC) return 0;
        if (hi == CFDNN_PBC_HALO) return N;
        if (hi == CFDNN_PBC_DIRICHLET) s = -s;
        return N - 1;
    }
    return i;
}
// Residual ghosts: periodic wrap, else zero (apply_bc_to_residual :556-823). ok=false => 0.
__device__ __forceinline__ int rbc_map(int i, int N, int lo, int hi, bool& ok) {
    if (i < 0) {
        if (lo == CFDNN_PBC_PERIODIC) return N - 1;
        if (lo == CFDNN_PBC_HALO) return -1;
        ok = false;
        return 0;
    }
    if (i >= N) {
        if (hi == CFDNN_PBC_PERIODIC) return 0;
        if (hi == CFDNN_PBC_HALO) return N;
        ok = false;
        return N - 1;
    }
    return i;
}

__device__ __forceinline__ size_t lidx(const LevelGeom& g, int i, int j, int k) {
    return (size_t)(k + g.Hz) * (size_t)g.plane + (size_t)j * g.Px + i;
}

// ------------------------------------------------------------------ smoother sweep
// One weighted-Jacobi sweep  out = (1-w) u + w * ((uE+uW)/dx2 + (uN+uS)/dy2 + (uF+uB)/dz2 - f)/coeff
// (smooth_chebyshev :970-983, smooth_jacobi :1118-1131).  Each thread owns one (i,j) column
// and marches KZ planes keeping u(k-1), u(k), u(k+1) in registers: per cell 1 compulsory
// u load + 1 f load + 1 store; x/y neighbours come from L1/L2.
template <int KZ, bool NONUNIF>
__global__ void __launch_bounds__(256)
k_sweep(LevelGeom g, const double* __restrict__ u, const double* __restrict__ f,
        double* __restrict__ out, double omega, double om1) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k0 = blockIdx.z * KZ;
    if (i >= g.Nx || j >= g.Ny) return;

    double sxm = 1.0, sxp = 1.0, sym = 1.0, syp = 1.0;
    const int im = bc_map(i - 1, g.Nx, g.bc[0], g.bc[1], sxm);
    const int ip = bc_map(i + 1, g.Nx, g.bc[0], g.bc[1], sxp);
    const int jm = bc_map(j - 1, g.Ny, g.bc[2], g.bc[3], sym);
    const int jp = bc_map(j + 1, g.Ny, g.bc[2], g.bc[3], syp);
    const size_t rc = (size_t)j * g.Px, rm = (size_t)jm * g.Px, rp = (size_t)jp * g.Px;

    double aS = 0.0, aN = 0.0, diag = 0.0;
    if (NONUNIF) {
        aS = g.aS[j + g.joff];
        aN = g.aN[j + g.joff];
        diag = g.inv_dx2_2 + (aS + aN) + g.inv_dz2_2;   // :959
    }

    const int kend = min(k0 + KZ, g.Nz);
    double s = 1.0;
    int kb = bc_map(k0 - 1, g.Nz, g.bc[4], g.bc[5], s);
    double ub = s * u[(size_t)(kb + g.Hz) * g.plane + rc + i];
    double uc = u[(size_t)(k0 + g.Hz) * g.plane + rc + i];
    for (int k = k0; k < kend; ++k) {
        const size_t pc = (size_t)(k + g.Hz) * g.plane;
        double sf = 1.0;
        const int kf = bc_map(k + 1, g.Nz, g.bc[4], g.bc[5], sf);
        const double uf = sf * u[(size_t)(kf + g.Hz) * g.plane + rc + i];
        const double uE = sxp * u[pc + rc + ip];
        const double uW = sxm * u[pc + rc + im];
        const double uN = syp * u[pc + rp + i];
        const double uS = sym * u[pc + rm + i];
        const double fv = f[pc + rc + i];
        double uj;
        if (NONUNIF) {
            const double lap_x = (uE + uW) / g.dx2;
            const double lap_y = aS * uS + aN * uN;
            const double lap_z = (uf + ub) / g.dz2;
            uj = (lap_x + lap_y + lap_z - fv) / diag;
        } else {
            uj = ((uE + uW) / g.dx2 + (uN + uS) / g.dy2 + (uf + ub) / g.dz2 - fv) / g.coeff;
        }
        out[pc + rc + i] = om1 * uc + omega * uj;
        ub = uc;
        uc = uf * sf;   // sf is +-1: undo the sign so the raw stored value moves on
    }
}

// ------------------------------------------------------------------ residual
// r = f - ((uE-2u+uW)/dx2 + (uN-2u+uS)/dy2 + (uF-2u+uB)/dz2)   (compute_residual :1807-1816)
// NORMS: also per-block max|r| and sum r^2 (compute_residual_and_norms :1943-1958).
// r == nullptr: norms only (the convergence check; nothing reads r before the next residual pass).
template <int KZ, bool NONUNIF, bool NORMS>
__global__ void __launch_bounds__(256)
k_residual(LevelGeom g, const double* __restrict__ u, const double* __restrict__ f,
           double* __restrict__ r, double* __restrict__ part_max, double* __restrict__ part_sq) {
    __shared__ double sh[32];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k0 = blockIdx.z * KZ;
    double mx = 0.0, sq = 0.0;
    if (i < g.Nx && j < g.Ny) {
        double sxm = 1.0, sxp = 1.0, sym = 1.0, syp = 1.0;
        const int im = bc_map(i - 1, g.Nx, g.bc[0], g.bc[1], sxm);
        const int ip = bc_map(i + 1, g.Nx, g.bc[0], g.bc[1], sxp);
        const int jm = bc_map(j - 1, g.Ny, g.bc[2], g.bc[3], sym);
        const int jp = bc_map(j + 1, g.Ny, g.bc[2], g.bc[3], syp);
        const size_t rc = (size_t)j * g.Px, rm = (size_t)jm * g.Px, rp = (size_t)jp * g.Px;
        double aS = 0.0, aN = 0.0, aP = 0.0;
        if (NONUNIF) { aS = g.aS[j + g.joff]; aN = g.aN[j + g.joff]; aP = g.aP[j + g.joff]; }
        const int kend = min(k0 + KZ, g.Nz);
        double s = 1.0;
        int kb = bc_map(k0 - 1, g.Nz, g.bc[4], g.bc[5], s);
        double ub = s * u[(size_t)(kb + g.Hz) * g.plane + rc + i];
        double uc = u[(size_t)(k0 + g.Hz) * g.plane + rc + i];
        for (int k = k0; k < kend; ++k) {
            const size_t pc = (size_t)(k + g.Hz) * g.plane;
            double sf = 1.0;
            const int kf = bc_map(k + 1, g.Nz, g.bc[4], g.bc[5], sf);
            const double uf = sf * u[(size_t)(kf + g.Hz) * g.plane + rc + i];
            const double uE = sxp * u[pc + rc + ip];
            const double uW = sxm * u[pc + rc + im];
            const double uN = syp * u[pc + rp + i];
            const double uS = sym * u[pc + rm + i];
            const double two_c = 2.0 * uc;
            double lap;
            if (NONUNIF) {
                const double lap_x = (uE - two_c + uW) / g.dx2;
                const double lap_y = aS * uS + aP * uc + aN * uN;
                const double lap_z = (uf - two_c + ub) / g.dz2;
                lap = lap_x + lap_y + lap_z;
            } else {
                lap = (uE - two_c + uW) / g.dx2 + (uN - two_c + uS) / g.dy2 + (uf - two_c + ub) / g.dz2;
            }
            const double rv = f[pc + rc + i] - lap;
            if (r) r[pc + rc + i] = rv;
            if (NORMS) {
                mx = fmax(mx, fabs(rv));
                sq += rv * rv;
            }
            ub = uc;
            uc = uf * sf;
        }
    }
    if (NORMS) {
        const int tid = threadIdx.y * blockDim.x + threadIdx.x;
        const int lane = tid & 31, w = tid >> 5, nw = (blockDim.x * blockDim.y + 31) >> 5;
        const int bid = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        double m2 = warp_max(mx);
        if (lane == 0) sh[w] = m2;
        __syncthreads();
        if (w == 0) { m2 = lane < nw ? sh[lane] : 0.0; m2 = warp_max(m2); if (lane == 0) part_max[bid] = m2; }
        __syncthreads();
        double s2 = warp_sum(sq);
        if (lane == 0) sh[w] = s2;
        __syncthreads();
        if (w == 0) { s2 = lane < nw ? sh[lane] : 0.0; s2 = warp_sum(s2); if (lane == 0) part_sq[bid] = s2; }
    }
}

// ------------------------------------------------------------------ restriction
// 27-point full weighting centred on fine cell 2*ic (restrict_residual :2035-2062); residual
// ghosts: periodic wrap else 0.  One thread per coarse cell; also emits per-block partial sums
// of f_c for make_rhs_mean_free (:3479-3548).
static __global__ void __launch_bounds__(256)
k_restrict(LevelGeom gf, LevelGeom gc, const double* __restrict__ r, double* __restrict__ fc,
           int kc_off, double* __restrict__ part_sum) {
    __shared__ double sh[32];
    const int ic = blockIdx.x * blockDim.x + threadIdx.x;
    const int jc = blockIdx.y * blockDim.y + threadIdx.y;
    const int kc = blockIdx.z;                 // local coarse plane of this slab
    double val = 0.0;
    const int NzcLocal = gf.Nz >> 1;
    if (ic < gc.Nx && jc < gc.Ny && kc < NzcLocal) {
        const int i0 = 2 * ic, j0 = 2 * jc, k0 = 2 * kc;
        bool okx = true, oky = true, okz = true;
        const int im = rbc_map(i0 - 1, gf.Nx, gf.bc[0], gf.bc[1], okx);
        const int jm = rbc_map(j0 - 1, gf.Ny, gf.bc[2], gf.bc[3], oky);
        const int km = rbc_map(k0 - 1, gf.Nz, gf.bc[4], gf.bc[5], okz);
        const int ix[3] = {im, i0, i0 + 1};
        const int jx[3] = {jm, j0, j0 + 1};
        const int kx[3] = {km, k0, k0 + 1};
        double v[3][3][3];   // [dk][dj][di]
#pragma unroll
        for (int c = 0; c < 3; ++c)
#pragma unroll
            for (int b = 0; b < 3; ++b)
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    const bool ok = (a > 0 || okx) && (b > 0 || oky) && (c > 0 || okz);
                    v[c][b][a] = ok ? r[lidx(gf, ix[a], jx[b], kx[c])] : 0.0;
                }
#define RV(di, dj, dk) v[(dk) + 1][(dj) + 1][(di) + 1]
        double sum = 0.0;
        sum += 0.125 * RV(0, 0, 0);
        sum += 0.0625 * (RV(-1, 0, 0) + RV(1, 0, 0) + RV(0, -1, 0) + RV(0, 1, 0) + RV(0, 0, -1) + RV(0, 0, 1));
        sum += 0.03125 * (RV(-1, -1, 0) + RV(1, -1, 0) + RV(-1, 1, 0) + RV(1, 1, 0)
                        + RV(-1, 0, -1) + RV(1, 0, -1) + RV(-1, 0, 1) + RV(1, 0, 1)
                        + RV(0, -1, -1) + RV(0, 1, -1) + RV(0, -1, 1) + RV(0, 1, 1));
        sum += 0.015625 * (RV(-1, -1, -1) + RV(1, -1, -1) + RV(-1, 1, -1) + RV(1, 1, -1)
                         + RV(-1, -1, 1) + RV(1, -1, 1) + RV(-1, 1, 1) + RV(1, 1, 1));
#undef RV
        fc[lidx(gc, ic, jc, kc + kc_off)] = sum;
        val = sum;
    }
    if (part_sum) {
        const int tid = threadIdx.y * blockDim.x + threadIdx.x;
        const int lane = tid & 31, w = tid >> 5, nw = (blockDim.x * blockDim.y + 31) >> 5;
        const int bid = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        double s2 = warp_sum(val);
        if (lane == 0) sh[w] = s2;
        __syncthreads();
        if (w == 0) { s2 = lane < nw ? sh[lane] : 0.0; s2 = warp_sum(s2); if (lane == 0) part_sum[bid] = s2; }
    }
}

// ------------------------------------------------------------------ prolongation
// u_f += trilinear(u_c) with the reference's one-sided owner-computes weights
// (prolongate_correction :2230-2260): ic = i_f/2, weight 1 (even) or 1/2,1/2 toward ic+1 (odd).
static __global__ void __launch_bounds__(256)
k_prolong(LevelGeom gf, LevelGeom gc, const double* __restrict__ uc, double* __restrict__ uf,
          int kc_off) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i >= gf.Nx || j >= gf.Ny || k >= gf.Nz) return;
    const int ic = i >> 1, jc = j >> 1, kc = (k >> 1) + kc_off;
    const double wx1 = 0.5 * (i & 1), wx0 = 1.0 - wx1;
    const double wy1 = 0.5 * (j & 1), wy0 = 1.0 - wy1;
    const double wz1 = 0.5 * (k & 1), wz0 = 1.0 - wz1;
    double sx = 1.0, sy = 1.0, sz = 1.0;
    const int ip = bc_map(ic + 1, gc.Nx, gc.bc[0], gc.bc[1], sx);
    const int jp = bc_map(jc + 1, gc.Ny, gc.bc[2], gc.bc[3], sy);
    const int kp = bc_map(kc + 1, gc.Nz, gc.bc[4], gc.bc[5], sz);
    const double c000 = uc[lidx(gc, ic, jc, kc)];
    const double c100 = sx * uc[lidx(gc, ip, jc, kc)];
    const double c010 = sy * uc[lidx(gc, ic, jp, kc)];
    const double c110 = (sx * sy) * uc[lidx(gc, ip, jp, kc)];
    const double c001 = sz * uc[lidx(gc, ic, jc, kp)];
    const double c101 = (sx * sz) * uc[lidx(gc, ip, jc, kp)];
    const double c011 = (sy * sz) * uc[lidx(gc, ic, jp, kp)];
    const double c111 = (sx * sy * sz) * uc[lidx(gc, ip, jp, kp)];
    const double corr = wx0 * wy0 * wz0 * c000
                      + wx1 * wy0 * wz0 * c100
                      + wx0 * wy1 * wz0 * c010
                      + wx1 * wy1 * wz0 * c110
                      + wx0 * wy0 * wz1 * c001
                      + wx1 * wy0 * wz1 * c101
                      + wx0 * wy1 * wz1 * c011
                      + wx1 * wy1 * wz1 * c111;
    const size_t o = lidx(gf, i, j, k);
    uf[o] += corr;
}

// Same operator, one thread per COARSE cell updating its 8 fine children: the 8 coarse values are
// loaded once per thread instead of once per fine cell, fine rows are read/written as aligned
// double2.  For child (a,b,c) only the terms with (da<=a, db<=b, dc<=c) have non-zero weight; they
// are summed in the reference's order (c000,c100,c010,c110,c001,...), so the result is bitwise that
// of k_prolong (weights are exact powers of two).  Needs even fine sizes.
static __global__ void __launch_bounds__(256)
k_prolong8(LevelGeom gf, LevelGeom gc, const double* __restrict__ uc, double* __restrict__ uf, int kc_off) {
    const int ic = blockIdx.x * blockDim.x + threadIdx.x;
    const int jc = blockIdx.y * blockDim.y + threadIdx.y;
    const int kl = blockIdx.z;                       // local coarse plane of this slab
    if (2 * ic >= gf.Nx || 2 * jc >= gf.Ny || 2 * kl >= gf.Nz) return;
    const int kc = kl + kc_off;
    double sx = 1.0, sy = 1.0, sz = 1.0;
    const int ip = bc_map(ic + 1, gc.Nx, gc.bc[0], gc.bc[1], sx);
    const int jp = bc_map(jc + 1, gc.Ny, gc.bc[2], gc.bc[3], sy);
    const int kp = bc_map(kc + 1, gc.Nz, gc.bc[4], gc.bc[5], sz);
    double cv[2][2][2];                              // [dc][db][da]
    cv[0][0][0] = uc[lidx(gc, ic, jc, kc)];
    cv[0][0][1] = sx * uc[lidx(gc, ip, jc, kc)];
    cv[0][1][0] = sy * uc[lidx(gc, ic, jp, kc)];
    cv[0][1][1] = (sx * sy) * uc[lidx(gc, ip, jp, kc)];
    cv[1][0][0] = sz * uc[lidx(gc, ic, jc, kp)];
    cv[1][0][1] = (sx * sz) * uc[lidx(gc, ip, jc, kp)];
    cv[1][1][0] = (sy * sz) * uc[lidx(gc, ic, jp, kp)];
    cv[1][1][1] = (sx * sy * sz) * uc[lidx(gc, ip, jp, kp)];
#pragma unroll
    for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            double corr[2];
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                const double w = (a ? 0.5 : 1.0) * (b ? 0.5 : 1.0) * (c ? 0.5 : 1.0);
                double acc = w * cv[0][0][0];
#pragma unroll
                for (int dc = 0; dc <= c; ++dc)
#pragma unroll
                    for (int db = 0; db <= b; ++db)
#pragma unroll
                        for (int da = 0; da <= a; ++da)
                            if (dc | db | da) acc += w * cv[dc][db][da];
                corr[a] = acc;
            }
            double2* row = reinterpret_cast<double2*>(uf + lidx(gf, 2 * ic, 2 * jc + b, 2 * kl + c));
            double2 v = *row;
            v.x += corr[0];
            v.y += corr[1];
            *row = v;
        }
}

// ------------------------------------------------------------------ small helpers
// out[0] = sum(part[0..n)) (fixed order); if count>0 also out[1] = out[0]/count (the mean).
static __global__ void k_final_sum(const double* __restrict__ part, int n, double* __restrict__ out,
                            double count) {
    __shared__ double sh[32];
    double s = 0.0;
    for (int t = threadIdx.x; t < n; t += blockDim.x) s += part[t];
    s = block_sum(s, sh);
    if (threadIdx.x == 0) {
        out[0] = s;
        if (count > 0.0) out[1] = s / count;
    }
}
static __global__ void k_final_max(const double* __restrict__ part, int n, double* __restrict__ out) {
    __shared__ double sh[32];
    double s = 0.0;
    for (int t = threadIdx.x; t < n; t += blockDim.x) s = fmax(s, part[t]);
    s = block_max(s, sh);
    if (threadIdx.x == 0) out[0] = s;
}
// interior sum of a level array -> per-block partials
static __global__ void __launch_bounds__(256)
k_sum_interior(LevelGeom g, const double* __restrict__ a, double* __restrict__ part) {
    __shared__ double sh[32];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    double v = 0.0;
    if (i < g.Nx && j < g.Ny && k < g.Nz) v = a[lidx(g, i, j, k)];
    const int tid = threadIdx.y * blockDim.x + threadIdx.x;
    const int lane = tid & 31, w = tid >> 5, nw = (blockDim.x * blockDim.y + 31) >> 5;
    const int bid = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    double s2 = warp_sum(v);
    if (lane == 0) sh[w] = s2;
    __syncthreads();
    if (w == 0) { s2 = lane < nw ? sh[lane] : 0.0; s2 = warp_sum(s2); if (lane == 0) part[bid] = s2; }
}
// a -= *mean on interior cells (subtract_mean / make_rhs_mean_free second loops)
static __global__ void __launch_bounds__(256)
k_sub_scalar(LevelGeom g, double* __restrict__ a, const double* __restrict__ mean) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i < g.Nx && j < g.Ny && k < g.Nz) a[lidx(g, i, j, k)] -= mean[0];
}

// max|a| and sum a^2 over interior cells -> per-block partials (||b||_inf, ||b||_2 :3660-3682)
static __global__ void __launch_bounds__(256)
k_norms_interior(LevelGeom g, const double* __restrict__ a, int kz, double* __restrict__ part_max,
                 double* __restrict__ part_sq) {
    __shared__ double sh[32];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k0 = blockIdx.z * kz;
    double mx = 0.0, sq = 0.0;
    if (i < g.Nx && j < g.Ny) {
        const int kend = min(k0 + kz, g.Nz);
        for (int k = k0; k < kend; ++k) {
            const double v = a[lidx(g, i, j, k)];
            mx = fmax(mx, fabs(v));
            sq += v * v;
        }
    }
    const int tid = threadIdx.y * blockDim.x + threadIdx.x;
    const int lane = tid & 31, w = tid >> 5, nw = (blockDim.x * blockDim.y + 31) >> 5;
    const int bid = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    double m2 = warp_max(mx);
    if (lane == 0) sh[w] = m2;
    __syncthreads();
    if (w == 0) { m2 = lane < nw ? sh[lane] : 0.0; m2 = warp_max(m2); if (lane == 0) part_max[bid] = m2; }
    __syncthreads();
    double s2 = warp_sum(sq);
    if (lane == 0) sh[w] = s2;
    __syncthreads();
    if (w == 0) { s2 = lane < nw ? sh[lane] : 0.0; s2 = warp_sum(s2); if (lane == 0) part_sq[bid] = s2; }
}

// level-0 transfer between the reference's ghosted cell layout and the level layout.
// in : f <- rhs interior, u <- p interior                           (solve :3558-3574)
static __global__ void __launch_bounds__(256)
k_copy_in(LevelGeom g, int Ng, const double* __restrict__ rhs, const double* __restrict__ p,
          double* __restrict__ f, double* __restrict__ u) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i >= g.Nx || j >= g.Ny || k >= g.Nz) return;
    const size_t sx = g.Nx + 2 * Ng, sy = g.Ny + 2 * Ng;
    const size_t src = ((size_t)(k + Ng) * sy + (j + Ng)) * sx + (i + Ng);
    const size_t o = lidx(g, i, j, k);
    f[o] = rhs[src];
    u[o] = p[src];
}
// out: p <- u on ALL cells, ghosts from the BCs in x,y,z pass order    (solve :3769-3783).
// mean != nullptr: p <- u - mean[0], the interior-mean removal of fix_nullspace (:3404-3477) applied on
// the way out instead of in a separate read-modify-write pass over u (only with a null space, i.e. no
// Dirichlet face, so every ghost is a copy of an interior value and gets the same subtraction).
static __global__ void __launch_bounds__(256)
k_copy_out(LevelGeom g, int Ng, const double* __restrict__ u, double* __restrict__ p,
           const double* __restrict__ mean) {
    const int I = blockIdx.x * blockDim.x + threadIdx.x;   // ghosted indices
    const int J = blockIdx.y * blockDim.y + threadIdx.y;
    const int K = blockIdx.z;
    const int sx = g.Nx + 2 * Ng, sy = g.Ny + 2 * Ng, sz = g.Nz + 2 * Ng;
    if (I >= sx || J >= sy || K >= sz) return;
    double s = 1.0;
    const int i = bc_map(I - Ng, g.Nx, g.bc[0], g.bc[1], s);
    const int j = bc_map(J - Ng, g.Ny, g.bc[2], g.bc[3], s);
    const int k = bc_map(K - Ng, g.Nz, g.bc[4], g.bc[5], s);
    double v = u[lidx(g, i, j, k)];
    if (mean) v -= mean[0];
    p[((size_t)K * sy + J) * sx + I] = s * v;
}

// ------------------------------------------------------------------ coarsest level
// `iters` damped-Jacobi iterations with BCs re-applied every iteration, whole level in one CTA
// (smooth_jacobi :1003-1245 as used by vcycle :3193-3195: 100 or 50 iterations, w = 0.67).
// The reference spends ~4 loops per iteration on an 8^3 grid; here the level lives in shared
// memory and one launch does all iterations.
template <int CPT>
__global__ void __launch_bounds__(1024)
k_coarse_jacobi(LevelGeom g, double* __restrict__ u, const double* __restrict__ f, int iters,
                double omega, double om1) {
    CFDNN_DYN_SMEM(double, smem);
    const int n = g.Nx * g.Ny * g.Nz;
    double* a = smem;
    double* b = smem + n;
    int nb[CPT][6];
    double sg[CPT][6];
    double fv[CPT];
    int cell[CPT];
#pragma unroll
    for (int q = 0; q < CPT; ++q) {
        const int c = threadIdx.x + q * blockDim.x;
        cell[q] = c;
        if (c < n) {
            const int i = c % g.Nx, j = (c / g.Nx) % g.Ny, k = c / (g.Nx * g.Ny);
#pragma unroll
            for (int d = 0; d < 6; ++d) sg[q][d] = 1.0;
            const int ip = bc_map(i + 1, g.Nx, g.bc[0], g.bc[1], sg[q][0]);
            const int im = bc_map(i - 1, g.Nx, g.bc[0], g.bc[1], sg[q][1]);
            const int jp = bc_map(j + 1, g.Ny, g.bc[2], g.bc[3], sg[q][2]);
            const int jm = bc_map(j - 1, g.Ny, g.bc[2], g.bc[3], sg[q][3]);
            const int kp = bc_map(k + 1, g.Nz, g.bc[4], g.bc[5], sg[q][4]);
            const int km = bc_map(k - 1, g.Nz, g.bc[4], g.bc[5], sg[q][5]);
            nb[q][0] = (k * g.Ny + j) * g.Nx + ip;
            nb[q][1] = (k * g.Ny + j) * g.Nx + im;
            nb[q][2] = (k * g.Ny + jp) * g.Nx + i;
            nb[q][3] = (k * g.Ny + jm) * g.Nx + i;
            nb[q][4] = (kp * g.Ny + j) * g.Nx + i;
            nb[q][5] = (km * g.Ny + j) * g.Nx + i;
            const size_t o = lidx(g, i, j, k);
            fv[q] = f[o];
            a[c] = u[o];
        }
    }
    __syncthreads();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int q = 0; q < CPT; ++q) {
            const int c = cell[q];
            if (c < n) {
                const double uE = sg[q][0] * a[nb[q][0]], uW = sg[q][1] * a[nb[q][1]];
                const double uN = sg[q][2] * a[nb[q][2]], uS = sg[q][3] * a[nb[q][3]];
                const double uF = sg[q][4] * a[nb[q][4]], uB = sg[q][5] * a[nb[q][5]];
                const double uj = ((uE + uW) / g.dx2 + (uN + uS) / g.dy2 + (uF + uB) / g.dz2 - fv[q]) / g.coeff;
                b[c] = om1 * a[c] + omega * uj;
            }
        }
        __syncthreads();
        double* t = a; a = b; b = t;
    }
#pragma unroll
    for (int q = 0; q < CPT; ++q) {
        const int c = cell[q];
        if (c < n) {
            const int i = c % g.Nx, j = (c / g.Nx) % g.Ny, k = c / (g.Nx * g.Ny);
            u[lidx(g, i, j, k)] = a[c];
        }
    }
}

}  // namespace cfdnn
