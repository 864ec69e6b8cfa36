// solver_kernels.cuh — FP64 device kernels of the fractional-step path (HBM-bound stencils).
// Arithmetic follows the reference CPU path expression by expression (no FMA contraction:
// the file is compiled with -fmad=false; divisions go through the correctly rounded fdiv()).
// file:line citations are to the reference tree.
#pragma once
#include "solver.cuh"

namespace cfdnn {

#define UIDX(i, j, k) ((size_t)(k) * g.up + (size_t)(j) * g.us + (i))
#define VIDX(i, j, k) ((size_t)(k) * g.vp + (size_t)(j) * g.vs + (i))
#define WIDX(i, j, k) ((size_t)(k) * g.wp + (size_t)(j) * g.ws + (i))
#define CIDX(i, j, k) ((size_t)(k) * g.cp + (size_t)(j) * g.cs + (i))

// Index space shared by the face kernels: (ix,iy,iz) in [0,Nx] x [0,Ny] x [0,Nz];
// a thread owns u-face (i,j,k) when iy<Ny && iz<Nz, v-face when ix<Nx && iz<Nz,
// w-face when ix<Nx && iy<Ny  (loops over n_u_faces / n_v_faces / n_w_faces in the reference).
#define FACE_THREAD_PROLOGUE                                               \
    const int ix = blockIdx.x * blockDim.x + threadIdx.x;                  \
    const int iy = blockIdx.y * blockDim.y + threadIdx.y;                  \
    const int iz = blockIdx.z;                                             \
    if (ix > g.Nx || iy > g.Ny || iz > g.Nz) return;                       \
    const int i = ix + g.Ng, j = iy + g.Ng, k = iz + g.Ng;                 \
    const bool has_u = iy < g.Ny && iz < g.Nz;                             \
    const bool has_v = ix < g.Nx && iz < g.Nz;                             \
    const bool has_w = ix < g.Nx && iy < g.Ny;

// ------------------------------------------------------------------ convection
// SKEW = true : convective_{u,v,w}_face_kernel_skew_3d   (include/solver_kernels.hpp:1377-1569)
// SKEW = false: convective_{u,v,w}_face_kernel_staggered_3d with use_central (:845-1022)
template <bool SKEW>
__device__ __forceinline__ double conv_u(const Geom& g, const double* __restrict__ u, const double* __restrict__ v,
                                         const double* __restrict__ w, int i, int j, int k) {
    const double uu = u[UIDX(i, j, k)];
    const double u_ip1 = u[UIDX(i + 1, j, k)], u_im1 = u[UIDX(i - 1, j, k)];
    const double u_jp1 = u[UIDX(i, j + 1, k)], u_jm1 = u[UIDX(i, j - 1, k)];
    const double u_kp1 = u[UIDX(i, j, k + 1)], u_km1 = u[UIDX(i, j, k - 1)];
    const double v_bl = v[VIDX(i - 1, j, k)], v_br = v[VIDX(i, j, k)];
    const double v_tl = v[VIDX(i - 1, j + 1, k)], v_tr = v[VIDX(i, j + 1, k)];
    const double vv = 0.25 * (v_bl + v_br + v_tl + v_tr);
    const double w_bl = w[WIDX(i - 1, j, k)], w_br = w[WIDX(i, j, k)];
    const double w_fl = w[WIDX(i - 1, j, k + 1)], w_fr = w[WIDX(i, j, k + 1)];
    const double ww = 0.25 * (w_bl + w_br + w_fl + w_fr);
    double r;
    if (SKEW) {
        const double adv_x = uu * (u_ip1 - u_im1) / g.dx_2;
        const double uu_left = 0.5 * (u_im1 + uu), uu_right = 0.5 * (uu + u_ip1);
        const double cons_x = (uu_right * uu_right - uu_left * uu_left) / g.dx;
        const double skew_x = 0.5 * (adv_x + cons_x);
        const double adv_y = vv * (u_jp1 - u_jm1) / g.dy_2;
        const double vv_bot = 0.5 * (v_bl + v_br), vv_top = 0.5 * (v_tl + v_tr);
        const double u_bot = 0.5 * (uu + u_jm1), u_top = 0.5 * (uu + u_jp1);
        const double cons_y = (vv_top * u_top - vv_bot * u_bot) / g.dy;
        const double skew_y = 0.5 * (adv_y + cons_y);
        const double adv_z = ww * (u_kp1 - u_km1) / g.dz_2;
        const double ww_back = 0.5 * (w_bl + w_br), ww_front = 0.5 * (w_fl + w_fr);
        const double u_back = 0.5 * (uu + u_km1), u_front = 0.5 * (uu + u_kp1);
        const double cons_z = (ww_front * u_front - ww_back * u_back) / g.dz;
        const double skew_z = 0.5 * (adv_z + cons_z);
        r = skew_x + skew_y + skew_z;
    } else {
        const double dudx = (u_ip1 - u_im1) / g.dx_2;
        const double dudy = (u_jp1 - u_jm1) / g.dy_2;
        const double dudz = (u_kp1 - u_km1) / g.dz_2;
        r = uu * dudx + vv * dudy + ww * dudz;
    }
    return r;
}
template <bool SKEW>
__device__ __forceinline__ double conv_v(const Geom& g, const double* __restrict__ u, const double* __restrict__ v,
                                         const double* __restrict__ w, int i, int j, int k) {
    const double vv = v[VIDX(i, j, k)];
    const double v_ip1 = v[VIDX(i + 1, j, k)], v_im1 = v[VIDX(i - 1, j, k)];
    const double v_jp1 = v[VIDX(i, j + 1, k)], v_jm1 = v[VIDX(i, j - 1, k)];
    const double v_kp1 = v[VIDX(i, j, k + 1)], v_km1 = v[VIDX(i, j, k - 1)];
    const double u_bl = u[UIDX(i, j - 1, k)], u_br = u[UIDX(i + 1, j - 1, k)];
    const double u_tl = u[UIDX(i, j, k)], u_tr = u[UIDX(i + 1, j, k)];
    const double uu = 0.25 * (u_bl + u_br + u_tl + u_tr);
    const double w_bl = w[WIDX(i, j - 1, k)], w_tl = w[WIDX(i, j, k)];
    const double w_bf = w[WIDX(i, j - 1, k + 1)], w_tf = w[WIDX(i, j, k + 1)];
    const double ww = 0.25 * (w_bl + w_tl + w_bf + w_tf);
    double r;
    if (SKEW) {
        const double adv_x = uu * (v_ip1 - v_im1) / g.dx_2;
        const double uu_left = 0.5 * (u_bl + u_tl), uu_right = 0.5 * (u_br + u_tr);
        const double v_left = 0.5 * (v_im1 + vv), v_right = 0.5 * (vv + v_ip1);
        const double cons_x = (uu_right * v_right - uu_left * v_left) / g.dx;
        const double skew_x = 0.5 * (adv_x + cons_x);
        const double adv_y = vv * (v_jp1 - v_jm1) / g.dy_2;
        const double vv_bot = 0.5 * (v_jm1 + vv), vv_top = 0.5 * (vv + v_jp1);
        const double cons_y = (vv_top * vv_top - vv_bot * vv_bot) / g.dy;
        const double skew_y = 0.5 * (adv_y + cons_y);
        const double adv_z = ww * (v_kp1 - v_km1) / g.dz_2;
        const double ww_back = 0.5 * (w_bl + w_tl), ww_front = 0.5 * (w_bf + w_tf);
        const double v_back = 0.5 * (vv + v_km1), v_front = 0.5 * (vv + v_kp1);
        const double cons_z = (ww_front * v_front - ww_back * v_back) / g.dz;
        const double skew_z = 0.5 * (adv_z + cons_z);
        r = skew_x + skew_y + skew_z;
    } else {
        const double dvdx = (v_ip1 - v_im1) / g.dx_2;
        const double dvdy = (v_jp1 - v_jm1) / g.dy_2;
        const double dvdz = (v_kp1 - v_km1) / g.dz_2;
        r = uu * dvdx + vv * dvdy + ww * dvdz;
    }
    return r;
}
template <bool SKEW>
__device__ __forceinline__ double conv_w(const Geom& g, const double* __restrict__ u, const double* __restrict__ v,
                                         const double* __restrict__ w, int i, int j, int k) {
    const double ww = w[WIDX(i, j, k)];
    const double w_ip1 = w[WIDX(i + 1, j, k)], w_im1 = w[WIDX(i - 1, j, k)];
    const double w_jp1 = w[WIDX(i, j + 1, k)], w_jm1 = w[WIDX(i, j - 1, k)];
    const double w_kp1 = w[WIDX(i, j, k + 1)], w_km1 = w[WIDX(i, j, k - 1)];
    const double u_bl = u[UIDX(i, j, k - 1)], u_br = u[UIDX(i + 1, j, k - 1)];
    const double u_fl = u[UIDX(i, j, k)], u_fr = u[UIDX(i + 1, j, k)];
    const double uu = 0.25 * (u_bl + u_br + u_fl + u_fr);
    const double v_bl = v[VIDX(i, j, k - 1)], v_tl = v[VIDX(i, j + 1, k - 1)];
    const double v_bf = v[VIDX(i, j, k)], v_tf = v[VIDX(i, j + 1, k)];
    const double vv = 0.25 * (v_bl + v_tl + v_bf + v_tf);
    double r;
    if (SKEW) {
        const double adv_x = uu * (w_ip1 - w_im1) / g.dx_2;
        const double uu_left = 0.5 * (u_bl + u_fl), uu_right = 0.5 * (u_br + u_fr);
        const double w_left = 0.5 * (w_im1 + ww), w_right = 0.5 * (ww + w_ip1);
        const double cons_x = (uu_right * w_right - uu_left * w_left) / g.dx;
        const double skew_x = 0.5 * (adv_x + cons_x);
        const double adv_y = vv * (w_jp1 - w_jm1) / g.dy_2;
        const double vv_bot = 0.5 * (v_bl + v_bf), vv_top = 0.5 * (v_tl + v_tf);
        const double w_bot = 0.5 * (ww + w_jm1), w_top = 0.5 * (ww + w_jp1);
        const double cons_y = (vv_top * w_top - vv_bot * w_bot) / g.dy;
        const double skew_y = 0.5 * (adv_y + cons_y);
        const double adv_z = ww * (w_kp1 - w_km1) / g.dz_2;
        const double ww_back = 0.5 * (w_km1 + ww), ww_front = 0.5 * (ww + w_kp1);
        const double cons_z = (ww_front * ww_front - ww_back * ww_back) / g.dz;
        const double skew_z = 0.5 * (adv_z + cons_z);
        r = skew_x + skew_y + skew_z;
    } else {
        const double dwdx = (w_ip1 - w_im1) / g.dx_2;
        const double dwdy = (w_jp1 - w_jm1) / g.dy_2;
        const double dwdz = (w_kp1 - w_km1) / g.dz_2;
        r = uu * dwdx + vv * dwdy + ww * dwdz;
    }
    return r;
}
template <bool SKEW>
__global__ void __launch_bounds__(256)
k_convective(Geom g, const double* __restrict__ u, const double* __restrict__ v,
             const double* __restrict__ w, double* __restrict__ cu, double* __restrict__ cv,
             double* __restrict__ cw) {
    FACE_THREAD_PROLOGUE
    if (has_u) cu[UIDX(i, j, k)] = conv_u<SKEW>(g, u, v, w, i, j, k);
    if (has_v) cv[VIDX(i, j, k)] = conv_v<SKEW>(g, u, v, w, i, j, k);
    if (has_w) cw[WIDX(i, j, k)] = conv_w<SKEW>(g, u, v, w, i, j, k);
}

// ------------------------------------------------------------------ diffusion
// diffusive_{u,v,w}_face_kernel_staggered_3d (include/solver_kernels.hpp:2199-2337)
__device__ __forceinline__ double diff_u(const Geom& g, const double* __restrict__ u, const double* __restrict__ v,
                                         const double* __restrict__ w, const double* __restrict__ nu, int i, int j, int k) {
    const double uc = u[UIDX(i, j, k)];
    const double nu_left = nu[CIDX(i - 1, j, k)], nu_right = nu[CIDX(i, j, k)];
    const double d2x = (nu_right * (u[UIDX(i + 1, j, k)] - uc) - nu_left * (uc - u[UIDX(i - 1, j, k)])) / g.dx2;
    const double nu_n = 0.25 * (nu_left + nu_right + nu[CIDX(i - 1, j + 1, k)] + nu[CIDX(i, j + 1, k)]);
    const double nu_s = 0.25 * (nu_left + nu_right + nu[CIDX(i - 1, j - 1, k)] + nu[CIDX(i, j - 1, k)]);
    const double d2y = (nu_n * (u[UIDX(i, j + 1, k)] - uc) - nu_s * (uc - u[UIDX(i, j - 1, k)])) / g.dy2;
    const double nu_f = 0.25 * (nu_left + nu_right + nu[CIDX(i - 1, j, k + 1)] + nu[CIDX(i, j, k + 1)]);
    const double nu_b = 0.25 * (nu_left + nu_right + nu[CIDX(i - 1, j, k - 1)] + nu[CIDX(i, j, k - 1)]);
    const double d2z = (nu_f * (u[UIDX(i, j, k + 1)] - uc) - nu_b * (uc - u[UIDX(i, j, k - 1)])) / g.dz2;
    return d2x + d2y + d2z;
}
__device__ __forceinline__ double diff_v(const Geom& g, const double* __restrict__ u, const double* __restrict__ v,
                                         const double* __restrict__ w, const double* __restrict__ nu, int i, int j, int k) {
    const double vc = v[VIDX(i, j, k)];
    const double nu_bottom = nu[CIDX(i, j - 1, k)], nu_top = nu[CIDX(i, j, k)];
    const double nu_e = 0.25 * (nu_bottom + nu_top + nu[CIDX(i + 1, j - 1, k)] + nu[CIDX(i + 1, j, k)]);
    const double nu_w = 0.25 * (nu_bottom + nu_top + nu[CIDX(i - 1, j - 1, k)] + nu[CIDX(i - 1, j, k)]);
    const double d2x = (nu_e * (v[VIDX(i + 1, j, k)] - vc) - nu_w * (vc - v[VIDX(i - 1, j, k)])) / g.dx2;
    const double d2y = (nu_top * (v[VIDX(i, j + 1, k)] - vc) - nu_bottom * (vc - v[VIDX(i, j - 1, k)])) / g.dy2;
    const double nu_f = 0.25 * (nu_bottom + nu_top + nu[CIDX(i, j - 1, k + 1)] + nu[CIDX(i, j, k + 1)]);
    const double nu_b = 0.25 * (nu_bottom + nu_top + nu[CIDX(i, j - 1, k - 1)] + nu[CIDX(i, j, k - 1)]);
    const double d2z = (nu_f * (v[VIDX(i, j, k + 1)] - vc) - nu_b * (vc - v[VIDX(i, j, k - 1)])) / g.dz2;
    return d2x + d2y + d2z;
}
__device__ __forceinline__ double diff_w(const Geom& g, const double* __restrict__ u, const double* __restrict__ v,
                                         const double* __restrict__ w, const double* __restrict__ nu, int i, int j, int k) {
    const double wc = w[WIDX(i, j, k)];
    const double nu_back = nu[CIDX(i, j, k - 1)], nu_front = nu[CIDX(i, j, k)];
    const double nu_e = 0.25 * (nu_back + nu_front + nu[CIDX(i + 1, j, k - 1)] + nu[CIDX(i + 1, j, k)]);
    const double nu_w = 0.25 * (nu_back + nu_front + nu[CIDX(i - 1, j, k - 1)] + nu[CIDX(i - 1, j, k)]);
    const double d2x = (nu_e * (w[WIDX(i + 1, j, k)] - wc) - nu_w * (wc - w[WIDX(i - 1, j, k)])) / g.dx2;
    const double nu_n = 0.25 * (nu_back + nu_front + nu[CIDX(i, j + 1, k - 1)] + nu[CIDX(i, j + 1, k)]);
    const double nu_s = 0.25 * (nu_back + nu_front + nu[CIDX(i, j - 1, k - 1)] + nu[CIDX(i, j - 1, k)]);
    const double d2y = (nu_n * (w[WIDX(i, j + 1, k)] - wc) - nu_s * (wc - w[WIDX(i, j - 1, k)])) / g.dy2;
    const double d2z = (nu_front * (w[WIDX(i, j, k + 1)] - wc) - nu_back * (wc - w[WIDX(i, j, k - 1)])) / g.dz2;
    return d2x + d2y + d2z;
}
static __global__ void __launch_bounds__(256)
k_diffusive(Geom g, const double* __restrict__ u, const double* __restrict__ v,
            const double* __restrict__ w, const double* __restrict__ nu,
            double* __restrict__ du, double* __restrict__ dv, double* __restrict__ dw) {
    FACE_THREAD_PROLOGUE
    if (has_u) du[UIDX(i, j, k)] = diff_u(g, u, v, w, nu, i, j, k);
    if (has_v) dv[VIDX(i, j, k)] = diff_v(g, u, v, w, nu, i, j, k);
    if (has_w) dw[WIDX(i, j, k)] = diff_w(g, u, v, w, nu, i, j, k);
}

// ------------------------------------------------------------------ time advance / blends
// out = in + dt * (-conv + diff + f) on interior faces (euler_substep, src/solver_time.cpp:484-516;
// Euler predictor src/solver.cpp:1589-1667)
static __global__ void __launch_bounds__(256)
k_advance(Geom g, Vel in, Vel out, Vel conv, Vel diff, double dt, double fx, double fy, double fz) {
    FACE_THREAD_PROLOGUE
    if (has_u) { const size_t o = UIDX(i, j, k); out.u[o] = in.u[o] + dt * (-conv.u[o] + diff.u[o] + fx); }
    if (has_v) { const size_t o = VIDX(i, j, k); out.v[o] = in.v[o] + dt * (-conv.v[o] + diff.v[o] + fy); }
    if (has_w) { const size_t o = WIDX(i, j, k); out.w[o] = in.w[o] + dt * (-conv.w[o] + diff.w[o] + fz); }
}
// Fused momentum update: out = in + dt * (-conv(vel) + diff(vel, nu) + f) on interior faces in ONE pass
// (compute_convective_term + compute_diffusive_term + the advance loop of euler_substep,
// src/solver_time.cpp:269-516, same expressions in the same order).  Raw index space (I,J,K) over the
// ghost-inclusive arrays.  copy_ghosts = 1: every non-interior entry of out is set to vel's, so
// that a caller whose `out` aliases `vel` in the reference (RK stages 2-3) can pass a scratch
// buffer and swap it with vel afterwards — bit-identical to the in-place reference semantics.
template <bool SKEW>
__global__ void __launch_bounds__(256)
k_momentum(Geom g, Vel vel, const double* __restrict__ nu, Vel in, Vel out, double dt, double fx, double fy,
           double fz, int copy_ghosts) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    const int Ng = g.Ng;
    const int ex = g.Nx + 2 * Ng, ey = g.Ny + 2 * Ng, ez = g.Nz + 2 * Ng;   // cell-array extents
    if (i > ex || j > ey || k > ez) return;
    const bool ix_c = i >= Ng && i < Ng + g.Nx, iy_c = j >= Ng && j < Ng + g.Ny, iz_c = k >= Ng && k < Ng + g.Nz;
    const bool ix_f = i >= Ng && i <= Ng + g.Nx, iy_f = j >= Ng && j <= Ng + g.Ny, iz_f = k >= Ng && k <= Ng + g.Nz;
    if (j < ey && k < ez) {                       // u array entry (i <= ex always holds)
        const size_t o = UIDX(i, j, k);
        if (ix_f && iy_c && iz_c) {
            const double c = conv_u<SKEW>(g, vel.u, vel.v, vel.w, i, j, k);
            const double d = diff_u(g, vel.u, vel.v, vel.w, nu, i, j, k);
            out.u[o] = in.u[o] + dt * (-c + d + fx);
        } else if (copy_ghosts) {
            out.u[o] = vel.u[o];
        }
    }
    if (i < ex && k < ez) {                       // v array entry
        const size_t o = VIDX(i, j, k);
        if (ix_c && iy_f && iz_c) {
            const double c = conv_v<SKEW>(g, vel.u, vel.v, vel.w, i, j, k);
            const double d = diff_v(g, vel.u, vel.v, vel.w, nu, i, j, k);
            out.v[o] = in.v[o] + dt * (-c + d + fy);
        } else if (copy_ghosts) {
            out.v[o] = vel.v[o];
        }
    }
    if (i < ex && j < ey) {                       // w array entry
        const size_t o = WIDX(i, j, k);
        if (ix_c && iy_c && iz_f) {
            const double c = conv_w<SKEW>(g, vel.u, vel.v, vel.w, i, j, k);
            const double d = diff_w(g, vel.u, vel.v, vel.w, nu, i, j, k);
            out.w[o] = in.w[o] + dt * (-c + d + fz);
        } else if (copy_ghosts) {
            out.w[o] = vel.w[o];
        }
    }
}
// star <- vel (interior faces), vel <- vel - dt * grad p', p += p' in one pass
// (the velocity_ -> velocity_star_ copy of project_velocity, src/solver_time.cpp:972-995, fused with
// correct_velocity, src/solver_operators.cpp:1259-1604; both are pointwise in the velocity arrays)
static __global__ void __launch_bounds__(256)
k_copy_correct(Geom g, Vel star, Vel vel, const double* __restrict__ pc, double* __restrict__ p, double dt) {
    FACE_THREAD_PROLOGUE
    if (has_u) {
        const size_t o = UIDX(i, j, k);
        const double us = vel.u[o];
        const double dp = (pc[CIDX(i, j, k)] - pc[CIDX(i - 1, j, k)]) / g.dx;
        star.u[o] = us;
        vel.u[o] = us - dt * dp;
    }
    if (has_v) {
        const size_t o = VIDX(i, j, k);
        const double vs = vel.v[o];
        const double d = pc[CIDX(i, j, k)] - pc[CIDX(i, j - 1, k)];
        const double dp = g.y_stretched ? fdiv(d, g.dyc[j], g.rdyc[j]) : d / g.dy;
        star.v[o] = vs;
        vel.v[o] = vs - dt * dp;
    }
    if (has_w) {
        const size_t o = WIDX(i, j, k);
        const double ws = vel.w[o];
        const double dp = (pc[CIDX(i, j, k)] - pc[CIDX(i, j, k - 1)]) / g.dz;
        star.w[o] = ws;
        vel.w[o] = ws - dt * dp;
    }
    if (ix < g.Nx && iy < g.Ny && iz < g.Nz) {
        const size_t o = CIDX(i, j, k);
        p[o] += pc[o];
    }
}
// b <- a and c <- a on interior faces in one pass (velocity_old_ and velocity_rk_ copies at the top
// of step() / ssprk3_step: src/solver.cpp:1210-1237, src/solver_time.cpp:1313-1330)
static __global__ void __launch_bounds__(256)
k_copy2(Geom g, Vel a, Vel b, Vel c) {
    FACE_THREAD_PROLOGUE
    if (has_u) { const size_t o = UIDX(i, j, k); const double x = a.u[o]; b.u[o] = x; c.u[o] = x; }
    if (has_v) { const size_t o = VIDX(i, j, k); const double x = a.v[o]; b.v[o] = x; c.v[o] = x; }
    if (has_w) { const size_t o = WIDX(i, j, k); const double x = a.w[o]; b.w[o] = x; c.w[o] = x; }
}
// c = a1*a + a2*b on interior faces (blend_3d_uvw / blend_to_3d_uvw / copy_3d_uvw,
// src/solver_time_kernels_blend.cpp:38-158; the copy is literally 1.0*a + 0.0*b)
static __global__ void __launch_bounds__(256)
k_blend(Geom g, Vel a, Vel b, Vel c, double a1, double a2) {
    FACE_THREAD_PROLOGUE
    if (has_u) { const size_t o = UIDX(i, j, k); c.u[o] = a1 * a.u[o] + a2 * b.u[o]; }
    if (has_v) { const size_t o = VIDX(i, j, k); c.v[o] = a1 * a.v[o] + a2 * b.v[o]; }
    if (has_w) { const size_t o = WIDX(i, j, k); c.w[o] = a1 * a.w[o] + a2 * b.w[o]; }
}
// Periodic seam of the duplicated faces Ng and Ng+N.  AVG = false: far <- near (euler_substep,
// src/solver_time.cpp:518-559).  AVG = true: both <- 0.5*(near+far) (Euler predictor
// src/solver.cpp:1616-1633, correct_velocity src/solver_operators.cpp:1347-1485).
// Index space: (a, b, dir) with dir = blockIdx.z in {0:x,1:y,2:z}.
template <bool AVG>
__global__ void __launch_bounds__(256)
k_seam(Geom g, Vel vel, int px, int py, int pz) {
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y * blockDim.y + threadIdx.y;
    const int dir = blockIdx.z;
    size_t lo, hi;
    double* f;
    if (dir == 0) {          // u: a = j, b = k
        if (!px || a >= g.Ny || b >= g.Nz) return;
        f = vel.u; lo = UIDX(g.Ng, a + g.Ng, b + g.Ng); hi = UIDX(g.Ng + g.Nx, a + g.Ng, b + g.Ng);
    } else if (dir == 1) {   // v: a = i, b = k
        if (!py || a >= g.Nx || b >= g.Nz) return;
        f = vel.v; lo = VIDX(a + g.Ng, g.Ng, b + g.Ng); hi = VIDX(a + g.Ng, g.Ng + g.Ny, b + g.Ng);
    } else {                 // w: a = i, b = j
        if (!pz || a >= g.Nx || b >= g.Ny) return;
        f = vel.w; lo = WIDX(a + g.Ng, b + g.Ng, g.Ng); hi = WIDX(a + g.Ng, b + g.Ng, g.Ng + g.Nz);
    }
    if (AVG) {
        const double avg = 0.5 * (f[lo] + f[hi]);
        f[lo] = avg; f[hi] = avg;
    } else {
        f[hi] = f[lo];
    }
}

// out = 0.5 * (a + b) over one plane (periodic w seam across the end ranks of a slab run)
static __global__ void __launch_bounds__(256)
k_avg_plane(double* out, const double* a, const double* b, size_t n) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) out[t] = 0.5 * (a[t] + b[t]);
}

// ------------------------------------------------------------------ divergence / RHS / correction
// divergence_cell_kernel_staggered_3d (include/solver_kernels.hpp:568-591); stretched y divides
// by dyv[j] (src/solver_operators.cpp:1157-1175).  Also emits per-block partial sums of div for
// the mean (src/solver_time.cpp:812-866).
static __global__ void __launch_bounds__(256)
k_divergence(Geom g, Vel vel, double* __restrict__ div, double* __restrict__ part) {
    __shared__ double sh[32];
    const int ix = blockIdx.x * blockDim.x + threadIdx.x;
    const int iy = blockIdx.y * blockDim.y + threadIdx.y;
    const int iz = blockIdx.z;
    double d = 0.0;
    if (ix < g.Nx && iy < g.Ny && iz < g.Nz) {
        const int i = ix + g.Ng, j = iy + g.Ng, k = iz + g.Ng;
        const double dudx = (vel.u[UIDX(i + 1, j, k)] - vel.u[UIDX(i, j, k)]) / g.dx;
        const double dv_ = vel.v[VIDX(i, j + 1, k)] - vel.v[VIDX(i, j, k)];
        const double dvdy = g.y_stretched ? fdiv(dv_, g.dyv[j], g.rdyv[j]) : dv_ / g.dy;
        const double dwdz = (vel.w[WIDX(i, j, k + 1)] - vel.w[WIDX(i, j, k)]) / g.dz;
        d = dudx + dvdy + dwdz;
        div[CIDX(i, j, k)] = d;
    }
    const int tid = threadIdx.y * blockDim.x + threadIdx.x;
    const int lane = tid & 31, wid = tid >> 5, nw = (blockDim.x * blockDim.y + 31) >> 5;
    const int bid = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    double s = warp_sum(d);
    if (lane == 0) sh[wid] = s;
    __syncthreads();
    if (wid == 0) { s = lane < nw ? sh[lane] : 0.0; s = warp_sum(s); if (lane == 0) part[bid] = s; }
}
// Interior sum of a cell field in the reference's own order (k, j, i ascending, one running double:
// compute_mean_divergence_3d, src/solver_time_kernels_projection.cpp:31-52); out[0] = sum,
// out[1] = sum / count.  Used on y-stretched meshes only: their semi-coarsening Poisson solver
// amplifies a last-bit change of the right-hand side to ~1e-4 (measured on the reference, see
// DESIGN.md), so the mean must carry the same bits.  One CTA: all threads stage a chunk in shared
// memory, thread 0 adds it in order.
static __global__ void __launch_bounds__(256)
k_seq_sum_cells(Geom g, const double* __restrict__ a, double* __restrict__ out, double count) {
    constexpr int CH = 4096;
    __shared__ double buf[CH];
    const int n = g.Nx * g.Ny * g.Nz;
    double acc = 0.0;
    for (int base = 0; base < n; base += CH) {
        const int m = min(CH, n - base);
        __syncthreads();
        for (int t = threadIdx.x; t < m; t += blockDim.x) {
            const int q = base + t;
            const int i = q % g.Nx, j = (q / g.Nx) % g.Ny, k = q / (g.Nx * g.Ny);
            buf[t] = a[CIDX(i + g.Ng, j + g.Ng, k + g.Ng)];
        }
        __syncthreads();
        if (threadIdx.x == 0)
            for (int t = 0; t < m; ++t) acc += buf[t];
    }
    if (threadIdx.x == 0) { out[0] = acc; out[1] = acc / count; }
}
// rhs = (div - mean) * dt_inv on interior cells (src/solver_time.cpp:855-864)
static __global__ void __launch_bounds__(256)
k_build_rhs(Geom g, const double* __restrict__ div, const double* __restrict__ mean, double dt_inv,
            double* __restrict__ rhs) {
    const int ix = blockIdx.x * blockDim.x + threadIdx.x;
    const int iy = blockIdx.y * blockDim.y + threadIdx.y;
    const int iz = blockIdx.z;
    if (ix >= g.Nx || iy >= g.Ny || iz >= g.Nz) return;
    const size_t o = CIDX(ix + g.Ng, iy + g.Ng, iz + g.Ng);
    rhs[o] = (div[o] - mean[0]) * dt_inv;
}
// sum of squares of rhs over interior cells -> partials (rhs_rms, src/solver.cpp:2135-2176)
static __global__ void __launch_bounds__(256)
k_sumsq_cells(Geom g, const double* __restrict__ a, double* __restrict__ part) {
    __shared__ double sh[32];
    const int ix = blockIdx.x * blockDim.x + threadIdx.x;
    const int iy = blockIdx.y * blockDim.y + threadIdx.y;
    const int iz = blockIdx.z;
    double d = 0.0;
    if (ix < g.Nx && iy < g.Ny && iz < g.Nz) {
        const double x = a[CIDX(ix + g.Ng, iy + g.Ng, iz + g.Ng)];
        d = x * x;
    }
    const int tid = threadIdx.y * blockDim.x + threadIdx.x;
    const int lane = tid & 31, wid = tid >> 5, nw = (blockDim.x * blockDim.y + 31) >> 5;
    const int bid = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    double s = warp_sum(d);
    if (lane == 0) sh[wid] = s;
    __syncthreads();
    if (wid == 0) { s = lane < nw ? sh[lane] : 0.0; s = warp_sum(s); if (lane == 0) part[bid] = s; }
}
// u = u* - dt * grad p' on interior faces; p += p' on interior cells
// (correct_{u,v,w}_face_kernel_staggered_3d include/solver_kernels.hpp:667-715; stretched v uses
// dyc[j] src/solver_operators.cpp:1381-1400; pressure update :1490-1498)
static __global__ void __launch_bounds__(256)
k_correct(Geom g, Vel star, Vel vel, const double* __restrict__ pc, double* __restrict__ p, double dt) {
    FACE_THREAD_PROLOGUE
    if (has_u) {
        const size_t o = UIDX(i, j, k);
        const double dp = (pc[CIDX(i, j, k)] - pc[CIDX(i - 1, j, k)]) / g.dx;
        vel.u[o] = star.u[o] - dt * dp;
    }
    if (has_v) {
        const size_t o = VIDX(i, j, k);
        const double d = pc[CIDX(i, j, k)] - pc[CIDX(i, j - 1, k)];
        const double dp = g.y_stretched ? fdiv(d, g.dyc[j], g.rdyc[j]) : d / g.dy;
        vel.v[o] = star.v[o] - dt * dp;
    }
    if (has_w) {
        const size_t o = WIDX(i, j, k);
        const double dp = (pc[CIDX(i, j, k)] - pc[CIDX(i, j, k - 1)]) / g.dz;
        vel.w[o] = star.w[o] - dt * dp;
    }
    if (ix < g.Nx && iy < g.Ny && iz < g.Nz) {
        const size_t o = CIDX(i, j, k);
        p[o] += pc[o];
    }
}

// ------------------------------------------------------------------ velocity BCs
// One kernel per pass of apply_velocity_bc (src/solver_operators.cpp:43-491), same pass order.
// Only Ng == 1 (g == 0).  bc codes: CFDNN_VBC_*.
struct BcFlags { int xlo, xhi, ylo, yhi, zlo, zhi; };

// apply_normal_bc_staggered (include/solver_kernels.hpp:44-95): ptr viewed as 1-D line with
// element stride `es`, N interior cells
__device__ __forceinline__ void bc_normal_line(double* line, size_t es, int N, int Ng, int lo, int hi) {
    const bool lo_per = lo == CFDNN_VBC_PERIODIC, hi_per = hi == CFDNN_VBC_PERIODIC;
    const bool lo_ns = lo == CFDNN_VBC_NOSLIP, hi_ns = hi == CFDNN_VBC_NOSLIP;
    if (lo_per && hi_per) line[(size_t)(Ng + N) * es] = line[(size_t)Ng * es];
    if (lo_ns) {
        line[(size_t)Ng * es] = 0.0;
        line[(size_t)(Ng - 1) * es] = 0.0;
    } else if (lo_per) {
        line[(size_t)(Ng - 1) * es] = line[(size_t)(Ng + N - 1) * es];
    }
    if (hi_ns) {
        line[(size_t)(Ng + N) * es] = 0.0;
        line[(size_t)(Ng + N + 1) * es] = 0.0;
    } else if (hi_per) {
        line[(size_t)(Ng + N + 1) * es] = line[(size_t)(Ng + 1) * es];
    }
}
// apply_tangential_bc_staggered (:110-148)
__device__ __forceinline__ void bc_tangential_line(double* line, size_t es, int N, int Ng, int lo, int hi) {
    if (lo == CFDNN_VBC_NOSLIP) line[(size_t)(Ng - 1) * es] = -line[(size_t)Ng * es];
    else if (lo == CFDNN_VBC_PERIODIC) line[(size_t)(Ng - 1) * es] = line[(size_t)(N + Ng - 1) * es];
    if (hi == CFDNN_VBC_NOSLIP) line[(size_t)(N + Ng) * es] = -line[(size_t)(N + Ng - 1) * es];
    else if (hi == CFDNN_VBC_PERIODIC) line[(size_t)(N + Ng) * es] = line[(size_t)Ng * es];
}

// pass ids (in reference order)
enum { BC_U_X = 0, BC_U_Y, BC_V_X, BC_V_Y, BC_U_X2, BC_V_Y2, BC_OUTFLOW_UV, BC_INFLOW_UV,
       BC_U_Z, BC_V_Z, BC_W_Z, BC_W_X, BC_W_X_INFLOW, BC_W_X_OUTFLOW, BC_W_Y };

static __global__ void __launch_bounds__(256)
k_velocity_bc(Geom g, Vel vel, BcFlags bc, int pass, int na, int nb) {
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y * blockDim.y + threadIdx.y;
    if (a >= na || b >= nb) return;
    const int Ng = g.Ng;
    switch (pass) {
    case BC_U_X: case BC_U_X2:      // a = j (all rows), b = k (all planes)
        bc_normal_line(vel.u + UIDX(0, a, b), 1, g.Nx, Ng, bc.xlo, bc.xhi); break;
    case BC_U_Y:                     // a = i (all), b = k (all)
        bc_tangential_line(vel.u + UIDX(a, 0, b), g.us, g.Ny, Ng, bc.ylo, bc.yhi); break;
    case BC_V_X:                     // a = j (all v rows), b = k
        bc_tangential_line(vel.v + VIDX(0, a, b), 1, g.Nx, Ng, bc.xlo, bc.xhi); break;
    case BC_V_Y: case BC_V_Y2:      // a = i (all), b = k
        bc_normal_line(vel.v + VIDX(a, 0, b), g.vs, g.Ny, Ng, bc.ylo, bc.yhi); break;
    case BC_OUTFLOW_UV: {            // x_hi outflow (:196-226): a = j, b = k; u rows and v rows
        if (a < g.Ny + 2 * Ng) vel.u[UIDX(Ng + g.Nx + 1, a, b)] = vel.u[UIDX(Ng + g.Nx - 1, a, b)];
        vel.v[VIDX(Ng + g.Nx, a, b)] = vel.v[VIDX(Ng + g.Nx - 1, a, b)];
        break; }
    case BC_INFLOW_UV: {             // x_lo inflow (:229-257)
        if (a < g.Ny + 2 * Ng) vel.u[UIDX(Ng - 1, a, b)] = vel.u[UIDX(Ng, a, b)];
        vel.v[VIDX(Ng - 1, a, b)] = vel.v[VIDX(Ng, a, b)];
        break; }
    case BC_U_Z: {                   // a = i (all u), b = j (all) (:301-325)
        if (bc.zlo == CFDNN_VBC_PERIODIC && bc.zhi == CFDNN_VBC_PERIODIC) {
            vel.u[UIDX(a, b, Ng - 1)] = vel.u[UIDX(a, b, Ng + g.Nz - 1)];
            vel.u[UIDX(a, b, Ng + g.Nz)] = vel.u[UIDX(a, b, Ng)];
        } else {
            if (bc.zlo == CFDNN_VBC_NOSLIP) vel.u[UIDX(a, b, Ng - 1)] = -vel.u[UIDX(a, b, Ng)];
            if (bc.zhi == CFDNN_VBC_NOSLIP) vel.u[UIDX(a, b, Ng + g.Nz)] = -vel.u[UIDX(a, b, Ng + g.Nz - 1)];
        }
        break; }
    case BC_V_Z: {                   // a = i, b = j (all v rows) (:327-351)
        if (bc.zlo == CFDNN_VBC_PERIODIC && bc.zhi == CFDNN_VBC_PERIODIC) {
            vel.v[VIDX(a, b, Ng - 1)] = vel.v[VIDX(a, b, Ng + g.Nz - 1)];
            vel.v[VIDX(a, b, Ng + g.Nz)] = vel.v[VIDX(a, b, Ng)];
        } else {
            if (bc.zlo == CFDNN_VBC_NOSLIP) vel.v[VIDX(a, b, Ng - 1)] = -vel.v[VIDX(a, b, Ng)];
            if (bc.zhi == CFDNN_VBC_NOSLIP) vel.v[VIDX(a, b, Ng + g.Nz)] = -vel.v[VIDX(a, b, Ng + g.Nz - 1)];
        }
        break; }
    case BC_W_Z: {                   // a = i, b = j (:353-388): normal component in z
        if (bc.zlo == CFDNN_VBC_PERIODIC && bc.zhi == CFDNN_VBC_PERIODIC) {
            vel.w[WIDX(a, b, Ng + g.Nz)] = vel.w[WIDX(a, b, Ng)];
            vel.w[WIDX(a, b, Ng - 1)] = vel.w[WIDX(a, b, Ng + g.Nz - 1)];
            vel.w[WIDX(a, b, Ng + g.Nz + 1)] = vel.w[WIDX(a, b, Ng + 1)];
        } else {
            if (bc.zlo == CFDNN_VBC_NOSLIP) {
                vel.w[WIDX(a, b, Ng)] = 0.0;
                vel.w[WIDX(a, b, Ng - 1)] = -vel.w[WIDX(a, b, Ng + 1)];
            }
            if (bc.zhi == CFDNN_VBC_NOSLIP) {
                vel.w[WIDX(a, b, Ng + g.Nz)] = 0.0;
                vel.w[WIDX(a, b, Ng + g.Nz + 1)] = -vel.w[WIDX(a, b, Ng + g.Nz - 1)];
            }
        }
        break; }
    case BC_W_X: {                   // a = j (all), b = k (all w planes) (:391-411)
        if (bc.xlo == CFDNN_VBC_PERIODIC && bc.xhi == CFDNN_VBC_PERIODIC) {
            vel.w[WIDX(Ng - 1, a, b)] = vel.w[WIDX(Ng + g.Nx - 1, a, b)];
            vel.w[WIDX(Ng + g.Nx, a, b)] = vel.w[WIDX(Ng, a, b)];
        } else {
            if (bc.xlo == CFDNN_VBC_NOSLIP) vel.w[WIDX(Ng - 1, a, b)] = -vel.w[WIDX(Ng, a, b)];
            if (bc.xhi == CFDNN_VBC_NOSLIP) vel.w[WIDX(Ng + g.Nx, a, b)] = -vel.w[WIDX(Ng + g.Nx - 1, a, b)];
        }
        break; }
    case BC_W_X_INFLOW: vel.w[WIDX(Ng - 1, a, b)] = vel.w[WIDX(Ng, a, b)]; break;            // :413-425
    case BC_W_X_OUTFLOW: vel.w[WIDX(Ng + g.Nx, a, b)] = vel.w[WIDX(Ng + g.Nx - 1, a, b)]; break;  // :427-441
    case BC_W_Y: {                   // a = i (all), b = k (all w planes) (:443-463)
        if (bc.ylo == CFDNN_VBC_PERIODIC && bc.yhi == CFDNN_VBC_PERIODIC) {
            vel.w[WIDX(a, Ng - 1, b)] = vel.w[WIDX(a, Ng + g.Ny - 1, b)];
            vel.w[WIDX(a, Ng + g.Ny, b)] = vel.w[WIDX(a, Ng, b)];
        } else {
            if (bc.ylo == CFDNN_VBC_NOSLIP) vel.w[WIDX(a, Ng - 1, b)] = -vel.w[WIDX(a, Ng, b)];
            if (bc.yhi == CFDNN_VBC_NOSLIP) vel.w[WIDX(a, Ng + g.Ny, b)] = -vel.w[WIDX(a, Ng + g.Ny - 1, b)];
        }
        break; }
    }
}

// ------------------------------------------------------------------ nu_eff, reductions
// nu_eff = nu everywhere; interior += nu_t when a closure is attached (src/solver.cpp:1326-1410)
static __global__ void __launch_bounds__(256)
k_nu_eff(Geom g, double nu, const double* __restrict__ nut, int has_model, double* __restrict__ nueff) {
    const int I = blockIdx.x * blockDim.x + threadIdx.x;
    const int J = blockIdx.y * blockDim.y + threadIdx.y;
    const int K = blockIdx.z;
    if (I >= g.Nx + 2 * g.Ng || J >= g.Ny + 2 * g.Ng) return;
    const size_t o = CIDX(I, J, K);
    const bool interior = I >= g.Ng && I < g.Ng + g.Nx && J >= g.Ng && J < g.Ng + g.Ny && K >= g.Ng && K < g.Ng + g.Nz;
    nueff[o] = (has_model && interior) ? nu + nut[o] : nu;
}

// max |new - old| over interior faces of u, v, w; NaN -> 1e30 (src/solver.cpp:2659-2775)
static __global__ void __launch_bounds__(256)
k_max_change(Geom g, Vel a, Vel b, double* __restrict__ part) {
    __shared__ double sh[32];
    const int ix = blockIdx.x * blockDim.x + threadIdx.x;
    const int iy = blockIdx.y * blockDim.y + threadIdx.y;
    const int iz = blockIdx.z;
    double m = 0.0;
    if (ix <= g.Nx && iy <= g.Ny && iz <= g.Nz) {
        const int i = ix + g.Ng, j = iy + g.Ng, k = iz + g.Ng;
        if (iy < g.Ny && iz < g.Nz) { double d = fabs(a.u[UIDX(i, j, k)] - b.u[UIDX(i, j, k)]); if (!(d == d)) d = 1.0e30; m = fmax(m, d); }
        if (ix < g.Nx && iz < g.Nz) { double d = fabs(a.v[VIDX(i, j, k)] - b.v[VIDX(i, j, k)]); if (!(d == d)) d = 1.0e30; m = fmax(m, d); }
        if (ix < g.Nx && iy < g.Ny) { double d = fabs(a.w[WIDX(i, j, k)] - b.w[WIDX(i, j, k)]); if (!(d == d)) d = 1.0e30; m = fmax(m, d); }
    }
    const int tid = threadIdx.y * blockDim.x + threadIdx.x;
    const int lane = tid & 31, wid = tid >> 5, nw = (blockDim.x * blockDim.y + 31) >> 5;
    const int bid = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    double s = warp_max(m);
    if (lane == 0) sh[wid] = s;
    __syncthreads();
    if (wid == 0) { s = lane < nw ? sh[lane] : 0.0; s = warp_max(s); if (lane == 0) part[bid] = s; }
}

// compute_adaptive_dt maxima (src/solver.cpp:3369-3729): per-block partial maxima of
// |u| (cell-adjacent faces), |v|, |w|, |v|/dy_local (dyv indexed with the 0-based j — the
// reference's own quirk, :3440,3579,3649) and nu + nu_t.  part layout: [5][nblocks].
// A CTA strides over (j, k) rows and its threads over i, so the five block reductions (warp shuffles + one
// shared-memory exchange) happen once per CTA and a launch of 148 x 8 CTAs leaves 1184 partials per quantity
// instead of one per 256 cells.  Maxima are order-independent: any partition gives the same bits.
static __global__ void __launch_bounds__(256)
k_cfl_maxima(Geom g, Vel vel, const double* __restrict__ nut, double nu, int has_model,
             double* __restrict__ part, int nblocks) {
    __shared__ double sh[5][8];
    double m[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
    const int nrows = g.Ny * g.Nz;
    for (int row = blockIdx.x; row < nrows; row += gridDim.x) {
        const int iy = row % g.Ny, iz = row / g.Ny;
        const int j = iy + g.Ng, k = iz + g.Ng;
        for (int ix = threadIdx.x; ix < g.Nx; ix += blockDim.x) {
            const int i = ix + g.Ng;
            const double mu = fmax(fabs(vel.u[UIDX(i, j, k)]), fabs(vel.u[UIDX(i + 1, j, k)]));
            const double mv = fmax(fabs(vel.v[VIDX(i, j, k)]), fabs(vel.v[VIDX(i, j + 1, k)]));
            const double mw = fmax(fabs(vel.w[WIDX(i, j, k)]), fabs(vel.w[WIDX(i, j, k + 1)]));
            m[0] = fmax(m[0], mu); m[1] = fmax(m[1], mv); m[2] = fmax(m[2], mw);
            m[3] = fmax(m[3], g.y_stretched ? mv / g.dyv[iy] : mv / g.dy.d);
            if (has_model) m[4] = fmax(m[4], nu + nut[CIDX(i, j, k)]);
        }
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int q = 0; q < 5; ++q) {
        const double s = warp_max(m[q]);
        if (lane == 0) sh[q][wid] = s;
    }
    __syncthreads();
    if (wid == 0) {
#pragma unroll
        for (int q = 0; q < 5; ++q) {
            double s = lane < nw ? sh[q][lane] : 0.0;
            s = warp_max(s);
            if (lane == 0) part[(size_t)q * nblocks + blockIdx.x] = s;
        }
    }
}
// final max over several partial rows: out[q] = max(init[q], max(part[q][:]))
static __global__ void k_final_max_rows(const double* __restrict__ part, int n, int rows, double* __restrict__ out) {
    __shared__ double sh[32];
    for (int q = 0; q < rows; ++q) {
        double s = 0.0;
        for (int t = threadIdx.x; t < n; t += blockDim.x) s = fmax(s, part[(size_t)q * n + t]);
        s = block_max(s, sh);
        if (threadIdx.x == 0) out[q] = s;
        __syncthreads();
    }
}
// Trip-region forcing on the provisional velocity (Euler predictor, src/solver.cpp:1724-1810): v* += dt A env_x[i] g_y[j]
// Fz_v[k] on the v faces, w* += dt A_w env_x[i] (g_y[j] + g_y[j+1]) / 2 Fz_w[k] on the w faces; the tables come from the
// caller (initialize_trip_forcing :818-923).  fz == nullptr: 2-D, Fz = 1.  plane0: first z plane, nk planes.
static __global__ void k_trip_v(Geom g, double* __restrict__ v, const double* __restrict__ env_x, const double* __restrict__ g_y,
                                const double* __restrict__ fz, double A, double dt, int plane0, int nk) {
    const int ix = blockIdx.x * blockDim.x + threadIdx.x, iy = blockIdx.y * blockDim.y + threadIdx.y, iz = blockIdx.z;
    if (ix >= g.Nx || iy > g.Ny || iz >= nk) return;
    const int i = ix + g.Ng, j = iy + g.Ng, k = plane0 + iz;
    const double Fz = fz ? fz[k] : 1.0;
    const double f_trip = A * env_x[i] * g_y[j] * Fz;
    v[VIDX(i, j, k)] += dt * f_trip;
}
static __global__ void k_trip_w(Geom g, double* __restrict__ w, const double* __restrict__ env_x, const double* __restrict__ g_y,
                                const double* __restrict__ fz, double A_w, double dt) {
    const int ix = blockIdx.x * blockDim.x + threadIdx.x, iy = blockIdx.y * blockDim.y + threadIdx.y, iz = blockIdx.z;
    if (ix >= g.Nx || iy >= g.Ny || iz > g.Nz) return;
    const int i = ix + g.Ng, j = iy + g.Ng, k = iz + g.Ng;
    const double gy = 0.5 * (g_y[j] + g_y[j + 1]);              // j > 0 and j < size - 1 always hold for interior rows
    const double f_trip = A_w * env_x[i] * gy * fz[k];
    w[WIDX(i, j, k)] += dt * f_trip;
}

// Explicit Laplacian velocity filter (apply_velocity_filter, src/solver.cpp:4065-4162): q += alpha (Lx + Lz) + alpha_y Ly
// with second differences of the UNFILTERED field `old` (a copy of q made by the caller); the y term is dropped on the
// wall-adjacent rows (u, w: first and last cell row; v: the two wall faces).  COMP 0/1/2 = u/v/w.
template <int COMP>
static __global__ void k_velocity_filter(Geom g, double* __restrict__ q, const double* __restrict__ old, double alpha,
                                         double alpha_y) {
    const int ix = blockIdx.x * blockDim.x + threadIdx.x, iy = blockIdx.y * blockDim.y + threadIdx.y, iz = blockIdx.z;
    const int nx = g.Nx + (COMP == 0), ny = g.Ny + (COMP == 1), nz = g.Nz + (COMP == 2);
    if (ix >= nx || iy >= ny || iz >= nz) return;
    const int i = ix + g.Ng, j = iy + g.Ng, k = iz + g.Ng;
    const size_t sj = COMP == 0 ? g.us : COMP == 1 ? g.vs : g.ws;
    const size_t sk = COMP == 0 ? g.up : COMP == 1 ? g.vp : g.wp;
    const size_t c = (size_t)k * sk + (size_t)j * sj + i;
    const double Lx = old[c - 1] - 2.0 * old[c] + old[c + 1];
    const double Lz = old[c - sk] - 2.0 * old[c] + old[c + sk];
    double Ly = 0.0;
    const bool inner = COMP == 1 ? (j > g.Ng && j < g.Ny + g.Ng) : (j > g.Ng && j < g.Ny + g.Ng - 1);
    if (inner) Ly = old[c - sj] - 2.0 * old[c] + old[c + sj];
    q[c] = old[c] + (alpha * (Lx + Lz) + alpha_y * Ly);
}

// non-finite flag over a flat array: x - x != 0 catches NaN and Inf (check_for_nan_inf :3195-3350)
static __global__ void __launch_bounds__(256)
k_nonfinite(const double* __restrict__ a, size_t n, int* __restrict__ flag) {
    int bad = 0;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (size_t)gridDim.x * blockDim.x) {
        const double x = a[t];
        if (x - x != 0.0) bad = 1;
    }
    if (bad) *flag = 1;
}

static __global__ void k_flag_to_double(const int* __restrict__ flag, double* __restrict__ out) {
    out[0] = flag[0] ? 1.0 : 0.0;
}

}  // namespace cfdnn
