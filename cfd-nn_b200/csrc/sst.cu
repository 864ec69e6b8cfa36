// sst.cu — the reference's 2-D RANS closures, device-resident inside step(): SST k-omega transport + SST closure (the
// file's namesake), standard k-omega + Boussinesq, SST + EARSM (Wallin-Johansson / Gatski-Speziale / Pope), and the
// algebraic models (mixing length, algebraic k-omega, GEP).  Selected with cfdnn_solver_set_rans_model(CFDNN_RANS_*).
//
// Reference: SSTKOmegaTransport (src/turbulence_transport.cpp:315-958, constants include/turbulence_transport.hpp:32-58)
// with its default SSTClosure (:234-313).  Parity is defined against the reference's CPU path, which advances k and
// omega IN PLACE, one sweep over the cells in (j, i) order (:818-860): a cell sees the new values of its west and south
// neighbours and the old values of its east and north ones.  That order is a data dependence, not an accident of the
// loop, so the GPU version keeps it: cells on an anti-diagonal i + j = d depend only on diagonal d - 1, and one CTA
// walks the diagonals with a barrier between them (k_sst_sweep).  A 2-D RANS grid has 10^4 .. 10^5 cells; the sweep is
// latency-bound by construction (Nx + Ny - 1 barriers) and costs tens of microseconds.  The reference's own GPU path
// runs the same cell function as a Jacobi update (every neighbour old) and therefore differs from its CPU path.
//
// The model is 2-D in the reference (fields indexed (i, j) only); 3-D meshes are not supported here.
#include "sst.cuh"
#include <cmath>
#include <vector>

namespace cfdnn {
namespace {

struct SstGrid {                    // grid constants of the sweep, computed on the host like the reference does (:772-781)
    double dx, dy, dx2, dy2, inv_2dx, inv_2dy;
};

__device__ __forceinline__ double hi_of(double a, double b) { return (a > b) ? a : b; }     // (a > b) ? a : b
__device__ __forceinline__ double lo_of(double a, double b) { return (a < b) ? a : b; }     // (a < b) ? a : b

// The values a cell update reads: k and omega at the cell and its four neighbours (raw, ghosts included), the
// previous nu_t, the wall distance, the MAC velocity differences and the cell-centred velocity.
struct SstCellIn {
    double kC, kE, kW, kN, kS;
    double oC, oE, oW, oN, oS;
    double nut, ywall;
    double dudx, dudy, dvdx, dvdy, uc, vc;
};

// One point-implicit SST update (sst_transport_cell_kernel, src/turbulence_transport.cpp:56-190): F1 blending from the
// cross-diffusion term, limited production, first-order upwind advection, central diffusion with nu + sigma nu_t,
// destruction treated implicitly, clipping.  Expression order is the reference's (parity to the last bits on the
// CPU test double; tanh / sqrt are the only library calls).
__device__ __forceinline__ void sst_cell_update(const SstParams& p, const SstGrid& g, double dt, const SstCellIn& c,
                                                double& k_out, double& omega_out) {
    const double kc = hi_of(c.kC, p.k_min);
    const double oc = hi_of(c.oC, p.omega_min);
    const double y = hi_of(c.ywall, 1e-10);
    const double nut = hi_of(c.nut, 0.0);

    const double Sxy = 0.5 * (c.dudy + c.dvdx);
    const double S2 = 2.0 * (c.dudx * c.dudx + c.dvdy * c.dvdy + 2.0 * Sxy * Sxy);

    const double dkdx = (c.kE - c.kW) * g.inv_2dx, dkdy = (c.kN - c.kS) * g.inv_2dy;
    const double dodx = (c.oE - c.oW) * g.inv_2dx, dody = (c.oN - c.oS) * g.inv_2dy;
    const double grad_dot = dkdx * dodx + dkdy * dody;

    const double CDkw = hi_of(2.0 * p.sigma_omega2 / oc * grad_dot, p.CD_omega_min);
    const double a_visc = 500.0 * p.nu / (y * y * oc);
    const double a_turb = sqrt(kc) / (p.beta_star * oc * y);
    const double a_cd = 4.0 * p.sigma_omega2 * kc / (CDkw * y * y);
    const double arg1 = lo_of(hi_of(a_turb, a_visc), a_cd);
    const double F1 = tanh(arg1 * arg1 * arg1 * arg1);
    const double G1 = 1.0 - F1;

    const double beta = F1 * p.beta1 + G1 * p.beta2;
    const double alpha = F1 * p.alpha1 + G1 * p.alpha2;
    const double nu_k = p.nu + (F1 * p.sigma_k1 + G1 * p.sigma_k2) * nut;
    const double nu_o = p.nu + (F1 * p.sigma_omega1 + G1 * p.sigma_omega2) * nut;

    const double Pk = lo_of(2.0 * nut * S2, 10.0 * p.beta_star * kc * oc);

    double adv_k, adv_o;
    if (c.uc >= 0) { adv_k = c.uc * (kc - c.kW) / g.dx; adv_o = c.uc * (oc - c.oW) / g.dx; }
    else           { adv_k = c.uc * (c.kE - kc) / g.dx; adv_o = c.uc * (c.oE - oc) / g.dx; }
    if (c.vc >= 0) { adv_k += c.vc * (kc - c.kS) / g.dy; adv_o += c.vc * (oc - c.oS) / g.dy; }
    else           { adv_k += c.vc * (c.kN - kc) / g.dy; adv_o += c.vc * (c.oN - oc) / g.dy; }

    const double lap_k = (c.kE - 2.0 * kc + c.kW) / g.dx2 + (c.kN - 2.0 * kc + c.kS) / g.dy2;
    const double lap_o = (c.oE - 2.0 * oc + c.oW) / g.dx2 + (c.oN - 2.0 * oc + c.oS) / g.dy2;
    const double dif_k = nu_k * lap_k, dif_o = nu_o * lap_o;

    const double CD = hi_of(2.0 * G1 * p.sigma_omega2 / oc * grad_dot, 0.0);

    const double src_k = Pk + dif_k - adv_k;
    const double src_o = alpha * (oc / kc) * Pk + dif_o - adv_o + CD;
    double kn = (kc + dt * src_k) / (1.0 + dt * (p.beta_star * oc));
    double on = (oc + dt * src_o) / (1.0 + dt * (beta * oc));
    kn = lo_of(hi_of(kn, p.k_min), p.k_max);
    on = lo_of(hi_of(on, p.omega_min), p.omega_max);
    k_out = kn;
    omega_out = on;
}

// Standard k-omega (Wilcox 1988), KOmegaTransport::advance_turbulence host path (src/turbulence_transport.cpp:1062-1118):
// no blending, no cross-diffusion, the advecting velocity is the FACE value u(i, j), v(i, j) (not the cell average), the
// strain comes from compute_gradients_from_mac (stretched-aware), every k / omega entering a difference is raw.
struct KwParams { double nu, sigma_k, sigma_omega, beta, beta_star, alpha, k_min, k_max, omega_min, omega_max; };

__device__ __forceinline__ void kw_cell_update(const KwParams& p, const SstGrid& g, double dt, const SstCellIn& c,
                                               double& k_out, double& omega_out) {
    const double k_old = (p.k_min < c.kC) ? c.kC : p.k_min;            // std::max(k_min, k)
    const double o_old = (p.omega_min < c.oC) ? c.oC : p.omega_min;
    const double nut = (0.0 < c.nut) ? c.nut : 0.0;                     // std::max(0.0, nu_t)
    const double Sxy = 0.5 * (c.dudy + c.dvdx);
    const double S2 = 2.0 * (c.dudx * c.dudx + c.dvdy * c.dvdy + 2.0 * Sxy * Sxy);
    const double Pk = 2.0 * nut * S2;
    const double nu_k = p.nu + p.sigma_k * nut, nu_o = p.nu + p.sigma_omega * nut;
    const double dkdx = (c.uc >= 0) ? (c.kC - c.kW) / g.dx : (c.kE - c.kC) / g.dx;
    const double dkdy = (c.vc >= 0) ? (c.kC - c.kS) / g.dy : (c.kN - c.kC) / g.dy;
    const double dodx = (c.uc >= 0) ? (c.oC - c.oW) / g.dx : (c.oE - c.oC) / g.dx;
    const double dody = (c.vc >= 0) ? (c.oC - c.oS) / g.dy : (c.oN - c.oC) / g.dy;
    const double adv_k = c.uc * dkdx + c.vc * dkdy, adv_o = c.uc * dodx + c.vc * dody;
    const double dif_k = nu_k * ((c.kE - 2.0 * c.kC + c.kW) / g.dx2 + (c.kN - 2.0 * c.kC + c.kS) / g.dy2);
    const double dif_o = nu_o * ((c.oE - 2.0 * c.oC + c.oW) / g.dx2 + (c.oN - 2.0 * c.oC + c.oS) / g.dy2);
    const double src_k = Pk + dif_k - adv_k;
    const double src_o = p.alpha * (o_old / k_old) * Pk + dif_o - adv_o;
    const double kn = (k_old + dt * src_k) / (1.0 + dt * (p.beta_star * o_old));
    const double on = (o_old + dt * src_o) / (1.0 + dt * (p.beta * o_old));
    const double kl = (kn < p.k_min) ? p.k_min : kn, ol = (on < p.omega_min) ? p.omega_min : on;   // std::max(new, min)
    k_out = (p.k_max < kl) ? p.k_max : kl;                               // std::min(., max)
    omega_out = (p.omega_max < ol) ? p.omega_max : ol;
}

// The in-place sweep as a wavefront: ONE CTA, diagonals d = ix + iy in increasing order, a barrier after each.
// k and omega are read and written through plain pointers (no read-only path: a neighbour written on the previous
// diagonal must be seen).  u, v, nu_t, wall are read-only.
// SST = true: SSTKOmegaTransport; false: KOmegaTransport (dyc != nullptr: stretched-aware y differences of the strain).
template <bool SST>
__global__ void __launch_bounds__(1024)
k_sst_sweep(Geom g, SstParams p, KwParams kw, SstGrid sg, double dt, const double* __restrict__ u, const double* __restrict__ v,
            double* k, double* omega, const double* __restrict__ nu_t, const double* __restrict__ wall,
            const double* __restrict__ dyc) {
    const int Nx = g.Nx, Ny = g.Ny, Ng = g.Ng;
    const int cs = g.cs, us = g.us, vs = g.vs;
    for (int d = 0; d < Nx + Ny - 1; ++d) {
        const int iy_lo = d - (Nx - 1) > 0 ? d - (Nx - 1) : 0, iy_hi = d < Ny - 1 ? d : Ny - 1;
        for (int iy = iy_lo + (int)threadIdx.x; iy <= iy_hi; iy += blockDim.x) {
            const int i = d - iy + Ng, j = iy + Ng;
            const int o = j * cs + i;
            SstCellIn c;
            c.kC = k[o]; c.kE = k[o + 1]; c.kW = k[o - 1]; c.kN = k[o + cs]; c.kS = k[o - cs];
            c.oC = omega[o]; c.oE = omega[o + 1]; c.oW = omega[o - 1]; c.oN = omega[o + cs]; c.oS = omega[o - cs];
            c.nut = nu_t[o];
            c.ywall = wall[j];
            const int ou = j * us + i, ov = j * vs + i;
            const double wy = (!SST && dyc) ? 1.0 / (dyc[j] + dyc[j + 1]) : sg.inv_2dy;
            c.dudx = (u[ou + 1] - u[ou - 1]) * sg.inv_2dx;
            c.dudy = (u[ou + us] - u[ou - us]) * wy;
            c.dvdx = (v[ov + 1] - v[ov - 1]) * sg.inv_2dx;
            c.dvdy = (v[ov + vs] - v[ov - vs]) * wy;
            c.uc = SST ? 0.5 * (u[ou] + u[ou + 1]) : u[ou];
            c.vc = SST ? 0.5 * (v[ov] + v[ov + vs]) : v[ov];
            double kn, on;
            if (SST) sst_cell_update(p, sg, dt, c, kn, on);
            else kw_cell_update(kw, sg, dt, c, kn, on);
            k[o] = kn;
            omega[o] = on;
        }
        __syncthreads();
    }
}

// Wall BCs after the sweep (apply_wall_bc_k :547-567: ghost = -first/last interior row, every column, ghost columns
// included; apply_wall_bc_omega :569-603: ghost = 2 omega_wall - interior, interior columns only)
__global__ void k_sst_wall_bc(Geom g, SstParams p, double* __restrict__ k, double* __restrict__ omega,
                              const double* __restrict__ wall) {
    const int I = blockIdx.x * blockDim.x + threadIdx.x;
    const int Ng = g.Ng, cs = g.cs, jb = Ng, jt = g.Ny + Ng - 1;
    if (I >= g.Nx + 2 * Ng) return;
    k[(jb - 1) * cs + I] = -k[jb * cs + I];
    k[(jt + 1) * cs + I] = -k[jt * cs + I];
    if (I < Ng || I >= g.Nx + Ng) return;
    {
        const double y1 = hi_of(wall[jb], 1e-10);                       // std::max(wall_distance, 1e-10)
        double ow = 10.0 * 6.0 * p.nu / (p.beta1 * y1 * y1);
        ow = (p.omega_max < ow) ? p.omega_max : ow;                     // std::min(omega_wall, omega_max)
        omega[(jb - 1) * cs + I] = 2.0 * ow - omega[jb * cs + I];
    }
    {
        const double y1 = hi_of(wall[jt], 1e-10);
        double ow = 10.0 * 6.0 * p.nu / (p.beta1 * y1 * y1);
        ow = (p.omega_max < ow) ? p.omega_max : ow;
        omega[(jt + 1) * cs + I] = 2.0 * ow - omega[jt * cs + I];
    }
}

// SSTClosure::compute_nu_t (:259-313): nu_t = a1 k / max(a1 omega, |S| F2), F2 = tanh(max(2 sqrt k / (beta* omega y),
// 500 nu / (y^2 omega))^2), gradients from the MAC faces (compute_gradients_from_mac, src/gpu_kernels.cpp:17-130;
// stretched meshes divide by dyc[j] + dyc[j+1]), clipped to [0, 1000 nu]
__global__ void k_sst_closure(Geom g, SstParams p, double inv_2dx, double inv_2dy, const double* __restrict__ dyc,
                              const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ k,
                              const double* __restrict__ omega, const double* __restrict__ wall, double* __restrict__ nu_t) {
    const int ix = blockIdx.x * blockDim.x + threadIdx.x, iy = blockIdx.y * blockDim.y + threadIdx.y;
    if (ix >= g.Nx || iy >= g.Ny) return;
    const int i = ix + g.Ng, j = iy + g.Ng;
    const int o = j * g.cs + i, ou = j * g.us + i, ov = j * g.vs + i;
    const double wy = dyc ? 1.0 / (dyc[j] + dyc[j + 1]) : inv_2dy;
    const double dudx = (u[ou + 1] - u[ou - 1]) * inv_2dx;
    const double dudy = (u[ou + g.us] - u[ou - g.us]) * wy;
    const double dvdx = (v[ov + 1] - v[ov - 1]) * inv_2dx;
    const double dvdy = (v[ov + g.vs] - v[ov - g.vs]) * wy;
    const double kl = (p.k_min < k[o]) ? k[o] : p.k_min;                // std::max(k_min, k)
    const double ol = (p.omega_min < omega[o]) ? omega[o] : p.omega_min;
    const double Sxy = 0.5 * (dudy + dvdx);
    const double Smag = sqrt(2.0 * (dudx * dudx + dvdy * dvdy + 2.0 * Sxy * Sxy));
    // compute_F2 (:244-257) clamps its arguments once more
    const double ks = (p.k_min < kl) ? kl : p.k_min;
    const double os = (p.omega_min < ol) ? ol : p.omega_min;
    const double ys = (1e-10 < wall[j]) ? wall[j] : 1e-10;
    const double t1 = 2.0 * sqrt(ks) / (p.beta_star * os * ys);
    const double t2 = 500.0 * p.nu / (ys * ys * os);
    const double arg2 = (t1 < t2) ? t2 : t1;
    const double F2 = tanh(arg2 * arg2);
    const double den_a = p.a1 * ol, den_b = Smag * F2;
    const double den = (den_a < den_b) ? den_b : den_a;
    const double mag = (fabs(den) < 1e-10) ? 1e-10 : fabs(den);        // numerics::safe_divide with K_FLOOR
    double r = p.a1 * kl / ((den < 0.0) ? -mag : mag);
    const double cap = 1000.0 * p.nu;
    r = (r < 0.0) ? 0.0 : ((cap < r) ? cap : r);                        // std::clamp
    nu_t[o] = r;
}
// BoussinesqClosure::compute_nu_t (:202-228): nu_t = k / omega with the numerics:: floors, |.| capped at 1000 nu
__global__ void k_kw_closure(Geom g, double nu, const double* __restrict__ k, const double* __restrict__ omega,
                             double* __restrict__ nu_t) {
    const int ix = blockIdx.x * blockDim.x + threadIdx.x, iy = blockIdx.y * blockDim.y + threadIdx.y;
    if (ix >= g.Nx || iy >= g.Ny) return;
    const int o = (iy + g.Ng) * g.cs + ix + g.Ng;
    const double kl = (1e-10 < k[o]) ? k[o] : 1e-10;                    // std::max(K_FLOOR, k)
    const double ol = (1e-10 < omega[o]) ? omega[o] : 1e-10;            // std::max(OMEGA_FLOOR, omega)
    const double mag = (fabs(ol) < 1e-10) ? 1e-10 : fabs(ol);          // bounded_ratio -> safe_divide(k, omega, OMEGA_FLOOR)
    const double r = kl / ((ol < 0.0) ? -mag : mag);
    const double cap = 1000.0 * nu;
    nu_t[o] = (r < -cap) ? -cap : ((cap < r) ? cap : r);                // std::clamp(ratio, -ceiling, ceiling)
}
// ---- the two algebraic 2-D RANS baselines (src/turbulence_baseline.cpp): both scale with a friction velocity taken from
// the mean |du/dy| of the first interior row, summed over i in order like the reference's loop (one thread: Nx values).
// ML = true : MixingLengthModel (:330-352) — one-sided difference u(i, j+1) - u(i, j) over the local centre spacing
// ML = false: AlgebraicKOmegaModel (:433-447) — the MAC central difference dudy(i, j) of compute_gradients_from_mac
template <bool ML>
__global__ void k_rans_utau(Geom g, double nu, double inv_2dy, double dy_local, const double* __restrict__ dyc,
                            const double* __restrict__ u, double* __restrict__ out) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    const int j = g.Ng;
    double acc = 0.0;
    for (int ix = 0; ix < g.Nx; ++ix) {
        const int ou = j * g.us + ix + g.Ng;
        if (ML) acc += fabs((u[ou + g.us] - u[ou]) / dy_local);
        else acc += fabs((u[ou + g.us] - u[ou - g.us]) * (dyc ? 1.0 / (dyc[j] + dyc[j + 1]) : inv_2dy));
    }
    acc /= g.Nx;
    const double ut = sqrt(nu * acc);
    out[0] = (ut < 1e-10) ? 1e-10 : ut;                                 // std::max(u_tau, 1e-10)
}

// MixingLengthModel: l = kappa y (1 - exp(-y+ / A+)) capped at delta / 2, nu_t = l^2 |S| capped at 1000 nu, under-relaxed
// with the previous nu_t (mixing_length_cell_kernel :36-76)
__global__ void k_mixing_length(Geom g, double nu, double kappa, double A_plus, double delta, double inv_2dx, double inv_2dy,
                                const double* __restrict__ dyc, const double* __restrict__ u, const double* __restrict__ v,
                                const double* __restrict__ wall, const double* __restrict__ utau, double* __restrict__ nu_t) {
    const int ix = blockIdx.x * blockDim.x + threadIdx.x, iy = blockIdx.y * blockDim.y + threadIdx.y;
    if (ix >= g.Nx || iy >= g.Ny) return;
    const int i = ix + g.Ng, j = iy + g.Ng;
    const int o = j * g.cs + i, ou = j * g.us + i, ov = j * g.vs + i;
    const double wy = dyc ? 1.0 / (dyc[j] + dyc[j + 1]) : inv_2dy;
    const double dudx = (u[ou + 1] - u[ou - 1]) * inv_2dx, dudy = (u[ou + g.us] - u[ou - g.us]) * wy;
    const double dvdx = (v[ov + 1] - v[ov - 1]) * inv_2dx, dvdy = (v[ov + g.vs] - v[ov - g.vs]) * wy;
    const double Sxy = 0.5 * (dudy + dvdx);
    const double Smag = sqrt(2.0 * (dudx * dudx + dvdy * dvdy + 2.0 * Sxy * Sxy));
    const double yw = wall[j];
    const double y_plus = yw * utau[0] / nu;
    const double damping = 1.0 - exp(-y_plus / A_plus);
    double l_mix = kappa * yw * damping;
    if (l_mix > 0.5 * delta) l_mix = 0.5 * delta;
    double nut_new = l_mix * l_mix * Smag;
    const double cap = 1000.0 * nu;
    if (nut_new > cap) nut_new = cap;
    nu_t[o] = 0.5 * nut_new + 0.5 * nu_t[o];
}

// AlgebraicKOmegaModel (:449-470): k and omega estimated from u_tau and the wall distance, nu_t = k / omega
__global__ void k_algebraic_komega(Geom g, double nu, double C_mu, const double* __restrict__ wall,
                                   const double* __restrict__ utau, double* __restrict__ nu_t) {
    const int ix = blockIdx.x * blockDim.x + threadIdx.x, iy = blockIdx.y * blockDim.y + threadIdx.y;
    if (ix >= g.Nx || iy >= g.Ny) return;
    const int j = iy + g.Ng, o = j * g.cs + ix + g.Ng;
    const double ut = utau[0], yw = wall[j];
    const double y_plus = yw * ut / nu;
    const double f_mu = 1.0 - exp(-y_plus / 26.0);                                      // numerics::A_PLUS
    const double k_est = (ut * ut / sqrt(C_mu)) * f_mu * f_mu;
    const double yf = (yw < 1e-10) ? 1e-10 : yw;                                        // std::max(y_wall, Y_WALL_FLOOR)
    double o_est = ut / (0.41 * yf * f_mu + 1e-10);                                     // numerics::KAPPA
    if (y_plus < 5.0) o_est = 6.0 * nu / (0.075 * yw * yw + 1e-20);
    const double od = (o_est < 1e-10) ? 1e-10 : o_est;
    const double r = k_est / od;
    const double r0 = (0.0 < r) ? r : 0.0;                                              // std::max(0.0, .)
    const double cap = 1000.0 * nu;
    nu_t[o] = (cap < r0) ? cap : r0;                                                    // std::min(., 1000 nu)
}
// TurbulenceGEP (src/turbulence_gep.cpp:21-83, gep_cell_kernel): an algebraic mixing-length form whose damping is one of
// three gene-expression correction functions (variant 0 channel, 1 periodic hill, else plain van Driest squared);
// y+ is approximated with |S| y / nu
__global__ void k_gep_closure(Geom g, double nu, int variant, double inv_2dx, double inv_2dy, const double* __restrict__ dyc,
                              const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ wall,
                              double* __restrict__ nu_t) {
    const int ix = blockIdx.x * blockDim.x + threadIdx.x, iy = blockIdx.y * blockDim.y + threadIdx.y;
    if (ix >= g.Nx || iy >= g.Ny) return;
    const int i = ix + g.Ng, j = iy + g.Ng;
    const int o = j * g.cs + i, ou = j * g.us + i, ov = j * g.vs + i;
    const double wy = dyc ? 1.0 / (dyc[j] + dyc[j + 1]) : inv_2dy;
    const double dudx = (u[ou + 1] - u[ou - 1]) * inv_2dx, dudy = (u[ou + g.us] - u[ou - g.us]) * wy;
    const double dvdx = (v[ov + 1] - v[ov - 1]) * inv_2dx, dvdy = (v[ov + g.vs] - v[ov - g.vs]) * wy;
    const double Sxy = 0.5 * (dudy + dvdx), Oxy = 0.5 * (dudy - dvdx);
    const double Smag = sqrt(2.0 * (dudx * dudx + dvdy * dvdy + 2.0 * Sxy * Sxy));
    const double Omag = sqrt(2.0 * Oxy * Oxy);
    const double yw = hi_of(wall[j], 1e-10);
    const double y_plus = Smag * yw / (nu + 1e-20);
    double f_damp = 1.0 - exp(-y_plus / 26.0);
    f_damp = f_damp * f_damp;
    double f = f_damp;
    if (variant == 0) {
        const double ratio = (Smag > 1e-10) ? Omag / Smag : 1.0;
        f = f_damp * (1.0 / (1.0 + 0.1 * ratio * ratio));
    } else if (variant == 1) {
        const double ratio = (Smag > 1e-10) ? Omag / Smag : 1.0;
        const double f_sep = 1.0 / (1.0 + 0.2 * ratio * ratio);
        f = tanh(y_plus / 50.0) * f_sep;
    }
    const double l = 0.41 * yw * f;
    double r = l * l * Smag;
    r = hi_of(r, 0.0);
    nu_t[o] = lo_of(r, 1000.0 * nu);
}
// ---- explicit algebraic Reynolds-stress closures on the SST transport (SSTWithEARSM, src/turbulence_earsm.cpp:205-346,
// 349-529): b_ij = sum_n G_n(eta, zeta) T^n_ij on the TBNN tensor basis (TensorBasis::compute, src/features.cpp:192-292,
// built from the stretched-aware MAC gradients and the RAW k, omega like FeatureComputer::compute_tbnn_features :361-396),
// the nonlinear terms faded in with the turbulence Reynolds number, nu_t recovered from b_xy.
// VARIANT 0: Wallin-Johansson 2000, 1: Gatski-Speziale 1993, 2: Pope 1975 quadratic.
struct EarsmParams {
    double nu, k_min, omega_min, Re_t_center, Re_t_width;     // EARSMClosure thresholds (include/turbulence_earsm.hpp:84-92)
    double pope_C1, pope_C2;                                  // PopeQuadraticEARSM (:255)
};

__device__ __forceinline__ double clampd(double v, double lo, double hi) {       // std::max(lo, std::min(v, hi))
    const double m = (hi < v) ? hi : v;
    return (lo < m) ? m : lo;
}
__device__ __forceinline__ bool finite_d(double x) { return x - x == 0.0; }

template <int VARIANT>
__device__ __forceinline__ void earsm_G(const EarsmParams& p, double eta, double zeta, double G[4]) {
    if (VARIANT == 0) {                                        // WallinJohanssonEARSM::compute_G / solve_for_N (:349-449)
        const double C2 = 0.36, C3 = 1.25, C4 = 0.4, C5 = 1.88;                      // WJConstants (:51-66)
        const double A1 = 4.0 / 5.0 - C2 / 2.0, A2 = 2.0 - C4 / 2.0, A3 = 2.0 - C3 / 2.0, A4 = 2.0 * C5 - 1.0;
        const double II_S = eta * eta, II_O = -zeta * zeta;
        double den = 1.0 + A3 * II_S + A4 * fabs(II_O);
        const double dm = (fabs(den) < 0.1) ? 0.1 : fabs(den);
        den = dm * (den >= 0 ? 1.0 : -1.0);
        const double N = clampd(-A1 / den, -10.0, 10.0);
        double d2 = A1 + N;
        if (fabs(d2) < 0.01) d2 = 0.01 * (d2 >= 0 ? 1.0 : -1.0);
        G[0] = -N / d2;
        G[1] = (fabs(II_O) > 1e-10) ? A2 * N * N / (d2 * d2) : 0.0;
        G[2] = (II_S > 1e-10) ? A3 * N / d2 : 0.0;
        G[3] = 0.0;
        for (int n = 0; n < 4; ++n) G[n] = clampd(G[n], -10.0, 10.0);
    } else if (VARIANT == 1) {                                 // GatskiSpezialeEARSM::compute_G (:452-494)
        const double C_mu0 = 0.09, C1 = 1.8, C2 = 0.6, eta_max = 10.0;               // GSConstants (:69-79)
        const double reg = 1.0 + (eta * eta) / (eta_max * eta_max);
        const double Ce = C_mu0 / reg;
        double rot = 1.0;
        if (eta > 1e-10) { const double ratio = zeta / eta; rot = 1.0 / (1.0 + 0.1 * ratio * ratio); }
        G[0] = -Ce * rot; G[1] = C1 * Ce * Ce; G[2] = C2 * Ce; G[3] = 0.0;
        for (int n = 0; n < 4; ++n) G[n] = clampd(G[n], -5.0, 5.0);
    } else {                                                   // PopeQuadraticEARSM::compute_G (:497-528)
        const double es = (finite_d(eta) && eta < 100.0) ? eta : 100.0;
        const double Ce = 0.09 / (1.0 + 0.01 * es * es);
        G[0] = -Ce; G[1] = p.pope_C2 * es; G[2] = p.pope_C1 * es; G[3] = 0.0;
    }
}

template <int VARIANT>
__global__ void k_earsm_closure(Geom g, EarsmParams p, double inv_2dx, double inv_2dy, const double* __restrict__ dyc,
                                const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ k,
                                const double* __restrict__ omega, double* __restrict__ nu_t, double* __restrict__ txx,
                                double* __restrict__ txy, double* __restrict__ tyy) {
    const int ix = blockIdx.x * blockDim.x + threadIdx.x, iy = blockIdx.y * blockDim.y + threadIdx.y;
    if (ix >= g.Nx || iy >= g.Ny) return;
    const int i = ix + g.Ng, j = iy + g.Ng;
    const int o = j * g.cs + i, ou = j * g.us + i, ov = j * g.vs + i;
    const double dux = u[ou + 1] - u[ou - 1], duy = u[ou + g.us] - u[ou - g.us];
    const double dvx = v[ov + 1] - v[ov - 1], dvy = v[ov + g.vs] - v[ov - g.vs];
    // tensor basis: stretched-aware gradients, raw k / omega (TensorBasis::compute)
    double T[3][3];
    {
        const double wy = dyc ? 1.0 / (dyc[j] + dyc[j + 1]) : inv_2dy;
        const double dudx = dux * inv_2dx, dudy = duy * wy, dvdx = dvx * inv_2dx, dvdy = dvy * wy;
        const double kv = k[o], eps = 0.09 * kv * omega[o];
        const double tau = ((kv < 1e-10) ? 1e-10 : kv) / ((eps < 1e-20) ? 1e-20 : eps);
        const double Sxx = dudx * tau, Syy = dvdy * tau, Sxy = (0.5 * (dudy + dvdx)) * tau, Oxy = (0.5 * (dudy - dvdx)) * tau;
        T[0][0] = Sxx; T[0][1] = Sxy; T[0][2] = Syy;
        T[1][0] = -2.0 * Sxy * Oxy; T[1][1] = (Sxx - Syy) * Oxy; T[1][2] = 2.0 * Sxy * Oxy;
        const double S2xx = Sxx * Sxx + Sxy * Sxy, S2yy = Sxy * Sxy + Syy * Syy, S2xy = Sxy * (Sxx + Syy);
        const double trS2 = S2xx + S2yy;
        T[2][0] = S2xx - 0.5 * trS2; T[2][1] = S2xy; T[2][2] = S2yy - 0.5 * trS2;
    }
    const double kl = (p.k_min < k[o]) ? k[o] : p.k_min;                 // std::max(k_min, k)
    const double ol = (p.omega_min < omega[o]) ? omega[o] : p.omega_min;
    double Re_t;
    {                                                                    // safe_divide(k, nu omega, OMEGA_FLOOR nu)
        const double den = p.nu * ol, fl = 1e-10 * p.nu;
        const double mag = (fabs(den) < fl) ? fl : fabs(den);
        Re_t = kl / ((den < 0.0) ? -mag : mag);
    }
    const double alpha = 0.5 * (1.0 + tanh((Re_t - p.Re_t_center) / p.Re_t_width));
    double tau;
    {                                                                    // safe_divide(k, eps, K_FLOOR)
        const double eps = 0.09 * kl * ol;
        const double mag = (fabs(eps) < 1e-10) ? 1e-10 : fabs(eps);
        tau = kl / ((eps < 0.0) ? -mag : mag);
    }
    // uniform-spacing gradients for the invariants (:252-258)
    const double dudx = dux * inv_2dx, dudy = duy * inv_2dy, dvdx = dvx * inv_2dx, dvdy = dvy * inv_2dy;
    const double Sxy = 0.5 * (dudy + dvdx), Oxy = 0.5 * (dudy - dvdx);
    const double S_mag = sqrt(dudx * dudx + dvdy * dvdy + 2.0 * Sxy * Sxy);
    const double O_mag = sqrt(2.0 * Oxy * Oxy);
    double eta = tau * S_mag, zeta = tau * O_mag;
    if (!finite_d(eta)) eta = 100.0;
    if (!finite_d(zeta)) zeta = 100.0;
    double G[4];
    earsm_G<VARIANT>(p, eta, zeta, G);
    for (int n = 0; n < 4; ++n) if (!finite_d(G[n])) G[n] = 0.0;
    G[1] *= alpha; G[2] *= alpha; G[3] *= alpha;
    double b_xx = 0.0, b_xy = 0.0, b_yy = 0.0;
    for (int n = 0; n < 3; ++n) { b_xx += G[n] * T[n][0]; b_xy += G[n] * T[n][1]; b_yy += G[n] * T[n][2]; }
    b_xx += G[3] * 0.0; b_xy += G[3] * 0.0; b_yy += G[3] * 0.0;          // T4 = 0 in 2-D (keeps a NaN G[3] visible)
    if (!finite_d(b_xx)) b_xx = 0.0;
    if (!finite_d(b_xy)) b_xy = 0.0;
    if (!finite_d(b_yy)) b_yy = 0.0;
    if (txx) {                                                           // anisotropy_to_reynolds_stress :310-326
        const double ks = (kl < 0.0) ? 0.0 : kl;
        txx[o] = 2.0 * ks * (b_xx + 1.0 / 3.0);
        txy[o] = 2.0 * ks * b_xy;
        tyy[o] = 2.0 * ks * (b_yy + 1.0 / 3.0);
    }
    double r = 0.0;
    if (fabs(Sxy) > 1e-10) {
        const double mag = (fabs(Sxy) < 1e-10) ? 1e-10 : fabs(Sxy);
        r = fabs((-b_xy * kl) / ((Sxy < 0.0) ? -mag : mag));
    } else if (S_mag > 1e-10) {
        const double b_mag = sqrt(b_xx * b_xx + 2.0 * b_xy * b_xy + b_yy * b_yy);
        r = (kl * b_mag) / ((S_mag < 1e-10) ? 1e-10 : S_mag);
    }
    r = (0.0 < r) ? r : 0.0;
    const double cap = 100.0 * p.nu;
    r = (cap < r) ? cap : r;
    if (!finite_d(r)) r = 0.0;
    nu_t[o] = r;
}
}  // namespace

SstModel::SstModel(const cfdnn_mesh_desc& m, double nu, int kind) : kind_(kind), dx_(m.dx), dy_(m.dy) {
    if (kind < CFDNN_RANS_SST || kind > CFDNN_RANS_EARSM_POPE)
        throw std::invalid_argument("2-D RANS model: kind must be one of CFDNN_RANS_SST ... CFDNN_RANS_EARSM_POPE");
    require_device();
    if (m.Nz > 1)
        throw std::invalid_argument("SST k-omega: the reference model is two-dimensional (fields indexed (i, j)); 3-D meshes "
                                    "are not supported");
    if (m.Ng != 1) throw std::invalid_argument("SST k-omega: only Nghost == 1 is supported");
    prm.nu = nu;
    const int n = m.Ny + 2 * m.Ng;
    std::vector<double> w(n);
    for (int j = 0; j < n; ++j) {
        const double y = m.yc ? m.yc[j] : m.y_lo + (j - m.Ng + 0.5) * m.dy;                          // src/mesh.cpp:66-68
        const double lo = std::abs(y - m.y_lo), hi = std::abs(y - m.y_hi);
        w[j] = std::min(lo, hi);
    }
    wall_.alloc(n, false);
    CFDNN_CUDA(cudaMemcpy(wall_.p, w.data(), n * sizeof(double), cudaMemcpyHostToDevice));
    dy_wall_ = m.dy;                                 // centre spacing at the lower wall row, src/turbulence_baseline.cpp:337-341
    if (m.y_stretched && m.dyv) dy_wall_ = 0.5 * (m.dyv[m.Ng] + m.dyv[m.Ng + 1]);
    utau_.alloc(1, true);
    if (m.y_stretched && m.dyc) {
        dyc_.alloc(n + 1, false);
        CFDNN_CUDA(cudaMemcpy(dyc_.p, m.dyc, (n + 1) * sizeof(double), cudaMemcpyHostToDevice));
    }
}

void SstModel::advance(const Geom& g, const Vel& vel, double* k, double* omega, const double* nu_t, double dt, int plane,
                       cudaStream_t st) {
    SstGrid sg;
    sg.dx = dx_; sg.dy = dy_; sg.dx2 = dx_ * dx_; sg.dy2 = dy_ * dy_;
    sg.inv_2dx = 1.0 / (2.0 * dx_); sg.inv_2dy = 1.0 / (2.0 * dy_);
    const size_t oc = (size_t)plane * g.cp, ou = (size_t)plane * g.up, ov = (size_t)plane * g.vp;
    const bool earsm = kind_ >= CFDNN_RANS_EARSM_WJ && kind_ <= CFDNN_RANS_EARSM_POPE;         // SSTWithEARSM: SST transport
    if (kind_ != CFDNN_RANS_SST && kind_ != CFDNN_RANS_KOMEGA && !earsm) return;                // algebraic: no transport
    KwParams kw;                                    // KOmegaConstants, include/turbulence_transport.hpp:61-72
    kw.nu = prm.nu; kw.sigma_k = 0.5; kw.sigma_omega = 0.5; kw.beta = 0.075; kw.beta_star = 0.09; kw.alpha = 5.0 / 9.0;
    kw.k_min = 1e-10; kw.k_max = 100.0; kw.omega_min = 1e-10; kw.omega_max = 1e8;
    if (kind_ == CFDNN_RANS_SST || earsm) {
        CFDNN_LAUNCH_COOP((k_sst_sweep<true>), 1, 1024, 0, st, g, prm, kw, sg, dt, vel.u + ou, vel.v + ov, k + oc, omega + oc,
                          nu_t + oc, wall_.p, dyc_.p);
        CFDNN_LAUNCH(k_sst_wall_bc, (g.Nx + 2 * g.Ng + 127) / 128, 128, 0, st, g, prm, k + oc, omega + oc, wall_.p);
    } else {                                        // the standard model applies no wall BCs after its sweep
        CFDNN_LAUNCH_COOP((k_sst_sweep<false>), 1, 1024, 0, st, g, prm, kw, sg, dt, vel.u + ou, vel.v + ov, k + oc, omega + oc,
                          nu_t + oc, wall_.p, dyc_.p);
    }
}

void SstModel::closure(const Geom& g, const Vel& vel, const double* k, const double* omega, double* nu_t, double* txx,
                       double* txy, double* tyy, int plane, cudaStream_t st) {
    const size_t oc = (size_t)plane * g.cp, ou = (size_t)plane * g.up, ov = (size_t)plane * g.vp;
    const dim3 blk(64, 4);
    if (kind_ >= CFDNN_RANS_EARSM_WJ && kind_ <= CFDNN_RANS_EARSM_POPE) {
        EarsmParams ep;
        ep.nu = prm.nu; ep.k_min = 1e-10; ep.omega_min = 1e-10; ep.Re_t_center = 10.0; ep.Re_t_width = 5.0;
        ep.pope_C1 = pope_C1; ep.pope_C2 = pope_C2;
        const dim3 grd((g.Nx + 63) / 64, (g.Ny + 3) / 4);
#define CFDNN_EARSM(V) CFDNN_LAUNCH((k_earsm_closure<V>), grd, blk, 0, st, g, ep, 1.0 / (2.0 * dx_), 1.0 / (2.0 * dy_), dyc_.p, \
                                    vel.u + ou, vel.v + ov, k + oc, omega + oc, nu_t + oc, txx ? txx + oc : nullptr,             \
                                    txy ? txy + oc : nullptr, tyy ? tyy + oc : nullptr)
        if (kind_ == CFDNN_RANS_EARSM_WJ) CFDNN_EARSM(0);
        else if (kind_ == CFDNN_RANS_EARSM_GS) CFDNN_EARSM(1);
        else CFDNN_EARSM(2);
#undef CFDNN_EARSM
        return;
    }
    if (kind_ == CFDNN_RANS_MIXING_LENGTH) {
        CFDNN_LAUNCH((k_rans_utau<true>), 1, 32, 0, st, g, prm.nu, 1.0 / (2.0 * dy_), dy_wall_, dyc_.p, vel.u + ou, utau_.p);
        CFDNN_LAUNCH(k_mixing_length, dim3((g.Nx + 63) / 64, (g.Ny + 3) / 4), blk, 0, st, g, prm.nu, ml_kappa, ml_A_plus, ml_delta,
                     1.0 / (2.0 * dx_), 1.0 / (2.0 * dy_), dyc_.p, vel.u + ou, vel.v + ov, wall_.p, utau_.p, nu_t + oc);
        return;
    }
    if (kind_ == CFDNN_RANS_GEP) {
        CFDNN_LAUNCH(k_gep_closure, dim3((g.Nx + 63) / 64, (g.Ny + 3) / 4), blk, 0, st, g, gep_nu, gep_variant, 1.0 / (2.0 * dx_),
                     1.0 / (2.0 * dy_), dyc_.p, vel.u + ou, vel.v + ov, wall_.p, nu_t + oc);
        return;
    }
    if (kind_ == CFDNN_RANS_ALGEBRAIC_KOMEGA) {
        CFDNN_LAUNCH((k_rans_utau<false>), 1, 32, 0, st, g, prm.nu, 1.0 / (2.0 * dy_), dy_wall_, dyc_.p, vel.u + ou, utau_.p);
        CFDNN_LAUNCH(k_algebraic_komega, dim3((g.Nx + 63) / 64, (g.Ny + 3) / 4), blk, 0, st, g, prm.nu, 0.09, wall_.p, utau_.p,
                     nu_t + oc);
        return;
    }
    if (kind_ == CFDNN_RANS_KOMEGA) {
        CFDNN_LAUNCH(k_kw_closure, dim3((g.Nx + 63) / 64, (g.Ny + 3) / 4), blk, 0, st, g, prm.nu, k + oc, omega + oc, nu_t + oc);
        return;
    }
    CFDNN_LAUNCH(k_sst_closure, dim3((g.Nx + 63) / 64, (g.Ny + 3) / 4), blk, 0, st, g, prm, 1.0 / (2.0 * dx_), 1.0 / (2.0 * dy_),
                 dyc_.p, vel.u + ou, vel.v + ov, k + oc, omega + oc, wall_.p, nu_t + oc);
}

}  // namespace cfdnn
