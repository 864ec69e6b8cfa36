// les.cu — static LES subgrid-scale closures: Smagorinsky, WALE, Vreman, Sigma.
//
// Stands in for LESModel::update and its four static subclasses (reference include/turbulence_les.hpp:24-115,
// src/turbulence_les.cpp:158-335) with the velocity-gradient tensor of GradientComputer::compute
// (src/velocity_gradient.cpp:20-119) — the CPU path, which is the oracle: dv/dy over the face spacing
// yf[j+1] - yf[j], du/dy and dw/dy over yc[j+1] - yc[j-1], filter width cbrt(dx dy_local dz) (sqrt(dx dy_local) in
// 2-D).  One fused kernel: the nine gradients are formed in registers from the staggered velocities and go straight
// into nu_sgs; nothing but nu_t is written (the reference stores nine gradient arrays first).  HBM-bound: u, v, w in
// (24 B per cell), nu_t out (8 B).  The dynamic Smagorinsky model (test filter + plane averages) is not built.
#include "les.cuh"
#include <cmath>
#include <vector>

namespace cfdnn {
namespace {

// (Cs delta)^2 |S|, |S| = sqrt(2 Sij Sij)                          (src/turbulence_les.cpp:158-167)
__device__ __forceinline__ double nu_smagorinsky(const double g[9], double Cs, double delta) {
    const double S11 = g[0], S22 = g[4], S33 = g[8];
    const double S12 = 0.5 * (g[1] + g[3]);
    const double S13 = 0.5 * (g[2] + g[6]);
    const double S23 = 0.5 * (g[5] + g[7]);
    const double S_mag = sqrt(2.0 * (S11 * S11 + S22 * S22 + S33 * S33 + 2.0 * (S12 * S12 + S13 * S13 + S23 * S23)));
    return (Cs * delta) * (Cs * delta) * S_mag;
}

// Nicoud & Ducros 1999                                             (:170-221)
__device__ __forceinline__ double nu_wale(const double g[9], double Cw, double delta) {
    const double S11 = g[0], S22 = g[4], S33 = g[8];
    const double S12 = 0.5 * (g[1] + g[3]);
    const double S13 = 0.5 * (g[2] + g[6]);
    const double S23 = 0.5 * (g[5] + g[7]);
    const double SijSij = S11 * S11 + S22 * S22 + S33 * S33 + 2.0 * (S12 * S12 + S13 * S13 + S23 * S23);
    double g2[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            double a = 0.0;
#pragma unroll
            for (int kk = 0; kk < 3; ++kk) a += g[i * 3 + kk] * g[kk * 3 + j];
            g2[i * 3 + j] = a;
        }
    const double trace_g2 = g2[0] + g2[4] + g2[8];
    double SdSd = 0.0;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            double sd = 0.5 * (g2[i * 3 + j] + g2[j * 3 + i]);
            if (i == j) sd -= trace_g2 / 3.0;
            SdSd += sd * sd;
        }
    const double numerator = SdSd * sqrt(SdSd);
    const double denom_a = SijSij * SijSij * sqrt(SijSij);
    const double denom_b = SdSd * sqrt(sqrt(SdSd));
    const double denominator = denom_a + denom_b;
    if (denominator < 1e-30) return 0.0;
    return (Cw * delta) * (Cw * delta) * numerator / denominator;
}

// Vreman 2004                                                      (:224-248)
__device__ __forceinline__ double nu_vreman(const double g[9], double Cv, double delta) {
    const double delta2 = delta * delta;
    double alpha2 = 0.0;
#pragma unroll
    for (int i = 0; i < 9; ++i) alpha2 += g[i] * g[i];
    if (alpha2 < 1e-30) return 0.0;
    double beta[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            double a = 0.0;
#pragma unroll
            for (int m = 0; m < 3; ++m) a += g[m * 3 + i] * g[m * 3 + j];
            beta[i * 3 + j] = a * delta2;
        }
    double B_beta = beta[0] * beta[4] - beta[1] * beta[1] + beta[0] * beta[8] - beta[2] * beta[2] +
                    beta[4] * beta[8] - beta[5] * beta[5];
    if (B_beta < 0.0) B_beta = 0.0;
    return Cv * sqrt(B_beta / alpha2);
}

// Nicoud et al. 2011: singular values from the invariants of G = g^T g   (:251-306)
__device__ __forceinline__ double nu_sigma(const double g[9], double Cs, double delta) {
    double G[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            double a = 0.0;
#pragma unroll
            for (int kk = 0; kk < 3; ++kk) a += g[kk * 3 + i] * g[kk * 3 + j];
            G[i * 3 + j] = a;
        }
    const double I1 = G[0] + G[4] + G[8];
    const double I2 = G[0] * G[4] + G[0] * G[8] + G[4] * G[8] - G[1] * G[1] - G[2] * G[2] - G[5] * G[5];
    const double I3 = G[0] * (G[4] * G[8] - G[5] * G[5]) - G[1] * (G[1] * G[8] - G[5] * G[2]) +
                      G[2] * (G[1] * G[5] - G[4] * G[2]);
    if (I1 < 1e-30) return 0.0;
    double p = (I1 * I1 - 3.0 * I2) / 9.0;
    const double q = (2.0 * I1 * I1 * I1 - 9.0 * I1 * I2 + 27.0 * I3) / 54.0;
    if (p < 0.0) p = 0.0;
    const double sp = sqrt(p);
    double arg = (sp > 1e-30) ? q / (p * sp) : 0.0;
    if (arg < -1.0) arg = -1.0;
    if (arg > 1.0) arg = 1.0;
    const double theta = acos(arg) / 3.0;
    double lam1 = I1 / 3.0 + 2.0 * sp * cos(theta);
    double lam2 = I1 / 3.0 + 2.0 * sp * cos(theta - 2.0 * M_PI / 3.0);
    double lam3 = I1 / 3.0 + 2.0 * sp * cos(theta + 2.0 * M_PI / 3.0);
    if (lam1 < lam2) { const double t = lam1; lam1 = lam2; lam2 = t; }
    if (lam1 < lam3) { const double t = lam1; lam1 = lam3; lam3 = t; }
    if (lam2 < lam3) { const double t = lam2; lam2 = lam3; lam3 = t; }
    const double sig1 = sqrt(lam1 > 0.0 ? lam1 : 0.0);
    const double sig2 = sqrt(lam2 > 0.0 ? lam2 : 0.0);
    const double sig3 = sqrt(lam3 > 0.0 ? lam3 : 0.0);
    if (sig1 < 1e-30) return 0.0;
    const double D_sigma = sig3 * (sig1 - sig2) * (sig2 - sig3) / (sig1 * sig1);
    if (D_sigma < 0.0) return 0.0;
    return (Cs * delta) * (Cs * delta) * D_sigma;
}

// One thread per interior cell.  yc: Ny+2Ng cell centres, yf: Ny+2Ng+1 faces, delta: filter width per row
// (computed on the host with the reference's own expression).  plane0: z-plane of the first cell layer.
template <int MODEL>
__global__ void __launch_bounds__(256)
k_les_sgs(Geom g, double dx, double dz, const double* __restrict__ yc, const double* __restrict__ yf,
          const double* __restrict__ delta, const double* __restrict__ u, const double* __restrict__ v,
          const double* __restrict__ w, double C, int is2d, int plane0, double* __restrict__ nu_t) {
    const int ix = blockIdx.x * blockDim.x + threadIdx.x;
    const int iy = blockIdx.y * blockDim.y + threadIdx.y;
    if (ix >= g.Nx || iy >= g.Ny) return;
    const int ig = ix + g.Ng, jg = iy + g.Ng, kg = plane0 + (int)blockIdx.z;
#define LU(I, J, K) u[(size_t)(K) * g.up + (size_t)(J) * g.us + (I)]
#define LV(I, J, K) v[(size_t)(K) * g.vp + (size_t)(J) * g.vs + (I)]
#define LW(I, J, K) w[(size_t)(K) * g.wp + (size_t)(J) * g.ws + (I)]
    const double dy_central = yc[jg + 1] - yc[jg - 1];
    const double dy_face = yf[jg + 1] - yf[jg];
    double G[9];
    G[0] = (LU(ig + 1, jg, kg) - LU(ig, jg, kg)) / dx;
    {
        const double u_jp = 0.5 * (LU(ig, jg + 1, kg) + LU(ig + 1, jg + 1, kg));
        const double u_jm = 0.5 * (LU(ig, jg - 1, kg) + LU(ig + 1, jg - 1, kg));
        G[1] = (u_jp - u_jm) / dy_central;
    }
    {
        const double v_ip = 0.5 * (LV(ig + 1, jg, kg) + LV(ig + 1, jg + 1, kg));
        const double v_im = 0.5 * (LV(ig - 1, jg, kg) + LV(ig - 1, jg + 1, kg));
        G[3] = (v_ip - v_im) / (2.0 * dx);
    }
    G[4] = (LV(ig, jg + 1, kg) - LV(ig, jg, kg)) / dy_face;
    if (is2d) {
        G[2] = 0.0; G[5] = 0.0; G[6] = 0.0; G[7] = 0.0; G[8] = 0.0;
    } else {
        {
            const double u_kp = 0.5 * (LU(ig, jg, kg + 1) + LU(ig + 1, jg, kg + 1));
            const double u_km = 0.5 * (LU(ig, jg, kg - 1) + LU(ig + 1, jg, kg - 1));
            G[2] = (u_kp - u_km) / (2.0 * dz);
        }
        {
            const double v_kp = 0.5 * (LV(ig, jg, kg + 1) + LV(ig, jg + 1, kg + 1));
            const double v_km = 0.5 * (LV(ig, jg, kg - 1) + LV(ig, jg + 1, kg - 1));
            G[5] = (v_kp - v_km) / (2.0 * dz);
        }
        {
            const double w_ip = 0.5 * (LW(ig + 1, jg, kg) + LW(ig + 1, jg, kg + 1));
            const double w_im = 0.5 * (LW(ig - 1, jg, kg) + LW(ig - 1, jg, kg + 1));
            G[6] = (w_ip - w_im) / (2.0 * dx);
        }
        {
            const double w_jp = 0.5 * (LW(ig, jg + 1, kg) + LW(ig, jg + 1, kg + 1));
            const double w_jm = 0.5 * (LW(ig, jg - 1, kg) + LW(ig, jg - 1, kg + 1));
            G[7] = (w_jp - w_jm) / dy_central;
        }
        G[8] = (LW(ig, jg, kg + 1) - LW(ig, jg, kg)) / dz;
    }
#undef LU
#undef LV
#undef LW
    const double d = delta[jg];
    double r;
    if (MODEL == CFDNN_LES_SMAGORINSKY) r = nu_smagorinsky(G, C, d);
    else if (MODEL == CFDNN_LES_WALE) r = nu_wale(G, C, d);
    else if (MODEL == CFDNN_LES_VREMAN) r = nu_vreman(G, C, d);
    else r = nu_sigma(G, C, d);
    nu_t[(size_t)kg * g.cp + (size_t)jg * g.cs + ig] = r;
}
}  // namespace

double les_default_constant(int model) {      // include/turbulence_les.hpp:56,73,89,105
    switch (model) {
        case CFDNN_LES_SMAGORINSKY: return 0.17;
        case CFDNN_LES_WALE: return 0.325;
        case CFDNN_LES_VREMAN: return 0.07;
        case CFDNN_LES_SIGMA: return 1.35;
    }
    throw std::invalid_argument("unknown LES model (1 = Smagorinsky, 2 = WALE, 3 = Vreman, 4 = Sigma)");
}

LesMetrics::LesMetrics(const cfdnn_mesh_desc& m) {
    const int Ng = m.Ng, n = m.Ny + 2 * Ng;
    std::vector<double> hyc(n), hyf(n + 1), hd(n, 0.0);
    for (int j = 0; j < n; ++j) hyc[j] = m.yc ? m.yc[j] : m.y_lo + (j - Ng + 0.5) * m.dy;          // src/mesh.cpp:66-68
    if (m.y_stretched) {
        if (!m.yf) throw std::invalid_argument("LES closure on a y-stretched mesh needs mesh.yf");
        for (int j = 0; j <= n; ++j) hyf[j] = m.yf[j];
    } else {
        for (int j = 0; j <= n; ++j) hyf[j] = m.y_lo + (j - Ng) * m.dy;                            // src/mesh.cpp:84-86
    }
    const bool is2d = m.Nz <= 1;
    for (int j = 0; j < n; ++j) {                                                                   // filter_width :308-314
        const double dy_local = hyf[j + 1] - hyf[j];
        hd[j] = is2d ? std::sqrt(m.dx * dy_local) : std::cbrt(m.dx * dy_local * m.dz);
    }
    yc.alloc(n, false); yf.alloc(n + 1, false); delta.alloc(n, false);
    CFDNN_CUDA(cudaMemcpy(yc.p, hyc.data(), n * sizeof(double), cudaMemcpyHostToDevice));
    CFDNN_CUDA(cudaMemcpy(yf.p, hyf.data(), (n + 1) * sizeof(double), cudaMemcpyHostToDevice));
    CFDNN_CUDA(cudaMemcpy(delta.p, hd.data(), n * sizeof(double), cudaMemcpyHostToDevice));
}

void les_update_device(const Geom& g, const cfdnn_mesh_desc& m, const LesMetrics& met, int model, double C, const Vel& vel,
                       double* nu_t, int plane0, int nplanes, bool is2d, cudaStream_t st) {
    if (C <= 0.0) C = les_default_constant(model);
    const dim3 blk(64, 4, 1);
    const dim3 grd((g.Nx + 63) / 64, (g.Ny + 3) / 4, nplanes);
#define CFDNN_LES(M) CFDNN_LAUNCH((k_les_sgs<M>), grd, blk, 0, st, g, m.dx, m.dz, met.yc.p, met.yf.p, met.delta.p, vel.u, vel.v, \
                                  vel.w, C, is2d ? 1 : 0, plane0, nu_t)
    switch (model) {
        case CFDNN_LES_SMAGORINSKY: CFDNN_LES(CFDNN_LES_SMAGORINSKY); break;
        case CFDNN_LES_WALE: CFDNN_LES(CFDNN_LES_WALE); break;
        case CFDNN_LES_VREMAN: CFDNN_LES(CFDNN_LES_VREMAN); break;
        case CFDNN_LES_SIGMA: CFDNN_LES(CFDNN_LES_SIGMA); break;
        default: throw std::invalid_argument("unknown LES model (1 = Smagorinsky, 2 = WALE, 3 = Vreman, 4 = Sigma)");
    }
#undef CFDNN_LES
}

// LESModel::update on host arrays (2-D mesh: plane 0 holds the data, include/fields.hpp:18-19)
void les_update_host(const cfdnn_mesh_desc& m, int model, double C, const double* u, const double* v, const double* w,
                     double* nu_t) {
    require_device();
    if (m.Ng != 1) throw std::invalid_argument("les_update: only Nghost == 1 is supported");
    if (!u || !v || !nu_t) throw std::invalid_argument("les_update: null field pointer");
    const bool is2d = m.Nz <= 1;
    if (!is2d && !w) throw std::invalid_argument("les_update: 3-D mesh needs w");
    (void)les_default_constant(model);
    cfdnn_mesh_desc mm = m;
    if (is2d && !(mm.dz > 0.0)) mm.dz = 1.0;
    Geom g = make_geom(mm);
    const int Ng = m.Ng, nzt = m.Nz + 2 * Ng;
    const size_t su = (size_t)g.up * nzt, sv = (size_t)g.vp * nzt, sw = (size_t)g.wp * (nzt + 1), sc = (size_t)g.cp * nzt;
    LesMetrics met(mm);
    DevBuf<double> du, dv, dw, dn;
    du.alloc(su, false); dv.alloc(sv, false); dw.alloc(is2d ? 1 : sw, false); dn.alloc(sc, false);
    cudaStream_t st = nullptr;
    CFDNN_CUDA(cudaMemcpyAsync(du.p, u, su * sizeof(double), cudaMemcpyHostToDevice, st));
    CFDNN_CUDA(cudaMemcpyAsync(dv.p, v, sv * sizeof(double), cudaMemcpyHostToDevice, st));
    if (!is2d) CFDNN_CUDA(cudaMemcpyAsync(dw.p, w, sw * sizeof(double), cudaMemcpyHostToDevice, st));
    CFDNN_CUDA(cudaMemcpyAsync(dn.p, nu_t, sc * sizeof(double), cudaMemcpyHostToDevice, st));
    Vel vel; vel.u = du.p; vel.v = dv.p; vel.w = dw.p;
    les_update_device(g, mm, met, model, C, vel, dn.p, is2d ? 0 : Ng, is2d ? 1 : m.Nz, is2d, st);
    CFDNN_CUDA(cudaMemcpyAsync(nu_t, dn.p, sc * sizeof(double), cudaMemcpyDeviceToHost, st));
    CFDNN_CUDA(cudaStreamSynchronize(st));
    check_launch("les_update");
}

}  // namespace cfdnn
