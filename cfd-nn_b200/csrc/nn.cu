// nn.cu — NN closures: feature extraction -> MLP -> tensor-basis contraction -> nu_t, fused per cell.
#include "nn.cuh"
#include <cmath>
#include <algorithm>
#include <cstdlib>

namespace cfdnn {

namespace {
constexpr double C_MU = 0.09;      // reference include/numerics.hpp:86
constexpr double KAPPA = 0.41;     // :90

__device__ __forceinline__ double activate(double x, int act) {
    if (act == 1) return tanh(x);
    if (act == 2) return x > 0.0 ? x : 0.0;
    return x;
}

// MLP::forward (src/nn_core.cpp:144-204): z-score, then per layer y = W x + b (sum starts at
// the bias, j ascending, :36-47) and the activation.  One thread evaluates one sample.
__device__ void mlp_forward(const NNDev& net, const double* x_in, double* out) {
    double a[NN_MAX_WIDTH], c[NN_MAX_WIDTH];
    const int in0 = net.dims[0];
    for (int i = 0; i < in0; ++i) {
        double x = x_in[i];
        if (net.means && i < net.n_scale) x = (x - net.means[i]) / net.stds[i];
        a[i] = x;
    }
    double* cur = a;
    double* nxt = c;
    for (int l = 0; l < net.n_layers; ++l) {
        const int in = net.dims[l], od = net.dims[l + 1];
        const double* W = net.W + net.woff[l];
        const double* b = net.b + net.boff[l];
        for (int i = 0; i < od; ++i) {
            double sum = b[i];
            const double* Wr = W + (size_t)i * in;
            for (int j = 0; j < in; ++j) sum += Wr[j] * cur[j];
            nxt[i] = activate(sum, net.act[l]);
        }
        double* t = cur; cur = nxt; nxt = t;
    }
    const int od = net.dims[net.n_layers];
    for (int i = 0; i < od; ++i) out[i] = cur[i];
}

__global__ void __launch_bounds__(128)
k_mlp_forward(NNDev net, const double* __restrict__ x, double* __restrict__ y, size_t n) {
    const size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    double out[NN_MAX_WIDTH];
    mlp_forward(net, x + s * net.dims[0], out);
    const int od = net.dims[net.n_layers];
    for (int i = 0; i < od; ++i) y[s * od + i] = out[i];
}

// k "provided" test per plane: any interior k_in > k_min (src/turbulence_nn_tbnn.cpp:659-667)
__global__ void __launch_bounds__(256)
k_k_provided(Geom g, const double* __restrict__ k_in, double k_min, int plane0, int* __restrict__ flag) {
    const int ix = blockIdx.x * blockDim.x + threadIdx.x;
    const int iy = blockIdx.y * blockDim.y + threadIdx.y;
    const int pl = blockIdx.z;
    if (ix >= g.Nx || iy >= g.Ny) return;
    const size_t o = (size_t)(plane0 + pl) * g.cp + (size_t)(iy + g.Ng) * g.cs + (ix + g.Ng);
    if (k_in[o] > k_min) flag[pl] = 1;
}

__device__ __forceinline__ double wall_distance(double y, double y_lo, double y_hi) {   // src/mesh.cpp:6-12
    return fmin(fabs(y - y_lo), fabs(y - y_hi));
}

// Per-cell context carried from the feature stage to the nu_t stage.
struct CellCtx {
    double T[4][3];
    double k_val, Sxy_, S_mag;
    double Oxy;              // tau-scaled rotation rate the basis was built from (with T[0] it rebuilds T[1..3])
};

// TensorBasis::compute (src/features.cpp:192-292) from the normalised strain / rotation components:
// T1 = S, T2 = [S, Omega], T3 = dev(S^2) (half trace in 2-D), T4 = 0
__device__ __forceinline__ void tensor_basis(double Sxx, double Sxy, double Syy, double Oxy, CellCtx& c) {
    c.T[0][0] = Sxx; c.T[0][1] = Sxy; c.T[0][2] = Syy;
    c.T[1][0] = -2.0 * Sxy * Oxy; c.T[1][1] = (Sxx - Syy) * Oxy; c.T[1][2] = 2.0 * Sxy * Oxy;
    const double S2xx = Sxx * Sxx + Sxy * Sxy, S2yy = Sxy * Sxy + Syy * Syy, S2xy = Sxy * (Sxx + Syy);
    const double trS2 = S2xx + S2yy;
    c.T[2][0] = S2xx - 0.5 * trS2; c.T[2][1] = S2xy; c.T[2][2] = S2yy - 0.5 * trS2;
    c.T[3][0] = 0.0; c.T[3][1] = 0.0; c.T[3][2] = 0.0;
    c.Oxy = Oxy;
}

// Features of one cell (i, j) of plane kk: TBNN 5 invariants (src/features.cpp:121-190) or the 6
// scalar-nu_t features (:67-119); for TBNN also the tensor basis (:192-292) and k / omega.
template <int KIND>
__device__ __forceinline__ void cell_features(const Geom& g, const NNParams& prm, const double* __restrict__ yc,
                                              double y_lo, double y_hi, const double* __restrict__ u,
                                              const double* __restrict__ v, const double* __restrict__ k_in,
                                              const double* __restrict__ omega_in, int kprovided, int i, int j,
                                              size_t oc, double* feat, CellCtx& c) {
#define U2(I, J) u[(size_t)(J) * g.us + (I)]
#define V2(I, J) v[(size_t)(J) * g.vs + (I)]
    // uniform-spacing MAC gradients (features.cpp:132-140, turbulence_nn_tbnn.cpp:838-845)
    const double inv_2dx = g.dx_2.r;     // RN(1 / (2 dx)), the same double the reference's 1.0 / (2.0 * dx) gives
    const double inv_2dy = g.dy_2.r;
    const double dudx = (U2(i + 1, j) - U2(i - 1, j)) * inv_2dx;
    const double dudy = (U2(i, j + 1) - U2(i, j - 1)) * inv_2dy;
    const double dvdx = (V2(i + 1, j) - V2(i - 1, j)) * inv_2dx;
    const double dvdy = (V2(i, j + 1) - V2(i, j - 1)) * inv_2dy;
    const double Sxx_ = dudx, Syy_ = dvdy, Sxy_ = 0.5 * (dudy + dvdx), Oxy_ = 0.5 * (dudy - dvdx);
    const double S_mag = sqrt(Sxx_ * Sxx_ + Syy_ * Syy_ + 2.0 * Sxy_ * Sxy_);
    const double Omega_mag = sqrt(2.0 * Oxy_ * Oxy_);
    const double y_wall = wall_distance(yc[j], y_lo, y_hi);
    c.Sxy_ = Sxy_; c.S_mag = S_mag;

    if (KIND == CFDNN_TURB_NN_MLP) {
        // compute_features_scalar_nut (src/features.cpp:67-119)
        const double uc = 0.5 * (U2(i, j) + U2(i + 1, j));
        const double vc = 0.5 * (V2(i, j) + V2(i, j + 1));
        const double u_mag = sqrt(uc * uc + vc * vc);
        const double u_ref = fmax(u_mag, 1e-10);
        feat[0] = S_mag * prm.delta / u_ref;
        feat[1] = Omega_mag * prm.delta / u_ref;
        feat[2] = y_wall / prm.delta;
        feat[3] = (S_mag > 1e-10) ? Omega_mag / S_mag : 0.0;
        feat[4] = S_mag * prm.delta * prm.delta / prm.nu;
        feat[5] = u_mag / u_ref;
        c.k_val = 0.0;
        return;
    }

    // ---- TBNN: k / omega (turbulence_nn_tbnn.cpp:659-684, estimate_k :274-309)
    double k_val, omega_val;
    if (kprovided) {
        k_val = k_in[oc];
        omega_val = omega_in[oc];
    } else {
        const double u_tau = 1e-6;       // the reference's u_tau loop never counts a row (:280-296) -> floor
        const double y_plus = y_wall * u_tau / (prm.nu + 1e-20);
        const double f_mu = 1.0 - exp(-fmin(y_plus / 26.0, 20.0));
        const double k_est = (u_tau * u_tau / sqrt(C_MU)) * f_mu * f_mu;
        k_val = fmax(prm.k_min, fmin(k_est, 10.0 * u_tau * u_tau));
        omega_val = sqrt(k_val) / (KAPPA * fmax(y_wall, 1e-10));
    }
    c.k_val = k_val;
    // compute_features_tbnn (src/features.cpp:121-190)
    const double eps = C_MU * k_val * omega_val;
    const double tau = fmax(k_val, 1e-10) / fmax(eps, 1e-20);
    {
        const double S_norm = S_mag * tau, Omega_norm = Omega_mag * tau;
        feat[0] = S_norm * S_norm;
        feat[1] = Omega_norm * Omega_norm;
        const double Sxx = Sxx_ * tau, Syy = Syy_ * tau, Sxy = Sxy_ * tau, Oxy = Oxy_ * tau;
        feat[2] = Sxx * Sxx + Syy * Syy + 2.0 * Sxy * Sxy;
        feat[3] = 2.0 * Oxy * Oxy;
        feat[4] = y_wall / prm.delta;
    }
    // tensor basis on the stretched-aware MAC gradients (src/gpu_kernels.cpp:103-131, features.cpp:192-292)
    double gdudy = dudy, gdvdy = dvdy;
    if (g.y_stretched) {
        const double inv_dy_local = 1.0 / (g.dyc[j] + g.dyc[j + 1]);
        gdudy = (U2(i, j + 1) - U2(i, j - 1)) * inv_dy_local;
        gdvdy = (V2(i, j + 1) - V2(i, j - 1)) * inv_dy_local;
    }
    {
        const double Sxx = dudx * tau, Syy = gdvdy * tau;
        const double Sxy = (0.5 * (gdudy + dvdx)) * tau, Oxy = (0.5 * (gdudy - dvdx)) * tau;
        tensor_basis(Sxx, Sxy, Syy, Oxy, c);
    }
#undef U2
#undef V2
}

// MLP output of one cell -> nu_t (and tau_ij for TBNN).
template <int KIND>
__device__ __forceinline__ void cell_finish(const NNParams& prm, const CellCtx& c, const double* out, int nout,
                                            size_t oc, double* __restrict__ nu_t, double* __restrict__ txx,
                                            double* __restrict__ txy, double* __restrict__ tyy) {
    if (KIND == CFDNN_TURB_NN_MLP) {
        double nn = out[0];                                   // turbulence_nn_mlp.cpp:262-278
        if (nn - nn != 0.0) nn = 0.0;
        else { nn = fmax(0.0, nn); nn = fmin(nn, prm.nu_t_max); }
        nu_t[oc] = nn;
        return;
    }
    double b_xx = 0.0, b_xy = 0.0, b_yy = 0.0;                // construct_anisotropy :294-306
#pragma unroll
    for (int n = 0; n < 4; ++n) {
        const double G = n < nout ? out[n] : 0.0;
        b_xx += G * c.T[n][0]; b_xy += G * c.T[n][1]; b_yy += G * c.T[n][2];
    }
    if (txx) {                                                // anisotropy_to_reynolds_stress :308-326
        const double k_safe = fmax(c.k_val, 0.0);
        txx[oc] = 2.0 * k_safe * (b_xx + 1.0 / 3.0);
        txy[oc] = 2.0 * k_safe * b_xy;
        tyy[oc] = 2.0 * k_safe * (b_yy + 1.0 / 3.0);
    }
    double nt;                                                // turbulence_nn_tbnn.cpp:846-864
    if (fabs(c.Sxy_) > 1e-10) {
        nt = fabs(-b_xy * c.k_val / c.Sxy_);
    } else if (c.S_mag > 1e-10) {
        const double b_mag = sqrt(b_xx * b_xx + 2.0 * b_xy * b_xy + b_yy * b_yy);
        nt = c.k_val * b_mag / c.S_mag;
    } else {
        nt = 0.0;
    }
    nt = fmax(0.0, fmin(nt, 10.0 * prm.nu));
    if (nt - nt != 0.0) nt = 0.0;
    nu_t[oc] = nt;
}

// ---- scalar (SIMT) path: one thread per cell of one z-plane.  Used for networks whose weight
// fragments do not fit the tensor-core kernel's shared memory, by CFDNN_NN_SIMT=1 (A/B checks)
// and by the host-side test double.  KIND: CFDNN_TURB_NN_TBNN or CFDNN_TURB_NN_MLP.
template <int KIND>
__global__ void __launch_bounds__(128)
k_nn_update(Geom g, NNDev net, NNParams prm, const double* __restrict__ yc, double y_lo, double y_hi,
            const double* __restrict__ ub, const double* __restrict__ vb, const double* __restrict__ k_in,
            const double* __restrict__ omega_in, const int* __restrict__ kflag, int plane0,
            double* __restrict__ nu_t, double* __restrict__ txx, double* __restrict__ txy,
            double* __restrict__ tyy) {
    const int ix = blockIdx.x * blockDim.x + threadIdx.x;
    const int iy = blockIdx.y * blockDim.y + threadIdx.y;
    const int pl = blockIdx.z;
    if (ix >= g.Nx || iy >= g.Ny) return;
    const int i = ix + g.Ng, j = iy + g.Ng, kk = plane0 + pl;
    const size_t oc = (size_t)kk * g.cp + (size_t)j * g.cs + i;
    double feat[8];
    double out[NN_MAX_WIDTH];
    CellCtx c;
    cell_features<KIND>(g, prm, yc, y_lo, y_hi, ub + (size_t)kk * g.up, vb + (size_t)kk * g.vp, k_in, omega_in,
                        KIND == CFDNN_TURB_NN_TBNN ? kflag[pl] : 0, i, j, oc, feat, c);
    mlp_forward(net, feat, out);
    cell_finish<KIND>(prm, c, out, net.dims[net.n_layers], oc, nu_t, txx, txy, tyy);
}

#ifndef CFDNN_DEVSIM
// ---------------------------------------------------------------- tensor-core path (FP64 DMMA)
// The hidden-layer GEMMs  Y[cells x out] = act(X[cells x in] W^T + b)  run on the FP64 tensor
// cores with mma.sync.m8n8k4.f64.  One warp owns 32 consecutive cells of a row: each lane builds
// the features of its own cell, then the warp pushes the cells through the network 16 at a time
// (two 8-row M tiles share every weight fragment).  The D fragment of layer l (row = lane/4,
// columns 2q, 2q+1 of each 8-wide N tile, q = lane%4) is used directly as the A fragments of
// layer l+1 by enumerating K in the order (tile, parity, q) instead of ascending — the weights
// are stored pre-permuted in that fragment order (NNClosure::build_fragments), so no shuffles or
// shared-memory round trips sit between layers.  Summation order inside a dot product differs
// from the reference's ascending-j loop (and DMMA fuses), which moves results by O(1e-16) —
// the closure's gate is 1e-6.
__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// tanh to ~2e-13 absolute: t = exp(-2|x|) (Cody-Waite + degree-10 Taylor on |r| <= ln2/2), then
// (1-t)/(1+t) with a Newton-refined reciprocal.  ~22 FP64 instructions instead of libdevice's
// branchy sequence; explicit fma intrinsics (the library is compiled with -fmad=false).
__device__ __forceinline__ double fast_tanh(double x) {
    const double ax = fmin(fabs(x), 20.0);
    const double y = -2.0 * ax;
    const double MAGIC = 6755399441055744.0;                       // 1.5 * 2^52
    double fn = __fma_rn(y, 1.4426950408889634, MAGIC);
    const int n = __double2loint(fn);
    fn = fn - MAGIC;
    double r = __fma_rn(fn, -6.93147180369123816490e-01, y);
    r = __fma_rn(fn, -1.90821492927058770002e-10, r);
    double p = 2.7557319223985893e-07;                              // 1/10!
    p = __fma_rn(p, r, 2.7557319223985888e-06);                     // 1/9!
    p = __fma_rn(p, r, 2.4801587301587302e-05);
    p = __fma_rn(p, r, 1.9841269841269841e-04);
    p = __fma_rn(p, r, 1.3888888888888889e-03);
    p = __fma_rn(p, r, 8.3333333333333332e-03);
    p = __fma_rn(p, r, 4.1666666666666664e-02);
    p = __fma_rn(p, r, 1.6666666666666666e-01);
    p = __fma_rn(p, r, 0.5);
    p = __fma_rn(p, r, 1.0);
    p = __fma_rn(p, r, 1.0);
    const double t = __hiloint2double(__double2hiint(p) + n * 1048576, __double2loint(p));   // p * 2^n, n in [-58, 0]
    const double num = 1.0 - t, den = 1.0 + t;                     // den in [1, 2]
    double rc;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(rc) : "d"(den));
    double e = __fma_rn(-den, rc, 1.0);
    rc = __fma_rn(rc, e, rc);
    e = __fma_rn(-den, rc, 1.0);
    rc = __fma_rn(rc, e, rc);
    const double q = num * rc;
    return x != x ? x : copysign(q, x);                            // NaN in -> NaN out
}

__device__ __forceinline__ double act_mma(double s, int act) {
    if (act == 1) return fast_tanh(s);
    if (act == 2) return s > 0.0 ? s : 0.0;
    return s;
}

// tanh for the shape-specialised kernel.  The FP64 FMA pipe and the DMMA pipe are ONE resource on sm_100a
// (scripts/fp64_peaks.cu: 37.1 TFLOP/s of DMMA alone, 33.4 of DFMA alone, 32-35 combined in any mix), so every FP64
// instruction of the activation is taken from the GEMMs.  Same scheme as fast_tanh, t = exp(-2|x|),
// tanh = (1 - t) / (1 + t), with the work moved off that pipe where possible: the argument reduction is to
// 1/32 of an octave, exp = 2^m * 2^(j/32) * P4(r) with 2^(j/32) from a 32-entry shared-memory table (stored 16 times
// so that the lanes of a half-warp read 16 different banks whatever j they need) and a degree-4 polynomial on
// |r| <= ln2/64; |x| is clamped, the 2^m scaling, the sign and the NaN test are integer operations on the high word;
// one Newton step on the 20-bit reciprocal seed.  12 FP64 instructions instead of 26; absolute error <= 1e-12.
constexpr int TANH_TAB_DOUBLES = 32 * 16;
__device__ __forceinline__ double tanh_tab(double x, const double* __restrict__ tab /* + (lane & 15) */) {
    const int hx = __double2hiint(x);
    const int habs = hx & 0x7fffffff;
    const int hcl = min(habs, 0x40340000);                          // |x| clamped just above 20: tanh = 1 - 8e-18 there
    const double ax = __hiloint2double(hcl, __double2loint(x));
    const double MAGIC = 6755399441055744.0;                        // 1.5 * 2^52
    double fn = __fma_rn(ax, -92.33248261689366, MAGIC);            // n = round(-2|x| * 32 / ln2) in [-1847, 0]
    const int n = __double2loint(fn);
    fn = fn - MAGIC;
    // -2|x| = n ln2/32 + r, |r| <= ln2/64; the polynomial is written in s = -r/2 = |x| + n ln2/64 so that the
    // product -2|x| never has to be formed: exp(r) = 1 - 2s + 2s^2 - 4/3 s^3 + 2/3 s^4
    const double sr = __fma_rn(fn, 0.010830424696249145, ax);       // ln2 / 64
    double p = __fma_rn(sr, 6.6666666666666663e-01, -1.3333333333333333e+00);
    p = __fma_rn(p, sr, 2.0);
    p = __fma_rn(p, sr, -2.0);
    p = __fma_rn(p, sr, 1.0);
    // -2 * 2^(n/32): table value with the exponent moved by m + 1 and the sign bit set (integer pipe)
    const double T = tab[(n & 31) * 16];
    const double Tm2 = __hiloint2double((__double2hiint(T) + (((n >> 5) + 1) << 20)) | 0x80000000, __double2loint(T));
    const double tm2 = Tm2 * p;                                     // -2 t,  t = exp(-2|x|)
    const double den = __fma_rn(tm2, -0.5, 1.0);                    // 1 + t in [1, 2]
    double rc;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(rc) : "d"(den));
    const double e = __fma_rn(-den, rc, 1.0);
    rc = __fma_rn(rc, e, rc);
    const double q = __fma_rn(tm2, rc, 1.0);                        // (1 - t) / (1 + t) = 1 - 2t / (1 + t)
    const double qs = __hiloint2double(__double2hiint(q) | (hx & 0x80000000), __double2loint(q));
    return habs > 0x7ff00000 ? x : qs;                              // NaN in -> NaN out; +-inf -> +-1
}

constexpr int MMA_WARPS = 4;

template <int KIND>
__global__ void __launch_bounds__(MMA_WARPS * 32)
k_nn_update_mma(Geom g, NNFrag net, NNParams prm, const double* __restrict__ yc, double y_lo, double y_hi,
                const double* __restrict__ ub, const double* __restrict__ vb, const double* __restrict__ k_in,
                const double* __restrict__ omega_in, const int* __restrict__ kflag, int plane0,
                double* __restrict__ nu_t, double* __restrict__ txx, double* __restrict__ txy,
                double* __restrict__ tyy) {
    extern __shared__ double2 smem2[];
    double2* Wsm = smem2;                                               // fragments
    double* bsm = reinterpret_cast<double*>(Wsm + net.w_total);         // biases
    double* xsm = bsm + ((net.b_total + 1) & ~1);                       // [warp][32 cells][8] features
    double* gsm = xsm + MMA_WARPS * 32 * 8;                             // [warp][32 cells][8] outputs
    for (int t = threadIdx.x; t < net.w_total; t += blockDim.x) Wsm[t] = net.W[t];
    for (int t = threadIdx.x; t < net.b_total; t += blockDim.x) bsm[t] = net.b[t];
    __syncthreads();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int grp = lane >> 2, q = lane & 3;
    double* xw = xsm + warp * 256;
    double* gw = gsm + warp * 256;
    const int pl = blockIdx.z;
    const int kk = plane0 + pl;
    const double* u = ub + (size_t)kk * g.up;
    const double* v = vb + (size_t)kk * g.vp;
    const int kprov = KIND == CFDNN_TURB_NN_TBNN ? kflag[pl] : 0;
    const int ix = blockIdx.x * 32 + lane;
    const int ixc = min(ix, g.Nx - 1);                                  // clamp: tail lanes recompute the last cell
    const int nrows = (g.Ny + MMA_WARPS - 1) / MMA_WARPS;

    for (int rg = blockIdx.y; rg < nrows; rg += gridDim.y) {
        const int iy = rg * MMA_WARPS + warp;
        if (iy >= g.Ny) continue;                                       // warp-uniform
        const int i = ixc + g.Ng, j = iy + g.Ng;
        const size_t oc = (size_t)kk * g.cp + (size_t)j * g.cs + i;
        CellCtx c;
        {
            double feat[8];
#pragma unroll
            for (int f = 0; f < 8; ++f) feat[f] = 0.0;
            cell_features<KIND>(g, prm, yc, y_lo, y_hi, u, v, k_in, omega_in, kprov, i, j, oc, feat, c);
            if (net.means) {
#pragma unroll
                for (int f = 0; f < 8; ++f)
                    if (f < net.n_in) feat[f] = (feat[f] - net.means[f]) / net.stds[f];
            }
#pragma unroll
            for (int f = 0; f < 8; f += 2)
                *reinterpret_cast<double2*>(xw + lane * 8 + f) = make_double2(feat[f], feat[f + 1]);
        }
        __syncwarp();

#pragma unroll 1
        for (int half = 0; half < 2; ++half) {
            // A fragments: a[mt][tile] = columns (tile*8 + 2q, +1) of cell row half*16 + mt*8 + grp
            double2 a[2][8];
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) {
                a[mt][0] = *reinterpret_cast<const double2*>(xw + (half * 16 + mt * 8 + grp) * 8 + 2 * q);
#pragma unroll
                for (int t = 1; t < 8; ++t) a[mt][t] = make_double2(0.0, 0.0);
            }
#pragma unroll 1
            for (int l = 0; l < net.n_layers; ++l) {
                const double2* Wl = Wsm + net.woff[l];
                const double* bl = bsm + net.boff[l];
                const int nti = net.nti[l], nto = net.nto[l], act = net.act[l];
                double2 o[2][8];
                // two N tiles x two M tiles per step: four independent accumulator chains in flight
#pragma unroll
                for (int no = 0; no < 8; no += 2) {
                    if (no < nto) {
                        const bool two = no + 1 < nto;                  // warp-uniform
                        const double2 bias0 = *reinterpret_cast<const double2*>(bl + no * 8 + 2 * q);
                        const double2 bias1 = two ? *reinterpret_cast<const double2*>(bl + (no + 1) * 8 + 2 * q)
                                                  : make_double2(0.0, 0.0);
                        double2 c00 = bias0, c10 = bias0, c01 = bias1, c11 = bias1;   // c[mt][n]
                        const double2* W0 = Wl + (no * nti) * 32 + lane;
                        const double2* W1 = Wl + ((two ? no + 1 : no) * nti) * 32 + lane;
#pragma unroll
                        for (int ni = 0; ni < 8; ++ni) {
                            if (ni < nti) {
                                const double2 w0 = W0[ni * 32];
                                const double2 w1 = W1[ni * 32];
                                dmma(c00.x, c00.y, a[0][ni].x, w0.x);
                                dmma(c10.x, c10.y, a[1][ni].x, w0.x);
                                dmma(c01.x, c01.y, a[0][ni].x, w1.x);
                                dmma(c11.x, c11.y, a[1][ni].x, w1.x);
                                dmma(c00.x, c00.y, a[0][ni].y, w0.y);
                                dmma(c10.x, c10.y, a[1][ni].y, w0.y);
                                dmma(c01.x, c01.y, a[0][ni].y, w1.y);
                                dmma(c11.x, c11.y, a[1][ni].y, w1.y);
                            }
                        }
                        o[0][no] = make_double2(act_mma(c00.x, act), act_mma(c00.y, act));
                        o[1][no] = make_double2(act_mma(c10.x, act), act_mma(c10.y, act));
                        if (two) {
                            o[0][no + 1] = make_double2(act_mma(c01.x, act), act_mma(c01.y, act));
                            o[1][no + 1] = make_double2(act_mma(c11.x, act), act_mma(c11.y, act));
                        } else {
                            o[0][no + 1] = make_double2(0.0, 0.0);
                            o[1][no + 1] = make_double2(0.0, 0.0);
                        }
                    } else {
                        o[0][no] = make_double2(0.0, 0.0); o[1][no] = make_double2(0.0, 0.0);
                        o[0][no + 1] = make_double2(0.0, 0.0); o[1][no + 1] = make_double2(0.0, 0.0);
                    }
                }
#pragma unroll
                for (int t = 0; t < 8; ++t) { a[0][t] = o[0][t]; a[1][t] = o[1][t]; }
            }
            // outputs (<= 8 wide, tile 0): row = cell, columns 2q, 2q+1
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
                *reinterpret_cast<double2*>(gw + (half * 16 + mt * 8 + grp) * 8 + 2 * q) = a[mt][0];
        }
        __syncwarp();
        if (ix < g.Nx) {
            double out[4];
#pragma unroll
            for (int n = 0; n < 4; ++n) out[n] = gw[lane * 8 + n];
            cell_finish<KIND>(prm, c, out, net.n_out, oc, nu_t, txx, txy, tyy);
        }
        __syncwarp();
    }
}
// ---- shape-specialised tensor-core kernel (the product path for the shipped presets).
// Networks of the form  in(<=8) -> H*8 -> ... -> H*8 -> out(<=8)  with tanh hidden layers and a linear
// output (5-64-64-64-4 TBNN: H = 8; 6-32-32-1 MLP: H = 4).  Differences to the generic kernel above:
//   * every tile loop has compile-time bounds and there is no branch inside a layer, so ptxas sees one
//     straight-line block per layer and interleaves the tanh of one group of N tiles (FP64 FMA pipe)
//     with the DMMAs of the next group (tensor pipe);
//   * a warp pushes ONE 8-cell M tile through the network at a time with FOUR N tiles in flight
//     (four independent accumulator chains, as before) — half the live activations, so the kernel fits
//     128 registers and 16 warps per SM (4 per scheduler) instead of 8, which is what hides the DMMA
//     and tanh latencies;
//   * the per-cell context of the nu_t stage (7 doubles) waits in shared memory, not in registers;
//   * the single-N-tile output layer splits K over four chains instead of one 16-deep dependent chain.
constexpr int MMAH_WARPS = 8;

template <int NTI, int NTO, bool TANH>
__device__ __forceinline__ void mma_layer(const double2* __restrict__ Wl, const double* __restrict__ bl, int lane,
                                          int q, const double* __restrict__ tab, const double2 (&a)[NTI],
                                          double2 (&o)[NTO]) {
    if constexpr (NTO == 1) {
        // one N tile: K tiles dealt to four chains, summed pairwise at the end
        double2 c[4];
        c[0] = *reinterpret_cast<const double2*>(bl + 2 * q);
#pragma unroll
        for (int j = 1; j < 4; ++j) c[j] = make_double2(0.0, 0.0);
#pragma unroll
        for (int ni = 0; ni < NTI; ++ni) {
            const double2 w = Wl[ni * 32 + lane];
            dmma(c[ni & 3].x, c[ni & 3].y, a[ni].x, w.x);
            dmma(c[ni & 3].x, c[ni & 3].y, a[ni].y, w.y);
        }
        double2 r;
        r.x = (c[0].x + c[1].x) + (c[2].x + c[3].x);
        r.y = (c[0].y + c[1].y) + (c[2].y + c[3].y);
        o[0] = TANH ? make_double2(tanh_tab(r.x, tab), tanh_tab(r.y, tab)) : r;
    } else {
        static_assert(NTO % 4 == 0, "N tiles are processed four at a time");
#pragma unroll
        for (int ng = 0; ng < NTO; ng += 4) {
            double2 c[4];
#pragma unroll
            for (int n = 0; n < 4; ++n) c[n] = *reinterpret_cast<const double2*>(bl + (ng + n) * 8 + 2 * q);
#pragma unroll
            for (int ni = 0; ni < NTI; ++ni) {
                double2 w[4];
#pragma unroll
                for (int n = 0; n < 4; ++n) w[n] = Wl[((ng + n) * NTI + ni) * 32 + lane];
#pragma unroll
                for (int n = 0; n < 4; ++n) dmma(c[n].x, c[n].y, a[ni].x, w[n].x);
#pragma unroll
                for (int n = 0; n < 4; ++n) dmma(c[n].x, c[n].y, a[ni].y, w[n].y);
            }
#pragma unroll
            for (int n = 0; n < 4; ++n)
                o[ng + n] = TANH ? make_double2(tanh_tab(c[n].x, tab), tanh_tab(c[n].y, tab)) : c[n];
        }
    }
}

template <int KIND, int H>
__global__ void __launch_bounds__(MMAH_WARPS * 32, 2)
k_nn_update_mmah(Geom g, NNFrag net, NNParams prm, const double* __restrict__ yc, double y_lo, double y_hi,
                 const double* __restrict__ ub, const double* __restrict__ vb, const double* __restrict__ k_in,
                 const double* __restrict__ omega_in, const int* __restrict__ kflag, int plane0,
                 double* __restrict__ nu_t, double* __restrict__ txx, double* __restrict__ txy,
                 double* __restrict__ tyy) {
    extern __shared__ double2 smem2[];
    double2* Wsm = smem2;                                               // fragments
    double* bsm = reinterpret_cast<double*>(Wsm + net.w_total);         // biases
    double* xsm = bsm + ((net.b_total + 1) & ~1);                       // [warp][32 cells][8]: features in, outputs back
    double* csm = xsm + MMAH_WARPS * 32 * 8;                            // [warp][8][32 cells]: context of the nu_t stage
    double* tsm = csm + MMAH_WARPS * 32 * 8;                            // [32 entries][16 copies]: 2^(j/32) for tanh_tab
    for (int t = threadIdx.x; t < TANH_TAB_DOUBLES; t += blockDim.x) tsm[t] = exp2((double)(t >> 4) * 0.03125);
    for (int t = threadIdx.x; t < net.w_total; t += blockDim.x) Wsm[t] = net.W[t];
    for (int t = threadIdx.x; t < net.b_total; t += blockDim.x) bsm[t] = net.b[t];
    __syncthreads();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int grp = lane >> 2, q = lane & 3;
    double* xw = xsm + warp * 256;
    double* cw = csm + warp * 256;
    const double* tab = tsm + (lane & 15);
    const int pl = blockIdx.z;
    const int kk = plane0 + pl;
    const double* u = ub + (size_t)kk * g.up;
    const double* v = vb + (size_t)kk * g.vp;
    const int kprov = KIND == CFDNN_TURB_NN_TBNN ? kflag[pl] : 0;
    const int ix = blockIdx.x * 32 + lane;
    const int ixc = min(ix, g.Nx - 1);                                  // clamp: tail lanes recompute the last cell
    const int nrows = (g.Ny + MMAH_WARPS - 1) / MMAH_WARPS;
    const int nhid = net.n_layers - 2;                                  // hidden -> hidden layers

    for (int rg = blockIdx.y; rg < nrows; rg += gridDim.y) {
        const int iy = rg * MMAH_WARPS + warp;
        if (iy >= g.Ny) continue;                                       // warp-uniform
        const int i = ixc + g.Ng, j = iy + g.Ng;
        const size_t oc = (size_t)kk * g.cp + (size_t)j * g.cs + i;
        {
            CellCtx c;
            double feat[8];
#pragma unroll
            for (int f = 0; f < 8; ++f) feat[f] = 0.0;
            cell_features<KIND>(g, prm, yc, y_lo, y_hi, u, v, k_in, omega_in, kprov, i, j, oc, feat, c);
            if (net.means) {
#pragma unroll
                for (int f = 0; f < 8; ++f)
                    if (f < net.n_in) feat[f] = (feat[f] - net.means[f]) / net.stds[f];
            }
#pragma unroll
            for (int f = 0; f < 8; f += 2)
                *reinterpret_cast<double2*>(xw + lane * 8 + f) = make_double2(feat[f], feat[f + 1]);
            // T1 = (Sxx, Sxy, Syy) and Oxy rebuild the whole tensor basis (features.cpp:192-292)
            cw[0 * 32 + lane] = c.T[0][0]; cw[1 * 32 + lane] = c.T[0][1]; cw[2 * 32 + lane] = c.T[0][2];
            cw[3 * 32 + lane] = c.Oxy; cw[4 * 32 + lane] = c.k_val; cw[5 * 32 + lane] = c.Sxy_;
            cw[6 * 32 + lane] = c.S_mag;
        }
        __syncwarp();

#pragma unroll 1
        for (int mt = 0; mt < 4; ++mt) {
            double* xrow = xw + (mt * 8 + grp) * 8 + 2 * q;             // this lane's column pair of its cell row
            double2 a[H], o[H];
            {
                double2 a0[1];
                a0[0] = *reinterpret_cast<const double2*>(xrow);
                mma_layer<1, H, true>(Wsm + net.woff[0], bsm + net.boff[0], lane, q, tab, a0, a);
            }
#pragma unroll 1
            for (int l = 1; l <= nhid; ++l) {
                mma_layer<H, H, true>(Wsm + net.woff[l], bsm + net.boff[l], lane, q, tab, a, o);
#pragma unroll
                for (int t = 0; t < H; ++t) a[t] = o[t];
            }
            double2 r[1];
            mma_layer<H, 1, false>(Wsm + net.woff[nhid + 1], bsm + net.boff[nhid + 1], lane, q, tab, a, r);
            *reinterpret_cast<double2*>(xrow) = r[0];                   // outputs 2q, 2q+1 of cell mt*8 + grp
        }
        __syncwarp();
        if (ix < g.Nx) {
            double out[4];
#pragma unroll
            for (int n = 0; n < 4; ++n) out[n] = xw[lane * 8 + n];
            CellCtx c;
            if (KIND == CFDNN_TURB_NN_TBNN) {
                const double Sxx = cw[0 * 32 + lane], Sxy = cw[1 * 32 + lane], Syy = cw[2 * 32 + lane];
                tensor_basis(Sxx, Sxy, Syy, cw[3 * 32 + lane], c);
            }
            c.k_val = cw[4 * 32 + lane]; c.Sxy_ = cw[5 * 32 + lane]; c.S_mag = cw[6 * 32 + lane];
            cell_finish<KIND>(prm, c, out, net.n_out, oc, nu_t, txx, txy, tyy);
        }
        __syncwarp();
    }
}
#endif  // !CFDNN_DEVSIM
}  // namespace

// ---------------------------------------------------------------- host side
NNClosure::NNClosure(int kind, int n_layers, const int* dims, const int* act, const double* const* W,
                     const double* const* b, const double* means, const double* stds) : kind_(kind) {
    require_device();
    if (kind != CFDNN_TURB_NN_MLP && kind != CFDNN_TURB_NN_TBNN) throw std::invalid_argument("nn_create: bad closure kind");
    if (n_layers < 1 || n_layers > NN_MAX_LAYERS) throw std::invalid_argument("nn_create: 1..8 layers supported");
    net_.n_layers = n_layers;
    size_t nw = 0, nb = 0;
    for (int l = 0; l <= n_layers; ++l) {
        if (dims[l] < 1 || dims[l] > NN_MAX_WIDTH) throw std::invalid_argument("nn_create: layer width must be 1..64");
        net_.dims[l] = dims[l];
    }
    if (net_.dims[0] > 8) throw std::invalid_argument("nn_create: at most 8 input features");
    for (int l = 0; l < n_layers; ++l) {
        net_.act[l] = act ? act[l] : (l + 1 < n_layers ? 1 : 0);   // tanh hidden, linear output (nn_core.cpp:532-543)
        if (net_.act[l] < 0 || net_.act[l] > 2) throw std::invalid_argument("nn_create: activation must be 0..2");
        net_.woff[l] = (int)nw; net_.boff[l] = (int)nb;
        nw += (size_t)dims[l] * dims[l + 1]; nb += dims[l + 1];
    }
    std::vector<double> hw(nw), hb(nb);
    for (int l = 0; l < n_layers; ++l) {
        if (!W[l] || !b[l]) throw std::invalid_argument("nn_create: null weight pointer");
        std::copy(W[l], W[l] + (size_t)dims[l] * dims[l + 1], hw.begin() + net_.woff[l]);
        std::copy(b[l], b[l] + dims[l + 1], hb.begin() + net_.boff[l]);
    }
    for (double x : hw) if (!std::isfinite(x)) throw Error(CFDNN_ERR_RUNTIME, "Neural network weights must be finite values");
    for (double x : hb) if (!std::isfinite(x)) throw Error(CFDNN_ERR_RUNTIME, "Neural network biases must be finite values");
    W_.alloc(nw); b_.alloc(nb);
    CFDNN_CUDA(cudaMemcpy(W_.p, hw.data(), nw * sizeof(double), cudaMemcpyHostToDevice));
    CFDNN_CUDA(cudaMemcpy(b_.p, hb.data(), nb * sizeof(double), cudaMemcpyHostToDevice));
    net_.W = W_.p; net_.b = b_.p;
    if (means && stds) {
        means_.alloc(dims[0]); stds_.alloc(dims[0]);
        CFDNN_CUDA(cudaMemcpy(means_.p, means, dims[0] * sizeof(double), cudaMemcpyHostToDevice));
        CFDNN_CUDA(cudaMemcpy(stds_.p, stds, dims[0] * sizeof(double), cudaMemcpyHostToDevice));
        net_.means = means_.p; net_.stds = stds_.p; net_.n_scale = dims[0];
    }
    build_fragments(hw, hb);
    CFDNN_CUDA(cudaStreamCreateWithFlags(&st_, cudaStreamNonBlocking));
    CFDNN_CUDA(cudaEventCreate(&ev0_));
    CFDNN_CUDA(cudaEventCreate(&ev1_));
}

NNClosure::~NNClosure() {
    if (ev0_) cudaEventDestroy(ev0_);
    if (ev1_) cudaEventDestroy(ev1_);
    if (st_) cudaStreamDestroy(st_);
}

// Weights in DMMA fragment order (see k_nn_update_mma): layer l holds [nto][nti][32 lanes] double2,
// lane (grp = lane/4, q = lane%4) of tile pair (no, ni) gets W_l[no*8 + grp][ni*8 + 2q + {0,1}],
// zero where the index is outside the layer (K and N padded to multiples of 8).
void NNClosure::build_fragments(const std::vector<double>& hw, const std::vector<double>& hb) {
    frag_ = NNFrag{};
    mma_ok_ = false;
    const int nl = net_.n_layers;
    size_t nw = 0, nb = 0;
    for (int l = 0; l < nl; ++l) {
        frag_.nti[l] = (net_.dims[l] + 7) / 8;
        frag_.nto[l] = (net_.dims[l + 1] + 7) / 8;
        frag_.act[l] = net_.act[l];
        frag_.woff[l] = (int)nw; frag_.boff[l] = (int)nb;
        nw += (size_t)frag_.nti[l] * frag_.nto[l] * 32;
        nb += (size_t)frag_.nto[l] * 8;
    }
    frag_.n_layers = nl; frag_.w_total = (int)nw; frag_.b_total = (int)nb;
    frag_.n_in = net_.dims[0]; frag_.n_out = net_.dims[nl];
    mma_smem_ = nw * sizeof(double2) + ((nb + 1) & ~(size_t)1) * sizeof(double) + 2 * 4 * 32 * 8 * sizeof(double);
    // tensor-core path: inputs and outputs fit one 8-wide tile, fragments fit shared memory
    if (net_.dims[0] > 8 || net_.dims[nl] > 8 || mma_smem_ > 200 * 1024) return;
    std::vector<double2> fw(nw);
    std::vector<double> fb(nb, 0.0);
    for (int l = 0; l < nl; ++l) {
        const int in = net_.dims[l], od = net_.dims[l + 1];
        const double* W = hw.data() + net_.woff[l];
        for (int no = 0; no < frag_.nto[l]; ++no)
            for (int ni = 0; ni < frag_.nti[l]; ++ni)
                for (int lane = 0; lane < 32; ++lane) {
                    const int row = no * 8 + (lane >> 2), col = ni * 8 + 2 * (lane & 3);
                    double2 w;
                    w.x = (row < od && col < in) ? W[(size_t)row * in + col] : 0.0;
                    w.y = (row < od && col + 1 < in) ? W[(size_t)row * in + col + 1] : 0.0;
                    fw[frag_.woff[l] + ((size_t)no * frag_.nti[l] + ni) * 32 + lane] = w;
                }
        for (int i = 0; i < od; ++i) fb[frag_.boff[l] + i] = hb[net_.boff[l] + i];
    }
    Wf_.alloc(nw, false); bf_.alloc(nb, false);
    CFDNN_CUDA(cudaMemcpy(Wf_.p, fw.data(), nw * sizeof(double2), cudaMemcpyHostToDevice));
    CFDNN_CUDA(cudaMemcpy(bf_.p, fb.data(), nb * sizeof(double), cudaMemcpyHostToDevice));
    frag_.W = Wf_.p; frag_.b = bf_.p;
#ifndef CFDNN_DEVSIM
    CFDNN_CUDA(cudaFuncSetAttribute(k_nn_update_mma<CFDNN_TURB_NN_TBNN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mma_smem_));
    CFDNN_CUDA(cudaFuncSetAttribute(k_nn_update_mma<CFDNN_TURB_NN_MLP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mma_smem_));
    const char* e = std::getenv("CFDNN_NN_SIMT");
    mma_ok_ = !(e && std::atoi(e) != 0);
    // shape-specialised kernel: in(<=8) -> 8H -> ... -> 8H -> out(<=8), tanh hidden layers, linear output
    mmah_ = 0;
    if (nl >= 2 && net_.act[nl - 1] == 0) {
        const int hw = net_.dims[1];
        bool ok = hw == 64 || hw == 32;
        for (int l = 1; l < nl; ++l) ok = ok && net_.dims[l] == hw;
        for (int l = 0; l + 1 < nl; ++l) ok = ok && net_.act[l] == 1;
        if (ok) mmah_ = hw / 8;
    }
    if (const char* g = std::getenv("CFDNN_NN_GENERIC")) if (std::atoi(g) != 0) mmah_ = 0;   // A/B: generic DMMA kernel
    mmah_smem_ = nw * sizeof(double2) + ((nb + 1) & ~(size_t)1) * sizeof(double) +
                 (2 * MMAH_WARPS * 32 * 8 + TANH_TAB_DOUBLES) * sizeof(double);
    if (mmah_smem_ > 113 * 1024) mmah_ = 0;             // two CTAs per SM
    if (mmah_) {
        CFDNN_CUDA(cudaFuncSetAttribute(k_nn_update_mmah<CFDNN_TURB_NN_TBNN, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mmah_smem_));
        CFDNN_CUDA(cudaFuncSetAttribute(k_nn_update_mmah<CFDNN_TURB_NN_MLP, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mmah_smem_));
        CFDNN_CUDA(cudaFuncSetAttribute(k_nn_update_mmah<CFDNN_TURB_NN_TBNN, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mmah_smem_));
        CFDNN_CUDA(cudaFuncSetAttribute(k_nn_update_mmah<CFDNN_TURB_NN_MLP, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mmah_smem_));
    }
#endif
}

void NNClosure::set_params(double nu, double delta, double u_ref, double k_min, double nu_t_max) {
    prm_ = NNParams{nu, delta, u_ref, k_min, nu_t_max};
}

void NNClosure::update_device(const Geom& g, const double* yc, double y_lo, double y_hi, const Vel& vel,
                              const double* k, const double* omega, double* nu_t, double* txx, double* txy,
                              double* tyy, cudaStream_t st, int data_plane) {
    const bool is2d = g.Nz == 1;
    // 2-D accessors address plane k = 0 (include/fields.hpp:18-19); the solver keeps 2-D data in plane Ng
    const int plane0 = data_plane >= 0 ? data_plane : (is2d ? 0 : g.Ng);
    const int nplanes = is2d ? 1 : g.Nz;
    if ((size_t)nplanes > kflag_.n) kflag_.alloc(nplanes);
    const dim3 blk(32, 4, 1);
    const dim3 grd((g.Nx + 31) / 32, (g.Ny + 3) / 4, nplanes);
    if (kind_ == CFDNN_TURB_NN_TBNN && net_.dims[0] != 5) throw std::invalid_argument("TBNN closure expects 5 input features");
    if (kind_ == CFDNN_TURB_NN_MLP && net_.dims[0] != 6) throw std::invalid_argument("MLP closure expects 6 input features");
#ifndef CFDNN_DEVSIM
    if (mma_ok_) {
        // tensor-core kernel: 4 warps x 32 cells per CTA pass; CTAs stride over row groups so the
        // weight fragments are staged into shared memory a few times per SM, not once per 128 cells
        frag_.means = net_.means; frag_.stds = net_.stds;
        const int xt = (g.Nx + 31) / 32, rows = (g.Ny + MMA_WARPS - 1) / MMA_WARPS;
        const long long want = 148LL * 8;
        int gy = (int)std::max(1LL, std::min<long long>(rows, (want + (long long)xt * nplanes - 1) / ((long long)xt * nplanes)));
        const dim3 mg(xt, gy, nplanes);
        if (mmah_) {
            // 8 warps per CTA, two CTAs per SM; CTAs stride over groups of 8 rows
            const int rows8 = (g.Ny + MMAH_WARPS - 1) / MMAH_WARPS;
            const long long want8 = 148LL * 2 * 4;
            const int gy8 = (int)std::max(1LL, std::min<long long>(rows8, (want8 + (long long)xt * nplanes - 1) / ((long long)xt * nplanes)));
            const dim3 hg(xt, gy8, nplanes);
            if (kind_ == CFDNN_TURB_NN_TBNN) {
                CFDNN_CUDA(cudaMemsetAsync(kflag_.p, 0, nplanes * sizeof(int), st));
                CFDNN_LAUNCH(k_k_provided, grd, blk, 0, st, g, k, prm_.k_min, plane0, kflag_.p);
            }
#define CFDNN_MMAH(KIND, HH)                                                                                       \
            CFDNN_LAUNCH((k_nn_update_mmah<KIND, HH>), hg, MMAH_WARPS * 32, mmah_smem_, st, g, frag_, prm_, yc, y_lo, \
                         y_hi, vel.u, vel.v, k, omega, kflag_.p, plane0, nu_t, txx, txy, tyy)
            if (kind_ == CFDNN_TURB_NN_TBNN) { if (mmah_ == 8) CFDNN_MMAH(CFDNN_TURB_NN_TBNN, 8); else CFDNN_MMAH(CFDNN_TURB_NN_TBNN, 4); }
            else { if (mmah_ == 8) CFDNN_MMAH(CFDNN_TURB_NN_MLP, 8); else CFDNN_MMAH(CFDNN_TURB_NN_MLP, 4); }
#undef CFDNN_MMAH
            return;
        }
        if (kind_ == CFDNN_TURB_NN_TBNN) {
            CFDNN_CUDA(cudaMemsetAsync(kflag_.p, 0, nplanes * sizeof(int), st));
            CFDNN_LAUNCH(k_k_provided, grd, blk, 0, st, g, k, prm_.k_min, plane0, kflag_.p);
            CFDNN_LAUNCH((k_nn_update_mma<CFDNN_TURB_NN_TBNN>), mg, MMA_WARPS * 32, mma_smem_, st, g, frag_, prm_, yc, y_lo,
                         y_hi, vel.u, vel.v, k, omega, kflag_.p, plane0, nu_t, txx, txy, tyy);
        } else {
            CFDNN_LAUNCH((k_nn_update_mma<CFDNN_TURB_NN_MLP>), mg, MMA_WARPS * 32, mma_smem_, st, g, frag_, prm_, yc, y_lo,
                         y_hi, vel.u, vel.v, k, omega, kflag_.p, plane0, nu_t, txx, txy, tyy);
        }
        return;
    }
#endif
    if (kind_ == CFDNN_TURB_NN_TBNN) {
        if (net_.dims[0] != 5) throw std::invalid_argument("TBNN closure expects 5 input features");
        CFDNN_CUDA(cudaMemsetAsync(kflag_.p, 0, nplanes * sizeof(int), st));
        CFDNN_LAUNCH(k_k_provided, grd, blk, 0, st, g, k, prm_.k_min, plane0, kflag_.p);
        CFDNN_LAUNCH((k_nn_update<CFDNN_TURB_NN_TBNN>), grd, blk, 0, st, g, net_, prm_, yc, y_lo, y_hi, vel.u, vel.v,
                     k, omega, kflag_.p, plane0, nu_t, txx, txy, tyy);
    } else {
        if (net_.dims[0] != 6) throw std::invalid_argument("MLP closure expects 6 input features");
        CFDNN_LAUNCH((k_nn_update<CFDNN_TURB_NN_MLP>), grd, blk, 0, st, g, net_, prm_, yc, y_lo, y_hi, vel.u, vel.v,
                     k, omega, kflag_.p, plane0, nu_t, txx, txy, tyy);
    }
}

void NNClosure::update_host(const cfdnn_mesh_desc& m, const double* u, const double* v, const double* k,
                            const double* omega, double* nu_t, double* txx, double* txy, double* tyy) {
    if (m.Ng != 1) throw std::invalid_argument("nn_update: only Nghost == 1 is supported");
    Geom g = make_geom(m);
    const int Ng = m.Ng, nzt = m.Nz + 2 * Ng;
    const size_t su = (size_t)g.up * nzt, sv = (size_t)g.vp * nzt, sc = (size_t)g.cp * nzt;
    DevBuf<double> du, dv, dk, dom, dnut, dt0, dt1, dt2, dyc, ddyc;
    du.alloc(su, false); dv.alloc(sv, false); dk.alloc(sc); dom.alloc(sc); dnut.alloc(sc);
    const bool want_tau = txx && txy && tyy && kind_ == CFDNN_TURB_NN_TBNN;
    if (want_tau) { dt0.alloc(sc); dt1.alloc(sc); dt2.alloc(sc); }
    std::vector<double> yc(m.Ny + 2 * Ng);
    for (int j = 0; j < m.Ny + 2 * Ng; ++j) yc[j] = m.yc ? m.yc[j] : m.y_lo + (j - Ng + 0.5) * m.dy;   // src/mesh.cpp:66-68
    dyc.alloc(yc.size(), false);
    CFDNN_CUDA(cudaMemcpyAsync(dyc.p, yc.data(), yc.size() * sizeof(double), cudaMemcpyHostToDevice, st_));
    if (m.y_stretched) {
        if (!m.dyc) throw std::invalid_argument("nn_update: stretched mesh needs dyc");
        ddyc.alloc(m.Ny + 2 * Ng + 1, false);
        CFDNN_CUDA(cudaMemcpyAsync(ddyc.p, m.dyc, (m.Ny + 2 * Ng + 1) * sizeof(double), cudaMemcpyHostToDevice, st_));
        g.dyc = ddyc.p;
    }
    CFDNN_CUDA(cudaMemcpyAsync(du.p, u, su * sizeof(double), cudaMemcpyHostToDevice, st_));
    CFDNN_CUDA(cudaMemcpyAsync(dv.p, v, sv * sizeof(double), cudaMemcpyHostToDevice, st_));
    if (k) CFDNN_CUDA(cudaMemcpyAsync(dk.p, k, sc * sizeof(double), cudaMemcpyHostToDevice, st_));
    if (omega) CFDNN_CUDA(cudaMemcpyAsync(dom.p, omega, sc * sizeof(double), cudaMemcpyHostToDevice, st_));
    CFDNN_CUDA(cudaMemcpyAsync(dnut.p, nu_t, sc * sizeof(double), cudaMemcpyHostToDevice, st_));
    Vel vel; vel.u = du.p; vel.v = dv.p; vel.w = nullptr;
    CFDNN_CUDA(cudaEventRecord(ev0_, st_));
    update_device(g, dyc.p, m.y_lo, m.y_hi, vel, dk.p, dom.p, dnut.p, want_tau ? dt0.p : nullptr,
                  want_tau ? dt1.p : nullptr, want_tau ? dt2.p : nullptr, st_);
    CFDNN_CUDA(cudaEventRecord(ev1_, st_));
    CFDNN_CUDA(cudaMemcpyAsync(nu_t, dnut.p, sc * sizeof(double), cudaMemcpyDeviceToHost, st_));
    if (want_tau) {
        CFDNN_CUDA(cudaMemcpyAsync(txx, dt0.p, sc * sizeof(double), cudaMemcpyDeviceToHost, st_));
        CFDNN_CUDA(cudaMemcpyAsync(txy, dt1.p, sc * sizeof(double), cudaMemcpyDeviceToHost, st_));
        CFDNN_CUDA(cudaMemcpyAsync(tyy, dt2.p, sc * sizeof(double), cudaMemcpyDeviceToHost, st_));
    }
    CFDNN_CUDA(cudaStreamSynchronize(st_));
    check_launch("nn_update");
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev0_, ev1_);
    last_ms = ms;
}

void NNClosure::update_dev_ptrs(const cfdnn_mesh_desc& m, const double* u, const double* v, const double* k,
                                const double* omega, double* nu_t, double* txx, double* txy, double* tyy) {
    if (m.Ng != 1) throw std::invalid_argument("nn_update_dev: only Nghost == 1 is supported");
    if (!u || !v || !nu_t) throw std::invalid_argument("nn_update_dev: u, v and nu_t device pointers are required");
    Geom g = make_geom(m);
    const int Ng = m.Ng, nzt = m.Nz + 2 * Ng;
    const size_t sc = (size_t)g.cp * nzt;
    DevBuf<double> dyc, ddyc, zk, zo;
    std::vector<double> yc(m.Ny + 2 * Ng);
    for (int j = 0; j < m.Ny + 2 * Ng; ++j) yc[j] = m.yc ? m.yc[j] : m.y_lo + (j - Ng + 0.5) * m.dy;   // src/mesh.cpp:66-68
    dyc.alloc(yc.size(), false);
    CFDNN_CUDA(cudaMemcpyAsync(dyc.p, yc.data(), yc.size() * sizeof(double), cudaMemcpyHostToDevice, st_));
    if (m.y_stretched) {
        if (!m.dyc) throw std::invalid_argument("nn_update_dev: stretched mesh needs dyc");
        ddyc.alloc(m.Ny + 2 * Ng + 1, false);
        CFDNN_CUDA(cudaMemcpyAsync(ddyc.p, m.dyc, (m.Ny + 2 * Ng + 1) * sizeof(double), cudaMemcpyHostToDevice, st_));
        g.dyc = ddyc.p;
    }
    if (!k) { zk.alloc(sc); k = zk.p; }            // absent k / omega == all zero (estimated by the closure)
    if (!omega) { zo.alloc(sc); omega = zo.p; }
    const bool want_tau = txx && txy && tyy && kind_ == CFDNN_TURB_NN_TBNN;
    Vel vel; vel.u = const_cast<double*>(u); vel.v = const_cast<double*>(v); vel.w = nullptr;
    CFDNN_CUDA(cudaEventRecord(ev0_, st_));
    update_device(g, dyc.p, m.y_lo, m.y_hi, vel, k, omega, nu_t, want_tau ? txx : nullptr, want_tau ? txy : nullptr,
                  want_tau ? tyy : nullptr, st_);
    CFDNN_CUDA(cudaEventRecord(ev1_, st_));
    CFDNN_CUDA(cudaStreamSynchronize(st_));
    check_launch("nn_update_dev");
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev0_, ev1_);
    last_ms = ms;
}

void NNClosure::forward_host(const double* x, double* y, size_t n) {
    if (n == 0) return;
    const int in = net_.dims[0], od = net_.dims[net_.n_layers];
    DevBuf<double> dx, dy;
    dx.alloc(n * in, false); dy.alloc(n * od, false);
    CFDNN_CUDA(cudaMemcpyAsync(dx.p, x, n * in * sizeof(double), cudaMemcpyHostToDevice, st_));
    CFDNN_CUDA(cudaEventRecord(ev0_, st_));
    CFDNN_LAUNCH(k_mlp_forward, (unsigned)((n + 127) / 128), 128, 0, st_, net_, dx.p, dy.p, n);
    CFDNN_CUDA(cudaEventRecord(ev1_, st_));
    CFDNN_CUDA(cudaMemcpyAsync(y, dy.p, n * od * sizeof(double), cudaMemcpyDeviceToHost, st_));
    CFDNN_CUDA(cudaStreamSynchronize(st_));
    check_launch("nn_forward");
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev0_, ev1_);
    last_ms = ms;
}

}  // namespace cfdnn
