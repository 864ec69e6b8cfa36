// nn.cu — NN closures: feature extraction -> MLP -> tensor-basis contraction -> nu_t, fused per cell.
#include "nn.cuh"
#include <cmath>
#include <algorithm>

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

// One thread per cell of one z-plane.  KIND: CFDNN_TURB_NN_TBNN or CFDNN_TURB_NN_MLP.
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
    const double* u = ub + (size_t)kk * g.up;
    const double* v = vb + (size_t)kk * g.vp;
    const size_t oc = (size_t)kk * g.cp + (size_t)j * g.cs + i;
#define U2(I, J) u[(size_t)(J) * g.us + (I)]
#define V2(I, J) v[(size_t)(J) * g.vs + (I)]
    // uniform-spacing MAC gradients (features.cpp:132-140, turbulence_nn_tbnn.cpp:838-845)
    const double inv_2dx = 1.0 / (2.0 * g.dx.d);
    const double inv_2dy = 1.0 / (2.0 * g.dy.d);
    const double dudx = (U2(i + 1, j) - U2(i - 1, j)) * inv_2dx;
    const double dudy = (U2(i, j + 1) - U2(i, j - 1)) * inv_2dy;
    const double dvdx = (V2(i + 1, j) - V2(i - 1, j)) * inv_2dx;
    const double dvdy = (V2(i, j + 1) - V2(i, j - 1)) * inv_2dy;
    const double Sxx_ = dudx, Syy_ = dvdy, Sxy_ = 0.5 * (dudy + dvdx), Oxy_ = 0.5 * (dudy - dvdx);
    const double S_mag = sqrt(Sxx_ * Sxx_ + Syy_ * Syy_ + 2.0 * Sxy_ * Sxy_);
    const double Omega_mag = sqrt(2.0 * Oxy_ * Oxy_);
    const double y_wall = wall_distance(yc[j], y_lo, y_hi);
    double feat[8];
    double out[NN_MAX_WIDTH];

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
        mlp_forward(net, feat, out);
        double nn = out[0];                                   // turbulence_nn_mlp.cpp:262-278
        if (nn - nn != 0.0) nn = 0.0;
        else { nn = fmax(0.0, nn); nn = fmin(nn, prm.nu_t_max); }
        nu_t[oc] = nn;
        return;
    }

    // ---- TBNN: k / omega (turbulence_nn_tbnn.cpp:659-684, estimate_k :274-309)
    double k_val, omega_val;
    if (kflag[pl]) {
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
    double T[4][3];
    {
        const double Sxx = dudx * tau, Syy = gdvdy * tau;
        const double Sxy = (0.5 * (gdudy + dvdx)) * tau, Oxy = (0.5 * (gdudy - dvdx)) * tau;
        T[0][0] = Sxx; T[0][1] = Sxy; T[0][2] = Syy;
        T[1][0] = -2.0 * Sxy * Oxy; T[1][1] = (Sxx - Syy) * Oxy; T[1][2] = 2.0 * Sxy * Oxy;
        const double S2xx = Sxx * Sxx + Sxy * Sxy, S2yy = Sxy * Sxy + Syy * Syy, S2xy = Sxy * (Sxx + Syy);
        const double trS2 = S2xx + S2yy;
        T[2][0] = S2xx - 0.5 * trS2; T[2][1] = S2xy; T[2][2] = S2yy - 0.5 * trS2;
        T[3][0] = 0.0; T[3][1] = 0.0; T[3][2] = 0.0;
    }
    mlp_forward(net, feat, out);
    const int nout = net.dims[net.n_layers];
    double b_xx = 0.0, b_xy = 0.0, b_yy = 0.0;                // construct_anisotropy :294-306
    for (int n = 0; n < 4; ++n) {
        const double G = n < nout ? out[n] : 0.0;
        b_xx += G * T[n][0]; b_xy += G * T[n][1]; b_yy += G * T[n][2];
    }
    if (txx) {                                                // anisotropy_to_reynolds_stress :308-326
        const double k_safe = fmax(k_val, 0.0);
        txx[oc] = 2.0 * k_safe * (b_xx + 1.0 / 3.0);
        txy[oc] = 2.0 * k_safe * b_xy;
        tyy[oc] = 2.0 * k_safe * (b_yy + 1.0 / 3.0);
    }
    double nt;                                                // turbulence_nn_tbnn.cpp:846-864
    if (fabs(Sxy_) > 1e-10) {
        nt = fabs(-b_xy * k_val / Sxy_);
    } else if (S_mag > 1e-10) {
        const double b_mag = sqrt(b_xx * b_xx + 2.0 * b_xy * b_xy + b_yy * b_yy);
        nt = k_val * b_mag / S_mag;
    } else {
        nt = 0.0;
    }
    nt = fmax(0.0, fmin(nt, 10.0 * prm.nu));
    if (nt - nt != 0.0) nt = 0.0;
    nu_t[oc] = nt;
#undef U2
#undef V2
}
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
    CFDNN_CUDA(cudaStreamCreateWithFlags(&st_, cudaStreamNonBlocking));
    CFDNN_CUDA(cudaEventCreate(&ev0_));
    CFDNN_CUDA(cudaEventCreate(&ev1_));
}

NNClosure::~NNClosure() {
    if (ev0_) cudaEventDestroy(ev0_);
    if (ev1_) cudaEventDestroy(ev1_);
    if (st_) cudaStreamDestroy(st_);
}

void NNClosure::set_params(double nu, double delta, double u_ref, double k_min, double nu_t_max) {
    prm_ = NNParams{nu, delta, u_ref, k_min, nu_t_max};
}

void NNClosure::update_device(const Geom& g, const double* yc, double y_lo, double y_hi, const Vel& vel,
                              const double* k, const double* omega, double* nu_t, double* txx, double* txy,
                              double* tyy, cudaStream_t st) {
    const bool is2d = g.Nz == 1;
    const int plane0 = is2d ? 0 : g.Ng;          // 2-D accessors address plane k = 0 (include/fields.hpp:18-19)
    const int nplanes = is2d ? 1 : g.Nz;
    if ((size_t)nplanes > kflag_.n) kflag_.alloc(nplanes);
    const dim3 blk(32, 4, 1);
    const dim3 grd((g.Nx + 31) / 32, (g.Ny + 3) / 4, nplanes);
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
