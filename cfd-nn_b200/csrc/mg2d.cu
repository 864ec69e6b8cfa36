// mg2d.cu — 2-D multigrid, whole V-cycle sequences in one single-CTA launch (see mg2d.cuh).
// file:line citations are to the reference's src/poisson_solver_multigrid.cpp (CPU path).
#include "mg2d.cuh"
#include <cmath>
#include <algorithm>

namespace cfdnn {

namespace {

#ifndef CFDNN_DEVSIM
constexpr int MG2D_CTA = 1024;
#else
constexpr int MG2D_CTA = 64;          // host test double: fewer fibers per barrier
#endif
constexpr int SEQ_CAP = 2048;         // staging buffer of the in-order sums (doubles)

// ------------------------------------------------------------------ ghost fill
// apply_bc 2-D (:340-455): x pass over all rows, y pass over all columns, and when any side is
// periodic both passes once more (corner consistency).  Residual mode (:590-640, :700-745):
// periodic wrap else 0, x pass then y pass.
__device__ __forceinline__ void bc_x(const MG2DParams& P, const Lvl2D& L, double* a, int j, bool res) {
    double* row = a + (size_t)j * L.S;
    const int blo = P.bc[0], bhi = P.bc[1];
    if (blo == CFDNN_PBC_PERIODIC) row[0] = row[L.Nx];
    else if (res) row[0] = 0.0;
    else if (blo == CFDNN_PBC_NEUMANN) row[0] = row[1];
    else row[0] = 2.0 * 0.0 - row[1];
    if (bhi == CFDNN_PBC_PERIODIC) row[L.Nx + 1] = row[1];
    else if (res) row[L.Nx + 1] = 0.0;
    else if (bhi == CFDNN_PBC_NEUMANN) row[L.Nx + 1] = row[L.Nx];
    else row[L.Nx + 1] = 2.0 * 0.0 - row[L.Nx];
}
__device__ __forceinline__ void bc_y(const MG2DParams& P, const Lvl2D& L, double* a, int i, bool res) {
    const size_t S = (size_t)L.S;
    const int blo = P.bc[2], bhi = P.bc[3];
    if (blo == CFDNN_PBC_PERIODIC) a[i] = a[L.Ny * S + i];
    else if (res) a[i] = 0.0;
    else if (blo == CFDNN_PBC_NEUMANN) a[i] = a[S + i];
    else a[i] = 2.0 * 0.0 - a[S + i];
    if (bhi == CFDNN_PBC_PERIODIC) a[(L.Ny + 1) * S + i] = a[S + i];
    else if (res) a[(L.Ny + 1) * S + i] = 0.0;
    else if (bhi == CFDNN_PBC_NEUMANN) a[(L.Ny + 1) * S + i] = a[L.Ny * S + i];
    else a[(L.Ny + 1) * S + i] = 2.0 * 0.0 - a[L.Ny * S + i];
}
__device__ void cta_bc(const MG2DParams& P, const Lvl2D& L, double* a, bool res) {
    const bool corner = !res && (P.bc[0] == 2 || P.bc[1] == 2 || P.bc[2] == 2 || P.bc[3] == 2);
    const int passes = corner ? 2 : 1;
    __syncthreads();
    for (int p = 0; p < passes; ++p) {
        for (int j = threadIdx.x; j < L.Ny + 2; j += blockDim.x) bc_x(P, L, a, j, res);
        __syncthreads();
        for (int i = threadIdx.x; i < L.Nx + 2; i += blockDim.x) bc_y(P, L, a, i, res);
        __syncthreads();
    }
}

// ------------------------------------------------------------------ one weighted-Jacobi sweep u -> u
// (smooth_chebyshev :930-958, smooth_jacobi :1085-1112): tmp = (1-w) u + w * jacobi(u), then u <- tmp
// and apply_bc(u)
__device__ void cta_sweep(const MG2DParams& P, const Lvl2D& L, double omega) {
    const int n = L.Nx * L.Ny;
    const double om1 = 1.0 - omega;
    const size_t S = (size_t)L.S;
    for (int t = threadIdx.x; t < n; t += blockDim.x) {
        const int i = t % L.Nx + 1, j = t / L.Nx + 1;
        const size_t idx = (size_t)j * S + i;
        const double u_old = L.u[idx];
        double u_jacobi;
        if (P.semi) {
            const double lap_x = (L.u[idx + 1] + L.u[idx - 1]) / L.dx2;
            const double lap_y = P.aS[j] * L.u[idx - S] + P.aN[j] * L.u[idx + S];
            const double diag_j = L.two_inv_dx2 + (P.aS[j] + P.aN[j]);
            u_jacobi = (lap_x + lap_y - L.f[idx]) / diag_j;
        } else {
            u_jacobi = ((L.u[idx + 1] + L.u[idx - 1]) / L.dx2 + (L.u[idx + S] + L.u[idx - S]) / L.dy2 - L.f[idx]) / L.coeff;
        }
        L.tmp[idx] = om1 * u_old + omega * u_jacobi;
    }
    __syncthreads();
    for (int t = threadIdx.x; t < n; t += blockDim.x) {
        const size_t idx = (size_t)(t / L.Nx + 1) * S + (t % L.Nx + 1);
        L.u[idx] = L.tmp[idx];
    }
    cta_bc(P, L, L.u, false);
}

// r = f - L u on the interior (compute_residual :1759-1795)
__device__ void cta_residual(const MG2DParams& P, const Lvl2D& L) {
    const int n = L.Nx * L.Ny;
    const size_t S = (size_t)L.S;
    for (int t = threadIdx.x; t < n; t += blockDim.x) {
        const int i = t % L.Nx + 1, j = t / L.Nx + 1;
        const size_t idx = (size_t)j * S + i;
        const double uc = L.u[idx];
        double lap;
        if (P.semi) {
            const double lap_x = (L.u[idx + 1] - 2.0 * uc + L.u[idx - 1]) / L.dx2;
            const double lap_y = P.aS[j] * L.u[idx - S] + P.aP[j] * uc + P.aN[j] * L.u[idx + S];
            lap = lap_x + lap_y;
        } else {
            lap = (L.u[idx + 1] - 2.0 * uc + L.u[idx - 1]) / L.dx2 + (L.u[idx + S] - 2.0 * uc + L.u[idx - S]) / L.dy2;
        }
        L.r[idx] = L.f[idx] - lap;
    }
    __syncthreads();
}

// interior sum in the reference's order (i fastest, one running double)
__device__ double cta_seq_sum(const Lvl2D& L, const double* a, double* buf, double* slot) {
    const int n = L.Nx * L.Ny;
    double acc = 0.0;
    for (int base = 0; base < n; base += SEQ_CAP) {
        const int m = min(SEQ_CAP, n - base);
        __syncthreads();
        for (int t = threadIdx.x; t < m; t += blockDim.x) {
            const int q = base + t;
            buf[t] = a[(size_t)(q / L.Nx + 1) * L.S + (q % L.Nx + 1)];
        }
        __syncthreads();
        if (threadIdx.x == 0)
            for (int t = 0; t < m; ++t) acc += buf[t];
    }
    __syncthreads();
    if (threadIdx.x == 0) *slot = acc;
    __syncthreads();
    return *slot;
}
// make_rhs_mean_free / subtract_mean (:3479-3548, :3411-3477)
__device__ void cta_mean_free(const Lvl2D& L, double* a, double* buf, double* slot) {
    const double mean = cta_seq_sum(L, a, buf, slot) / (L.Nx * L.Ny);
    const int n = L.Nx * L.Ny;
    for (int t = threadIdx.x; t < n; t += blockDim.x) a[(size_t)(t / L.Nx + 1) * L.S + (t % L.Nx + 1)] -= mean;
    __syncthreads();
}

// restriction (:1995-2022 full weighting; :2100-2122 x-only for semi-coarsening)
__device__ void cta_restrict(const MG2DParams& P, const Lvl2D& F, const Lvl2D& C) {
    const int n = C.Nx * C.Ny;
    const size_t Sf = (size_t)F.S;
    for (int t = threadIdx.x; t < n; t += blockDim.x) {
        const int ic = t % C.Nx + 1, jc = t / C.Nx + 1;
        const int i_f = 1 + 2 * (ic - 1);
        double v;
        if (P.semi) {
            const size_t idx = (size_t)jc * Sf + i_f;
            v = 0.25 * F.r[idx - 1] + 0.50 * F.r[idx] + 0.25 * F.r[idx + 1];
        } else {
            const size_t idx = (size_t)(1 + 2 * (jc - 1)) * Sf + i_f;
            v = 0.25 * F.r[idx] + 0.125 * (F.r[idx - 1] + F.r[idx + 1] + F.r[idx - Sf] + F.r[idx + Sf])
              + 0.0625 * (F.r[idx - 1 - Sf] + F.r[idx + 1 - Sf] + F.r[idx - 1 + Sf] + F.r[idx + 1 + Sf]);
        }
        C.f[(size_t)jc * C.S + ic] = v;
    }
    __syncthreads();
}

// prolongation (:2183-2219 bilinear toward i_c+1 / j_c+1; :2296-2324 x-only, alpha = 1)
__device__ void cta_prolong(const MG2DParams& P, const Lvl2D& F, const Lvl2D& C) {
    const int n = F.Nx * F.Ny;
    const size_t Sc = (size_t)C.S;
    for (int t = threadIdx.x; t < n; t += blockDim.x) {
        const int i_f = t % F.Nx + 1, j_f = t / F.Nx + 1;
        const int ic = (i_f - 1) / 2 + 1, di = (i_f - 1) & 1;
        const double wx1 = 0.5 * di, wx0 = 1.0 - wx1;
        const size_t o = (size_t)j_f * F.S + i_f;
        if (P.semi) {
            const size_t c = (size_t)j_f * Sc + ic;
            const double correction = wx0 * C.u[c] + wx1 * C.u[c + 1];
            const double alpha = 1.0;
            F.u[o] += alpha * correction;
        } else {
            const int jc = (j_f - 1) / 2 + 1, dj = (j_f - 1) & 1;
            const double wy1 = 0.5 * dj, wy0 = 1.0 - wy1;
            const size_t c = (size_t)jc * Sc + ic;
            const double correction = wx0 * wy0 * C.u[c] + wx1 * wy0 * C.u[c + 1] + wx0 * wy1 * C.u[c + Sc]
                                    + wx1 * wy1 * C.u[c + 1 + Sc];
            F.u[o] += correction;
        }
    }
    __syncthreads();
}

// do_smooth of vcycle (:3143-3172), 2-D branches
__device__ void cta_smooth(const MG2DParams& P, const Lvl2D& L, int nu, int degree) {
    for (int pass = 0; pass < nu; ++pass) {
        if (P.semi) {
            for (int it = 0; it < degree * 4; ++it) cta_sweep(P, L, 0.67);
        } else {
            cta_bc(P, L, L.u, false);                     // smooth_chebyshev applies the BCs first (:909)
            for (int k = 0; k < degree; ++k) cta_sweep(P, L, P.omega[k]);
        }
    }
}

// `ncycles` V-cycles (:3121-3322) in one launch
__global__ void __launch_bounds__(MG2D_CTA) k_mg2d_vcycles(MG2DParams P, int ncycles, int nu1, int nu2, int degree) {
    __shared__ double buf[SEQ_CAP];
    __shared__ double slot;
    for (int c = 0; c < ncycles; ++c) {
        for (int l = 0; l < P.nl - 1; ++l) {
            const Lvl2D& L = P.L[l];
            const Lvl2D& C = P.L[l + 1];
            cta_smooth(P, L, nu1, degree);
            cta_residual(P, L);
            cta_bc(P, L, L.r, true);
            cta_restrict(P, L, C);
            if (P.nullspace) cta_mean_free(C, C.f, buf, &slot);
            for (int t = threadIdx.x; t < C.total; t += blockDim.x) C.u[t] = 0.0;
            __syncthreads();
        }
        {
            const Lvl2D& L = P.L[P.nl - 1];                // coarsest (:3176-3197)
            const int iters = P.semi ? (P.nullspace ? 400 : 200) : (P.nullspace ? 100 : 50);
            for (int it = 0; it < iters; ++it) cta_sweep(P, L, 0.67);
        }
        for (int l = P.nl - 2; l >= 0; --l) {
            const Lvl2D& L = P.L[l];
            cta_prolong(P, L, P.L[l + 1]);
            cta_bc(P, L, L.u, false);
            if (nu2 > 0) cta_smooth(P, L, nu2, degree);
        }
    }
}

// level-0 helpers: copy in (+ apply_bc), residual + norms, ||b|| norms, nullspace fix, copy out
// mode 0: f,u <- interior of rhs,p (:3558-3566), apply_bc(0)
// mode 1: r = f - L u; out[0] = max|r|, out[1] = sum r^2   (compute_residual_and_norms :1833-1970)
// mode 2: out[0] = max|f|, out[1] = sum f^2                 (:3660-3670)
// mode 3: fix_nullspace(0) (:3404-3409) then p <- all cells of u (:3769-3775)
__global__ void __launch_bounds__(MG2D_CTA)
k_mg2d_level0(MG2DParams P, int mode, const double* rhs, double* p, double* out, int dup_planes, long long plane_stride) {
    __shared__ double buf[SEQ_CAP];
    __shared__ double slot;
    __shared__ double sh[32];
    const Lvl2D& L = P.L[0];
    const int n = L.Nx * L.Ny;
    if (mode == 0) {
        for (int t = threadIdx.x; t < n; t += blockDim.x) {
            const size_t idx = (size_t)(t / L.Nx + 1) * L.S + (t % L.Nx + 1);
            L.f[idx] = rhs[idx];
            L.u[idx] = p[idx];
        }
        cta_bc(P, L, L.u, false);
    } else if (mode == 1 || mode == 2) {
        if (mode == 1) cta_residual(P, L);
        const double* a = mode == 1 ? L.r : L.f;
        double mx = 0.0, sq = 0.0;
        for (int t = threadIdx.x; t < n; t += blockDim.x) {
            const double v = a[(size_t)(t / L.Nx + 1) * L.S + (t % L.Nx + 1)];
            mx = fmax(mx, fabs(v)); sq += v * v;
        }
        mx = block_max(mx, sh);
        sq = block_sum(sq, sh);
        if (threadIdx.x == 0) { out[0] = mx; out[1] = sq; }
    } else {
        if (P.nullspace) {
            cta_mean_free(L, L.u, buf, &slot);
            cta_bc(P, L, L.u, false);
        }
        __syncthreads();
        for (int t = threadIdx.x; t < L.total; t += blockDim.x) {
            const double v = L.u[t];
            p[t] = v;
            if (dup_planes) { p[t - plane_stride] = v; p[t + plane_stride] = v; }
        }
    }
}

}  // namespace

// ---------------------------------------------------------------- host side
// create_hierarchy 2-D branches (:176-178, :189-196, :211-218), compute_gershgorin_bounds (:232-305)
Multigrid2D::Multigrid2D(const cfdnn_mesh_desc& m) : mesh_(m) {
    require_device();
    if (m.Nz != 1) throw std::invalid_argument("Multigrid2D: 2-D meshes only (Nz == 1)");
    if (m.Ng != 1) throw std::invalid_argument("Multigrid2D: only Nghost == 1 is supported");
    P_.semi = m.y_stretched != 0;
    if (P_.semi) {
        if (!m.yLap_aS || !m.yLap_aN || !m.yLap_aP) throw std::invalid_argument("Multigrid2D: stretched mesh needs yLap_aS/aN/aP");
        const int ny2 = m.Ny + 2;
        aS_.alloc(ny2, false); aN_.alloc(ny2, false); aP_.alloc(ny2, false);
        CFDNN_CUDA(cudaMemcpy(aS_.p, m.yLap_aS, ny2 * sizeof(double), cudaMemcpyHostToDevice));
        CFDNN_CUDA(cudaMemcpy(aN_.p, m.yLap_aN, ny2 * sizeof(double), cudaMemcpyHostToDevice));
        CFDNN_CUDA(cudaMemcpy(aP_.p, m.yLap_aP, ny2 * sizeof(double), cudaMemcpyHostToDevice));
        P_.aS = aS_.p; P_.aN = aN_.p; P_.aP = aP_.p;
    }
    int Nx = m.Nx, Ny = m.Ny;
    double dx = m.dx, dy = m.dy;
    auto add = [&](int nx, int ny, double hx, double hy) {
        if (P_.nl >= MG2D_MAX_LEVELS) throw std::invalid_argument("Multigrid2D: too many levels");
        Lvl2D& L = P_.L[P_.nl++];
        L.Nx = nx; L.Ny = ny; L.S = nx + 2; L.total = L.S * (ny + 2);
        L.dx2 = hx * hx; L.dy2 = hy * hy;
        L.coeff = 2.0 / L.dx2 + 2.0 / L.dy2;
        L.two_inv_dx2 = 2.0 / L.dx2;
        for (double** q : {&L.u, &L.f, &L.r, &L.tmp}) {
            bufs_.emplace_back();
            bufs_.back().alloc(L.total);
            *q = bufs_.back().p;
        }
    };
    bufs_.reserve(4 * MG2D_MAX_LEVELS);
    add(Nx, Ny, dx, dy);
    constexpr int MIN_COARSE = 8;
    if (P_.semi) {
        while (Nx > MIN_COARSE) {
            if (Nx & 1) throw std::invalid_argument("Multigrid2D: Nx must stay even while coarsening");
            Nx /= 2; dx *= 2.0;
            add(Nx, Ny, dx, dy);
        }
    } else {
        while (Nx > MIN_COARSE && Ny > MIN_COARSE) {
            if ((Nx | Ny) & 1) throw std::invalid_argument("Multigrid2D: grid sizes must stay even while coarsening");
            Nx /= 2; Ny /= 2; dx *= 2.0; dy *= 2.0;
            add(Nx, Ny, dx, dy);
        }
    }
    scalars_.alloc(8);
    CFDNN_CUDA(cudaMallocHost(&h_scalars_, 8 * sizeof(double)));
    for (int d = 0; d < 4; ++d) P_.bc[d] = bc_[d];
    P_.nullspace = has_nullspace() ? 1 : 0;
}

Multigrid2D::~Multigrid2D() {
    if (h_scalars_) cudaFreeHost(h_scalars_);
}

void Multigrid2D::set_bc(const int bc[6]) {
    for (int d = 0; d < 4; ++d) {
        if (bc[d] < 0 || bc[d] > 2) throw std::invalid_argument("set_bc: bad PoissonBC value");
        bc_[d] = bc[d];
        P_.bc[d] = bc[d];
    }
    P_.nullspace = has_nullspace() ? 1 : 0;
}

bool Multigrid2D::has_nullspace() const {   // :3393-3402 (the z sides of a 2-D solver stay at their periodic default)
    for (int d = 0; d < 4; ++d) if (bc_[d] == CFDNN_PBC_DIRICHLET) return false;
    return true;
}

// Chebyshev weights (:877-881, :912-918): uniform levels all have lambda_max = max(1.8, 2 * 1.1)
void Multigrid2D::set_degree(int degree) {
    if (degree == degree_cached_) return;
    if (degree < 1 || degree > MG2D_MAX_DEGREE) throw std::invalid_argument("Multigrid2D: chebyshev_degree must be 1..8");
    const double lambda_max = std::max(1.8, (1.0 + 1.0) * 1.10);
    const double lambda_min = 0.1 * lambda_max;
    const double d = (lambda_max + lambda_min) / 2.0;
    const double c = (lambda_max - lambda_min) / 2.0;
    for (int k = 0; k < degree; ++k) {
        const double theta = M_PI * (2.0 * k + 1.0) / (2.0 * degree);
        P_.omega[k] = std::min(1.0 / (d - c * std::cos(theta)), 2.0);
    }
    degree_cached_ = degree;
}

int Multigrid2D::solve_dev(const double* rhs, double* p, const cfdnn_poisson_cfg& cfg, cudaStream_t st, int plane,
                           bool dup_planes) {
    const Lvl2D& L0 = P_.L[0];
    const long long ps = L0.total;                     // plane stride of the caller's arrays
    const double* rhs_pl = rhs + (size_t)plane * ps;
    double* p_pl = p + (size_t)plane * ps;
    if (dup_planes && plane != 1) throw std::invalid_argument("Multigrid2D: dup_planes needs the data in plane 1");
    const int degree = cfg.chebyshev_degree;
    set_degree(degree);
    stats = cfdnn_poisson_stats{};
    double* S = scalars_.p;
    CFDNN_LAUNCH_COOP(k_mg2d_level0, 1, MG2D_CTA, 0, st, P_, 0, rhs_pl, p_pl, S, 0, ps);
    int cycles_used = 0;
    if (cfg.fixed_cycles > 0) {
        const int nu1 = cfg.nu1 > 0 ? cfg.nu1 : 3;
        const int nu2 = cfg.nu2 > 0 ? cfg.nu2 : 1;
        CFDNN_LAUNCH_COOP(k_mg2d_vcycles, 1, MG2D_CTA, 0, st, P_, cfg.fixed_cycles, nu1, nu2, degree);
        cycles_used = cfg.fixed_cycles;
        stats.fixed_cycle_mode = 1;
        stats.status = CFDNN_PSTAT_FIXED;
    } else {
        if (cfg.max_vcycles <= 0) throw std::invalid_argument("PoissonConfig.max_vcycles must be positive");
        const int max_cycles = cfg.max_vcycles;
        const bool accurate = max_cycles > 5;                                  // :3647-3650
        const int nu1 = cfg.nu1 > 0 ? cfg.nu1 : (accurate ? 3 : 2);
        const int nu2 = cfg.nu2 > 0 ? cfg.nu2 : 1;
        const int check_interval = std::max(1, P_.semi ? std::max(4, cfg.check_interval) : cfg.check_interval);
        CFDNN_LAUNCH_COOP(k_mg2d_level0, 1, MG2D_CTA, 0, st, P_, 2, rhs_pl, p_pl, S + 0, 0, ps);
        CFDNN_LAUNCH_COOP(k_mg2d_level0, 1, MG2D_CTA, 0, st, P_, 1, rhs_pl, p_pl, S + 2, 0, ps);
        CFDNN_CUDA(cudaMemcpyAsync(h_scalars_, S, 4 * sizeof(double), cudaMemcpyDeviceToHost, st));
        CFDNN_CUDA(cudaStreamSynchronize(st));
        stats.b_inf = h_scalars_[0];
        stats.b_l2 = std::sqrt(h_scalars_[1]);
        stats.r0 = h_scalars_[2];
        stats.r0_l2 = std::sqrt(h_scalars_[3]);
        stats.residual = stats.r0;
        stats.residual_l2 = stats.r0_l2;
        int cycle = 0;
        while (cycle < max_cycles) {
            // run up to the next convergence check in one launch
            int next = (cycle / check_interval + 1) * check_interval;       // first cycle count with a check
            next = std::min(next, max_cycles);
            CFDNN_LAUNCH_COOP(k_mg2d_vcycles, 1, MG2D_CTA, 0, st, P_, next - cycle, nu1, nu2, degree);
            cycle = next;
            cycles_used = cycle;
            CFDNN_LAUNCH_COOP(k_mg2d_level0, 1, MG2D_CTA, 0, st, P_, 1, rhs_pl, p_pl, S + 2, 0, ps);
            CFDNN_CUDA(cudaMemcpyAsync(h_scalars_ + 2, S + 2, 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
            CFDNN_CUDA(cudaStreamSynchronize(st));
            stats.residual = h_scalars_[2];
            stats.residual_l2 = std::sqrt(h_scalars_[3]);
            const double r_norm = cfg.use_l2_norm ? stats.residual_l2 : stats.residual;
            const double b_norm = cfg.use_l2_norm ? stats.b_l2 : stats.b_inf;
            const double r0_norm = cfg.use_l2_norm ? stats.r0_l2 : stats.r0;
            bool conv = false;                                                  // :3728-3747
            if (cfg.tol_abs > 0.0 && r_norm <= cfg.tol_abs) conv = true;
            if (cfg.tol_rhs > 0.0 && r_norm <= cfg.tol_rhs * (b_norm + 1e-30)) conv = true;
            if (cfg.tol_rel > 0.0 && r_norm <= cfg.tol_rel * (r0_norm + 1e-30)) conv = true;
            if (cfg.tol > 0.0 && r_norm <= cfg.tol) conv = true;
            if (conv && cfg.use_l2_norm && cfg.linf_safety_factor > 0.0 && cfg.tol_rhs > 0.0) {
                const double linf_ratio = stats.residual / (stats.b_inf + 1e-30);
                if (linf_ratio > cfg.tol_rhs * cfg.linf_safety_factor) conv = false;
            }
            if (conv) { stats.converged = 1; break; }
        }
        stats.status = stats.converged ? CFDNN_PSTAT_CONVERGED : CFDNN_PSTAT_HIT_MAX;
    }
    stats.cycles = cycles_used;
    CFDNN_LAUNCH_COOP(k_mg2d_level0, 1, MG2D_CTA, 0, st, P_, 3, rhs_pl, p_pl, S, dup_planes ? 1 : 0, ps);
    return cycles_used;
}

}  // namespace cfdnn
