// mg_semi.cu — semi-coarsening multigrid for y-stretched meshes (see mg_semi.cuh).
// file:line citations are to the reference's src/poisson_solver_multigrid.cpp (CPU path).
#include "mg_semi.cuh"
#include <cmath>
#include <algorithm>
#include <cstring>

namespace cfdnn {

namespace {

#ifndef CFDNN_DEVSIM
constexpr int SEMI_CTA = 1024;        // threads per CTA of the serial-dependency kernels (y-lines, PCG)
constexpr int SEMI_CLUSTER = 8;       // CTAs of the y-line kernel's cluster (portable maximum)
#else
constexpr int SEMI_CTA = 64;          // host test double: fewer fibers per barrier
constexpr int SEMI_CLUSTER = 1;
#endif

__device__ __forceinline__ size_t sidx(const SemiGeom& g, int i, int j, int k) {
    return (size_t)k * (size_t)g.P + (size_t)j * g.S + i;
}

// ------------------------------------------------------------------ ghost fill
// apply_bc (:307-554): periodic u[g] = u[N+g], Neumann u[g] = u[first], Dirichlet u[g] = 2*0 - u[first];
// apply_bc_to_residual (:556-823): periodic wrap, else 0.  One pass = one direction over the FULL
// extent of the other two (ghosts included), passes in x, y, z order.
__device__ __forceinline__ void bc_pair(double* a, size_t lo, size_t lo_src_per, size_t first, size_t hi,
                                        size_t hi_src_per, size_t last, int blo, int bhi, bool res) {
    if (blo == CFDNN_PBC_PERIODIC) a[lo] = a[lo_src_per];
    else if (res) a[lo] = 0.0;
    else if (blo == CFDNN_PBC_NEUMANN) a[lo] = a[first];
    else a[lo] = 2.0 * 0.0 - a[first];
    if (bhi == CFDNN_PBC_PERIODIC) a[hi] = a[hi_src_per];
    else if (res) a[hi] = 0.0;
    else if (bhi == CFDNN_PBC_NEUMANN) a[hi] = a[last];
    else a[hi] = 2.0 * 0.0 - a[last];
}

// one direction, item t of that direction's face (device function: used by the grid kernel and
// by the single-CTA kernels)
template <int DIR>
__device__ __forceinline__ void bc_item(const SemiGeom& g, double* a, int t, bool res) {
    const int SX = g.Nx + 2, SY = g.Ny + 2;
    if (DIR == 0) {
        const int j = t % SY, k = t / SY;
        const size_t b = sidx(g, 0, j, k);
        bc_pair(a, b, b + g.Nx, b + 1, b + g.Nx + 1, b + 1, b + g.Nx, g.bc[0], g.bc[1], res);
    } else if (DIR == 1) {
        const int i = t % SX, k = t / SX;
        const size_t b = sidx(g, i, 0, k), s = (size_t)g.S;
        bc_pair(a, b, b + g.Ny * s, b + s, b + (g.Ny + 1) * s, b + s, b + g.Ny * s, g.bc[2], g.bc[3], res);
    } else {
        const int i = t % SX, j = t / SX;
        const size_t b = sidx(g, i, j, 0), s = (size_t)g.P;
        bc_pair(a, b, b + g.Nz * s, b + s, b + (g.Nz + 1) * s, b + s, b + g.Nz * s, g.bc[4], g.bc[5], res);
    }
}
__device__ __forceinline__ int bc_count(const SemiGeom& g, int dir) {
    return dir == 0 ? (g.Ny + 2) * (g.Nz + 2) : dir == 1 ? (g.Nx + 2) * (g.Nz + 2) : (g.Nx + 2) * (g.Ny + 2);
}

template <int DIR>
__global__ void __launch_bounds__(256) k_semi_bc(SemiGeom g, double* __restrict__ a, int res) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < bc_count(g, DIR)) bc_item<DIR>(g, a, t, res != 0);
}

// A team = the CTAs that run one serial-dependency kernel together: one CTA (PCG, host test
// double) or one thread-block cluster of `n` CTAs (y-line relaxation on the GPU).  team_sync is a
// barrier with release/acquire ordering of global memory across the whole team.
struct Team { int rank, n; };
__device__ __forceinline__ void team_sync(const Team& tm) {
#ifndef CFDNN_DEVSIM
    if (tm.n > 1) {
        asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
        return;
    }
#endif
    __syncthreads();
}

// whole apply_bc by one team (three passes, barrier between them)
__device__ void cta_apply_bc(const SemiGeom& g, double* a, const Team tm = Team{0, 1}) {
    const int t0 = tm.rank * blockDim.x + threadIdx.x, dt = tm.n * blockDim.x;
    team_sync(tm);
    for (int t = t0; t < bc_count(g, 0); t += dt) bc_item<0>(g, a, t, false);
    team_sync(tm);
    for (int t = t0; t < bc_count(g, 1); t += dt) bc_item<1>(g, a, t, false);
    team_sync(tm);
    for (int t = t0; t < bc_count(g, 2); t += dt) bc_item<2>(g, a, t, false);
    team_sync(tm);
}

// ------------------------------------------------------------------ y-line relaxation
// smooth_y_lines (:1420-1716), CPU branch: for k, for i: Thomas solve of the y-line (i,k) with the
// x/z neighbours frozen at their CURRENT values — in place, so line (i,k) sees the new values of
// lines (i-1,k) and (i,k-1) and the old values of (i+1,k), (i,k+1): the lines of one wavefront
// i + k = const are independent and wavefronts run in ascending order.  Ghost cells keep the
// values of the last apply_bc during a pass.  Per wavefront: phase A builds every right-hand side
// (all threads), phase B runs the two Thomas recurrences (one thread per line, c' and the pivots
// are level constants: SemiGeom::ta/tcp/tden).
// sm: [3*Ny coefficient copies][LB*Ny line workspace, layout j*LB + l]
__device__ void cta_y_lines(const SemiGeom& g, double* u, const double* f,
                            int iterations, double* sm, int LB, const Team tm = Team{0, 1}) {
    const int Nx = g.Nx, Ny = g.Ny, Nz = g.Nz;
    double* s_a = sm;
    double* s_cp = sm + Ny;
    double* s_den = sm + 2 * Ny;
    double* s_rden = sm + 3 * Ny;
    double* ws = sm + 4 * Ny;
    for (int j = threadIdx.x; j < Ny; j += blockDim.x) {
        s_a[j] = g.ta[j]; s_cp[j] = g.tcp[j]; s_den[j] = g.tden[j]; s_rden[j] = g.trden[j];
    }
    const bool dir_lo = g.bc[2] != CFDNN_PBC_NEUMANN, dir_hi = g.bc[3] != CFDNN_PBC_NEUMANN;
    const double aS1 = g.aS[1], aNN = g.aN[Ny];
    const size_t S = (size_t)g.S, P = (size_t)g.P;
    __syncthreads();
    for (int it = 0; it < iterations; ++it) {
        for (int d = 0; d <= Nx + Nz - 2; ++d) {
            const int k_lo = max(0, d - (Nx - 1)), k_hi = min(Nz - 1, d);
            const int nl = k_hi - k_lo + 1;
            // this CTA's lines of the wavefront: l = rank, rank + n, ...  (local index ll = l / n)
            const int mine = nl > tm.rank ? (nl - tm.rank + tm.n - 1) / tm.n : 0;
            for (int l0 = 0; l0 < mine; l0 += LB) {
                const int nb = min(LB, mine - l0);
                // phase A: right-hand sides  inv_dx2*(uE+uW) + inv_dz2*(uF+uB) - f  (+ Dirichlet ghost terms)
                for (int t = threadIdx.x; t < nb * Ny; t += blockDim.x) {
                    const int ll = t % nb, jl = t / nb;
                    const int k0 = k_lo + (l0 + ll) * tm.n + tm.rank, i0 = d - k0;
                    const size_t idx = sidx(g, i0 + 1, jl + 1, k0 + 1);
                    double rhs = g.inv_dx2 * (u[idx + 1] + u[idx - 1]) + g.inv_dz2 * (u[idx + P] + u[idx - P]) - f[idx];
                    if (jl == 0 && dir_lo) rhs += aS1 * u[idx - S];
                    if (jl == Ny - 1 && dir_hi) rhs += aNN * u[idx + S];
                    ws[jl * LB + ll] = rhs;
                }
                __syncthreads();
                // phase B: forward elimination, back substitution
                if (threadIdx.x < nb) {
                    const int ll = threadIdx.x;
                    const int k0 = k_lo + (l0 + ll) * tm.n + tm.rank, i0 = d - k0;
                    double dp = fdiv(ws[ll], s_den[0], s_rden[0]);
                    ws[ll] = dp;
                    for (int jl = 1; jl < Ny; ++jl) {
                        dp = fdiv(ws[jl * LB + ll] - s_a[jl] * dp, s_den[jl], s_rden[jl]);
                        ws[jl * LB + ll] = dp;
                    }
                    size_t idx = sidx(g, i0 + 1, Ny, k0 + 1);
                    double un = dp;
                    u[idx] = un;
                    for (int jl = Ny - 2; jl >= 0; --jl) {
                        idx -= S;
                        un = ws[jl * LB + ll] - s_cp[jl] * un;
                        u[idx] = un;
                    }
                }
                __syncthreads();
            }
            if (tm.n > 1) team_sync(tm);                 // next wavefront reads these lines' new values
        }
        cta_apply_bc(g, u, tm);
    }
}

// grid = ONE thread-block cluster (SEMI_CLUSTER CTAs; 1 CTA on the host test double)
__global__ void __launch_bounds__(SEMI_CTA) k_semi_ylines(SemiGeom g, double* u,
                                                           const double* f, int iterations, int LB) {
    CFDNN_DYN_SMEM(double, sm_y);
    cta_y_lines(g, u, f, iterations, sm_y, LB, Team{(int)blockIdx.x, (int)gridDim.x});
}

// ------------------------------------------------------------------ x-z red-black Gauss-Seidel
// smooth_xz_plane_rbgs (:1247-1418), CPU branch: colour (i+k)%2 over ghosted indices; inside one
// colour the loop runs k, j, i ascending IN PLACE, so u(i,j-1,k) (same colour) is already new:
// one thread per (i,k) column marches j upward carrying the new value.
__global__ void __launch_bounds__(128)
k_semi_rbgs(SemiGeom g, double* __restrict__ u, const double* __restrict__ f, const double* __restrict__ dg,
            const double* __restrict__ rdg, int colour) {
    const int half = (g.Nx + 1) / 2;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= half * g.Nz) return;
    const int k = 1 + t / half;
    const int first = 1 + ((colour + 1 + k) & 1);        // smallest i >= 1 with (i + k) % 2 == colour
    const int i = first + 2 * (t % half);
    if (i > g.Nx) return;
    const size_t S = (size_t)g.S, P = (size_t)g.P;
    size_t idx = sidx(g, i, 1, k);
    double us = u[idx - S];                               // ghost row below (value of the last apply_bc)
    for (int j = 1; j <= g.Ny; ++j, idx += S) {
        const double lap_x = (u[idx + 1] + u[idx - 1]) / g.dx2;
        const double lap_z = (u[idx + P] + u[idx - P]) / g.dz2;
        const double lap_y = g.aS[j] * us + g.aN[j] * u[idx + S];
        us = fdiv(lap_x + lap_y + lap_z - f[idx], dg[j], rdg[j]);
        u[idx] = us;
    }
}

// ------------------------------------------------------------------ residual (nonuniform y)
// compute_residual (:1797-1813), compute_residual_and_norms (:1833-1970)
template <bool NORMS>
__global__ void __launch_bounds__(256)
k_semi_residual(SemiGeom g, const double* __restrict__ u, const double* __restrict__ f, double* __restrict__ r,
                double* __restrict__ part_max, double* __restrict__ part_sq) {
    __shared__ double sh[32];
    const int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
    const int j = blockIdx.y + 1, k = blockIdx.z + 1;
    double mx = 0.0, sq = 0.0;
    if (i <= g.Nx) {
        const size_t idx = sidx(g, i, j, k), S = (size_t)g.S, P = (size_t)g.P;
        const double uc = u[idx];
        const double lap_x = (u[idx + 1] - 2.0 * uc + u[idx - 1]) / g.dx2;
        const double lap_y = g.aS[j] * u[idx - S] + g.aP[j] * uc + g.aN[j] * u[idx + S];
        const double lap_z = (u[idx + P] - 2.0 * uc + u[idx - P]) / g.dz2;
        const double rv = f[idx] - (lap_x + lap_y + lap_z);
        r[idx] = rv;
        mx = fabs(rv); sq = rv * rv;
    }
    if (NORMS) {
        const double bm = block_max(mx, sh);
        const double bs = block_sum(sq, sh);
        if (threadIdx.x == 0) {
            const size_t b = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
            part_max[b] = bm; part_sq[b] = bs;
        }
    }
}

// per-block max|a| and sum a^2 over the interior (||b|| norms, :3660-3682); sum over the interior
__global__ void __launch_bounds__(256)
k_semi_norms(SemiGeom g, const double* __restrict__ a, double* __restrict__ part_max, double* __restrict__ part_sq,
             double* __restrict__ part_sum) {
    __shared__ double sh[32];
    const int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
    const int j = blockIdx.y + 1, k = blockIdx.z + 1;
    double v = 0.0;
    if (i <= g.Nx) v = a[sidx(g, i, j, k)];
    const size_t b = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    if (part_max) { const double bm = block_max(fabs(v), sh); if (threadIdx.x == 0) part_max[b] = bm; }
    if (part_sq) { const double bs = block_sum(v * v, sh); if (threadIdx.x == 0) part_sq[b] = bs; }
    if (part_sum) { const double bs = block_sum(v, sh); if (threadIdx.x == 0) part_sum[b] = bs; }
}

// fixed-order final reductions (one CTA); out[0] = result, out[1] = result / count when count > 0
__global__ void __launch_bounds__(1024) k_semi_final_sum(const double* __restrict__ part, int n, double* out, double count) {
    __shared__ double sh[32];
    double s = 0.0;
    for (int t = threadIdx.x; t < n; t += blockDim.x) s += part[t];
    s = block_sum(s, sh);
    if (threadIdx.x == 0) { out[0] = s; if (count > 0.0) out[1] = s / count; }
}
__global__ void __launch_bounds__(1024) k_semi_final_max(const double* __restrict__ part, int n, double* out) {
    __shared__ double sh[32];
    double s = 0.0;
    for (int t = threadIdx.x; t < n; t += blockDim.x) s = fmax(s, part[t]);
    s = block_max(s, sh);
    if (threadIdx.x == 0) out[0] = s;
}
// Interior sum in the reference's order (i fastest, one running double): make_rhs_mean_free
// (:3479-3548) and subtract_mean (:3411-3477) feed the chaotic coarse solve, see cta_dot.
// out[0] = sum, out[1] = sum / count.
constexpr int SEQ_CHUNK = 4096;
__global__ void __launch_bounds__(256) k_semi_seq_sum(SemiGeom g, const double* __restrict__ a, double* out, double count) {
    __shared__ double buf[SEQ_CHUNK];
    const int n = g.Nx * g.Ny * g.Nz;
    double acc = 0.0;
    for (int base = 0; base < n; base += SEQ_CHUNK) {
        const int m = min(SEQ_CHUNK, n - base);
        __syncthreads();
        for (int t = threadIdx.x; t < m; t += blockDim.x) {
            const int q = base + t;
            const int i = q % g.Nx, j = (q / g.Nx) % g.Ny, k = q / (g.Nx * g.Ny);
            buf[t] = a[sidx(g, i + 1, j + 1, k + 1)];
        }
        __syncthreads();
        if (threadIdx.x == 0)
            for (int t = 0; t < m; ++t) acc += buf[t];
    }
    if (threadIdx.x == 0) { out[0] = acc; out[1] = acc / count; }
}
__global__ void __launch_bounds__(256) k_semi_sub(SemiGeom g, double* __restrict__ a, const double* __restrict__ mean) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
    const int j = blockIdx.y + 1, k = blockIdx.z + 1;
    if (i <= g.Nx) a[sidx(g, i, j, k)] -= *mean;
}

// ------------------------------------------------------------------ transfers
// restrict_residual_xz (:2128-2150): 9-point x-z full weighting at the same j
__global__ void __launch_bounds__(256)
k_semi_restrict(SemiGeom gf, SemiGeom gc, const double* __restrict__ r, double* __restrict__ fc) {
    const int ic = blockIdx.x * blockDim.x + threadIdx.x + 1;
    const int j = blockIdx.y + 1, kc = blockIdx.z + 1;
    if (ic > gc.Nx) return;
    const int i_f = 1 + 2 * (ic - 1), k_f = 1 + 2 * (kc - 1);
    const size_t idx = sidx(gf, i_f, j, k_f), P = (size_t)gf.P;
    double sum = 0.0;
    sum += 0.25 * r[idx];
    sum += 0.125 * (r[idx - 1] + r[idx + 1] + r[idx - P] + r[idx + P]);
    sum += 0.0625 * (r[idx - 1 - P] + r[idx + 1 - P] + r[idx - 1 + P] + r[idx + 1 + P]);
    fc[sidx(gc, ic, j, kc)] = sum;
}

// prolongate_correction_xz (:2326-2358): one-sided bilinear in x-z (toward i_c+1 / k_c+1), alpha = 1
__global__ void __launch_bounds__(256)
k_semi_prolong(SemiGeom gf, SemiGeom gc, const double* __restrict__ uc, double* __restrict__ uf) {
    const int i_f = blockIdx.x * blockDim.x + threadIdx.x + 1;
    const int j = blockIdx.y + 1, k_f = blockIdx.z + 1;
    if (i_f > gf.Nx) return;
    const int ic = (i_f - 1) / 2 + 1, kc = (k_f - 1) / 2 + 1;
    const int di = (i_f - 1) & 1, dk = (k_f - 1) & 1;
    const double wx1 = 0.5 * di, wx0 = 1.0 - wx1, wz1 = 0.5 * dk, wz0 = 1.0 - wz1;
    const size_t c = sidx(gc, ic, j, kc), P = (size_t)gc.P;
    const double correction = wx0 * wz0 * uc[c] + wx1 * wz0 * uc[c + 1] + wx0 * wz1 * uc[c + P] + wx1 * wz1 * uc[c + 1 + P];
    const double alpha = 1.0;
    uf[sidx(gf, i_f, j, k_f)] += alpha * correction;
}

// solve(): interior of rhs / p into f / u (:3558-3574); all cells of u back into p (:3769-3783).
// The caller's arrays and the level-0 arrays share the ghosted layout.
__global__ void __launch_bounds__(256)
k_semi_copy_in(SemiGeom g, const double* __restrict__ rhs, const double* __restrict__ p, double* __restrict__ f,
               double* __restrict__ u) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
    const int j = blockIdx.y + 1, k = blockIdx.z + 1;
    if (i > g.Nx) return;
    const size_t idx = sidx(g, i, j, k);
    f[idx] = rhs[idx];
    u[idx] = p[idx];
}

// ------------------------------------------------------------------ PCG on the coarsest level
// solve_coarse_pcg (:2490-3119): conjugate gradients preconditioned by 2 y-line passes, mean
// projection when the operator is singular, convergence looked at every 4th iteration, restart on
// |pAp| < 1e-30 — all of it inside ONE single-CTA launch (the coarsest level is a few thousand
// cells; the reference's host-side decisions become block-uniform branches).
__device__ __forceinline__ double cta_bcast(double v, double* slot) {
    __syncthreads();
    if (threadIdx.x == 0) *slot = v;
    __syncthreads();
    return *slot;
}
__device__ __forceinline__ size_t interior_idx(const SemiGeom& g, int t) {
    const int i = t % g.Nx, j = (t / g.Nx) % g.Ny, k = t / (g.Nx * g.Ny);
    return sidx(g, i + 1, j + 1, k + 1);
}
// y = A x on the interior (apply_laplacian_coarse :2455-2470); with f != nullptr: y = f - A x (residual)
__device__ void cta_laplacian(const SemiGeom& g, const double* x, const double* f,
                              double* y) {
    const int n = g.Nx * g.Ny * g.Nz;
    const size_t S = (size_t)g.S, P = (size_t)g.P;
    for (int t = threadIdx.x; t < n; t += blockDim.x) {
        const int j = (t / g.Nx) % g.Ny + 1;
        const size_t idx = interior_idx(g, t);
        const double xc = x[idx];
        const double lap_x = (x[idx + 1] - 2.0 * xc + x[idx - 1]) / g.dx2;
        const double lap_y = g.aS[j] * x[idx - S] + g.aP[j] * xc + g.aN[j] * x[idx + S];
        const double lap_z = (x[idx + P] - 2.0 * xc + x[idx - P]) / g.dz2;
        const double lap = lap_x + lap_y + lap_z;
        y[idx] = f ? f[idx] - lap : lap;
    }
    __syncthreads();
}
// Interior sum of a*b (b == nullptr: of a) in the reference's own order — k, j, i ascending, one
// running double — so alpha, beta, the projections and the convergence test of the PCG see the
// same bits as the CPU loops (:2611-2635, :2680-2760).  The coarse solve runs CG well past its
// attainable accuracy, where a last-bit change of a dot product changes the iterate at 1e-6
// (measured on the reference itself), so a tree reduction is not good enough here.  Products are
// staged in shared memory by all threads, thread 0 adds them in order.
__device__ double cta_dot(const SemiGeom& g, const double* a, const double* b, double* buf, int cap, double* slot) {
    const int n = g.Nx * g.Ny * g.Nz;
    double acc = 0.0;
    for (int base = 0; base < n; base += cap) {
        const int m = min(cap, n - base);
        __syncthreads();
        for (int t = threadIdx.x; t < m; t += blockDim.x) {
            const size_t idx = interior_idx(g, base + t);
            buf[t] = b ? a[idx] * b[idx] : a[idx];
        }
        __syncthreads();
        if (threadIdx.x == 0)
            for (int t = 0; t < m; ++t) acc += buf[t];
    }
    return cta_bcast(acc, slot);
}
// z = M^-1 r : f <- r, u <- 0 (unless keep_u_f), 2 y-line passes, mean projection (:2640-2672)
__device__ void cta_precondition(const SemiGeom& g, double* u, double* f, const double* r, bool reset, bool nullspace,
                                 double* smy, int LB, double* buf, int cap, double* slot) {
    if (reset) {
        for (long long t = threadIdx.x; t < g.total; t += blockDim.x) { f[t] = r[t]; u[t] = 0.0; }
    }
    __syncthreads();
    cta_y_lines(g, u, f, 2, smy, LB);
    if (nullspace) {
        const int n = g.Nx * g.Ny * g.Nz;
        const double z_sum = cta_dot(g, u, nullptr, buf, cap, slot);
        const double z_mean = z_sum / n;
        for (int t = threadIdx.x; t < n; t += blockDim.x) u[interior_idx(g, t)] -= z_mean;
    }
    __syncthreads();
}

__global__ void __launch_bounds__(SEMI_CTA)
k_semi_pcg(SemiGeom g, double* u, double* f, double* r, double* p,
           double* z, double* Ap, double* x, double* f_save,
           int max_iters, double tol, int nullspace, int LB, int cap) {
    CFDNN_DYN_SMEM(double, sm_y);
    __shared__ double slot;
    double* sh = sm_y + (size_t)4 * g.Ny + (size_t)LB * g.Ny;     // dot-product staging, cap doubles
    const long long total = g.total;
    const int n = g.Nx * g.Ny * g.Nz;
    for (long long t = threadIdx.x; t < total; t += blockDim.x) { f_save[t] = f[t]; x[t] = u[t]; }
    cta_apply_bc(g, u);
    cta_laplacian(g, u, f, r);
    const double r_norm_sq_init = cta_dot(g, r, r, sh, cap, &slot);
    cta_precondition(g, u, f, r, true, nullspace != 0, sm_y, LB, sh, cap, &slot);
    for (long long t = threadIdx.x; t < total; t += blockDim.x) { z[t] = u[t]; p[t] = u[t]; }
    __syncthreads();
    double rho = cta_dot(g, r, z, sh, cap, &slot);

    for (int iter = 0; iter < max_iters; ++iter) {
        for (long long t = threadIdx.x; t < total; t += blockDim.x) u[t] = p[t];
        cta_apply_bc(g, u);
        cta_laplacian(g, u, nullptr, Ap);
        const double pAp = cta_dot(g, u, Ap, sh, cap, &slot);
        if (fabs(pAp) < 1e-30) {                                 // restart with a fresh direction (:2700-2722)
            for (long long t = threadIdx.x; t < total; t += blockDim.x) u[t] = r[t];
            cta_apply_bc(g, u);
            cta_precondition(g, u, f, r, false, nullspace != 0, sm_y, LB, sh, cap, &slot);
            for (long long t = threadIdx.x; t < total; t += blockDim.x) { z[t] = u[t]; p[t] = u[t]; }
            __syncthreads();
            rho = cta_dot(g, r, z, sh, cap, &slot);
            continue;
        }
        double alpha = rho / pAp;
        if (fabs(alpha) > 1e10) alpha = (alpha > 0) ? 1e10 : -1e10;
        for (int t = threadIdx.x; t < n; t += blockDim.x) {
            const size_t idx = interior_idx(g, t);
            x[idx] += alpha * u[idx];
            r[idx] -= alpha * Ap[idx];
        }
        __syncthreads();
        if (iter % 4 == 3 || iter == max_iters - 1) {
            const double r_norm_sq = cta_dot(g, r, r, sh, cap, &slot);
            if (r_norm_sq < tol * tol * r_norm_sq_init) break;
        }
        cta_precondition(g, u, f, r, true, nullspace != 0, sm_y, LB, sh, cap, &slot);
        for (long long t = threadIdx.x; t < total; t += blockDim.x) z[t] = u[t];
        __syncthreads();
        const double rho_new = cta_dot(g, r, z, sh, cap, &slot);
        const double beta = (fabs(rho) > 1e-30) ? (rho_new / rho) : 0.0;
        rho = rho_new;
        for (long long t = threadIdx.x; t < total; t += blockDim.x) p[t] = z[t] + beta * p[t];
        __syncthreads();
    }
    __syncthreads();
    for (long long t = threadIdx.x; t < total; t += blockDim.x) { u[t] = x[t]; f[t] = f_save[t]; }
}

}  // namespace

// ---------------------------------------------------------------- construction
// create_hierarchy, semi-coarsening branch (:181-207): halve Nx and Nz while both exceed 8, keep Ny
MultigridSemi::MultigridSemi(const cfdnn_mesh_desc& m) : mesh_(m) {
    require_device();
    if (m.Nz <= 1) throw std::invalid_argument("MultigridSemi: 3-D meshes only");
    if (m.Ng != 1) throw std::invalid_argument("MultigridSemi: only Nghost == 1 is supported");
    if (!m.yLap_aS || !m.yLap_aN || !m.yLap_aP) throw std::invalid_argument("MultigridSemi: stretched mesh needs yLap_aS/aN/aP");
    if (m.Nz_global > 0 && m.Nz_global != m.Nz)
        throw std::invalid_argument("MultigridSemi: z-slab decomposition of the y-stretched solver is not built yet");
    const int ny2 = m.Ny + 2;
    h_aS_.assign(m.yLap_aS, m.yLap_aS + ny2);
    h_aN_.assign(m.yLap_aN, m.yLap_aN + ny2);
    aS_.alloc(ny2, false); aN_.alloc(ny2, false); aP_.alloc(ny2, false);
    CFDNN_CUDA(cudaMemcpy(aS_.p, m.yLap_aS, ny2 * sizeof(double), cudaMemcpyHostToDevice));
    CFDNN_CUDA(cudaMemcpy(aN_.p, m.yLap_aN, ny2 * sizeof(double), cudaMemcpyHostToDevice));
    CFDNN_CUDA(cudaMemcpy(aP_.p, m.yLap_aP, ny2 * sizeof(double), cudaMemcpyHostToDevice));

    int Nx = m.Nx, Nz = m.Nz;
    double dx = m.dx, dz = m.dz;
    auto add = [&](int nx, int nz, double hx, double hz) {
        auto L = std::make_unique<SemiLevel>();
        SemiGeom& g = L->g;
        g.Nx = nx; g.Ny = m.Ny; g.Nz = nz;
        g.S = nx + 2; g.P = (long long)g.S * (m.Ny + 2); g.total = g.P * (nz + 2);
        L->dx = hx; L->dz = hz;
        const double dx2 = hx * hx, dz2 = hz * hz;
        g.dx2 = make_div_host(dx2); g.dz2 = make_div_host(dz2);
        g.inv_dx2 = 1.0 / dx2; g.inv_dz2 = 1.0 / dz2;
        g.two_inv_dx2 = 2.0 / dx2; g.two_inv_dz2 = 2.0 / dz2;
        g.aS = aS_.p; g.aN = aN_.p; g.aP = aP_.p;
        L->u.alloc(g.total); L->f.alloc(g.total); L->r.alloc(g.total);
        L->ta.alloc(m.Ny, false); L->tcp.alloc(m.Ny, false); L->tden.alloc(m.Ny, false); L->trden.alloc(m.Ny, false);
        L->rbd.alloc(ny2, false); L->rbrd.alloc(ny2, false);
        g.ta = L->ta.p; g.tcp = L->tcp.p; g.tden = L->tden.p; g.trden = L->trden.p;
        levels_.push_back(std::move(L));
    };
    add(Nx, Nz, dx, dz);
    constexpr int MIN_COARSE = 8;
    while (Nx > MIN_COARSE && Nz > MIN_COARSE) {
        if ((Nx | Nz) & 1) throw std::invalid_argument("MultigridSemi: Nx and Nz must stay even while coarsening");
        Nx /= 2; Nz /= 2; dx *= 2.0; dz *= 2.0;
        add(Nx, Nz, dx, dz);
    }
    const SemiGeom& g0 = levels_[0]->g;
    partial_.alloc(3 * ((size_t)((g0.Nx + 255) / 256) * g0.Ny * g0.Nz) + 16);
    scalars_.alloc(16);
    CFDNN_CUDA(cudaMallocHost(&h_scalars_, 16 * sizeof(double)));
    const SemiGeom& gc = levels_.back()->g;
    pcg_p_.alloc(gc.total); pcg_z_.alloc(gc.total); pcg_Ap_.alloc(gc.total); pcg_x_.alloc(gc.total); pcg_fsave_.alloc(gc.total);
    CFDNN_CUDA(cudaFuncSetAttribute(k_semi_ylines, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CFDNN_CUDA(cudaFuncSetAttribute(k_semi_pcg, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    set_bc(bc_);
}

MultigridSemi::~MultigridSemi() {
    if (h_scalars_) cudaFreeHost(h_scalars_);
}

void MultigridSemi::set_bc(const int bc[6]) {
    int tmp[6];
    for (int d = 0; d < 6; ++d) {
        if (bc[d] < 0 || bc[d] > 2) throw std::invalid_argument("set_bc: bad PoissonBC value");
        tmp[d] = bc[d];
    }
    for (int d = 0; d < 6; ++d) bc_[d] = tmp[d];
    for (auto& L : levels_) for (int d = 0; d < 6; ++d) L->g.bc[d] = bc_[d];
    factor_lines();
}

bool MultigridSemi::has_nullspace() const {   // :3393-3402
    for (int d = 0; d < 6; ++d) if (bc_[d] == CFDNN_PBC_DIRICHLET) return false;
    return true;
}

// The Thomas recurrences for c' and the pivots (:1675-1690) depend on the level and the y BCs only,
// not on u: factor once per level (same arithmetic, same order as the reference's per-line loop).
void MultigridSemi::factor_lines() {
    const int Ny = mesh_.Ny;
    const bool neumann_lo = bc_[2] == CFDNN_PBC_NEUMANN, neumann_hi = bc_[3] == CFDNN_PBC_NEUMANN;
    constexpr double PIVOT_EPS = 1e-14;
    std::vector<double> a(Ny), cp(Ny), den(Ny), rden(Ny), rbd(Ny + 2, 1.0), rbrd(Ny + 2, 1.0);
    for (auto& L : levels_) {
        const double dx2 = L->dx * L->dx, dz2 = L->dz * L->dz;
        const double inv_dx2 = 1.0 / dx2, inv_dz2 = 1.0 / dz2;
        for (int jl = 0; jl < Ny; ++jl) {
            const int jm = jl + 1;
            const bool south = jl == 0, north = jl == Ny - 1;
            const double aS = h_aS_[jm], aN = h_aN_[jm];
            const double av = south ? 0.0 : -aS;
            const double cv = north ? 0.0 : -aN;
            const double b = ((south && neumann_lo) ? 0.0 : aS) + ((north && neumann_hi) ? 0.0 : aN)
                           + 2.0 * inv_dx2 + 2.0 * inv_dz2;
            double dsafe;
            if (jl == 0) {
                dsafe = (std::abs(b) > PIVOT_EPS) ? b : PIVOT_EPS;
            } else {
                const double denom = b - av * cp[jl - 1];
                dsafe = (std::abs(denom) > PIVOT_EPS) ? denom : (denom >= 0 ? PIVOT_EPS : -PIVOT_EPS);
            }
            a[jl] = av; cp[jl] = cv / dsafe; den[jl] = dsafe; rden[jl] = 1.0 / dsafe;
            rbd[jm] = 2.0 / dx2 + (aS + aN) + 2.0 / dz2;      // RBGS diagonal (:1380)
            rbrd[jm] = 1.0 / rbd[jm];
        }
        CFDNN_CUDA(cudaMemcpy(L->ta.p, a.data(), Ny * sizeof(double), cudaMemcpyHostToDevice));
        CFDNN_CUDA(cudaMemcpy(L->tcp.p, cp.data(), Ny * sizeof(double), cudaMemcpyHostToDevice));
        CFDNN_CUDA(cudaMemcpy(L->tden.p, den.data(), Ny * sizeof(double), cudaMemcpyHostToDevice));
        CFDNN_CUDA(cudaMemcpy(L->trden.p, rden.data(), Ny * sizeof(double), cudaMemcpyHostToDevice));
        CFDNN_CUDA(cudaMemcpy(L->rbd.p, rbd.data(), (Ny + 2) * sizeof(double), cudaMemcpyHostToDevice));
        CFDNN_CUDA(cudaMemcpy(L->rbrd.p, rbrd.data(), (Ny + 2) * sizeof(double), cudaMemcpyHostToDevice));
    }
}

// ---------------------------------------------------------------- launchers
static inline dim3 cell_grid(const SemiGeom& g) { return dim3((g.Nx + 255) / 256, g.Ny, g.Nz); }

void MultigridSemi::apply_bc(int level, double* a, bool res) {
    const SemiGeom& g = levels_[level]->g;
    const int nx = (g.Ny + 2) * (g.Nz + 2), ny = (g.Nx + 2) * (g.Nz + 2), nz = (g.Nx + 2) * (g.Ny + 2);
    CFDNN_LAUNCH(k_semi_bc<0>, (nx + 255) / 256, 256, 0, st_, g, a, res ? 1 : 0);
    CFDNN_LAUNCH(k_semi_bc<1>, (ny + 255) / 256, 256, 0, st_, g, a, res ? 1 : 0);
    CFDNN_LAUNCH(k_semi_bc<2>, (nz + 255) / 256, 256, 0, st_, g, a, res ? 1 : 0);
}

// lines per batch of the y-line workspace and the dynamic shared memory it needs
static void yline_smem(const SemiGeom& g, int& LB, size_t& bytes) {
    const int maxl = std::min(g.Nx, g.Nz);
    const size_t budget = 190 * 1024 - (size_t)4 * g.Ny * sizeof(double);
    long long lb = (long long)(budget / ((size_t)g.Ny * sizeof(double)));
    if (lb < 1) throw std::invalid_argument("MultigridSemi: Ny too large for the y-line workspace");
    if (lb >= 32) lb = lb / 32 * 32;
    LB = (int)std::min<long long>(lb, std::min(maxl, SEMI_CTA));
    bytes = ((size_t)4 * g.Ny + (size_t)LB * g.Ny) * sizeof(double);
}

void MultigridSemi::smooth_y_lines(int level, int iterations) {
    SemiLevel& L = *levels_[level];
    int LB; size_t bytes;
    yline_smem(L.g, LB, bytes);
#ifndef CFDNN_DEVSIM
    // one cluster of SEMI_CLUSTER CTAs: the lines of a wavefront are dealt round-robin to the CTAs,
    // a cluster barrier separates wavefronts
    cudaLaunchConfig_t lc{};
    lc.gridDim = dim3(SEMI_CLUSTER); lc.blockDim = dim3(SEMI_CTA); lc.dynamicSmemBytes = bytes; lc.stream = st_;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = SEMI_CLUSTER; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    lc.attrs = at; lc.numAttrs = 1;
    CFDNN_CUDA(cudaLaunchKernelEx(&lc, k_semi_ylines, L.g, L.u.p, (const double*)L.f.p, iterations, LB));
    g_launches.fetch_add(1, std::memory_order_relaxed);
#else
    CFDNN_LAUNCH_COOP(k_semi_ylines, 1, SEMI_CTA, bytes, st_, L.g, L.u.p, L.f.p, iterations, LB);
#endif
}

void MultigridSemi::smooth_xz_rbgs(int level, int iterations) {
    SemiLevel& L = *levels_[level];
    const int n = ((L.g.Nx + 1) / 2) * L.g.Nz;
    for (int it = 0; it < iterations; ++it) {
        CFDNN_LAUNCH(k_semi_rbgs, (n + 127) / 128, 128, 0, st_, L.g, L.u.p, L.f.p, L.rbd.p, L.rbrd.p, 0);
        CFDNN_LAUNCH(k_semi_rbgs, (n + 127) / 128, 128, 0, st_, L.g, L.u.p, L.f.p, L.rbd.p, L.rbrd.p, 1);
        apply_bc(level, L.u.p);
    }
}

// do_smooth of vcycle (:3143-3172), 3-D semi-coarsening branch
void MultigridSemi::do_smooth(int level, int nu) {
    const bool strong = has_nullspace();
    for (int pass = 0; pass < nu; ++pass) {
        smooth_y_lines(level, strong ? 4 : 2);
        smooth_xz_rbgs(level, (level == 0 && strong) ? 4 : 2);
    }
}

void MultigridSemi::residual(int level) {
    SemiLevel& L = *levels_[level];
    CFDNN_LAUNCH_COOP((k_semi_residual<false>), cell_grid(L.g), 256, 0, st_, L.g, L.u.p, L.f.p, L.r.p, nullptr, nullptr);
}

void MultigridSemi::residual_norms(int level, double* d_max, double* d_sumsq) {
    SemiLevel& L = *levels_[level];
    const dim3 gr = cell_grid(L.g);
    const int nb = gr.x * gr.y * gr.z;
    CFDNN_LAUNCH_COOP((k_semi_residual<true>), gr, 256, 0, st_, L.g, L.u.p, L.f.p, L.r.p, partial_.p, partial_.p + nb);
    CFDNN_LAUNCH_COOP(k_semi_final_max, 1, 1024, 0, st_, partial_.p, nb, d_max);
    CFDNN_LAUNCH_COOP(k_semi_final_sum, 1, 1024, 0, st_, partial_.p + nb, nb, d_sumsq, 0.0);
}

void MultigridSemi::mean_free(int level, double* field) {   // make_rhs_mean_free :3479-3548, subtract_mean :3411-3477
    SemiLevel& L = *levels_[level];
    const dim3 gr = cell_grid(L.g);
    CFDNN_LAUNCH_COOP(k_semi_seq_sum, 1, 256, 0, st_, L.g, field, scalars_.p + 8, (double)(L.g.Nx * L.g.Ny * L.g.Nz));
    CFDNN_LAUNCH(k_semi_sub, gr, 256, 0, st_, L.g, field, scalars_.p + 9);
}

void MultigridSemi::restrict_xz(int fine) {
    SemiLevel& F = *levels_[fine];
    SemiLevel& C = *levels_[fine + 1];
    CFDNN_LAUNCH(k_semi_restrict, cell_grid(C.g), 256, 0, st_, F.g, C.g, F.r.p, C.f.p);
}

void MultigridSemi::prolong_xz(int coarse) {
    SemiLevel& C = *levels_[coarse];
    SemiLevel& F = *levels_[coarse - 1];
    CFDNN_LAUNCH(k_semi_prolong, cell_grid(F.g), 256, 0, st_, F.g, C.g, C.u.p, F.u.p);
}

void MultigridSemi::solve_coarse_pcg(int max_iters, double tol) {
    SemiLevel& L = *levels_.back();
    int LB; size_t bytes;
    yline_smem(L.g, LB, bytes);
    const int cap = 4096;                                     // dot-product staging buffer (doubles)
    CFDNN_LAUNCH_COOP(k_semi_pcg, 1, SEMI_CTA, bytes + cap * sizeof(double), st_, L.g, L.u.p, L.f.p, L.r.p, pcg_p_.p, pcg_z_.p,
                      pcg_Ap_.p, pcg_x_.p, pcg_fsave_.p, max_iters, tol, has_nullspace() ? 1 : 0, LB, cap);
}

// ---------------------------------------------------------------- V-cycle (:3121-3322, semi-coarsening 3-D branches)
void MultigridSemi::vcycle(int level, int nu1, int nu2, int degree) {
    const int nl = (int)levels_.size();
    const bool nullsp = has_nullspace();
    if (level == nl - 1) {                                    // :3176-3186
        solve_coarse_pcg(nullsp ? 50 : 30, 1e-10);
        return;
    }
    auto zero_coarse = [&]() {
        SemiLevel& C = *levels_[level + 1];
        CFDNN_CUDA(cudaMemsetAsync(C.u.p, 0, (size_t)C.g.total * sizeof(double), st_));
    };
    SemiLevel& L = *levels_[level];
    if (level == nl - 2 && nullsp) {                          // K-cycle-lite: two coarse solves (:3203-3255)
        for (int rep = 0; rep < 2; ++rep) {
            if (rep == 0) do_smooth(level, nu1);
            residual(level);
            apply_bc(level, L.r.p, true);
            restrict_xz(level);
            zero_coarse();
            solve_coarse_pcg(50, 1e-10);
            prolong_xz(level + 1);
            apply_bc(level, L.u.p);
            do_smooth(level, nu2);
        }
        return;
    }
    do_smooth(level, nu1);
    residual(level);
    apply_bc(level, L.r.p, true);
    restrict_xz(level);
    if (nullsp) mean_free(level + 1, levels_[level + 1]->f.p);
    zero_coarse();
    vcycle(level + 1, nu1, nu2, degree);
    prolong_xz(level + 1);
    apply_bc(level, L.u.p);
    if (nu2 > 0) do_smooth(level, nu2);
}

// ---------------------------------------------------------------- solve (:3550-3786)
int MultigridSemi::solve_dev(const double* rhs, double* p, const cfdnn_poisson_cfg& cfg, cudaStream_t st) {
    st_ = st;
    SemiLevel& L0 = *levels_[0];
    const SemiGeom& g = L0.g;
    CFDNN_LAUNCH(k_semi_copy_in, cell_grid(g), 256, 0, st_, g, rhs, p, L0.f.p, L0.u.p);
    apply_bc(0, L0.u.p);
    stats = cfdnn_poisson_stats{};
    const int degree = cfg.chebyshev_degree;
    int cycles_used = 0;
    if (cfg.fixed_cycles > 0) {
        const int nu1 = cfg.nu1 > 0 ? cfg.nu1 : 3;
        const int nu2 = cfg.nu2 > 0 ? cfg.nu2 : 1;
        for (int c = 0; c < cfg.fixed_cycles; ++c) vcycle(0, nu1, nu2, degree);
        cycles_used = cfg.fixed_cycles;
        stats.fixed_cycle_mode = 1;
        stats.converged = 0;
        stats.status = CFDNN_PSTAT_FIXED;
    } else {
        if (cfg.max_vcycles <= 0) throw std::invalid_argument("PoissonConfig.max_vcycles must be positive");
        const int max_cycles = cfg.max_vcycles;
        const bool accurate = max_cycles > 5;                                  // :3647-3650
        const int nu1 = cfg.nu1 > 0 ? cfg.nu1 : (accurate ? 3 : 2);
        const int nu2 = cfg.nu2 > 0 ? cfg.nu2 : 1;
        const int check_interval = std::max(1, std::max(4, cfg.check_interval));   // :3653
        double* S = scalars_.p;
        {
            const dim3 gr = cell_grid(g);
            const int nb = gr.x * gr.y * gr.z;
            CFDNN_LAUNCH_COOP(k_semi_norms, gr, 256, 0, st_, g, L0.f.p, partial_.p, partial_.p + nb, nullptr);
            CFDNN_LAUNCH_COOP(k_semi_final_max, 1, 1024, 0, st_, partial_.p, nb, S + 0);
            CFDNN_LAUNCH_COOP(k_semi_final_sum, 1, 1024, 0, st_, partial_.p + nb, nb, S + 1, 0.0);
            residual_norms(0, S + 2, S + 3);
        }
        CFDNN_CUDA(cudaMemcpyAsync(h_scalars_, S, 4 * sizeof(double), cudaMemcpyDeviceToHost, st_));
        CFDNN_CUDA(cudaStreamSynchronize(st_));
        stats.b_inf = h_scalars_[0];
        stats.b_l2 = std::sqrt(h_scalars_[1]);
        stats.r0 = h_scalars_[2];
        stats.r0_l2 = std::sqrt(h_scalars_[3]);
        stats.residual = stats.r0;
        stats.residual_l2 = stats.r0_l2;
        for (int cycle = 0; cycle < max_cycles; ++cycle) {
            vcycle(0, nu1, nu2, degree);
            cycles_used = cycle + 1;
            if ((cycle % check_interval) == (check_interval - 1) || cycle == max_cycles - 1) {
                residual_norms(0, S + 2, S + 3);
                CFDNN_CUDA(cudaMemcpyAsync(h_scalars_ + 2, S + 2, 2 * sizeof(double), cudaMemcpyDeviceToHost, st_));
                CFDNN_CUDA(cudaStreamSynchronize(st_));
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
        }
        stats.status = stats.converged ? CFDNN_PSTAT_CONVERGED : CFDNN_PSTAT_HIT_MAX;
    }
    stats.cycles = cycles_used;
    if (has_nullspace()) {                                     // fix_nullspace(0) :3404-3409
        mean_free(0, L0.u.p);
        apply_bc(0, L0.u.p);
    }
    CFDNN_CUDA(cudaMemcpyAsync(p, L0.u.p, (size_t)g.total * sizeof(double), cudaMemcpyDeviceToDevice, st_));   // all cells (:3769-3783)
    return cycles_used;
}

}  // namespace cfdnn
