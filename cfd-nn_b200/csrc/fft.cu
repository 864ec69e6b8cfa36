// fft.cu — FFT-hybrid Poisson solver for channels: periodic x/z, Neumann walls in y, uniform or stretched y.
//
// Reference: FFTPoissonSolver, src/poisson_solver_fft.cpp (GPU build only: cuFFT 2-D plans over a packed
// [i][k][j] copy + cuSPARSE gtsv2StridedBatch).  Same algorithm (:960-1098) — volume-weighted mean subtraction,
// forward 2-D FFT over (x, z), one tridiagonal system in y per mode with the (0,0) mode pinned at j = 0, inverse
// FFT, 1/(Nx Nz), ghost fill — but designed for this backend's layout instead of a library's:
//   * no pack/unpack passes: the x transform reads the solver's ghosted rows and the inverse x transform writes
//     them (x ghosts included);
//   * the half spectrum is taken along x (the contiguous index), two real rows ride one complex transform;
//   * hand-written Stockham autosort FFTs in shared memory (radix 2/3/4/5, conflict-free swizzled [n][line] tiles):
//     x lines are contiguous rows, z lines are gathered as tiles of L consecutive kx so every global access is a
//     run of L complex numbers;
//   * the Thomas solve runs on the [kz][j][kx] spectrum in place, one thread per mode, coalesced in kx; c' depends
//     on the mesh only and is tabulated once.
// Algorithmic traffic per real cell: x fwd 8+8, z fwd 8+8, Thomas 8+8 (+8+4+8: second read, c', write), z inv 8+8,
// x inv 8+8 = 100 B — 6.7 GB at 512x256x512, against ~20 V-cycles of ~1 GB each for the multigrid.
#include "fft.cuh"
#include <cmath>
#include <cstring>

namespace cfdnn {
namespace {

constexpr int FFT_THREADS = 256;
constexpr int FFT_MAX_RAD = 12;

struct Radices {
    int n, r[FFT_MAX_RAD];
};

__device__ __forceinline__ double2 c2(double x, double y) { double2 r; r.x = x; r.y = y; return r; }
__device__ __forceinline__ double2 cadd(double2 a, double2 b) { return c2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ double2 csub(double2 a, double2 b) { return c2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ double2 cmul(double2 a, double2 b) { return c2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
__device__ __forceinline__ double2 cconj(double2 a) { return c2(a.x, -a.y); }

// element n of line l in a tile of L lines (L a power of two): lines are the fast index, rotated by n so that both
// "consecutive lines" (the FFT stages, z gathers) and "consecutive n of one line" (row loads) spread over all banks
__device__ __forceinline__ int sw(int n, int l, int L) { return n * L + ((l + n) & (L - 1)); }

// One Stockham stage of radix R on all L lines: with Ns = product of the radices already done, butterfly j takes
// x[j + r N/R] * w^(r (j mod Ns) N/(Ns R)), r = 0..R-1, through a length-R DFT to y[(j / Ns) Ns R + (j mod Ns) + q Ns].
// tw[n] = exp(-2 pi i n / N).
template <int R>
__device__ __forceinline__ void fft_stage(const double2* __restrict__ A, double2* __restrict__ B, int N, int Ns, int L, int logL,
                                          const double2* __restrict__ tw) {
    const int M = N / R, tstep = N / (Ns * R);
    for (int t = threadIdx.x; t < M * L; t += blockDim.x) {
        const int l = t & (L - 1), j = t >> logL;
        const int k = j % Ns, d0 = (j / Ns) * Ns * R + k;
        double2 v[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            v[r] = A[sw(j + r * M, l, L)];
            if (r > 0 && k > 0) v[r] = cmul(v[r], tw[r * k * tstep]);
        }
        if (R == 2) {
            B[sw(d0, l, L)] = cadd(v[0], v[1]);
            B[sw(d0 + Ns, l, L)] = csub(v[0], v[1]);
        } else if (R == 4) {
            const double2 t0 = cadd(v[0], v[2]), t1 = csub(v[0], v[2]), t2 = cadd(v[1], v[3]), e = csub(v[1], v[3]);
            const double2 t3 = c2(e.y, -e.x);                          // -i (v1 - v3)
            B[sw(d0, l, L)] = cadd(t0, t2);
            B[sw(d0 + Ns, l, L)] = cadd(t1, t3);
            B[sw(d0 + 2 * Ns, l, L)] = csub(t0, t2);
            B[sw(d0 + 3 * Ns, l, L)] = csub(t1, t3);
        } else {                                                       // radix 3 / 5: direct DFT, roots from the table
            const int q0 = N / R;
#pragma unroll
            for (int q = 0; q < R; ++q) {
                double2 acc = v[0];
#pragma unroll
                for (int r = 1; r < R; ++r) acc = cadd(acc, cmul(v[r], tw[((q * r) % R) * q0]));
                B[sw(d0 + q * Ns, l, L)] = acc;
            }
        }
    }
}

// Forward DFTs of all L lines of length N held in A; returns the buffer that holds the result
__device__ double2* fft_lines(double2* A, double2* B, int N, int L, int logL, const Radices& rad,
                              const double2* __restrict__ tw) {
    int Ns = 1;
    for (int s = 0; s < rad.n; ++s) {
        const int R = rad.r[s];
        if (R == 4) fft_stage<4>(A, B, N, Ns, L, logL, tw);
        else if (R == 2) fft_stage<2>(A, B, N, Ns, L, logL, tw);
        else if (R == 3) fft_stage<3>(A, B, N, Ns, L, logL, tw);
        else fft_stage<5>(A, B, N, Ns, L, logL, tw);
        __syncthreads();
        double2* T = A; A = B; B = T;
        Ns *= R;
    }
    return A;
}

// ---------------------------------------------------------------- volume-weighted sum of the right-hand side
// sum f dyv over the interior (kernel_pack_and_partial_sum, src/poisson_solver_fft.cpp:53-89): a CTA strides over
// (k, j) rows, per-block partials, fixed-order final pass (bitwise-reproducible)
__global__ void __launch_bounds__(256)
k_fft_wsum(const double* __restrict__ rhs, const double* __restrict__ dyv, int Nx, int Ny, int Nz, int Ng, int cs, int cp,
           double* __restrict__ partial) {
    __shared__ double red[256];
    double acc = 0.0;
    const int rows = Ny * Nz;
    for (int row = blockIdx.x; row < rows; row += gridDim.x) {
        const int k = row / Ny, j = row - k * Ny;
        const double* src = rhs + (size_t)(k + Ng) * cp + (size_t)(j + Ng) * cs + Ng;
        double a = 0.0;
        for (int i = threadIdx.x; i < Nx; i += blockDim.x) a += src[i];
        acc += a * dyv[j];
    }
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) partial[1 + blockIdx.x] = red[0];
}
__global__ void k_fft_mean(double* __restrict__ partial, int n, double inv_volume) {       // :91-110, :930-938
    double s = 0.0;
    for (int i = 0; i < n; ++i) s += partial[1 + i];
    partial[0] = s * inv_volume;
}

// ---------------------------------------------------------------- x transforms (rows of the ghosted cell layout)
// forward: line l of the CTA carries rows 2l (real part) and 2l+1 (imaginary part); row = k Ny + j
__global__ void __launch_bounds__(FFT_THREADS)
k_fft_x_fwd(const double* __restrict__ rhs, const double* __restrict__ mean_p, double2* __restrict__ spec, int Nx, int Ny,
            int Nz, int Ng, int cs, int cp, int Pxh, int L, int logL, Radices rad, const double2* __restrict__ tw) {
    CFDNN_DYN_SMEM(double2, sm);
    double2* A = sm;
    double2* B = sm + (size_t)Nx * L;
    const int rows = Ny * Nz;
    const double mean = mean_p[0];
    const int line0 = blockIdx.x * L;
    for (int t = threadIdx.x; t < Nx * L; t += blockDim.x) {
        const int l = t / Nx, n = t - l * Nx;
        const int ra = 2 * (line0 + l), rb = ra + 1;
        double a = 0.0, b = 0.0;
        if (ra < rows) {
            const int k = ra / Ny, j = ra - k * Ny;
            a = rhs[(size_t)(k + Ng) * cp + (size_t)(j + Ng) * cs + Ng + n] - mean;
        }
        if (rb < rows) {
            const int k = rb / Ny, j = rb - k * Ny;
            b = rhs[(size_t)(k + Ng) * cp + (size_t)(j + Ng) * cs + Ng + n] - mean;
        }
        A[sw(n, l, L)] = c2(a, b);
    }
    __syncthreads();
    const double2* Z = fft_lines(A, B, Nx, L, logL, rad, tw);
    const int Nxh = Nx / 2 + 1;
    for (int t = threadIdx.x; t < Nxh * L; t += blockDim.x) {
        const int l = t / Nxh, kx = t - l * Nxh;
        const int ra = 2 * (line0 + l), rb = ra + 1;
        const double2 z = Z[sw(kx, l, L)], zc = cconj(Z[sw(kx == 0 ? 0 : Nx - kx, l, L)]);
        const double2 s = cadd(z, zc), d = csub(z, zc);
        if (ra < rows) spec[(size_t)ra * Pxh + kx] = c2(0.5 * s.x, 0.5 * s.y);
        if (rb < rows) spec[(size_t)rb * Pxh + kx] = c2(0.5 * d.y, -0.5 * d.x);      // (z - conj z') / (2 i)
    }
}

// inverse: Z = A + i B on all Nx modes (Hermitian completion of both rows), z = conj(FFT(conj Z)); row a = Re z,
// row b = Im z; scaled by 1 / (Nx Nz) and written with the periodic x ghosts (unpack_and_apply_bc :772-804, x part)
__global__ void __launch_bounds__(FFT_THREADS)
k_fft_x_inv(const double2* __restrict__ spec, double* __restrict__ p, int Nx, int Ny, int Nz, int Ng, int cs, int cp, int Pxh,
            int L, int logL, Radices rad, const double2* __restrict__ tw, double norm) {
    CFDNN_DYN_SMEM(double2, sm);
    double2* A = sm;
    double2* B = sm + (size_t)Nx * L;
    const int rows = Ny * Nz;
    const int line0 = blockIdx.x * L;
    for (int t = threadIdx.x; t < Nx * L; t += blockDim.x) {
        const int l = t / Nx, n = t - l * Nx;
        const int ra = 2 * (line0 + l), rb = ra + 1;
        const bool upper = n > Nx / 2;
        const int kx = upper ? Nx - n : n;
        double2 a = c2(0.0, 0.0), b = c2(0.0, 0.0);
        if (ra < rows) a = spec[(size_t)ra * Pxh + kx];
        if (rb < rows) b = spec[(size_t)rb * Pxh + kx];
        if (upper) { a = cconj(a); b = cconj(b); }
        const double2 z = c2(a.x - b.y, a.y + b.x);                    // a + i b
        A[sw(n, l, L)] = cconj(z);
    }
    __syncthreads();
    const double2* Z = fft_lines(A, B, Nx, L, logL, rad, tw);
    for (int t = threadIdx.x; t < (Nx + 2) * L; t += blockDim.x) {
        const int l = t / (Nx + 2), i = t - l * (Nx + 2);              // i: ghosted column 0 .. Nx+1 (Ng = 1)
        const int n = (i + Nx - 1) % Nx;
        const int ra = 2 * (line0 + l), rb = ra + 1;
        const double2 z = Z[sw(n, l, L)];
        if (ra < rows) {
            const int k = ra / Ny, j = ra - k * Ny;
            p[(size_t)(k + Ng) * cp + (size_t)(j + Ng) * cs + i] = z.x * norm;
        }
        if (rb < rows) {
            const int k = rb / Ny, j = rb - k * Ny;
            p[(size_t)(k + Ng) * cp + (size_t)(j + Ng) * cs + i] = -z.y * norm;
        }
    }
}

// ---------------------------------------------------------------- z transform, in place on spec[kz][j][kx]
// a CTA owns row j and a tile of L consecutive kx: every global access is a run of L complex numbers
template <bool INV>
__global__ void __launch_bounds__(FFT_THREADS)
k_fft_z(double2* __restrict__ spec, int Ny, int Nz, int Pxh, int L, int logL, Radices rad, const double2* __restrict__ tw) {
    CFDNN_DYN_SMEM(double2, sm);
    double2* A = sm;
    double2* B = sm + (size_t)Nz * L;
    const int j = blockIdx.y, kx0 = blockIdx.x * L;
    const size_t zs = (size_t)Ny * Pxh;
    double2* base = spec + (size_t)j * Pxh + kx0;
    for (int t = threadIdx.x; t < Nz * L; t += blockDim.x) {
        const int l = t & (L - 1), n = t >> logL;
        const double2 v = base[(size_t)n * zs + l];
        A[sw(n, l, L)] = INV ? cconj(v) : v;
    }
    __syncthreads();
    const double2* Z = fft_lines(A, B, Nz, L, logL, rad, tw);
    for (int t = threadIdx.x; t < Nz * L; t += blockDim.x) {
        const int l = t & (L - 1), n = t >> logL;
        const double2 v = Z[sw(n, l, L)];
        base[(size_t)n * zs + l] = INV ? cconj(v) : v;
    }
}

// ---------------------------------------------------------------- tridiagonal systems in y, one thread per mode
// Thomas algorithm with the reference's rows (:606-641 for the matrix, :1027-1075 for the recurrences): diagonal
// -(aS + aN) - (lambda_x + lambda_z); the (0, 0) mode is pinned to 0 at j = 0.  d' overwrites the spectrum, c' is
// written when BUILD (mesh-only data, tabulated at the first solve) and read by the back substitution.
template <bool BUILD>
__global__ void __launch_bounds__(128)
k_fft_thomas(double2* __restrict__ spec, double* __restrict__ cprime, const double* __restrict__ lamx,
             const double* __restrict__ lamz, const double* __restrict__ lo, const double* __restrict__ up,
             const double* __restrict__ dbase, int Ny, int Nxh, int Pxh) {
    const int kx = blockIdx.x * blockDim.x + threadIdx.x, kz = blockIdx.y;
    if (kx >= Nxh) return;
    const double shift = lamx[kx] + lamz[kz];
    const bool zero_mode = kx == 0 && kz == 0;
    size_t idx = (size_t)kz * Ny * Pxh + kx;
    double c = 0.0;
    double2 d = c2(0.0, 0.0);
    for (int j = 0; j < Ny; ++j, idx += Pxh) {
        double diag = dbase[j] - shift;
        const double lower = j > 0 ? lo[j] : 0.0;
        double upper = j < Ny - 1 ? up[j] : 0.0;
        double2 r = spec[idx];
        if (zero_mode && j == 0) { diag = 1.0; upper = 0.0; r = c2(0.0, 0.0); }
        if (BUILD) {
            const double den = diag - lower * c;
            c = upper / den;
            d = c2((r.x - lower * d.x) / den, (r.y - lower * d.y) / den);
            cprime[idx] = c;
        } else {                        // c'[j] comes from the table: one division per row less
            const double cn = cprime[idx];
            const double den = diag - lower * c;
            d = c2((r.x - lower * d.x) / den, (r.y - lower * d.y) / den);
            c = cn;
        }
        spec[idx] = d;
    }
    idx -= Pxh;
    double2 x = d;
    for (int j = Ny - 2; j >= 0; --j) {
        idx -= Pxh;
        const double2 dj = spec[idx];
        const double cj = cprime[idx];
        x = c2(dj.x - cj * x.x, dj.y - cj * x.y);
        spec[idx] = x;
    }
}

// Neumann ghost rows in y over the interior planes, all columns (apply_bc_device :806-866, y part; the z planes are
// two plane copies issued by the host after this kernel)
__global__ void k_fft_ghost_y(double* __restrict__ p, int Nxt, int Ny, int Nz, int Ng, int cs, int cp) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x, k = blockIdx.y;
    if (i >= Nxt) return;
    double* pl = p + (size_t)(k + Ng) * cp;
    pl[i] = pl[(size_t)Ng * cs + i];
    pl[(size_t)(Ng + Ny) * cs + i] = pl[(size_t)(Ng + Ny - 1) * cs + i];
}

int ilog2(int v) { int l = 0; while ((1 << l) < v) ++l; return l; }

Radices factorize(int n) {              // 4s first, then one 2, then 3s and 5s
    Radices r; r.n = 0;
    for (int q = 0; q < FFT_MAX_RAD; ++q) r.r[q] = 0;
    while (n % 4 == 0) { r.r[r.n++] = 4; n /= 4; }
    while (n % 2 == 0) { r.r[r.n++] = 2; n /= 2; }
    while (n % 3 == 0) { r.r[r.n++] = 3; n /= 3; }
    while (n % 5 == 0) { r.r[r.n++] = 5; n /= 5; }
    if (n != 1) throw std::invalid_argument("FFT Poisson: grid sizes in x and z must factor into 2, 3 and 5");
    return r;
}

int lines_per_cta(int n) {              // 2 buffers of n L complex numbers within 64 KB
    int L = 8;
    while (L > 1 && (size_t)2 * n * L * sizeof(double2) > 65536) L >>= 1;
    return L;
}

void fill_twiddles(DevBuf<double2>& buf, int n) {
    std::vector<double2> h(n);
    for (int i = 0; i < n; ++i) {
        const long double a = -2.0L * 3.14159265358979323846264338327950288L * (long double)i / (long double)n;
        h[i].x = (double)cosl(a);
        h[i].y = (double)sinl(a);
    }
    buf.alloc(n, false);
    CFDNN_CUDA(cudaMemcpy(buf.p, h.data(), n * sizeof(double2), cudaMemcpyHostToDevice));
}

void upload(DevBuf<double>& buf, const std::vector<double>& h) {
    buf.alloc(h.size(), false);
    CFDNN_CUDA(cudaMemcpy(buf.p, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice));
}
}  // namespace

bool FftPoisson::size_supported(int n) {
    if (n < 2 || n > 1024) return false;
    for (int f : {2, 3, 5}) while (n % f == 0) n /= f;
    return n == 1;
}

FftPoisson::FftPoisson(const cfdnn_mesh_desc& m) : Nx_(m.Nx), Ny_(m.Ny), Nz_(m.Nz), Ng_(m.Ng) {
    require_device();
    if (m.Ng != 1) throw std::invalid_argument("FFT Poisson: only Nghost == 1 is supported");
    if (m.Nz_global > 0 && m.Nz_global != m.Nz)
        throw std::invalid_argument("FFT Poisson: single-GPU only (no z-slab decomposition)");
    if (Ny_ < 2) throw std::invalid_argument("FFT Poisson: Ny >= 2 required");
    if (!size_supported(Nx_) || !size_supported(Nz_))
        throw std::invalid_argument("FFT Poisson: Nx and Nz must be 2^a 3^b 5^c in [2, 1024]");
    if (m.y_stretched && (!m.yLap_aS || !m.yLap_aN || !m.dyv))
        throw std::invalid_argument("FFT Poisson: y-stretched mesh needs yLap_aS, yLap_aN and dyv");
    Nxh_ = Nx_ / 2 + 1;
    Pxh_ = (Nxh_ + 7) / 8 * 8;
    cs_ = Nx_ + 2 * Ng_;
    cp_ = cs_ * (Ny_ + 2 * Ng_);
    Lx_ = lines_per_cta(Nx_);
    Lz_ = lines_per_cta(Nz_);
    Radices rx = factorize(Nx_), rz = factorize(Nz_);
    nrx_ = rx.n; nrz_ = rz.n;
    std::memcpy(radx_, rx.r, sizeof(radx_));
    std::memcpy(radz_, rz.r, sizeof(radz_));
    setup_tables(m);
    const size_t ns = (size_t)Nz_ * Ny_ * Pxh_;
    spec_.alloc(ns, true);
    cprime_.alloc(ns, true);
    partial_.alloc(1 + 1184, true);
    CFDNN_CUDA(cudaFuncSetAttribute(k_fft_x_fwd, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    CFDNN_CUDA(cudaFuncSetAttribute(k_fft_x_inv, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    CFDNN_CUDA(cudaFuncSetAttribute(k_fft_z<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    CFDNN_CUDA(cudaFuncSetAttribute(k_fft_z<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
}

void FftPoisson::setup_tables(const cfdnn_mesh_desc& m) {
    fill_twiddles(twx_, Nx_);
    fill_twiddles(twz_, Nz_);
    // compute_eigenvalues, second-order branch (src/poisson_solver_fft.cpp:494-504)
    std::vector<double> lx(Nxh_), lz(Nz_);
    for (int k = 0; k < Nxh_; ++k) lx[k] = (2.0 - 2.0 * std::cos(2.0 * M_PI * k / Nx_)) / (m.dx * m.dx);
    for (int k = 0; k < Nz_; ++k) lz[k] = (2.0 - 2.0 * std::cos(2.0 * M_PI * k / Nz_)) / (m.dz * m.dz);
    upload(lamx_, lx);
    upload(lamz_, lz);
    // compute_tridiagonal_coeffs (:509-547) with Neumann walls, volume weights (:420-435)
    std::vector<double> lo(Ny_), up(Ny_), base(Ny_), w(Ny_);
    double Ly = 0.0;
    for (int j = 0; j < Ny_; ++j) {
        double aS, aN;
        if (m.y_stretched) {
            aS = m.yLap_aS[j + Ng_];
            aN = m.yLap_aN[j + Ng_];
        } else {
            const double inv = 1.0 / (m.dy * m.dy);
            aS = inv; aN = inv;
        }
        if (j == 0) aS = 0.0;
        if (j == Ny_ - 1) aN = 0.0;
        lo[j] = aS; up[j] = aN; base[j] = -(aS + aN);
        w[j] = m.y_stretched ? m.dyv[j + Ng_] : m.dy;
        Ly += w[j];
    }
    total_volume_ = (double)Nx_ * Nz_ * Ly;
    upload(lo_, lo); upload(up_, up); upload(base_, base); upload(dyv_, w);
}

void FftPoisson::set_bc(const int bc[6]) {      // set_bc :253-265 (runtime_error for non-periodic x / z)
    if (bc[0] != CFDNN_PBC_PERIODIC || bc[1] != CFDNN_PBC_PERIODIC)
        throw Error(CFDNN_ERR_RUNTIME, "FFTPoissonSolver requires periodic BC in x");
    if (bc[4] != CFDNN_PBC_PERIODIC || bc[5] != CFDNN_PBC_PERIODIC)
        throw Error(CFDNN_ERR_RUNTIME, "FFTPoissonSolver requires periodic BC in z");
    if (bc[2] != CFDNN_PBC_NEUMANN || bc[3] != CFDNN_PBC_NEUMANN)
        throw std::invalid_argument("FFT Poisson: only Neumann walls in y are supported");
    bc_set_ = true;
}

int FftPoisson::solve_dev(const double* rhs, double* p, cudaStream_t st) {          // solve_device :960-1098
    if (!bc_set_) throw Error(CFDNN_ERR_RUNTIME, "FFTPoissonSolver: set_bc has not been called");
    if (!rhs || !p) throw std::invalid_argument("FFT Poisson: null field pointer");
    Radices rx, rz;
    rx.n = nrx_; rz.n = nrz_;
    std::memcpy(rx.r, radx_, sizeof(rx.r));
    std::memcpy(rz.r, radz_, sizeof(rz.r));
    const int rows = Ny_ * Nz_;
    // 1. volume-weighted mean of the right-hand side
    const int nb = rows < 1184 ? rows : 1184;
    CFDNN_LAUNCH_COOP(k_fft_wsum, nb, 256, 0, st, rhs, dyv_.p, Nx_, Ny_, Nz_, Ng_, cs_, cp_, partial_.p);
    CFDNN_LAUNCH(k_fft_mean, 1, 1, 0, st, partial_.p, nb, 1.0 / total_volume_);
    // 2. forward transforms
    const int xlines = (rows + 1) / 2;
    const size_t smx = (size_t)2 * Nx_ * Lx_ * sizeof(double2), smz = (size_t)2 * Nz_ * Lz_ * sizeof(double2);
    CFDNN_LAUNCH_COOP(k_fft_x_fwd, (xlines + Lx_ - 1) / Lx_, FFT_THREADS, smx, st, rhs, partial_.p, spec_.p, Nx_, Ny_, Nz_, Ng_,
                      cs_, cp_, Pxh_, Lx_, ilog2(Lx_), rx, twx_.p);
    const dim3 gz(Pxh_ / Lz_, Ny_);
    CFDNN_LAUNCH_COOP((k_fft_z<false>), gz, FFT_THREADS, smz, st, spec_.p, Ny_, Nz_, Pxh_, Lz_, ilog2(Lz_), rz, twz_.p);
    // 3. tridiagonal solves
    const dim3 gt((Nxh_ + 127) / 128, Nz_);
    if (!cprime_ready_) {
        CFDNN_LAUNCH((k_fft_thomas<true>), gt, 128, 0, st, spec_.p, cprime_.p, lamx_.p, lamz_.p, lo_.p, up_.p, base_.p, Ny_,
                     Nxh_, Pxh_);
        cprime_ready_ = true;
    } else {
        CFDNN_LAUNCH((k_fft_thomas<false>), gt, 128, 0, st, spec_.p, cprime_.p, lamx_.p, lamz_.p, lo_.p, up_.p, base_.p, Ny_,
                     Nxh_, Pxh_);
    }
    // 4. inverse transforms, normalisation, ghosts (x in the kernel, then y rows, then the periodic z planes)
    CFDNN_LAUNCH_COOP((k_fft_z<true>), gz, FFT_THREADS, smz, st, spec_.p, Ny_, Nz_, Pxh_, Lz_, ilog2(Lz_), rz, twz_.p);
    CFDNN_LAUNCH_COOP(k_fft_x_inv, (xlines + Lx_ - 1) / Lx_, FFT_THREADS, smx, st, spec_.p, p, Nx_, Ny_, Nz_, Ng_, cs_, cp_, Pxh_,
                      Lx_, ilog2(Lx_), rx, twx_.p, 1.0 / ((double)Nx_ * (double)Nz_));
    CFDNN_LAUNCH(k_fft_ghost_y, dim3((cs_ + 127) / 128, Nz_), 128, 0, st, p, cs_, Ny_, Nz_, Ng_, cs_, cp_);
    CFDNN_CUDA(cudaMemcpyAsync(p, p + (size_t)Nz_ * cp_, (size_t)cp_ * sizeof(double), cudaMemcpyDeviceToDevice, st));
    CFDNN_CUDA(cudaMemcpyAsync(p + (size_t)(Nz_ + 1) * cp_, p + (size_t)cp_, (size_t)cp_ * sizeof(double),
                               cudaMemcpyDeviceToDevice, st));
    return 1;
}

int FftPoisson::solve_host(const double* rhs, double* p) {
    if (!rhs || !p) throw std::invalid_argument("FFT Poisson: null field pointer");
    const size_t n = field_elems();
    DevBuf<double> dr, dp;
    dr.alloc(n, false);
    dp.alloc(n, true);
    cudaStream_t st = nullptr;
    CFDNN_CUDA(cudaMemcpyAsync(dr.p, rhs, n * sizeof(double), cudaMemcpyHostToDevice, st));
    const int it = solve_dev(dr.p, dp.p, st);
    CFDNN_CUDA(cudaMemcpyAsync(p, dp.p, n * sizeof(double), cudaMemcpyDeviceToHost, st));
    CFDNN_CUDA(cudaStreamSynchronize(st));
    check_launch("fft_solve");
    return it;
}

}  // namespace cfdnn
