// fft.cu — FFT-hybrid Poisson solver for channels: periodic x/z, Neumann walls in y, uniform or stretched y.
//
// Reference: FFTPoissonSolver, src/poisson_solver_fft.cpp (GPU build only: cuFFT 2-D plans over a packed
// [i][k][j] copy + cuSPARSE gtsv2StridedBatch).  Same algorithm (:960-1098) — volume-weighted mean subtraction,
// forward 2-D FFT over (x, z), one tridiagonal system in y per mode with the (0,0) mode pinned at j = 0, inverse
// FFT, 1/(Nx Nz), ghost fill — but designed for this backend's layout instead of a library's:
//   * no pack/unpack passes: the x transform reads the solver's ghosted rows and the inverse x transform writes
//     them (x ghosts included);
//   * the half spectrum is taken along x (the contiguous index), two real rows ride one complex transform;
//   * hand-written Stockham autosort FFTs in shared memory (radix 8/4/2/3/5, conflict-free swizzled [n][line] tiles):
//     x lines are contiguous rows, z lines are gathered as tiles of L consecutive kx so every global access is a
//     run of L complex numbers;
//   * the Thomas solve runs on the [kz][j][kx] spectrum in place, one thread per mode, coalesced in kx; its
//     factorisation (c', 1/den) depends on the mesh only and is tabulated once; the mean removal is folded into the
//     zero mode;
//   * on z-slabs the spectrum is kept in per-rank blocks and the two transposes per solve are either NCCL all-to-alls or
//     direct stores of the transform kernels into the other GPUs' arrays (CUDA IPC), see setup_peer_access().
// Algorithmic traffic per real cell: x fwd 8+8, z fwd 8+8, Thomas 36, z inv 8+8, x inv 8+8 = 100 B — 6.7 GB at
// 512x256x512 (1.89 ms per solve measured), against ~20 V-cycles of ~1 GB each for the multigrid.
#include "fft.cuh"
#include <cmath>
#include <cstring>
#include <cstdlib>
#include <cstdio>

namespace cfdnn {
namespace {

constexpr int FFT_THREADS = 256;
constexpr int FFT_MAX_RAD = 12;

struct Radices {
    int n, r[FFT_MAX_RAD];
};

// Where the blocks of the decomposed spectrum live, as seen by the kernel that WRITES them: b[q] = base of the block
// that belongs to rank q.  One rank / NCCL transposes: offsets into this rank's own array.  Peer-to-peer transposes: the
// IPC-mapped arrays of the other GPUs, so the store of the transform IS the all-to-all (NVLink writes, no pack, no pass).
constexpr int FFT_MAX_RANKS = 16;
struct BlockPtrs {
    double2* b[FFT_MAX_RANKS];
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
    const int logNs = (Ns & (Ns - 1)) == 0 ? 31 - __clz(Ns) : -1;     // 4s and 2s come first: almost always a shift
    for (int t = threadIdx.x; t < M * L; t += blockDim.x) {
        const int l = t & (L - 1), j = t >> logL;
        const int k = logNs >= 0 ? (j & (Ns - 1)) : j % Ns;
        const int d0 = (j - k) * R + k;                                 // (j / Ns) Ns R + k
        double2 v[R];
#pragma unroll
        for (int r = 0; r < R; ++r) v[r] = A[sw(j + r * M, l, L)];
        if constexpr (R == 8) {
            if (k > 0) {                    // w^r, r = 1..7, from three table entries (w, w^2, w^4)
                const double2 w1 = tw[k * tstep], w2 = tw[2 * k * tstep], w4 = tw[4 * k * tstep];
                const double2 w3 = cmul(w1, w2), w5 = cmul(w4, w1), w6 = cmul(w4, w2);
                v[1] = cmul(v[1], w1); v[2] = cmul(v[2], w2); v[3] = cmul(v[3], w3); v[4] = cmul(v[4], w4);
                v[5] = cmul(v[5], w5); v[6] = cmul(v[6], w6); v[7] = cmul(v[7], cmul(w4, w3));
            }
            // decimation in frequency: X[2m] = DFT4(v[q] + v[q+4])[m], X[2m+1] = DFT4((v[q] - v[q+4]) W8^q)[m]
            const double h = 0.70710678118654752440;
            const double2 a0 = cadd(v[0], v[4]), a1 = cadd(v[1], v[5]), a2 = cadd(v[2], v[6]), a3 = cadd(v[3], v[7]);
            const double2 b0 = csub(v[0], v[4]), e1 = csub(v[1], v[5]), e2 = csub(v[2], v[6]), e3 = csub(v[3], v[7]);
            const double2 b1 = c2(h * (e1.x + e1.y), h * (e1.y - e1.x));      // e1 (1 - i) / sqrt 2
            const double2 b2 = c2(e2.y, -e2.x);                                // -i e2
            const double2 b3 = c2(h * (e3.y - e3.x), -h * (e3.x + e3.y));     // e3 (-1 - i) / sqrt 2
            {
                const double2 t0 = cadd(a0, a2), t1 = csub(a0, a2), t2 = cadd(a1, a3), e = csub(a1, a3);
                const double2 t3 = c2(e.y, -e.x);
                B[sw(d0, l, L)] = cadd(t0, t2);
                B[sw(d0 + 2 * Ns, l, L)] = cadd(t1, t3);
                B[sw(d0 + 4 * Ns, l, L)] = csub(t0, t2);
                B[sw(d0 + 6 * Ns, l, L)] = csub(t1, t3);
            }
            {
                const double2 t0 = cadd(b0, b2), t1 = csub(b0, b2), t2 = cadd(b1, b3), e = csub(b1, b3);
                const double2 t3 = c2(e.y, -e.x);
                B[sw(d0 + Ns, l, L)] = cadd(t0, t2);
                B[sw(d0 + 3 * Ns, l, L)] = cadd(t1, t3);
                B[sw(d0 + 5 * Ns, l, L)] = csub(t0, t2);
                B[sw(d0 + 7 * Ns, l, L)] = csub(t1, t3);
            }
            continue;
        }
        if (k > 0) {
#pragma unroll
            for (int r = 1; r < R; ++r) v[r] = cmul(v[r], tw[r * k * tstep]);
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
        if (R == 8) fft_stage<8>(A, B, N, Ns, L, logL, tw);
        else if (R == 4) fft_stage<4>(A, B, N, Ns, L, logL, tw);
        else if (R == 2) fft_stage<2>(A, B, N, Ns, L, logL, tw);
        else if (R == 3) fft_stage<3>(A, B, N, Ns, L, logL, tw);
        else fft_stage<5>(A, B, N, Ns, L, logL, tw);
        __syncthreads();
        double2* T = A; A = B; B = T;
        Ns *= R;
    }
    return A;
}

// ---------------------------------------------------------------- x transforms (rows of the ghosted cell layout)
// The half spectrum of the local rows is stored in blocks of Pq columns, S[q][row][c] with kx = q Pq + c: block q is what
// rank q receives in the all-to-all (one rank: a single block, i.e. the plain [row][kx] array).  Offset of (row 0, kx):
__device__ __forceinline__ size_t sblock(int kx, int Pq, int rows) {
    const int q = kx < Pq ? 0 : kx / Pq;
    return (size_t)q * rows * Pq + (size_t)(kx - q * Pq);
}

// forward: line l of the CTA carries rows 2l (real part) and 2l+1 (imaginary part); row = k Ny + j
__global__ void __launch_bounds__(FFT_THREADS)
k_fft_x_fwd(const double* __restrict__ rhs, BlockPtrs dst, int Nx, int Ny,
            int Nz, int Ng, int cs, int cp, int Pxh, int L, int logL, Radices rad, const double2* __restrict__ tw) {
    CFDNN_DYN_SMEM(double2, sm);
    double2* A = sm;
    double2* B = sm + (size_t)Nx * L;
    const int rows = Ny * Nz;
    const int line0 = blockIdx.x * L;
    __shared__ long long ro[16];                    // offset of the first interior cell of rows 2l, 2l+1 (-1: no such row)
    if ((int)threadIdx.x < 2 * L) {
        const int row = 2 * line0 + threadIdx.x;
        const int k = row / Ny, j = row - k * Ny;
        ro[threadIdx.x] = row < rows ? (long long)(k + Ng) * cp + (long long)(j + Ng) * cs + Ng : -1;
    }
    __syncthreads();
    for (int l0 = 0; l0 < L; l0 += 4) {             // 8 loads in flight per thread before the first shared-memory store
        for (int n = threadIdx.x; n < Nx; n += blockDim.x) {
            double a[4], b[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const long long oa = l0 + q < L ? ro[2 * (l0 + q)] : -1, ob = l0 + q < L ? ro[2 * (l0 + q) + 1] : -1;
                a[q] = oa >= 0 ? rhs[oa + n] : 0.0;
                b[q] = ob >= 0 ? rhs[ob + n] : 0.0;
            }
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (l0 + q < L) A[sw(n, l0 + q, L)] = c2(a[q], b[q]);
        }
    }
    __syncthreads();
    const double2* Z = fft_lines(A, B, Nx, L, logL, rad, tw);
    const int Nxh = Nx / 2 + 1;
    for (int l = 0; l < L; ++l) {
        const int ra = 2 * (line0 + l), rb = ra + 1;
        for (int kx = threadIdx.x; kx < Nxh; kx += blockDim.x) {
            const double2 z = Z[sw(kx, l, L)], zc = cconj(Z[sw(kx == 0 ? 0 : Nx - kx, l, L)]);
            const double2 s = cadd(z, zc), d = csub(z, zc);
            const int q = kx < Pxh ? 0 : kx / Pxh;                       // owner of this kx column
            double2* const blk = dst.b[q] + (kx - q * Pxh);
            if (ra < rows) blk[(size_t)ra * Pxh] = c2(0.5 * s.x, 0.5 * s.y);
            if (rb < rows) blk[(size_t)rb * Pxh] = c2(0.5 * d.y, -0.5 * d.x);        // (z - conj z') / (2 i)
        }
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
    __shared__ long long ro[16];                    // offset of column 0 (x ghost) of rows 2l, 2l+1 (-1: no such row)
    if ((int)threadIdx.x < 2 * L) {
        const int row = 2 * line0 + threadIdx.x;
        const int k = row / Ny, j = row - k * Ny;
        ro[threadIdx.x] = row < rows ? (long long)(k + Ng) * cp + (long long)(j + Ng) * cs : -1;
    }
    for (int l0 = 0; l0 < L; l0 += 4) {             // 8 loads in flight per thread before the first shared-memory store
        for (int n = threadIdx.x; n < Nx; n += blockDim.x) {
            const bool upper = n > Nx / 2;
            const int kx = upper ? Nx - n : n;
            double2 a[4], b[4];
            const size_t o = sblock(kx, Pxh, rows);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int ra = 2 * (line0 + l0 + q), rb = ra + 1;
                a[q] = (l0 + q < L && ra < rows) ? spec[o + (size_t)ra * Pxh] : c2(0.0, 0.0);
                b[q] = (l0 + q < L && rb < rows) ? spec[o + (size_t)rb * Pxh] : c2(0.0, 0.0);
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (l0 + q >= L) continue;
                if (upper) { a[q] = cconj(a[q]); b[q] = cconj(b[q]); }
                A[sw(n, l0 + q, L)] = c2(a[q].x - b[q].y, -(a[q].y + b[q].x));      // conj(a + i b)
            }
        }
    }
    __syncthreads();
    const double2* Z = fft_lines(A, B, Nx, L, logL, rad, tw);
    for (int l = 0; l < L; ++l) {
        const long long oa = ro[2 * l], ob = ro[2 * l + 1];
        for (int i = threadIdx.x; i < Nx + 2; i += blockDim.x) {        // ghosted column 0 .. Nx+1 (Ng = 1)
            const int n = i == 0 ? Nx - 1 : (i == Nx + 1 ? 0 : i - 1);
            const double2 z = Z[sw(n, l, L)];
            if (oa >= 0) p[oa + i] = z.x * norm;
            if (ob >= 0) p[ob + i] = -z.y * norm;
        }
    }
}

// ---------------------------------------------------------------- z transform, in place on spec[kz][j][kx]
// a CTA owns row j and a tile of L consecutive kx: every global access is a run of L complex numbers
template <bool INV>
__global__ void __launch_bounds__(FFT_THREADS)
k_fft_z(double2* __restrict__ spec, BlockPtrs out, int Nzl, int Ny, int Nz, int Pxh, int L, int logL, Radices rad,
        const double2* __restrict__ tw) {
    CFDNN_DYN_SMEM(double2, sm);
    double2* A = sm;
    double2* B = sm + (size_t)Nz * L;
    const int j = blockIdx.y, kx0 = blockIdx.x * L;
    const size_t zs = (size_t)Ny * Pxh;
    double2* base = spec + (size_t)j * Pxh + kx0;
    for (int t0 = threadIdx.x; t0 < Nz * L; t0 += 8 * blockDim.x) {      // 8 loads in flight per thread
        double2 v[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const int t = t0 + q * blockDim.x;
            if (t < Nz * L) v[q] = base[(size_t)(t >> logL) * zs + (t & (L - 1))];
        }
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const int t = t0 + q * blockDim.x;
            if (t < Nz * L) A[sw(t >> logL, t & (L - 1), L)] = INV ? cconj(v[q]) : v[q];
        }
    }
    __syncthreads();
    const double2* Z = fft_lines(A, B, Nz, L, logL, rad, tw);
    for (int t = threadIdx.x; t < Nz * L; t += blockDim.x) {
        const int l = t & (L - 1), n = t >> logL;
        const double2 v = Z[sw(n, l, L)];
        if (INV) {                                  // plane n belongs to the slab of rank n / Nzl: its block of S
            const int s = n / Nzl;
            out.b[s][((size_t)(n - s * Nzl) * Ny + j) * Pxh + kx0 + l] = cconj(v);
        } else {
            base[(size_t)n * zs + l] = v;
        }
    }
}

// ---------------------------------------------------------------- tridiagonal systems in y, one thread per mode
// Thomas algorithm with the reference's rows (:606-641 for the matrix, :1027-1075 for the recurrences): diagonal
// -(aS + aN) - (lambda_x + lambda_z); the (0, 0) mode is pinned to 0 at j = 0.  The factorisation depends on the mesh
// only: c'[j] = upper / den[j] and 1 / den[j], den[j] = diag[j] - lower[j] c'[j-1], are tabulated once per solver.
__global__ void __launch_bounds__(128)
k_fft_thomas_setup(double* __restrict__ cprime, double* __restrict__ invden, const double* __restrict__ lamx,
                   const double* __restrict__ lamz, const double* __restrict__ lo, const double* __restrict__ up,
                   const double* __restrict__ dbase, int Ny, int Nxh, int Pxh, int kx0) {
    const int c_ = blockIdx.x * blockDim.x + threadIdx.x, kz = blockIdx.y;      // column of this rank, kx = kx0 + column
    const int kx = kx0 + c_;
    if (c_ >= Pxh || kx >= Nxh) return;
    const double shift = lamx[kx] + lamz[kz];
    const bool zero_mode = kx == 0 && kz == 0;
    size_t idx = (size_t)kz * Ny * Pxh + c_;
    double c = 0.0;
    for (int j = 0; j < Ny; ++j, idx += Pxh) {
        double diag = dbase[j] - shift;
        const double lower = j > 0 ? lo[j] : 0.0;
        double upper = j < Ny - 1 ? up[j] : 0.0;
        if (zero_mode && j == 0) { diag = 1.0; upper = 0.0; }
        const double den = diag - lower * c;
        c = upper / den;
        cprime[idx] = c;
        invden[idx] = 1.0 / den;
    }
}

// Forward elimination d'[j] = (r[j] - lower[j] d'[j-1]) / den[j] and back substitution x[j] = d'[j] - c'[j] x[j+1],
// in place on the spectrum.  Rows are taken U at a time: the U loads of a chunk do not depend on the recurrence and
// are all in flight before the first dependent operation (one thread per mode gives too few warps to hide a
// load per row otherwise).
template <int U>
__global__ void __launch_bounds__(128)
k_fft_thomas(double2* __restrict__ spec, const double* __restrict__ cprime, const double* __restrict__ invden,
             const double* __restrict__ lo, const double* __restrict__ dyv, double inv_Ly, int Ny, int Nxh, int Pxh,
             int kx0) {
    const int c_ = blockIdx.x * blockDim.x + threadIdx.x, kz = blockIdx.y;      // column of this rank, kx = kx0 + column
    const int kx = kx0 + c_;
    if (c_ >= Pxh || kx >= Nxh) return;
    const bool zero_mode = kx == 0 && kz == 0;
    const size_t base = (size_t)kz * Ny * Pxh + c_;
    // Mean removal (solvability, kernel_pack_and_partial_sum + kernel_subtract_mean :53-130): subtracting the
    // volume-weighted mean m = sum f dyv / (Nx Nz Ly) from every cell changes nothing but the (0, 0) coefficient of each
    // row, F00[j] -= m Nx Nz, and sum_j dyv[j] F00[j] IS sum f dyv: the thread of the zero mode does it on its column.
    double mean_hat = 0.0;
    if (zero_mode) {
        double s = 0.0;
        for (int j = 0; j < Ny; ++j) s += spec[base + (size_t)j * Pxh].x * dyv[j];
        mean_hat = s * inv_Ly;
    }
    double2 d = c2(0.0, 0.0);
    for (int j0 = 0; j0 < Ny; j0 += U) {
        double2 r[U];
        double iv[U], lw[U];
#pragma unroll
        for (int q = 0; q < U; ++q) {
            const int j = j0 + q;
            if (j < Ny) {
                r[q] = spec[base + (size_t)j * Pxh];
                iv[q] = invden[base + (size_t)j * Pxh];
                lw[q] = j > 0 ? lo[j] : 0.0;
            }
        }
        if (zero_mode) {
#pragma unroll
            for (int q = 0; q < U; ++q) r[q].x -= mean_hat;
            if (j0 == 0) r[0] = c2(0.0, 0.0);
        }
#pragma unroll
        for (int q = 0; q < U; ++q) {
            const int j = j0 + q;
            if (j < Ny) {
                d = c2((r[q].x - lw[q] * d.x) * iv[q], (r[q].y - lw[q] * d.y) * iv[q]);
                spec[base + (size_t)j * Pxh] = d;
            }
        }
    }
    double2 x = d;                                   // row Ny - 1
    for (int j0 = Ny - 2; j0 >= 0; j0 -= U) {
        double2 dj[U];
        double cj[U];
#pragma unroll
        for (int q = 0; q < U; ++q) {
            const int j = j0 - q;
            if (j >= 0) {
                dj[q] = spec[base + (size_t)j * Pxh];
                cj[q] = cprime[base + (size_t)j * Pxh];
            }
        }
#pragma unroll
        for (int q = 0; q < U; ++q) {
            const int j = j0 - q;
            if (j >= 0) {
                x = c2(dj[q].x - cj[q] * x.x, dj[q].y - cj[q] * x.y);
                spec[base + (size_t)j * Pxh] = x;
            }
        }
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

Radices factorize(int n) {              // 8s first, then 4s, one 2, then 3s and 5s (power-of-two strides as long as possible)
    Radices r; r.n = 0;
    for (int q = 0; q < FFT_MAX_RAD; ++q) r.r[q] = 0;
    const bool r8 = std::getenv("CFDNN_FFT_RADIX8") == nullptr || std::atoi(std::getenv("CFDNN_FFT_RADIX8")) != 0;
    while (r8 && n % 8 == 0) { r.r[r.n++] = 8; n /= 8; }
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
    if (Ny_ < 2) throw std::invalid_argument("FFT Poisson: Ny >= 2 required");
    Nzg_ = m.Nz_global > 0 ? m.Nz_global : m.Nz;
    if (Nzg_ % Nz_ != 0) throw std::invalid_argument("FFT Poisson: Nz_global must be a multiple of the local Nz (equal z-slabs)");
    if (!size_supported(Nx_) || !size_supported(Nzg_))
        throw std::invalid_argument("FFT Poisson: Nx and Nz must be 2^a 3^b 5^c in [2, 1024]");
    if (m.y_stretched && (!m.yLap_aS || !m.yLap_aN || !m.dyv))
        throw std::invalid_argument("FFT Poisson: y-stretched mesh needs yLap_aS, yLap_aN and dyv");
    Nxh_ = Nx_ / 2 + 1;
    cs_ = Nx_ + 2 * Ng_;
    cp_ = cs_ * (Ny_ + 2 * Ng_);
    Lx_ = lines_per_cta(Nx_);
    Lz_ = lines_per_cta(Nzg_);
    auto tune = [](const char* name, int& L, int n) {       // measurement aid: lines per CTA (a power of two <= 8)
        if (const char* e = std::getenv(name)) {
            const int v = std::atoi(e);
            if ((v == 1 || v == 2 || v == 4 || v == 8) && (size_t)2 * n * v * sizeof(double2) <= 65536) L = v;
        }
    };
    tune("CFDNN_FFT_LX", Lx_, Nx_);
    tune("CFDNN_FFT_LZ", Lz_, Nzg_);
    Radices rx = factorize(Nx_), rz = factorize(Nzg_);
    nrx_ = rx.n; nrz_ = rz.n;
    std::memcpy(radx_, rx.r, sizeof(radx_));
    std::memcpy(radz_, rz.r, sizeof(radz_));
    setup_tables(m);
    CFDNN_CUDA(cudaFuncSetAttribute(k_fft_x_fwd, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    CFDNN_CUDA(cudaFuncSetAttribute(k_fft_x_inv, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    CFDNN_CUDA(cudaFuncSetAttribute(k_fft_z<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    CFDNN_CUDA(cudaFuncSetAttribute(k_fft_z<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    if (Nzg_ == Nz_) plan(nullptr);                 // single domain; a slab waits for set_comm
}

// Decomposition of the spectrum.  The x transform runs on this rank's z-slab and leaves S[q][k_local][j][c] — the kx
// range of rank q (c < Pq columns, kx = q Pq + c) as one contiguous block; an all-to-all turns the blocks into
// T[kz][j][c], ALL z for this rank's kx columns (source ranks arrive in global z order), on which the z transform and
// the tridiagonal systems are local; the way back is the same in reverse.  One rank: S and T are the same array.
void FftPoisson::set_comm(Comm* c) {
    const int P = c ? c->nranks : 1;
    if ((long long)Nz_ * P != Nzg_)
        throw std::invalid_argument("FFT Poisson: need Nz_global == Nz_local * nranks (Nz_global = " + std::to_string(Nzg_) +
                                    ", Nz_local = " + std::to_string(Nz_) + ", nranks = " + std::to_string(P) + ")");
    plan(P > 1 ? c : nullptr);
}

// Peer-to-peer transposes: every rank maps the S and T arrays of the others through CUDA IPC (one process per GPU), so the
// transform kernels store straight into the owner's array over NVLink.  All ranks must succeed, else everybody stays on
// the NCCL all-to-all.  CFDNN_FFT_P2P=0 turns it off.
void FftPoisson::release_peer_access() {
#ifndef CFDNN_DEVSIM
    for (void* p : peer_opened_) cudaIpcCloseMemHandle(p);
#endif
    peer_opened_.clear();
    p2p_ = false;
}

void FftPoisson::setup_peer_access() {
    release_peer_access();
#ifndef CFDNN_DEVSIM
    if (P_ <= 1 || P_ > FFT_MAX_RANKS || !comm_->device_transport()) return;
    // Default route, from the measurements in DESIGN 6: the direct stores win on 2 GPUs (config 5 Poisson 44.0 -> 38.4 ms
    // per step) and lose on 8 (13.7 -> 15.9 ms): there the inverse z transform scatters 32-byte runs to eight GPUs.
    // CFDNN_FFT_P2P=1 / 0 forces the route.
    bool want = P_ == 2;
    if (const char* e = std::getenv("CFDNN_FFT_P2P")) want = std::atoi(e) != 0;
    if (!want) return;
    const bool dbg = std::getenv("CFDNN_FFT_P2P_DEBUG") != nullptr;
    auto note = [&](const char* what, cudaError_t e) {
        if (dbg) std::fprintf(stderr, "[cfdnn fft p2p] rank %d: %s: %s\n", rank_, what, cudaGetErrorString(e));
    };
    constexpr int HD = 2 * (int)(sizeof(cudaIpcMemHandle_t) / sizeof(double));        // doubles per rank: two handles
    static_assert(sizeof(cudaIpcMemHandle_t) % sizeof(double) == 0, "IPC handle size");
    std::vector<double> host((size_t)HD * P_, 0.0);
    cudaIpcMemHandle_t hs, ht;
    cudaError_t e1 = cudaIpcGetMemHandle(&hs, spec_.p), e2 = cudaIpcGetMemHandle(&ht, spec2_.p);
    note("cudaIpcGetMemHandle(S)", e1); note("cudaIpcGetMemHandle(T)", e2);
    bool ok = e1 == cudaSuccess && e2 == cudaSuccess;
    if (ok) {
        std::memcpy(host.data() + (size_t)HD * rank_, &hs, sizeof(hs));
        std::memcpy(host.data() + (size_t)HD * rank_ + HD / 2, &ht, sizeof(ht));
    }
    (void)cudaGetLastError();
    DevBuf<double> dev;
    dev.alloc((size_t)HD * P_ + 1, true);
    CFDNN_CUDA(cudaMemcpy(dev.p + (size_t)HD * rank_, host.data() + (size_t)HD * rank_, HD * sizeof(double), cudaMemcpyHostToDevice));
    comm_->allgather_planes(dev.p, HD, nullptr);
    CFDNN_CUDA(cudaMemcpy(host.data(), dev.p, host.size() * sizeof(double), cudaMemcpyDeviceToHost));
    peer_S_.assign(P_, nullptr);
    peer_T_.assign(P_, nullptr);
    for (int q = 0; q < P_ && ok; ++q) {
        if (q == rank_) { peer_S_[q] = spec_.p; peer_T_[q] = spec2_.p; continue; }
        cudaIpcMemHandle_t a, b;
        std::memcpy(&a, host.data() + (size_t)HD * q, sizeof(a));
        std::memcpy(&b, host.data() + (size_t)HD * q + HD / 2, sizeof(b));
        void *ps = nullptr, *pt = nullptr;
        cudaError_t eo = cudaIpcOpenMemHandle(&ps, a, cudaIpcMemLazyEnablePeerAccess);
        note("cudaIpcOpenMemHandle(S)", eo);
        if (eo != cudaSuccess) { ok = false; break; }
        peer_opened_.push_back(ps);
        eo = cudaIpcOpenMemHandle(&pt, b, cudaIpcMemLazyEnablePeerAccess);
        note("cudaIpcOpenMemHandle(T)", eo);
        if (eo != cudaSuccess) { ok = false; break; }
        peer_opened_.push_back(pt);
        peer_S_[q] = static_cast<double2*>(ps);
        peer_T_[q] = static_cast<double2*>(pt);
    }
    (void)cudaGetLastError();
    const double flag = ok ? 1.0 : 0.0;                          // everybody or nobody
    CFDNN_CUDA(cudaMemcpy(dev.p + (size_t)HD * P_, &flag, sizeof(double), cudaMemcpyHostToDevice));
    comm_->allreduce_min(dev.p + (size_t)HD * P_, 1, nullptr);
    double all = 0.0;
    CFDNN_CUDA(cudaMemcpy(&all, dev.p + (size_t)HD * P_, sizeof(double), cudaMemcpyDeviceToHost));
    if (all > 0.5) p2p_ = true; else release_peer_access();
#endif
}

void FftPoisson::plan(Comm* c) {
    release_peer_access();
    comm_ = c;
    P_ = c ? c->nranks : 1;
    rank_ = c ? c->rank : 0;
    if (P_ > FFT_MAX_RANKS)
        throw std::invalid_argument("FFT Poisson: at most " + std::to_string(FFT_MAX_RANKS) + " z-slabs (block pointer table)");
    Pq_ = ((Nxh_ + P_ - 1) / P_ + 7) / 8 * 8;
    const size_t ns = (size_t)Nzg_ * Ny_ * Pq_;     // = P * Nz_local * Ny * Pq
    spec_.alloc(ns, true);
    if (P_ > 1) spec2_.alloc(ns, true); else spec2_.release();
    sync_.alloc(8, true);
    cprime_.alloc(ns, true);
    invden_.alloc(ns, true);
    setup_peer_access();
    const dim3 gt((Pq_ + 127) / 128, Nzg_);
    CFDNN_LAUNCH(k_fft_thomas_setup, gt, 128, 0, nullptr, cprime_.p, invden_.p, lamx_.p, lamz_.p, lo_.p, up_.p, base_.p, Ny_,
                 Nxh_, Pq_, rank_ * Pq_);
    CFDNN_CUDA(cudaDeviceSynchronize());
    check_launch("fft thomas setup");
    planned_ = true;
}

void FftPoisson::setup_tables(const cfdnn_mesh_desc& m) {
    fill_twiddles(twx_, Nx_);
    fill_twiddles(twz_, Nzg_);
    // compute_eigenvalues, second-order branch (src/poisson_solver_fft.cpp:494-504)
    std::vector<double> lx(Nxh_), lz(Nzg_);
    for (int k = 0; k < Nxh_; ++k) lx[k] = (2.0 - 2.0 * std::cos(2.0 * M_PI * k / Nx_)) / (m.dx * m.dx);
    for (int k = 0; k < Nzg_; ++k) lz[k] = (2.0 - 2.0 * std::cos(2.0 * M_PI * k / Nzg_)) / (m.dz * m.dz);
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
    inv_Ly_ = 1.0 / Ly;
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
    if (!planned_) throw Error(CFDNN_ERR_RUNTIME, "FFT Poisson: this mesh is a z-slab of a larger box; comm_init first");
    if (!rhs || !p) throw std::invalid_argument("FFT Poisson: null field pointer");
    Radices rx, rz;
    rx.n = nrx_; rz.n = nrz_;
    std::memcpy(rx.r, radx_, sizeof(rx.r));
    std::memcpy(rz.r, radz_, sizeof(rz.r));
    const int rows = Ny_ * Nz_;
    double2* S = spec_.p;                            // [q][k_local][j][c]
    double2* T = P_ > 1 ? spec2_.p : spec_.p;        // [kz][j][c]
    const size_t blk = (size_t)rows * Pq_ * 2;       // doubles per all-to-all block
    const size_t bel = (size_t)rows * Pq_;           // complex elements per block
    BlockPtrs to_T{}, to_S{};                        // where x_fwd / the inverse z transform store block q
    for (int q = 0; q < P_; ++q) {
        to_T.b[q] = p2p_ ? peer_T_[q] + bel * rank_ : S + bel * q;      // p2p: straight into the owner's T
        to_S.b[q] = p2p_ ? peer_S_[q] + bel * rank_ : T + bel * q;      // p2p: straight into the owner's S
    }
    auto barrier = [&] { comm_->allreduce_sum(sync_.p, 1, st); };      // stream-ordered: every rank's stores have landed
    // 1. forward transforms
    const int xlines = (rows + 1) / 2;
    const size_t smx = (size_t)2 * Nx_ * Lx_ * sizeof(double2), smz = (size_t)2 * Nzg_ * Lz_ * sizeof(double2);
    CFDNN_LAUNCH_COOP(k_fft_x_fwd, (xlines + Lx_ - 1) / Lx_, FFT_THREADS, smx, st, rhs, to_T, Nx_, Ny_, Nz_, Ng_, cs_, cp_, Pq_,
                      Lx_, ilog2(Lx_), rx, twx_.p);
    if (p2p_) barrier();
    else if (P_ > 1) comm_->alltoall(reinterpret_cast<const double*>(S), reinterpret_cast<double*>(T), blk, st);
    const dim3 gz(Pq_ / Lz_, Ny_);
    CFDNN_LAUNCH_COOP((k_fft_z<false>), gz, FFT_THREADS, smz, st, T, to_S, Nz_, Ny_, Nzg_, Pq_, Lz_, ilog2(Lz_), rz, twz_.p);
    // 2. tridiagonal solves (with the mean removal on the zero mode)
    const dim3 gt((Pq_ + 127) / 128, Nzg_);
    CFDNN_LAUNCH((k_fft_thomas<8>), gt, 128, 0, st, T, cprime_.p, invden_.p, lo_.p, dyv_.p, inv_Ly_, Ny_, Nxh_, Pq_,
                 rank_ * Pq_);
    // 3. inverse transforms, normalisation, ghosts (x in the kernel, then y rows, then the periodic z planes)
    CFDNN_LAUNCH_COOP((k_fft_z<true>), gz, FFT_THREADS, smz, st, T, to_S, Nz_, Ny_, Nzg_, Pq_, Lz_, ilog2(Lz_), rz, twz_.p);
    if (p2p_) barrier();
    else if (P_ > 1) comm_->alltoall(reinterpret_cast<const double*>(T), reinterpret_cast<double*>(S), blk, st);
    CFDNN_LAUNCH_COOP(k_fft_x_inv, (xlines + Lx_ - 1) / Lx_, FFT_THREADS, smx, st, S, p, Nx_, Ny_, Nz_, Ng_, cs_, cp_, Pq_,
                      Lx_, ilog2(Lx_), rx, twx_.p, 1.0 / ((double)Nx_ * (double)Nzg_));
    CFDNN_LAUNCH(k_fft_ghost_y, dim3((cs_ + 127) / 128, Nz_), 128, 0, st, p, cs_, Ny_, Nz_, Ng_, cs_, cp_);
    if (P_ > 1) {
        comm_->exchange_planes(p, (size_t)cp_, Nz_, Ng_, true, true, st);       // periodic ring of slabs
    } else {
        CFDNN_CUDA(cudaMemcpyAsync(p, p + (size_t)Nz_ * cp_, (size_t)cp_ * sizeof(double), cudaMemcpyDeviceToDevice, st));
        CFDNN_CUDA(cudaMemcpyAsync(p + (size_t)(Nz_ + 1) * cp_, p + (size_t)cp_, (size_t)cp_ * sizeof(double),
                                   cudaMemcpyDeviceToDevice, st));
    }
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
