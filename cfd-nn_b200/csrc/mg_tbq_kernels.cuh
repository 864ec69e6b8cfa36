// mg_tbq_kernels.cuh — the lean temporally blocked smoother (k_sweep_tbo, mg_tbo_kernels.cuh) with a 2 x 2
// cell patch per thread instead of a 4 x 1 quad.
//
// Same algorithm, tiles, z-march, register rings, split barrier and per-cell arithmetic as k_sweep_tbo: two
// Chebyshev sweeps of smooth_chebyshev (src/poisson_solver_multigrid.cpp:825-1001) per pass over HBM on levels
// with no physical face in reach.  k_sweep_tbo is bound by shared-memory / L1 transactions (l1tex 90 %, DRAM
// 57 % at 512^3, profiles/r1/ncu_k_sweep_tbo_512.txt): per plane and level a thread stored its quad (2 STS.128)
// and fetched the quads north and south of it (4 LDS.128), and its global loads / stores were 16-byte pieces at a
// 32-byte stride (every 32-byte sector requested twice).  With a 2 (x) x 2 (y) patch
//   * each cell has ONE of its north / south neighbours in the thread's own registers: 2 LDS.128 per plane and
//     level instead of 4 (the row below the patch and the row above it);
//   * a warp's 16-byte accesses are contiguous in shared memory (planes are stored in their natural
//     [row][column] layout) and in global memory (each LDG.128 / STG.128 of a warp covers one 512-byte run);
//   * east / west neighbours are the other column of the patch or one shuffle away (4 shuffles per plane and
//     level instead of 2).
// Per cell-sweep arithmetic is tbo_point(): bitwise the same results as k_sweep_tbo in both division modes.
#pragma once
#include "mg_tbo_kernels.cuh"

namespace cfdnn {

template <int BX, int BY>
struct TbqShape {
    static constexpr int T = 2;
    static constexpr int HX = 2;
    static constexpr int TX = BX + 2 * HX, TY = BY + 2 * T;
    static_assert(TX % 2 == 0 && TY % 2 == 0, "tile must split into 2 x 2 patches");
    static constexpr int NC = TX / 2, NR = TY / 2;        // patches per tile row / column
    static constexpr int NACTIVE = NC * NR;
    static constexpr int NTHREADS = (NACTIVE + 31) / 32 * 32;
    static constexpr int PL = TX * TY;                     // doubles per plane buffer (natural layout)
    static constexpr size_t SMEM = sizeof(double) * (size_t)PL * 2 * T;
};

// WALLY = true: y is NOT periodic (channel walls: Neumann or Dirichlet on both y faces; x periodic, z open as before).
// Tiles and tile rows start at even y, so only the lower row of a patch can be the first cell row (its south
// neighbour is the ghost = sign * itself, apply_bc :307-554) and only the upper row the last one; rows outside the
// domain are loaded from a clamped row and feed no stored cell.
template <int BX, int BY, bool FD, bool WALLY = false>
__global__ void __launch_bounds__(TbqShape<BX, BY>::NTHREADS, 1)
k_sweep_tbq(LevelGeom g, const double* __restrict__ u, const double* __restrict__ f, double* __restrict__ out,
            TbOmega<2> om, int ZC, int zoff, int zstep) {
    using S = TbqShape<BX, BY>;
    constexpr int T = 2, TX = S::TX, PL = S::PL;
    CFDNN_DYN_SMEM(double, sm);
    const int t = threadIdx.x;
    const int lane = t & 31;
    const bool active = t < S::NACTIVE;
    const int r = active ? t / S::NC : 0;                  // patch row / column in the tile
    const int c = active ? t % S::NC : 0;
    const int gx = (int)blockIdx.x * BX - S::HX + 2 * c;
    const int gya = (int)blockIdx.y * BY - T + 2 * r;
    const int z0 = (zoff + (int)blockIdx.z * zstep) * ZC;   // z chunk (zoff, zstep: a launch may cover a subset of chunks)
    const int z1 = min(z0 + ZC, g.Nz);
    const bool perz = g.bc[4] == CFDNN_PBC_PERIODIC;

    // per-thread in-plane offsets of the two patch rows (x and y wrap; Nx is even so a pair never straddles the seam)
    const int gxw = wrap_idx(gx, g.Nx);
    unsigned off[2];
    bool stok[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int gy = gya + h;
        off[h] = (unsigned)((WALLY ? min(max(gy, 0), g.Ny - 1) : wrap_idx(gy, g.Ny)) * g.Px + gxw);
        const int tx = 2 * c, ty = 2 * r + h;
        stok[h] = active && tx >= S::HX && tx < S::HX + BX && ty >= T && ty < T + BY &&
                  gx >= 0 && gx + 1 < g.Nx && gy >= 0 && gy < g.Ny;
    }
    // shared-memory offsets (doubles) in a plane buffer: own rows, the row below and the row above the patch
    const int oA = (2 * r) * TX + 2 * c, oB = oA + TX;
    const int oS = r > 0 ? oA - TX : oA;                   // tile edge: any valid address (feeds no stored cell)
    const int oN = (active && r < S::NR - 1) ? oB + TX : oB;
    const bool wsm = (lane == 0) && c > 0;                 // west neighbours must come from shared memory (warp seam)
    const bool esm = (lane == 31) && c < S::NC - 1;
    const bool selfS = WALLY && gya == 0, selfN = WALLY && gya + 1 == g.Ny - 1;
    const double sylo = g.bc[2] == CFDNN_PBC_DIRICHLET ? -1.0 : 1.0, syhi = g.bc[3] == CFDNN_PBC_DIRICHLET ? -1.0 : 1.0;

    TboCoef<FD> cf;
    cf.rx = g.dx2.r; cf.ry = g.dy2.r; cf.rz = g.dz2.r;
#pragma unroll
    for (int s = 0; s < T; ++s) { cf.wc[s] = om.w[s] * g.coeff.r; cf.om1[s] = 1.0 - om.w[s]; }

    // rings of 4 slots (slot = iteration mod 4); element order: [row a: x, x+1 | row b: x, x+1]
    double l0[4][4], l1[4][4], fr[4][4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int p = 0; p < 4; ++p) { l0[k][p] = 0.0; l1[k][p] = 0.0; fr[k][p] = 0.0; }

    const size_t plane = (size_t)g.plane;
    const int n_it = (z1 - z0) + 2 * T;
    int zu = z0 - T;                                       // logical z of the next u plane
    if (perz && zu < 0) zu += g.Nz;
    size_t zoff_u = (size_t)(zu + g.Hz) * plane;
    size_t zoff_f = 0;
    auto advance = [&](int& z, size_t& zo) {
        ++z; zo += plane;
        if (perz && z == g.Nz) { z = 0; zo = (size_t)g.Hz * plane; }
    };
    auto load4 = [&](const double* __restrict__ a, size_t zo, double* dst) {
        const double* base = a + zo;
        ld_pair(base + off[0], dst[0], dst[1]);
        ld_pair(base + off[1], dst[2], dst[3]);
    };
    load4(u, zoff_u, l0[0]);
    int zf_ = zu;
    zoff_f = zoff_u;
    advance(zu, zoff_u);

    __shared__ unsigned long long bar_storage;
    unsigned long long* bar = &bar_storage;
    unsigned phase = 0u;
    if (t == 0) tbo_bar_init(bar, S::NTHREADS / 32);
    __syncthreads();
    tbo_arrive(bar);                                       // phase 0: nothing to wait for in the first iteration
    const double* rb = sm + PL;                            // level-0 plane published one iteration ago
    double* wb = sm;
    int zo_store = z0 - T - T;

    // neighbours of the centre plane v (published in `buf` one iteration ago) for the four cells of the patch
    auto gather = [&](const double* __restrict__ buf, const double (&v)[4], double (&W)[2], double (&E)[2],
                      double (&Sv)[2], double (&Nv)[2]) {
        W[0] = __shfl_up_sync(0xffffffffu, v[1], 1);
        W[1] = __shfl_up_sync(0xffffffffu, v[3], 1);
        E[0] = __shfl_down_sync(0xffffffffu, v[0], 1);
        E[1] = __shfl_down_sync(0xffffffffu, v[2], 1);
        if (wsm) { W[0] = buf[oA - 1]; W[1] = buf[oB - 1]; }
        if (esm) { E[0] = buf[oA + 2]; E[1] = buf[oB + 2]; }
        ld_pair(buf + oS, Sv[0], Sv[1]);
        ld_pair(buf + oN, Nv[0], Nv[1]);
    };

    auto iter = [&](auto PH, bool prefetch) {
        constexpr int A = decltype(PH)::value, C = (A + 3) % 4, M = (A + 2) % 4, N = (A + 1) % 4;
        tbo_wait(bar, phase);                              // planes published in the previous iteration are visible
        const double* rb1 = rb + 2 * PL;
        double* wb1 = wb + 2 * PL;
        // ---- level-0 plane zi arrives: publish it
        st_pair(wb + oA, l0[A][0], l0[A][1]);
        st_pair(wb + oB, l0[A][2], l0[A][3]);
        // prefetch of the next plane, pinned after the arrival of this one (see mg_tbo_kernels.cuh)
        bool arrived;
        {
#ifndef CFDNN_DEVSIM
            unsigned lw = (unsigned)__double2loint(l0[A][0]);
            asm volatile("or.b32 %0, %0, 1;" : "+r"(lw));
            arrived = lw != 0u;
#else
            arrived = true;
#endif
        }
        if (prefetch) {
            if (arrived) {
                load4(u, zoff_u, l0[N]);
                load4(f, zoff_f, fr[N]);
            }
            advance(zu, zoff_u);
            advance(zf_, zoff_f);
        }
        // ---- everything that reads the previous iteration's planes, for both stages
        double w0[2], e0[2], s0[2], n0[2], w1[2], e1[2], s1[2], n1[2];
        gather(rb, l0[C], w0, e0, s0, n0);
        gather(rb1, l1[C], w1, e1, s1, n1);
        // ---- stage 0: level-1 plane zi - 1.  Cell order: (a, x) (a, x+1) (b, x) (b, x+1)
        {
            const double* v = l0[C];
            if (WALLY) {
                if (selfS) { s0[0] = sylo * v[0]; s0[1] = sylo * v[1]; }
                if (selfN) { n0[0] = syhi * v[2]; n0[1] = syhi * v[3]; }
            }
            l1[A][0] = tbo_point<FD>(g, cf, 0, v[0], w0[0], v[1], s0[0], v[2], l0[M][0], l0[A][0], fr[A][0], om.w[0]);
            l1[A][1] = tbo_point<FD>(g, cf, 0, v[1], v[0], e0[0], s0[1], v[3], l0[M][1], l0[A][1], fr[A][1], om.w[0]);
            l1[A][2] = tbo_point<FD>(g, cf, 0, v[2], w0[1], v[3], v[0], n0[0], l0[M][2], l0[A][2], fr[A][2], om.w[0]);
            l1[A][3] = tbo_point<FD>(g, cf, 0, v[3], v[2], e0[1], v[1], n0[1], l0[M][3], l0[A][3], fr[A][3], om.w[0]);
        }
        // ---- stage 1: level-1 plane zi - 1 arrived; level-2 plane zi - 2
        st_pair(wb1 + oA, l1[A][0], l1[A][1]);
        st_pair(wb1 + oB, l1[A][2], l1[A][3]);
        tbo_arrive(bar);                                   // this thread's shared-memory reads and writes of the iteration are done
        double o[4];
        {
            const double* v = l1[C];
            if (WALLY) {
                if (selfS) { s1[0] = sylo * v[0]; s1[1] = sylo * v[1]; }
                if (selfN) { n1[0] = syhi * v[2]; n1[1] = syhi * v[3]; }
            }
            o[0] = tbo_point<FD>(g, cf, 1, v[0], w1[0], v[1], s1[0], v[2], l1[M][0], l1[A][0], fr[C][0], om.w[1]);
            o[1] = tbo_point<FD>(g, cf, 1, v[1], v[0], e1[0], s1[1], v[3], l1[M][1], l1[A][1], fr[C][1], om.w[1]);
            o[2] = tbo_point<FD>(g, cf, 1, v[2], w1[1], v[3], v[0], n1[0], l1[M][2], l1[A][2], fr[C][2], om.w[1]);
            o[3] = tbo_point<FD>(g, cf, 1, v[3], v[2], e1[1], v[1], n1[1], l1[M][3], l1[A][3], fr[C][3], om.w[1]);
        }
        if (zo_store >= z0) {
            double* base = out + (size_t)(zo_store + g.Hz) * plane;
            if (stok[0]) st_pair(base + off[0], o[0], o[1]);
            if (stok[1]) st_pair(base + off[1], o[2], o[3]);
        }
        ++zo_store;
        phase ^= 1u;
        { const double* tmp = rb; rb = wb; wb = const_cast<double*>(tmp); }
    };

    for (int it = 0; it < n_it; it += 4) {
        iter(std::integral_constant<int, 0>{}, it + 1 < n_it);
        if (it + 1 >= n_it) break;
        iter(std::integral_constant<int, 1>{}, it + 2 < n_it);
        if (it + 2 >= n_it) break;
        iter(std::integral_constant<int, 2>{}, it + 3 < n_it);
        if (it + 3 >= n_it) break;
        iter(std::integral_constant<int, 3>{}, it + 4 < n_it);
    }
}

}  // namespace cfdnn

