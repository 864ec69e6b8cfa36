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

template <int BX, int BY, bool FD>
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
        off[h] = (unsigned)(wrap_idx(gy, g.Ny) * g.Px + gxw);
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

#ifndef CFDNN_DEVSIM
namespace cfdnn {

// ---------------------------------------------------------------------------------------------------------
// k_sweep_tbx — k_sweep_tbq with the level-0 planes delivered by the TMA unit.
//
// k_sweep_tbq still spends its time waiting for global loads (half of its stall samples are long-scoreboard,
// profiles/r2/ncu_k_sweep_tbq_512.txt): with one 512-thread CTA per SM and 122 registers there is room for ONE
// plane of register prefetch, less than the ~35 KB per SM that must be in flight to cover HBM latency at full
// bandwidth.  Here the u planes arrive through bulk asynchronous copies (cp.async.bulk, one per tile row, the x
// wrap of an edge tile as two copies) straight into a ring of NS level-0 plane buffers in the natural
// [row][column] layout k_sweep_tbq publishes them in anyway: NS - 2 planes (33-49 KB) are in flight per SM
// without a register, the loads and the level-0 publish stores leave the LSU path (16 of ~100 l1tex
// wavefronts per warp and plane), and completion is an mbarrier with a transaction count instead of a
// scoreboard.  f (read once, by its owning thread only) stays a register prefetch, now two planes ahead in the
// ring slot the u planes no longer need.  Arithmetic: tbo_point(), bitwise the results of k_sweep_tbq.
template <int BX, int BY, int NS>
struct TbxShape {
    using Q = TbqShape<BX, BY>;
    static constexpr size_t SMEM = sizeof(double) * (size_t)Q::PL * (NS + 2) + 128;   // NS level-0 stages + 2 level-1 buffers (+ alignment slack)
};

__device__ __forceinline__ void tbx_expect(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                 ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tbx_bulk_g2s(double* dst_smem, const double* src_gmem, unsigned bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"((unsigned)__cvta_generic_to_shared(dst_smem)), "l"(src_gmem), "r"(bytes),
                   "r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void tbx_wait(unsigned long long* bar, unsigned parity) {
    unsigned ok, spins = 0;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"((unsigned)__cvta_generic_to_shared(bar)), "r"(parity) : "memory");
        if (!ok && ++spins > (1u << 26)) __trap();        // a lost copy must fault, not hang the device
    } while (!ok);
}

template <int BX, int BY, bool FD, int NS>
__global__ void __launch_bounds__(TbqShape<BX, BY>::NTHREADS, 1)
k_sweep_tbx(LevelGeom g, const double* __restrict__ u, const double* __restrict__ f, double* __restrict__ out,
            TbOmega<2> om, int ZC, int zoff, int zstep) {
    using S = TbqShape<BX, BY>;
    constexpr int T = 2, TX = S::TX, TY = S::TY, PL = S::PL;
    static_assert(NS >= 3 && TY <= 32, "one lane of warp 0 per tile row");
    extern __shared__ unsigned char tbx_raw[];
    double* sm = reinterpret_cast<double*>((reinterpret_cast<size_t>(tbx_raw) + 127) & ~(size_t)127);
    double* const st0 = sm;                                // level-0 ring: NS planes
    double* const lv1 = sm + (size_t)NS * PL;              // level-1 double buffer
    __shared__ unsigned long long full[NS];
    __shared__ unsigned long long bar_storage;
    const int t = threadIdx.x;
    const int lane = t & 31;
    const bool active = t < S::NACTIVE;
    const int r = active ? t / S::NC : 0;
    const int c = active ? t % S::NC : 0;
    const int gx = (int)blockIdx.x * BX - S::HX + 2 * c;
    const int gya = (int)blockIdx.y * BY - T + 2 * r;
    const int z0 = (zoff + (int)blockIdx.z * zstep) * ZC;
    const int z1 = min(z0 + ZC, g.Nz);
    const bool perz = g.bc[4] == CFDNN_PBC_PERIODIC;
    const int gxw = wrap_idx(gx, g.Nx);
    unsigned off[2];
    bool stok[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int gy = gya + h;
        off[h] = (unsigned)(wrap_idx(gy, g.Ny) * g.Px + gxw);
        const int tx = 2 * c, ty = 2 * r + h;
        stok[h] = active && tx >= S::HX && tx < S::HX + BX && ty >= T && ty < T + BY &&
                  gx >= 0 && gx + 1 < g.Nx && gy >= 0 && gy < g.Ny;
    }
    const int oA = (2 * r) * TX + 2 * c, oB = oA + TX;
    const int oS = r > 0 ? oA - TX : oA;
    const int oN = (active && r < S::NR - 1) ? oB + TX : oB;
    const bool wsm = (lane == 0) && c > 0;
    const bool esm = (lane == 31) && c < S::NC - 1;

    TboCoef<FD> cf;
    cf.rx = g.dx2.r; cf.ry = g.dy2.r; cf.rz = g.dz2.r;
#pragma unroll
    for (int s = 0; s < T; ++s) { cf.wc[s] = om.w[s] * g.coeff.r; cf.om1[s] = 1.0 - om.w[s]; }

    double l0[4][4], l1[4][4], fr[4][4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int p = 0; p < 4; ++p) { l0[k][p] = 0.0; l1[k][p] = 0.0; fr[k][p] = 0.0; }

    const size_t plane = (size_t)g.plane;
    const int n_it = (z1 - z0) + 2 * T;
    // stored plane offset of iteration `it` (u plane z0 - T + it; periodic wrap, halo planes are stored planes)
    auto zoff_of = [&](int it) {
        int z = z0 - T + it;
        if (perz) z = z < 0 ? z + g.Nz : (z >= g.Nz ? z - g.Nz : z);
        return (size_t)(z + g.Hz) * plane;
    };
    // ---- producer (the first TY lanes of warp 0): one tile row each, the periodic x seam as two copies
    const int gx_lo = (int)blockIdx.x * BX - S::HX;
    int segA_src = 0, segA_len = TX, segB_len = 0;         // doubles; segment B starts at source x = 0
    if (gx_lo < 0) { segA_src = gx_lo + g.Nx; segA_len = -gx_lo; segB_len = TX + gx_lo; }
    else if (gx_lo + TX > g.Nx) { segA_src = gx_lo; segA_len = g.Nx - gx_lo; segB_len = gx_lo + TX - g.Nx; }
    else segA_src = gx_lo;
    const unsigned prow = (unsigned)(wrap_idx((int)blockIdx.y * BY - T + min(t, TY - 1), g.Ny) * g.Px);
    auto issue = [&](int it) {                             // called by warp 0 only
        const int s = it % NS;
        if (lane == 0) tbx_expect(&full[s], (unsigned)(PL * sizeof(double)));
        __syncwarp();
        if (lane < TY) {
            const double* src = u + zoff_of(it) + prow;
            double* dst = st0 + (size_t)s * PL + lane * TX;
            tbx_bulk_g2s(dst, src + segA_src, (unsigned)(segA_len * sizeof(double)), &full[s]);
            if (segB_len) tbx_bulk_g2s(dst + segA_len, src, (unsigned)(segB_len * sizeof(double)), &full[s]);
        }
    };
    if (t == 0) {
        tbo_bar_init(&bar_storage, S::NTHREADS / 32);
#pragma unroll
        for (int s = 0; s < NS; ++s) tbo_bar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (t < 32) {
#pragma unroll
        for (int s = 0; s < NS; ++s)
            if (s < n_it) issue(s);
    }
    // f: register prefetch two planes ahead (slots A+1 and A+2 of the ring; u does not use its slots N and M here)
    auto load_f = [&](int it, double* dst) {               // f plane of iteration it = u plane index - 1, i.e. zoff_of(it - 1)
        const double* base = f + zoff_of(it - 1);
        ld_pair(base + off[0], dst[0], dst[1]);
        ld_pair(base + off[1], dst[2], dst[3]);
    };
    load_f(1, fr[1]);                                      // iteration `it` fetches the plane of iteration it + 2

    unsigned long long* bar = &bar_storage;
    unsigned phase = 0u;
    tbo_arrive(bar);
    int zo_store = z0 - T - T;
    int par1 = 0;                                          // level-1 double buffer parity

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

    auto iter = [&](auto PH, const int it) {
        constexpr int A = decltype(PH)::value, C = (A + 3) % 4, M = (A + 2) % 4;
        tbo_wait(bar, phase);                              // every thread is done with iteration it - 1
        // stage (it - 2) % NS held the centre plane of iteration it - 1: free now -> refill it NS planes ahead
        if (t < 32 && it >= 2 && it - 2 + NS < n_it) issue(it - 2 + NS);
        if (it + 2 < n_it) load_f(it + 2, fr[M]);          // fr[M] is read next as slot A of iteration it + 2
        tbx_wait(&full[it % NS], (unsigned)((it / NS) & 1));
        const double* cur = st0 + (size_t)(it % NS) * PL;          // arriving level-0 plane
        const double* rb = st0 + (size_t)((it + NS - 1) % NS) * PL;   // centre plane (arrived one iteration ago)
        ld_pair(cur + oA, l0[A][0], l0[A][1]);
        ld_pair(cur + oB, l0[A][2], l0[A][3]);
        const double* rb1 = lv1 + (size_t)(par1 ^ 1) * PL;
        double* wb1 = lv1 + (size_t)par1 * PL;
        double w0[2], e0[2], s0[2], n0[2], w1[2], e1[2], s1[2], n1[2];
        gather(rb, l0[C], w0, e0, s0, n0);
        gather(rb1, l1[C], w1, e1, s1, n1);
        {
            const double* v = l0[C];
            l1[A][0] = tbo_point<FD>(g, cf, 0, v[0], w0[0], v[1], s0[0], v[2], l0[M][0], l0[A][0], fr[A][0], om.w[0]);
            l1[A][1] = tbo_point<FD>(g, cf, 0, v[1], v[0], e0[0], s0[1], v[3], l0[M][1], l0[A][1], fr[A][1], om.w[0]);
            l1[A][2] = tbo_point<FD>(g, cf, 0, v[2], w0[1], v[3], v[0], n0[0], l0[M][2], l0[A][2], fr[A][2], om.w[0]);
            l1[A][3] = tbo_point<FD>(g, cf, 0, v[3], v[2], e0[1], v[1], n0[1], l0[M][3], l0[A][3], fr[A][3], om.w[0]);
        }
        st_pair(wb1 + oA, l1[A][0], l1[A][1]);
        st_pair(wb1 + oB, l1[A][2], l1[A][3]);
        tbo_arrive(bar);
        double o[4];
        {
            const double* v = l1[C];
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
        par1 ^= 1;
    };

    for (int it = 0; it < n_it; it += 4) {
        iter(std::integral_constant<int, 0>{}, it);
        if (it + 1 >= n_it) break;
        iter(std::integral_constant<int, 1>{}, it + 1);
        if (it + 2 >= n_it) break;
        iter(std::integral_constant<int, 2>{}, it + 2);
        if (it + 3 >= n_it) break;
        iter(std::integral_constant<int, 3>{}, it + 3);
    }
}

}  // namespace cfdnn
#endif  // !CFDNN_DEVSIM
