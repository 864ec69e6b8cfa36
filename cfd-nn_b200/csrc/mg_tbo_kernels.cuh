// mg_tbo_kernels.cuh — the temporally blocked smoother on levels with no physical face in x or y
// and open z sides (periodic, or slab interfaces with T halo planes): the 3-D periodic boxes of
// the headline workload, on one GPU and on z-slabs.
//
// Same tiling, thread mapping, shared-memory layout and per-cell expressions as k_sweep_tb
// (mg_tb_kernels.cuh) — two Chebyshev sweeps of smooth_chebyshev
// (src/poisson_solver_multigrid.cpp:825-1001) per pass over HBM — with the per-plane overhead of
// the general kernel removed:
//   * no boundary selects and no per-thread validity predicates (every load is in range);
//   * global addresses advance by one plane per iteration (uniform offset + a per-thread 32-bit
//     offset) instead of being rebuilt from (z, wrap, halo) every plane;
//   * register rings of 4 slots with the z loop unrolled by 4: the prefetched plane lands in the
//     slot it is consumed from, so no register moves between iterations;
//   * FD = true (reciprocal multiplies, LevelGeom::fastdiv): the update is evaluated with fused
//     multiply-adds, 9 FP64 instructions per cell-sweep instead of 13 (22 with exact divisions).
//     Each partial result is within 1 ulp of the reference's; FD = false keeps the reference's
//     expression and rounding bit for bit;
//   * the register prefetch of plane z+1 is issued right AFTER the first use of plane z (pinned by an
//     always-true predicate computed from the arriving data in an opaque asm).  ptxas puts
//     all global loads of the loop on one hardware scoreboard, and a scoreboard wait covers every
//     load outstanding on it: with the prefetch at the top of the iteration the wait for plane z
//     also waited for plane z+1, i.e. the full DRAM latency was exposed once per plane (measured:
//     0.89 ms per launch at 512^3, 0.71 ms with the same loads hitting L2, 0.47 ms without loads;
//     0.70 ms after the move).  Deeper pipelines were tried and lost: cp.async (LDGSTS) stages 3
//     planes deep ran 30% slower, prefetch.global.L2 two or four planes ahead 15% slower
//     (profiles/r1/tbo_experiments.log).
#pragma once
#include "mg_tb_kernels.cuh"
#include <type_traits>

namespace cfdnn {

template <bool FD>
struct TboCoef {
    double rx, ry, rz;      // 1/dx2, 1/dy2, 1/dz2
    double wc[2];           // omega_s / coeff
    double om1[2];          // 1 - omega_s
};

template <bool FD>
__device__ __forceinline__ double tbo_point(const LevelGeom& g, const TboCoef<FD>& c, int s, double uc, double uW,
                                            double uE, double uS, double uN, double ub, double uf, double fv,
                                            double omega) {
    if (FD) {
        double a = (uE + uW) * c.rx;
        a = fma(uN + uS, c.ry, a);
        a = fma(uf + ub, c.rz, a);
        a = a - fv;
        return fma(c.wc[s], a, c.om1[s] * uc);
    } else {
        const double uj = ((uE + uW) / g.dx2 + (uN + uS) / g.dy2 + (uf + ub) / g.dz2 - fv) / g.coeff;
        return c.om1[s] * uc + omega * uj;
    }
}

// Split barrier (MODE 2): a thread ARRIVES once its shared-memory reads and writes of the iteration
// are done and only WAITS at the top of the next iteration, so the second stage's arithmetic and the
// global store of a fast warp overlap the slower warps instead of idling at a full barrier.  All
// reads of the previous iteration's planes happen before the arrive, so two buffers per level suffice.
#ifndef CFDNN_DEVSIM
__device__ __forceinline__ void tbo_bar_init(unsigned long long* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void tbo_arrive(unsigned long long* bar) {   // one arrival per warp
    __syncwarp();
    if ((threadIdx.x & 31) == 0)
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void tbo_wait(unsigned long long* bar, unsigned parity) {
    unsigned ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"((unsigned)__cvta_generic_to_shared(bar)), "r"(parity) : "memory");
    } while (!ok);
}
#else   // CPU test double: one full barrier per plane, at the wait
inline void tbo_bar_init(unsigned long long*, unsigned) {}
inline void tbo_arrive(unsigned long long*) {}
inline void tbo_wait(unsigned long long*, unsigned) { __syncthreads(); }
#endif

// MODE 0: neighbours gathered per stage, full barrier per plane; 1: both stages' neighbours gathered
// at the top of the iteration; 2: as 1 with the split barrier.
template <int BX, int BY, bool FD, int MODE>
__global__ void __launch_bounds__(TbShape<2, BX, BY>::NTHREADS, 1)
k_sweep_tbo(LevelGeom g, const double* __restrict__ u, const double* __restrict__ f, double* __restrict__ out,
            TbOmega<2> om, int ZC, int zoff, int zstep) {
    constexpr int T = 2;
    using S = TbShape<T, BX, BY>;
    CFDNN_DYN_SMEM(double, sm);
    constexpr int NT = S::NTHREADS;
    constexpr int PL = 4 * NT;                             // doubles per smem plane
    const int t = threadIdx.x;
    const int lane = t & 31;
    const bool active = t < S::NACTIVE;
    const int row = active ? t / S::NQ : 0;
    const int q = active ? t % S::NQ : 0;
    const int lx = 4 * q;
    const int gx0 = (int)blockIdx.x * BX - S::HX + lx;
    const int gy = (int)blockIdx.y * BY - T + row;
    const int z0 = (zoff + (int)blockIdx.z * zstep) * ZC;   // z chunk (zoff, zstep: a launch may cover a subset of chunks)
    const int z1 = min(z0 + ZC, g.Nz);
    const bool perz = g.bc[4] == CFDNN_PBC_PERIODIC;

    // per-thread in-plane offsets (x and y wrap; Nx is even so a pair never straddles the seam)
    const int gyw = wrap_idx(gy, g.Ny);
    unsigned off[2];
    bool stok[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int gxp = gx0 + 2 * h;
        off[h] = (unsigned)(gyw * g.Px + wrap_idx(gxp, g.Nx));
        const int tx = lx + 2 * h;
        stok[h] = active && tx >= S::HX && tx < S::HX + BX && row >= T && row < T + BY &&
                  gxp >= 0 && gxp + 1 < g.Nx && gy >= 0 && gy < g.Ny;
    }
    // smem plane = [2 pairs][NT threads] double2
    const int tS = row > 0 ? t - S::NQ : t;
    const int tN = (active && row < S::TY - 1) ? t + S::NQ : t;
    const bool wsm = (lane == 0) && q > 0;                 // W neighbour must come from smem
    const bool esm = (lane == 31) && q < S::NQ - 1;

    TboCoef<FD> c;
    c.rx = g.dx2.r; c.ry = g.dy2.r; c.rz = g.dz2.r;
#pragma unroll
    for (int s = 0; s < T; ++s) { c.wc[s] = om.w[s] * g.coeff.r; c.om1[s] = 1.0 - om.w[s]; }

    // rings of 4 slots (slot = iteration mod 4): level-0 / level-1 planes and f planes
    double l0[4][4], l1[4][4], fr[4][4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int p = 0; p < 4; ++p) { l0[k][p] = 0.0; l1[k][p] = 0.0; fr[k][p] = 0.0; }

    // plane cursor: zl = stored-plane index of the NEXT u plane to load (f lags by one plane)
    const size_t plane = (size_t)g.plane;
    const int n_it = (z1 - z0) + 2 * T;
    int zu = z0 - T;                                       // logical z of the next u plane
    if (perz && zu < 0) zu += g.Nz;
    size_t zoff_u = (size_t)(zu + g.Hz) * plane;
    size_t zoff_f = 0;                                     // valid from the first prefetch on
    auto advance = [&](int& z, size_t& zoff) {
        ++z; zoff += plane;
        if (perz && z == g.Nz) { z = 0; zoff = (size_t)g.Hz * plane; }
    };
    auto load4 = [&](const double* __restrict__ a, size_t zoff, double* dst) {
        const double* base = a + zoff;
        ld_pair(base + off[0], dst[0], dst[1]);
        ld_pair(base + off[1], dst[2], dst[3]);
    };
    load4(u, zoff_u, l0[0]);
    int zf_ = zu;                                          // f plane to load at the first prefetch = u plane just loaded
    zoff_f = zoff_u;
    advance(zu, zoff_u);

    __shared__ unsigned long long bar_storage;
    unsigned long long* bar = &bar_storage;
    unsigned phase = 0u;                                   // parity waited on at the top of the iteration
    if (MODE == 2) {
        if (t == 0) tbo_bar_init(bar, NT / 32);
        __syncthreads();
        tbo_arrive(bar);                                   // phase 0: nothing to wait for in the first iteration
    }
    const double* rb = sm + PL;                            // read buffer (published one iteration ago)
    double* wb = sm;                                       // write buffer
    int zo_store = z0 - T - T;                             // level-2 plane produced by stage 1 of this iteration

    auto iter = [&](auto PH, bool prefetch) {
        constexpr int A = decltype(PH)::value, C = (A + 3) % 4, M = (A + 2) % 4, N = (A + 1) % 4;
        if (MODE == 2) tbo_wait(bar, phase);               // planes published in the previous iteration are visible
        const double* rb1 = rb + 2 * PL;
        double* wb1 = wb + 2 * PL;
        // ---- level-0 plane zi arrives: publish it
        st_pair(wb + 2 * t, l0[A][0], l0[A][1]);
        st_pair(wb + 2 * NT + 2 * t, l0[A][2], l0[A][3]);
        // The prefetch must be issued AFTER the wait for the previous one: all global loads of the loop
        // share one hardware scoreboard, so a wait placed after the new loads also waits for those and
        // exposes the full DRAM latency every plane.  ptxas hoists independent loads freely, so the
        // order is pinned by a data dependency that never changes what is loaded: the loads are predicated
        // on (low word of an arriving value | 1) != 0, evaluated in an opaque asm so that ptxas can neither
        // fold it nor move the loads above the arrival — true for every bit pattern, NaN included.
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
        if (prefetch) {                                    // next iteration's planes
            if (arrived) {
                load4(u, zoff_u, l0[N]);
                load4(f, zoff_f, fr[N]);
            }
            advance(zu, zoff_u);
            advance(zf_, zoff_f);
        }
        // ---- everything that reads the previous iteration's planes: E/W by shuffle (warp seams from
        // smem), N/S rows from smem — for both stages when MODE >= 1, so stage 1 only waits for stage 0
        double n0[4], s0[4], n1[4], s1[4];
        double w0 = __shfl_up_sync(0xffffffffu, l0[C][3], 1);
        double e0 = __shfl_down_sync(0xffffffffu, l0[C][0], 1);
        double w1, e1;
        if (wsm) w0 = rb[2 * NT + 2 * (t - 1) + 1];
        if (esm) e0 = rb[2 * (t + 1)];
        ld_pair(rb + 2 * tN, n0[0], n0[1]); ld_pair(rb + 2 * NT + 2 * tN, n0[2], n0[3]);
        ld_pair(rb + 2 * tS, s0[0], s0[1]); ld_pair(rb + 2 * NT + 2 * tS, s0[2], s0[3]);
        auto gather1 = [&]() {
            w1 = __shfl_up_sync(0xffffffffu, l1[C][3], 1);
            e1 = __shfl_down_sync(0xffffffffu, l1[C][0], 1);
            if (wsm) w1 = rb1[2 * NT + 2 * (t - 1) + 1];
            if (esm) e1 = rb1[2 * (t + 1)];
            ld_pair(rb1 + 2 * tN, n1[0], n1[1]); ld_pair(rb1 + 2 * NT + 2 * tN, n1[2], n1[3]);
            ld_pair(rb1 + 2 * tS, s1[0], s1[1]); ld_pair(rb1 + 2 * NT + 2 * tS, s1[2], s1[3]);
        };
        if (MODE >= 1) gather1();
        // ---- stage 0: produce level-1 plane zi - 1
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const double uW = p == 0 ? w0 : l0[C][p - 1];
            const double uE = p == 3 ? e0 : l0[C][p + 1];
            l1[A][p] = tbo_point<FD>(g, c, 0, l0[C][p], uW, uE, s0[p], n0[p], l0[M][p], l0[A][p], fr[A][p], om.w[0]);
        }
        // ---- stage 1: level-1 plane zi - 1 arrived; produce level-2 plane zi - 2
        st_pair(wb1 + 2 * t, l1[A][0], l1[A][1]);
        st_pair(wb1 + 2 * NT + 2 * t, l1[A][2], l1[A][3]);
        if (MODE == 2) tbo_arrive(bar);                    // this thread's reads and writes of the iteration are done
        if (MODE == 0) gather1();
        double o[4];
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const double uW = p == 0 ? w1 : l1[C][p - 1];
            const double uE = p == 3 ? e1 : l1[C][p + 1];
            o[p] = tbo_point<FD>(g, c, 1, l1[C][p], uW, uE, s1[p], n1[p], l1[M][p], l1[A][p], fr[C][p], om.w[1]);
        }
        if (zo_store >= z0) {                              // (zo_store < z1 by construction of n_it)
            double* base = out + (size_t)(zo_store + g.Hz) * plane;
            if (stok[0]) st_pair(base + off[0], o[0], o[1]);
            if (stok[1]) st_pair(base + off[1], o[2], o[3]);
        }
        ++zo_store;
        if (MODE != 2) __syncthreads();
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
