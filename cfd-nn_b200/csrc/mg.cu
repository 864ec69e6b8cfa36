// mg.cu — host orchestration of the multigrid V-cycle (see mg.cuh for the layout).
#include "mg_kernels.cuh"
#include "mg_tbo_kernels.cuh"
#include "mg_tbq_kernels.cuh"
#include "mg_rr_kernels.cuh"
#include "comm.cuh"
#include <cstdlib>
#include <cmath>
#include <algorithm>
#include <cstring>

namespace cfdnn {

static inline dim3 grid2(int nx, int ny, int nz, dim3 b) {
    return dim3((nx + b.x - 1) / b.x, (ny + b.y - 1) / b.y, nz);
}
static const dim3 kBlk(64, 4, 1);
constexpr int KZ = 8;

// ---------------------------------------------------------------- construction
// create_hierarchy (:155-230) + compute_gershgorin_bounds (:232-305)
Multigrid::Multigrid(const cfdnn_mesh_desc& m) : mesh_(m) {
    require_device();
    if (m.Nz <= 1) {     // 2-D mesh (reference Mesh::is2D()): 5-point solver, whole V-cycles per launch
        mg2d_impl_ = std::make_unique<Multigrid2D>(m);
        return;
    }
    if (m.Ng != 1) throw std::invalid_argument("Multigrid: only Nghost == 1 (O2 stencils) is supported");
    ystretched_ = m.y_stretched != 0;
    semi_ = ystretched_;
    if (ystretched_) {   // stretched y: x/z semi-coarsening with y-line / x-z RBGS smoothing (:181-207)
        semi_impl_ = std::make_unique<MultigridSemi>(m);
        return;
    }
    const int NzG = m.Nz_global > 0 ? m.Nz_global : m.Nz;

    int Nx = m.Nx, Ny = m.Ny, Nz = m.Nz, NzGlob = NzG;
    double dx = m.dx, dy = m.dy, dz = m.dz;
    auto add = [&](int nx, int ny, int nz, int nzg, double hx, double hy, double hz) {
        auto L = std::make_unique<Level>();
        L->g.Nx = nx; L->g.Ny = ny; L->g.Nz = nz;
        L->g.Hz = (nz != nzg) ? 2 : 1;    // slab levels: room for the 2-deep halo of the blocked smoother
        L->g.Px = (nx + 15) / 16 * 16;
        L->g.plane = (long long)L->g.Px * ny;
        L->dx = hx; L->dy = hy; L->dz = hz; L->NzGlobal = nzg;
        const double dx2 = hx * hx, dy2 = hy * hy, dz2 = hz * hz;
        L->g.dx2 = make_div_host(dx2);
        L->g.dy2 = make_div_host(dy2);
        L->g.dz2 = make_div_host(dz2);
        L->g.coeff = make_div_host(2.0 / dx2 + 2.0 / dy2 + 2.0 / dz2);   // :845-846
        L->g.inv_dx2_2 = 2.0 / dx2;
        L->g.inv_dz2_2 = 2.0 / dz2;
        L->g.nonuniform_y = 0; L->g.joff = 0;
        { const char* e = std::getenv("CFDNN_MG_FASTDIV"); L->g.fastdiv = (e && std::atoi(e) == 0) ? 0 : 1; }
        L->g.aS = L->g.aN = L->g.aP = nullptr;
        L->size = (size_t)(nz + 2 * L->g.Hz) * L->g.plane;
        L->u.alloc(L->size); L->f.alloc(L->size); L->r.alloc(L->size); L->tmp.alloc(L->size);
        L->lambda_max = std::max(1.8, (1.0 + 1.0) * 1.10);               // :283-294 uniform
        levels_.push_back(std::move(L));
    };
    add(Nx, Ny, Nz, NzGlob, dx, dy, dz);
    constexpr int MIN_COARSE = 8;
    // z-slab runs: a coarse level is REPLICATED on every rank (the slabs are all-gathered at the
    // restriction into it, restrict_to) once it is small or its slab would get thinner than 2
    // planes — no halo exchanges below that level, every rank solves it redundantly.
    // Threshold: a blocked 2-sweep launch on <= 128^3 cells takes ~20 us whether it runs on the whole level
    // or on a slab of it (latency-bound), while every halo exchange of a slab level costs ~25 us of NCCL
    // point-to-point latency (measured, profiles/r1/SCALING.md) — so such levels are cheaper replicated.
    long long REPLICATE_CELLS = 128LL * 128 * 128;
    if (const char* e = std::getenv("CFDNN_MG_REPLICATE_CELLS")) REPLICATE_CELLS = std::atoll(e);
    bool replicated = (Nz == NzGlob);
    while (Nx > MIN_COARSE && Ny > MIN_COARSE && NzGlob > MIN_COARSE) {
        if ((Nx | Ny | Nz | NzGlob) & 1)
            throw std::invalid_argument("Multigrid: grid sizes must stay even while coarsening");
        Nx /= 2; Ny /= 2; Nz /= 2; NzGlob /= 2;
        dx *= 2.0; dy *= 2.0; dz *= 2.0;
        if (!replicated && ((long long)Nx * Ny * NzGlob <= REPLICATE_CELLS || Nz < 2 || (Nz & 1))) replicated = true;
        if (replicated) Nz = NzGlob;
        add(Nx, Ny, Nz, NzGlob, dx, dy, dz);
    }
    size_t maxblocks = 0;
    for (auto& L : levels_) {
        dim3 g1 = grid2(L->g.Nx, L->g.Ny, L->g.Nz, kBlk);
        maxblocks = std::max(maxblocks, (size_t)g1.x * g1.y * g1.z);
    }
    partial_.alloc(2 * maxblocks + 16);
    scalars_.alloc(16);
    CFDNN_CUDA(cudaMallocHost(&h_scalars_, 16 * sizeof(double)));
    // coarsest-level Jacobi keeps the whole level (two copies) in shared memory: up to 64 KB
    CFDNN_CUDA(cudaFuncSetAttribute(k_coarse_jacobi<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
#ifdef CFDNN_DEVSIM
    use_graph_ = false;   // the CPU test double does not model CUDA graphs
#endif
    tb_setup();
    apply_bc_codes();
}

Multigrid::~Multigrid() {
    if (graph_) cudaGraphExecDestroy(graph_);
    if (sub_graph_) cudaGraphExecDestroy(sub_graph_);
    if (comm_st_) { cudaStreamDestroy(comm_st_); cudaEventDestroy(ev_fork_); cudaEventDestroy(ev_join_); }
    if (h_scalars_) cudaFreeHost(h_scalars_);
    for (auto& e : evpool_) { if (e.a) cudaEventDestroy(e.a); if (e.b) cudaEventDestroy(e.b); }
}

void Multigrid::set_comm(Comm* c) {
    if ((semi_impl_ || mg2d_impl_) && c && c->nranks > 1)
        throw std::invalid_argument("Multigrid: z-slab decomposition of the y-stretched / 2-D solvers is not built");
    comm_ = c;
    apply_bc_codes();
}

void Multigrid::set_bc(const int bc[6]) {
    if (mg2d_impl_) { mg2d_impl_->set_bc(bc); for (int d = 0; d < 4; ++d) bc_[d] = bc[d]; return; }
    if (semi_impl_) { semi_impl_->set_bc(bc); for (int d = 0; d < 6; ++d) bc_[d] = bc[d]; return; }
    for (int d = 0; d < 6; ++d) {
        if (bc[d] < 0 || bc[d] > 2) throw std::invalid_argument("set_bc: bad PoissonBC value");
        bc_[d] = bc[d];
    }
    apply_bc_codes();
}

void Multigrid::apply_bc_codes() {
    for (auto& L : levels_) {
        for (int d = 0; d < 6; ++d) L->g.bc[d] = bc_[d];
        if (comm_ && comm_->nranks > 1 && L->g.Nz != L->NzGlobal) {
            // slab interfaces: interior ranks always; end ranks only when z is periodic
            const bool per = bc_[4] == CFDNN_PBC_PERIODIC && bc_[5] == CFDNN_PBC_PERIODIC;
            if (comm_->rank > 0 || per) L->g.bc[4] = CFDNN_PBC_HALO;
            if (comm_->rank < comm_->nranks - 1 || per) L->g.bc[5] = CFDNN_PBC_HALO;
        }
    }
    invalidate_graph();
}

void Multigrid::invalidate_graph() {
    if (graph_) { cudaGraphExecDestroy(graph_); graph_ = nullptr; }
    graph_key_[0] = graph_key_[1] = graph_key_[2] = -1;
    if (sub_graph_) { cudaGraphExecDestroy(sub_graph_); sub_graph_ = nullptr; }
    sub_key_[0] = sub_key_[1] = sub_key_[2] = sub_key_[3] = -1;
}

bool Multigrid::has_nullspace() const {   // :3393-3402
    for (int d = 0; d < 6; ++d) if (bc_[d] == CFDNN_PBC_DIRICHLET) return false;
    return true;
}

void Multigrid::halo(int level, double* field, int depth) {
    if (!comm_ || comm_->nranks <= 1) return;
    Level& L = *levels_[level];
    if (L.g.Nz == L.NzGlobal) return;   // agglomerated level: whole domain on every rank
    const size_t pl = (size_t)L.g.plane;
    const size_t cnt = pl * (size_t)depth;
    const int Hz = L.g.Hz, Nz = L.g.Nz;
    // interior planes live at [Hz, Hz+Nz); the `depth` planes next to each interface travel
    comm_->exchange_ptrs(field + (size_t)Hz * pl, cnt, field + (size_t)(Hz - depth) * pl, cnt,
                         field + (size_t)(Hz + Nz - depth) * pl, cnt, field + (size_t)(Hz + Nz) * pl, cnt,
                         L.g.bc[4] == CFDNN_PBC_HALO, L.g.bc[5] == CFDNN_PBC_HALO, st_);
}

// ---------------------------------------------------------------- kernels' launchers
void Multigrid::sweep(int level, const double* uin, double* uout, double omega) {
    Level& L = *levels_[level];
    dim3 g = grid2(L.g.Nx, L.g.Ny, (L.g.Nz + KZ - 1) / KZ, kBlk);
    CFDNN_LAUNCH((k_sweep<KZ, false>), g, kBlk, 0, st_, L.g, uin, L.f.p, uout, omega, 1.0 - omega);
}

// ---- temporally blocked sweeps (mg_tb_kernels.cuh)
namespace {
// tile shapes: BIG for levels that fill the chip with 64-wide tiles, SMALL for the rest
constexpr int TBO_MODE = 2;   // k_sweep_tbo variant of the exact-division build
constexpr int RR_CX = 32, RR_CY = 8;   // coarse tile of the fused residual + restriction (mg_rr_kernels.cuh)
constexpr int TB_BX_BIG = 64, TB_BY_BIG = 26, TB_BX_SMALL = 32, TB_BY_SMALL = 12;

// z chunks (one CTA per SM): pick the chunk count that maximises
//   (fill of the last wave of 148 SMs) x (useful planes / planes marched incl. the 2T warm-up),
// with chunks no thinner than 4*T planes.
dim3 tb_grid(const LevelGeom& g, int T, int BX, int BY, int* ZC_out) {
    const int tx = (g.Nx + BX - 1) / BX, ty = (g.Ny + BY - 1) / BY;
    const int nch_max = std::max(1, g.Nz / (4 * T));
    int nch = 1;
    double best = -1.0;
    for (int c = 1; c <= nch_max; ++c) {
        const int zc = (g.Nz + c - 1) / c;
        const int cc = (g.Nz + zc - 1) / zc;
        const double ctas = (double)tx * ty * cc;
        const double waves = ctas / 148.0;
        const double fill = waves / std::ceil(waves);
        const double score = fill * zc / (zc + 2.0 * T);
        if (score > best + 1e-9) { best = score; nch = cc; }
    }
    const int ZC = (g.Nz + nch - 1) / nch;
    nch = (g.Nz + ZC - 1) / ZC;
    *ZC_out = ZC;
    return dim3(tx, ty, nch);
}

template <int T, int BX, int BY, bool RES = false>
void tb_attr() {
    using S = TbShape<T, BX, BY>;
    CFDNN_CUDA(cudaFuncSetAttribute(k_sweep_tb<T, BX, BY, RES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::SMEM));
}
template <int T, int BX, int BY, bool RES = false>
void tb_launch(const LevelGeom& g, const double* in, const double* f, double* out, const double* w,
               cudaStream_t st) {
    using S = TbShape<T, BX, BY>;
    TbOmega<T> om;
    for (int s = 0; s < T; ++s) om.w[s] = w ? w[s] : 0.0;
    int ZC;
    dim3 grid = tb_grid(g, T, BX, BY, &ZC);
    CFDNN_LAUNCH_COOP((k_sweep_tb<T, BX, BY, RES>), grid, S::NTHREADS, S::SMEM, st, g, in, f, out, om, ZC);
}
// two sweeps on a level without physical faces in reach (mg_tbo_kernels.cuh).  part = 0: all z chunks;
// 1: the chunks that read no halo plane (all but the first and the last); 2: the first and the last chunk.
template <int BX, int BY>
void tbo_launch(const LevelGeom& g, const double* in, const double* f, double* out, const double* w,
                cudaStream_t st, int mode, int part = 0, bool wally = false) {
    using S = TbShape<2, BX, BY>;
    TbOmega<2> om;
    om.w[0] = w[0]; om.w[1] = w[1];
    int ZC;
    dim3 grid = tb_grid(g, 2, BX, BY, &ZC);
    int zoff = 0, zstep = 1;
    const int nch = (int)grid.z;
    if (part == 1) { zoff = 1; grid.z = nch - 2; }
    else if (part == 2) { zstep = nch - 1; grid.z = 2; }
    if (wally && mode < 3) mode = 3;   // y walls: only the 2 x 2 patch kernels carry the wall rows
    using Q = TbqShape<BX, BY>;
#define CFDNN_TBQ(FD_, WY_) CFDNN_LAUNCH_COOP((k_sweep_tbq<BX, BY, FD_, WY_>), grid, Q::NTHREADS, Q::SMEM, st, g, in, f, out, om, ZC, zoff, zstep)
    if (mode > 3) mode = 3;
    if (mode == 3) {      // 2 x 2 patch per thread (mg_tbq_kernels.cuh)
        if (g.fastdiv) { if (wally) CFDNN_TBQ(true, true); else CFDNN_TBQ(true, false); }
        else { if (wally) CFDNN_TBQ(false, true); else CFDNN_TBQ(false, false); }
        return;
    }
#undef CFDNN_TBQ
    if (!g.fastdiv) {
        CFDNN_LAUNCH_COOP((k_sweep_tbo<BX, BY, false, TBO_MODE>), grid, S::NTHREADS, S::SMEM, st, g, in, f, out, om, ZC, zoff, zstep);
    } else if (mode == 0) {
        CFDNN_LAUNCH_COOP((k_sweep_tbo<BX, BY, true, 0>), grid, S::NTHREADS, S::SMEM, st, g, in, f, out, om, ZC, zoff, zstep);
    } else if (mode == 1) {
        CFDNN_LAUNCH_COOP((k_sweep_tbo<BX, BY, true, 1>), grid, S::NTHREADS, S::SMEM, st, g, in, f, out, om, ZC, zoff, zstep);
    } else {
        CFDNN_LAUNCH_COOP((k_sweep_tbo<BX, BY, true, 2>), grid, S::NTHREADS, S::SMEM, st, g, in, f, out, om, ZC, zoff, zstep);
    }
}
// can the end chunks be split off?  (>= 3 chunks, and the last one at least T planes thick so that the
// chunk before it reads no halo plane.)  Only launches of a few waves are split: there the exchange
// (~25 us) is a large share of the pass; a 12-wave level-0 launch at 512^3 loses more to the extra tail
// of the second launch than the hidden exchange saves (measured, profiles/r1/SCALING.md).
template <int BX, int BY>
bool tbo_splittable(const LevelGeom& g) {
    int ZC;
    const dim3 gr = tb_grid(g, 2, BX, BY, &ZC);
    const int nch = (int)gr.z;
    return nch >= 3 && g.Nz - (nch - 1) * ZC >= 2 && gr.x * gr.y * gr.z <= 4 * 148;
}
template <int BX, int BY>
void tbo_attr() {
    using S = TbShape<2, BX, BY>;
    CFDNN_CUDA(cudaFuncSetAttribute(k_sweep_tbo<BX, BY, true, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::SMEM));
    CFDNN_CUDA(cudaFuncSetAttribute(k_sweep_tbo<BX, BY, true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::SMEM));
    CFDNN_CUDA(cudaFuncSetAttribute(k_sweep_tbo<BX, BY, true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::SMEM));
    CFDNN_CUDA(cudaFuncSetAttribute(k_sweep_tbo<BX, BY, false, TBO_MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::SMEM));
    using Q = TbqShape<BX, BY>;
    CFDNN_CUDA(cudaFuncSetAttribute(k_sweep_tbq<BX, BY, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Q::SMEM));
    CFDNN_CUDA(cudaFuncSetAttribute(k_sweep_tbq<BX, BY, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Q::SMEM));
    CFDNN_CUDA(cudaFuncSetAttribute(k_sweep_tbq<BX, BY, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Q::SMEM));
    CFDNN_CUDA(cudaFuncSetAttribute(k_sweep_tbq<BX, BY, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Q::SMEM));
}
}  // namespace

void Multigrid::tb_setup() {
    tb_attr<1, TB_BX_BIG, TB_BY_BIG>(); tb_attr<2, TB_BX_BIG, TB_BY_BIG>();
    tb_attr<1, TB_BX_SMALL, TB_BY_SMALL>(); tb_attr<2, TB_BX_SMALL, TB_BY_SMALL>();
    tb_attr<1, TB_BX_BIG, TB_BY_BIG, true>(); tb_attr<1, TB_BX_SMALL, TB_BY_SMALL, true>();
    CFDNN_CUDA(cudaFuncSetAttribute(k_sweep_tb<1, TB_BX_BIG, TB_BY_BIG, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)TbShape<1, TB_BX_BIG, TB_BY_BIG>::SMEM));
    CFDNN_CUDA(cudaFuncSetAttribute(k_sweep_tb<1, TB_BX_SMALL, TB_BY_SMALL, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)TbShape<1, TB_BX_SMALL, TB_BY_SMALL>::SMEM));
    tbo_attr<TB_BX_BIG, TB_BY_BIG>(); tbo_attr<TB_BX_SMALL, TB_BY_SMALL>();
    if (const char* e = std::getenv("CFDNN_MG_TB")) tb_max_ = std::max(0, std::min(2, std::atoi(e)));
    if (const char* e = std::getenv("CFDNN_MG_TBO")) tbo_ = std::atoi(e) != 0;
    CFDNN_CUDA(cudaFuncSetAttribute(k_residual_restrict<RR_CX, RR_CY, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)RrShape<RR_CX, RR_CY>::SMEM));
    CFDNN_CUDA(cudaFuncSetAttribute(k_residual_restrict<RR_CX, RR_CY, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)RrShape<RR_CX, RR_CY>::SMEM));
    if (const char* e = std::getenv("CFDNN_MG_OVERLAP")) overlap_ = std::atoi(e) != 0;
    if (const char* e = std::getenv("CFDNN_MG_NORMS_TB")) norms_tb_ = std::atoi(e) != 0;
    if (const char* e = std::getenv("CFDNN_MG_RR")) rr_ = std::atoi(e) != 0;
    if (const char* e = std::getenv("CFDNN_MG_RR_MIN")) rr_min_ = std::max(8, std::atoi(e));
    if (const char* e = std::getenv("CFDNN_MG_TBO_MODE")) tbo_mode_ = std::max(0, std::min(3, std::atoi(e)));
}

// n (<= 2) sweeps in one pass (w != nullptr), or the tiled residual r = f - L u (w == nullptr, n = 1);
// false when this level is not eligible
bool Multigrid::sweeps_tb(int level, const double* in, double* out, const double* w, int n) {
    Level& L = *levels_[level];
    const LevelGeom& g = L.g;
    if (tb_max_ <= 0 || n > tb_max_) return false;
    if (g.nonuniform_y || (g.Nx & 1) || g.Nx < 16 || g.Ny < 8 || g.Nz < 4) return false;
    const bool slab = g.bc[4] == CFDNN_PBC_HALO || g.bc[5] == CFDNN_PBC_HALO;
    if (slab && (g.Hz < n || g.Nz < n)) return false;
    const bool big = g.Nx >= 128 && g.Ny >= 56 && g.Nz >= 32;
    const bool open_xy0 = g.bc[0] == CFDNN_PBC_PERIODIC && g.bc[1] == CFDNN_PBC_PERIODIC &&
                          g.bc[2] == CFDNN_PBC_PERIODIC && g.bc[3] == CFDNN_PBC_PERIODIC;
    if (slab && overlap_ && w && tbo_ && n == 2 && open_xy0 && g.bc[4] == CFDNN_PBC_HALO && g.bc[5] == CFDNN_PBC_HALO &&
        (big ? tbo_splittable<TB_BX_BIG, TB_BY_BIG>(g) : tbo_splittable<TB_BX_SMALL, TB_BY_SMALL>(g))) {
        // Halo exchange hidden behind the chunks that do not read it: the exchange and the two end chunks
        // run on the communication stream, the inner chunks on the solver stream, joined by events.
        if (!comm_st_) {
            CFDNN_CUDA(cudaStreamCreateWithFlags(&comm_st_, cudaStreamNonBlocking));
            CFDNN_CUDA(cudaEventCreateWithFlags(&ev_fork_, cudaEventDisableTiming));
            CFDNN_CUDA(cudaEventCreateWithFlags(&ev_join_, cudaEventDisableTiming));
        }
        cudaStream_t main_st = st_;
        CFDNN_CUDA(cudaEventRecord(ev_fork_, main_st));
        CFDNN_CUDA(cudaStreamWaitEvent(comm_st_, ev_fork_, 0));
        st_ = comm_st_;
        halo(level, const_cast<double*>(in), n);
        st_ = main_st;
        if (big) {
            tbo_launch<TB_BX_BIG, TB_BY_BIG>(g, in, L.f.p, out, w, comm_st_, tbo_mode_, 2);
            tbo_launch<TB_BX_BIG, TB_BY_BIG>(g, in, L.f.p, out, w, main_st, tbo_mode_, 1);
        } else {
            tbo_launch<TB_BX_SMALL, TB_BY_SMALL>(g, in, L.f.p, out, w, comm_st_, tbo_mode_, 2);
            tbo_launch<TB_BX_SMALL, TB_BY_SMALL>(g, in, L.f.p, out, w, main_st, tbo_mode_, 1);
        }
        CFDNN_CUDA(cudaEventRecord(ev_join_, comm_st_));
        CFDNN_CUDA(cudaStreamWaitEvent(main_st, ev_join_, 0));
        return true;
    }
    if (slab) halo(level, const_cast<double*>(in), n);     // n-deep halo: one exchange per pass
    if (!w) {
        if (big) tb_launch<1, TB_BX_BIG, TB_BY_BIG, true>(g, in, L.f.p, out, nullptr, st_);
        else tb_launch<1, TB_BX_SMALL, TB_BY_SMALL, true>(g, in, L.f.p, out, nullptr, st_);
        return true;
    }
    const bool open_xy = g.bc[0] == CFDNN_PBC_PERIODIC && g.bc[1] == CFDNN_PBC_PERIODIC &&
                         g.bc[2] == CFDNN_PBC_PERIODIC && g.bc[3] == CFDNN_PBC_PERIODIC;
    const bool open_z = (g.bc[4] == CFDNN_PBC_PERIODIC && g.bc[5] == CFDNN_PBC_PERIODIC) ||
                        (g.bc[4] == CFDNN_PBC_HALO && g.bc[5] == CFDNN_PBC_HALO);
    if (tbo_ && n == 2 && open_xy && open_z) {             // no physical face in reach: the lean kernel
        if (big) tbo_launch<TB_BX_BIG, TB_BY_BIG>(g, in, L.f.p, out, w, st_, tbo_mode_);
        else tbo_launch<TB_BX_SMALL, TB_BY_SMALL>(g, in, L.f.p, out, w, st_, tbo_mode_);
        return true;
    }
    // channel levels: x periodic, z open, walls (Neumann / Dirichlet) on both y faces — the 2 x 2 patch kernels
    // carry the two wall rows with one select each instead of the general kernel's per-cell boundary path
    const bool wall_y = g.bc[2] != CFDNN_PBC_PERIODIC && g.bc[3] != CFDNN_PBC_PERIODIC && !(g.Ny & 1);
    if (tbo_ && tbo_mode_ >= 3 && n == 2 && open_z && wall_y && g.bc[0] == CFDNN_PBC_PERIODIC && g.bc[1] == CFDNN_PBC_PERIODIC) {
        if (big) tbo_launch<TB_BX_BIG, TB_BY_BIG>(g, in, L.f.p, out, w, st_, tbo_mode_, 0, true);
        else tbo_launch<TB_BX_SMALL, TB_BY_SMALL>(g, in, L.f.p, out, w, st_, tbo_mode_, 0, true);
        return true;
    }
    if (big) {
        if (n == 2) tb_launch<2, TB_BX_BIG, TB_BY_BIG>(g, in, L.f.p, out, w, st_);
        else tb_launch<1, TB_BX_BIG, TB_BY_BIG>(g, in, L.f.p, out, w, st_);
    } else {
        if (n == 2) tb_launch<2, TB_BX_SMALL, TB_BY_SMALL>(g, in, L.f.p, out, w, st_);
        else tb_launch<1, TB_BX_SMALL, TB_BY_SMALL>(g, in, L.f.p, out, w, st_);
    }
    return true;
}

// smooth_chebyshev (:825-1001): degree sweeps with Chebyshev weights; BCs are implicit.
void Multigrid::smooth_chebyshev(int level, int degree) {
    Level& L = *levels_[level];
    const double lambda_max = L.lambda_max;
    const double lambda_min = 0.1 * lambda_max;
    const double d = (lambda_max + lambda_min) / 2.0;
    const double c = (lambda_max - lambda_min) / 2.0;
    double* a = L.u.p;
    double* b = L.tmp.p;
    auto omega_of = [&](int k) {
        const double theta = M_PI * (2.0 * k + 1.0) / (2.0 * degree);
        return std::min(1.0 / (d - c * std::cos(theta)), 2.0);
    };
    for (int k = 0; k < degree;) {
        const int n = std::min(2, degree - k);
        const double w[2] = {omega_of(k), omega_of(k + 1)};
        if (n == 2 && sweeps_tb(level, a, b, w, 2)) { k += 2; std::swap(a, b); continue; }
        if (!sweeps_tb(level, a, b, w, 1)) {
            halo(level, a);
            sweep(level, a, b, w[0]);
        }
        std::swap(a, b);
        ++k;
    }
    if (a != L.u.p) std::swap(L.u.p, L.tmp.p);   // result lives in L.u
}

void Multigrid::smooth_jacobi(int level, int iters, double omega) {
    Level& L = *levels_[level];
    const int n = L.g.Nx * L.g.Ny * L.g.Nz;
    const bool local = !(comm_ && comm_->nranks > 1 && L.g.Nz != L.NzGlobal);
    if (local && n <= 4096) {
        const size_t smem = 2 * (size_t)n * sizeof(double);
        const int threads = std::min(1024, (n + 31) / 32 * 32);
        if (n <= 1024) {
            CFDNN_LAUNCH_COOP((k_coarse_jacobi<1>), 1, threads, smem, st_, L.g, L.u.p, L.f.p, iters, omega, 1.0 - omega);
        } else {
            CFDNN_LAUNCH_COOP((k_coarse_jacobi<4>), 1, 1024, smem, st_, L.g, L.u.p, L.f.p, iters, omega, 1.0 - omega);
        }
        return;
    }
    double* a = L.u.p;
    double* b = L.tmp.p;
    for (int it = 0; it < iters; ++it) {
        halo(level, a);
        sweep(level, a, b, omega);
        std::swap(a, b);
    }
    if (a != L.u.p) std::swap(L.u.p, L.tmp.p);
}

void Multigrid::residual(int level) {
    Level& L = *levels_[level];
    if (sweeps_tb(level, L.u.p, L.r.p, nullptr, 1)) return;   // tiled residual (exchanges the halo itself)
    halo(level, L.u.p);
    dim3 g = grid2(L.g.Nx, L.g.Ny, (L.g.Nz + KZ - 1) / KZ, kBlk);
    CFDNN_LAUNCH((k_residual<KZ, false, false>), g, kBlk, 0, st_, L.g, L.u.p, L.f.p, L.r.p, nullptr, nullptr);
}

void Multigrid::residual_norms(int level, double* d_max, double* d_sumsq) {
    Level& L = *levels_[level];
    halo(level, L.u.p);
    dim3 g = grid2(L.g.Nx, L.g.Ny, (L.g.Nz + KZ - 1) / KZ, kBlk);
    int nb = g.x * g.y * g.z;
    double* pm = partial_.p;
    double* ps = partial_.p + nb;
    const LevelGeom& lg = L.g;
    if (tb_max_ > 0 && norms_tb_ && !lg.nonuniform_y && !(lg.Nx & 1) && lg.Nx >= 16 && lg.Ny >= 8 && lg.Nz >= 4) {
        // tiled, norms only: u and f are read once (16 B per cell), nothing is written
        TbOmega<1> om; om.w[0] = 0.0;
        int ZC;
        if (lg.Nx >= 128 && lg.Ny >= 56 && lg.Nz >= 32) {
            using S = TbShape<1, TB_BX_BIG, TB_BY_BIG>;
            dim3 tg = tb_grid(lg, 1, TB_BX_BIG, TB_BY_BIG, &ZC);
            nb = tg.x * tg.y * tg.z;
            ps = partial_.p + nb;
            CFDNN_LAUNCH_COOP((k_sweep_tb<1, TB_BX_BIG, TB_BY_BIG, true, true>), tg, S::NTHREADS, S::SMEM, st_, lg, L.u.p,
                              L.f.p, (double*)nullptr, om, ZC, pm, ps);
        } else {
            using S = TbShape<1, TB_BX_SMALL, TB_BY_SMALL>;
            dim3 tg = tb_grid(lg, 1, TB_BX_SMALL, TB_BY_SMALL, &ZC);
            nb = tg.x * tg.y * tg.z;
            ps = partial_.p + nb;
            CFDNN_LAUNCH_COOP((k_sweep_tb<1, TB_BX_SMALL, TB_BY_SMALL, true, true>), tg, S::NTHREADS, S::SMEM, st_, lg, L.u.p,
                              L.f.p, (double*)nullptr, om, ZC, pm, ps);
        }
    } else
    CFDNN_LAUNCH_COOP((k_residual<KZ, false, true>), g, kBlk, 0, st_, L.g, L.u.p, L.f.p, (double*)nullptr, pm, ps);
    CFDNN_LAUNCH_COOP(k_final_max, 1, 1024, 0, st_, pm, nb, d_max);
    CFDNN_LAUNCH_COOP(k_final_sum, 1, 1024, 0, st_, ps, nb, d_sumsq, 0.0);
    if (comm_ && comm_->nranks > 1) {
        comm_->allreduce_max(d_max, 1, st_);
        comm_->allreduce_sum(d_sumsq, 1, st_);
    }
}

void Multigrid::mean_free(int level, double* field, bool subtract) {
    Level& L = *levels_[level];
    dim3 g = grid2(L.g.Nx, L.g.Ny, L.g.Nz, kBlk);
    const int nb = g.x * g.y * g.z;
    const bool dist = comm_ && comm_->nranks > 1 && L.g.Nz != L.NzGlobal;
    const double count = (double)L.g.Nx * L.g.Ny * L.NzGlobal;
    CFDNN_LAUNCH_COOP(k_sum_interior, g, kBlk, 0, st_, L.g, field, partial_.p);
    CFDNN_LAUNCH_COOP(k_final_sum, 1, 1024, 0, st_, partial_.p, nb, scalars_.p + 8, dist ? 0.0 : count);
    if (dist) {
        comm_->allreduce_sum(scalars_.p + 8, 1, st_);
        CFDNN_LAUNCH_COOP(k_final_sum, 1, 32, 0, st_, scalars_.p + 8, 1, scalars_.p + 8, count);
    }
    if (subtract) CFDNN_LAUNCH(k_sub_scalar, g, kBlk, 0, st_, L.g, field, scalars_.p + 9);
}

// fused = true: r is never formed in HBM — k_residual_restrict evaluates it tile by tile from u and f
bool Multigrid::rr_eligible(int fine) const {
    const LevelGeom& g = levels_[fine]->g;
    if (!rr_ || g.nonuniform_y || ((g.Nx | g.Ny | g.Nz) & 1) || g.Nx < rr_min_ || g.Ny < rr_min_ / 4 || g.Nz < 8)
        return false;
    const bool slab = g.bc[4] == CFDNN_PBC_HALO || g.bc[5] == CFDNN_PBC_HALO;
    return !slab || g.Hz >= 2;       // the residual plane below the slab needs u two planes deep
}

void Multigrid::restrict_to(int fine, bool fused) {
    Level& F = *levels_[fine];
    Level& C = *levels_[fine + 1];
    if (fused) halo(fine, F.u.p, std::min(2, F.g.Hz));
    else halo(fine, F.r.p);
    const int nzc_local = F.g.Nz / 2;
    dim3 g = grid2(C.g.Nx, C.g.Ny, nzc_local, kBlk);
    int KC = 16;                     // coarse planes per CTA of the fused kernel
    if (fused) {
        const int tiles = ((C.g.Nx + RR_CX - 1) / RR_CX) * ((C.g.Ny + RR_CY - 1) / RR_CY);
        while (KC > 4 && tiles * ((nzc_local + KC - 1) / KC) < 4 * 148) KC /= 2;
        g = dim3((C.g.Nx + RR_CX - 1) / RR_CX, (C.g.Ny + RR_CY - 1) / RR_CY, (nzc_local + KC - 1) / KC);
    }
    const int nb = g.x * g.y * g.z;
    int kc_off = 0;
    const bool agglomerate = comm_ && comm_->nranks > 1 && F.g.Nz != F.NzGlobal && C.g.Nz == C.NzGlobal;
    if (agglomerate) kc_off = comm_->rank * nzc_local;
    const bool nullsp = has_nullspace();
    double* part = (nullsp && !agglomerate) ? partial_.p : nullptr;
    if (fused) {
        using S = RrShape<RR_CX, RR_CY>;
        const bool open_xy = F.g.bc[0] == CFDNN_PBC_PERIODIC && F.g.bc[2] == CFDNN_PBC_PERIODIC &&
                             F.g.Nx % (2 * RR_CX) == 0 && F.g.Ny % (2 * RR_CY) == 0;   // and no tile overshoot
        if (open_xy)
            CFDNN_LAUNCH_COOP((k_residual_restrict<RR_CX, RR_CY, true>), g, S::NT, S::SMEM, st_, F.g, C.g, F.u.p, F.f.p,
                              C.f.p, kc_off, KC, part);
        else
            CFDNN_LAUNCH_COOP((k_residual_restrict<RR_CX, RR_CY, false>), g, S::NT, S::SMEM, st_, F.g, C.g, F.u.p, F.f.p,
                              C.f.p, kc_off, KC, part);
    } else {
        CFDNN_LAUNCH_COOP(k_restrict, g, kBlk, 0, st_, F.g, C.g, F.r.p, C.f.p, kc_off, part);
    }
    if (agglomerate)
        comm_->allgather_planes(C.f.p + (size_t)C.g.Hz * C.g.plane, (size_t)C.g.plane * nzc_local, st_);
    if (nullsp) {   // make_rhs_mean_free(level+1) :3274-3276
        const bool dist = comm_ && comm_->nranks > 1 && C.g.Nz != C.NzGlobal;
        if (agglomerate) {
            mean_free(fine + 1, C.f.p);
        } else {
            const double count = (double)C.g.Nx * C.g.Ny * C.NzGlobal;
            CFDNN_LAUNCH_COOP(k_final_sum, 1, 1024, 0, st_, partial_.p, nb, scalars_.p + 8, dist ? 0.0 : count);
            if (dist) {
                comm_->allreduce_sum(scalars_.p + 8, 1, st_);
                CFDNN_LAUNCH_COOP(k_final_sum, 1, 32, 0, st_, scalars_.p + 8, 1, scalars_.p + 8, count);
            }
            dim3 gs = grid2(C.g.Nx, C.g.Ny, C.g.Nz, kBlk);
            CFDNN_LAUNCH(k_sub_scalar, gs, kBlk, 0, st_, C.g, C.f.p, scalars_.p + 9);
        }
    }
    halo(fine + 1, C.f.p, C.g.Hz);   // slab coarse level: deep halo of its (now final) right-hand side
}

void Multigrid::prolong_from(int coarse) {
    Level& C = *levels_[coarse];
    Level& F = *levels_[coarse - 1];
    int kc_off = 0;
    const bool agglomerated = comm_ && comm_->nranks > 1 && F.g.Nz != F.NzGlobal && C.g.Nz == C.NzGlobal;
    if (agglomerated) kc_off = comm_->rank * (F.g.Nz / 2);
    else halo(coarse, C.u.p);
    LevelGeom gc = C.g;
    dim3 g = grid2(F.g.Nx, F.g.Ny, F.g.Nz, kBlk);
    if (!((F.g.Nx | F.g.Ny | F.g.Nz) & 1)) {
        dim3 g8 = grid2(F.g.Nx / 2, F.g.Ny / 2, F.g.Nz / 2, kBlk);
        CFDNN_LAUNCH(k_prolong8, g8, kBlk, 0, st_, F.g, gc, C.u.p, F.u.p, kc_off);
    } else {
        CFDNN_LAUNCH(k_prolong, g, kBlk, 0, st_, F.g, gc, C.u.p, F.u.p, kc_off);
    }
}

// ---------------------------------------------------------------- V-cycle (:3121-3322)
void Multigrid::vcycle(int level, int nu1, int nu2, int degree) {
    const int nl = (int)levels_.size();
    if (level == nl - 1) {
        smooth_jacobi(level, has_nullspace() ? 100 : 50, 0.67);   // :3193-3195
        return;
    }
    for (int p = 0; p < nu1; ++p) smooth_chebyshev(level, degree);
    if (rr_eligible(level)) {
        restrict_to(level, true);     // residual + restriction in one pass over u and f
    } else {
        residual(level);
        restrict_to(level, false);
    }
    Level& C = *levels_[level + 1];
    CFDNN_CUDA(cudaMemsetAsync(C.u.p, 0, C.size * sizeof(double), st_));   // :3279-3304
    descend(level + 1, nu1, nu2, degree);
    prolong_from(level + 1);
    for (int p = 0; p < nu2; ++p) smooth_chebyshev(level, degree);
}

// z-slab runs cannot replay the whole V-cycle as a CUDA graph (NCCL exchanges on the distributed levels),
// but everything below the first REPLICATED level is communication-free: that sub-V-cycle (~100 launches of
// a few microseconds each) is captured once and replayed, so the coarse levels do not wait on the host.
void Multigrid::descend(int level, int nu1, int nu2, int degree) {
    const bool multi = comm_ && comm_->nranks > 1;
    const bool first_replicated = multi && use_graph_ && sub_ok_ && level > 0 &&
                                  levels_[level]->g.Nz == levels_[level]->NzGlobal &&
                                  levels_[level - 1]->g.Nz != levels_[level - 1]->NzGlobal;
    if (!first_replicated) { vcycle(level, nu1, nu2, degree); return; }
    if (!sub_graph_ || sub_key_[0] != nu1 || sub_key_[1] != nu2 || sub_key_[2] != degree || sub_key_[3] != level) {
        if (sub_graph_) { cudaGraphExecDestroy(sub_graph_); sub_graph_ = nullptr; }
        std::vector<double*> before;
        for (auto& L : levels_) before.push_back(L->u.p);
        cudaGraph_t gr = nullptr;
        CFDNN_CUDA(cudaStreamBeginCapture(st_, cudaStreamCaptureModeThreadLocal));
        const unsigned long long l0 = g_launches.load();
        try {
            vcycle(level, nu1, nu2, degree);
        } catch (...) {
            cudaStreamEndCapture(st_, &gr);
            if (gr) cudaGraphDestroy(gr);
            throw;
        }
        sub_launches_ = g_launches.load() - l0;
        g_launches.store(l0);
        CFDNN_CUDA(cudaStreamEndCapture(st_, &gr));
        bool same = true;
        for (size_t l = 0; l < levels_.size(); ++l) same = same && before[l] == levels_[l]->u.p;
        if (!same) {   // odd sweep count at some level: a replay would address the wrong buffers
            for (size_t l = 0; l < levels_.size(); ++l)
                if (before[l] != levels_[l]->u.p) std::swap(levels_[l]->u.p, levels_[l]->tmp.p);
            cudaGraphDestroy(gr);
            sub_ok_ = false;
            vcycle(level, nu1, nu2, degree);
            return;
        }
        CFDNN_CUDA(cudaGraphInstantiate(&sub_graph_, gr, 0));
        cudaGraphDestroy(gr);
        sub_key_[0] = nu1; sub_key_[1] = nu2; sub_key_[2] = degree; sub_key_[3] = level;
    }
    CFDNN_CUDA(cudaGraphLaunch(sub_graph_, st_));
    g_launches.fetch_add(sub_launches_, std::memory_order_relaxed);
}

void Multigrid::run_vcycle(int nu1, int nu2, int degree) {
    const bool multi = comm_ && comm_->nranks > 1;
    if (!use_graph_ || multi) { vcycle(0, nu1, nu2, degree); return; }
    if (!graph_ || graph_key_[0] != nu1 || graph_key_[1] != nu2 || graph_key_[2] != degree) {
        invalidate_graph();
        // The smoothers ping-pong u<->tmp by swapping the Level's pointers; a graph replays fixed
        // addresses, so capture requires every level to end a V-cycle with u where it began.
        // 16 = (nu1+nu2)*degree sweeps at each level: verify by pointer identity after capture.
        std::vector<double*> before;
        for (auto& L : levels_) before.push_back(L->u.p);
        cudaGraph_t gr = nullptr;
        CFDNN_CUDA(cudaStreamBeginCapture(st_, cudaStreamCaptureModeThreadLocal));
        const unsigned long long l0 = g_launches.load();
        try {
            vcycle(0, nu1, nu2, degree);
        } catch (...) {
            cudaStreamEndCapture(st_, &gr);
            if (gr) cudaGraphDestroy(gr);
            throw;
        }
        graph_launches_ = g_launches.load() - l0;
        g_launches.store(l0);
        CFDNN_CUDA(cudaStreamEndCapture(st_, &gr));
        bool same = true;
        for (size_t l = 0; l < levels_.size(); ++l) same = same && before[l] == levels_[l]->u.p;
        if (!same) {
            // odd sweep count at some level: restore pointers, run without a graph
            for (size_t l = 0; l < levels_.size(); ++l)
                if (before[l] != levels_[l]->u.p) std::swap(levels_[l]->u.p, levels_[l]->tmp.p);
            cudaGraphDestroy(gr);
            use_graph_ = false;
            vcycle(0, nu1, nu2, degree);
            return;
        }
        CFDNN_CUDA(cudaGraphInstantiate(&graph_, gr, 0));
        cudaGraphDestroy(gr);
        graph_key_[0] = nu1; graph_key_[1] = nu2; graph_key_[2] = degree;
    }
    CFDNN_CUDA(cudaGraphLaunch(graph_, st_));
    g_launches.fetch_add(graph_launches_, std::memory_order_relaxed);
}

// ---------------------------------------------------------------- solve (:3550-3786)
int Multigrid::solve_dev(const double* rhs, double* p, const cfdnn_poisson_cfg& cfg, cudaStream_t st) {
    st_ = st;
    if (semi_impl_ || mg2d_impl_) {
        cur_ev_ = nullptr;
        if (time_vcycles && evused_ < 4096) {
            if (evused_ == evpool_.size()) {
                EvPair e;
                CFDNN_CUDA(cudaEventCreate(&e.a));
                CFDNN_CUDA(cudaEventCreate(&e.b));
                evpool_.push_back(e);
            }
            cur_ev_ = &evpool_[evused_++];
            CFDNN_CUDA(cudaEventRecord(cur_ev_->a, st_));
        }
        const int cyc = mg2d_impl_ ? mg2d_impl_->solve_dev(rhs, p, cfg, st, plane2d_, dup2d_) : semi_impl_->solve_dev(rhs, p, cfg, st);
        if (cur_ev_) { CFDNN_CUDA(cudaEventRecord(cur_ev_->b, st_)); cur_ev_->cycles = cyc; }
        stats = mg2d_impl_ ? mg2d_impl_->stats : semi_impl_->stats;
        vcycle_count = cyc;
        return cyc;
    }
    Level& L0 = *levels_[0];
    const int Ng = mesh_.Ng;
    {
        dim3 g = grid2(L0.g.Nx, L0.g.Ny, L0.g.Nz, kBlk);
        CFDNN_LAUNCH(k_copy_in, g, kBlk, 0, st_, L0.g, Ng, rhs, p, L0.f.p, L0.u.p);
        halo(0, L0.f.p, L0.g.Hz);   // slab runs: f is constant during the solve, its deep halo travels once
    }
    int cycles_used = 0;
    stats = cfdnn_poisson_stats{};
    const int degree = cfg.chebyshev_degree;
    cur_ev_ = nullptr;
    if (time_vcycles && evused_ < 4096) {
        if (evused_ == evpool_.size()) {
            EvPair e;
            CFDNN_CUDA(cudaEventCreate(&e.a));
            CFDNN_CUDA(cudaEventCreate(&e.b));
            evpool_.push_back(e);
        }
        cur_ev_ = &evpool_[evused_++];
        CFDNN_CUDA(cudaEventRecord(cur_ev_->a, st_));
    }

    if (cfg.fixed_cycles > 0) {
        const int nu1 = cfg.nu1 > 0 ? cfg.nu1 : 3;
        const int nu2 = cfg.nu2 > 0 ? cfg.nu2 : 1;
        for (int c = 0; c < cfg.fixed_cycles; ++c) run_vcycle(nu1, nu2, degree);
        cycles_used = cfg.fixed_cycles;
        stats.fixed_cycle_mode = 1;
        stats.converged = 0;
        stats.status = CFDNN_PSTAT_FIXED;
        if (cur_ev_) CFDNN_CUDA(cudaEventRecord(cur_ev_->b, st_));
    } else {
        if (cfg.max_vcycles <= 0) throw std::invalid_argument("PoissonConfig.max_vcycles must be positive");
        const int max_cycles = cfg.max_vcycles;
        const bool accurate = max_cycles > 5;                                  // :3647-3650
        const int nu1 = cfg.nu1 > 0 ? cfg.nu1 : (accurate ? 3 : 2);
        const int nu2 = cfg.nu2 > 0 ? cfg.nu2 : 1;
        const int check_interval = std::max(1, semi_ ? std::max(4, cfg.check_interval) : cfg.check_interval);

        // ||b||_inf, ||b||_2 over the interior of f (:3660-3682) — residual kernel with u := 0 is
        // not the same thing, so use dedicated reductions through the r scratch.
        double* S = scalars_.p;
        {
            dim3 g = grid2(L0.g.Nx, L0.g.Ny, (L0.g.Nz + KZ - 1) / KZ, kBlk);
            const int nb = g.x * g.y * g.z;
            // r0 = f - L u  and its norms
            residual_norms(0, S + 2, S + 3);
            // b norms: reuse k_residual with u = zero array? cheaper: norm kernel on f
            CFDNN_LAUNCH_COOP(k_norms_interior, g, kBlk, 0, st_, L0.g, L0.f.p, KZ, partial_.p, partial_.p + nb);
            CFDNN_LAUNCH_COOP(k_final_max, 1, 1024, 0, st_, partial_.p, nb, S + 0);
            CFDNN_LAUNCH_COOP(k_final_sum, 1, 1024, 0, st_, partial_.p + nb, nb, S + 1, 0.0);
            if (comm_ && comm_->nranks > 1) {
                comm_->allreduce_max(S + 0, 1, st_);
                comm_->allreduce_sum(S + 1, 1, st_);
            }
        }
        CFDNN_CUDA(cudaMemcpyAsync(h_scalars_, S, 4 * sizeof(double), cudaMemcpyDeviceToHost, st_));
        CFDNN_CUDA(cudaStreamSynchronize(st_));
        stats.b_inf = h_scalars_[0];
        stats.b_l2 = std::sqrt(h_scalars_[1]);
        stats.r0 = h_scalars_[2];
        stats.r0_l2 = std::sqrt(h_scalars_[3]);
        stats.residual = stats.r0;
        stats.residual_l2 = stats.r0_l2;
        stats.converged = 0;
        stats.fixed_cycle_mode = 0;

        for (int cycle = 0; cycle < max_cycles; ++cycle) {
            run_vcycle(nu1, nu2, degree);
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
        if (cur_ev_) CFDNN_CUDA(cudaEventRecord(cur_ev_->b, st_));
    }
    stats.cycles = cycles_used;

    const bool nullsp = has_nullspace();
    if (nullsp) mean_free(0, L0.u.p, /*subtract=*/false);   // fix_nullspace(0) :3404-3409: the mean; subtracted in k_copy_out
    halo(0, L0.u.p);
    {
        dim3 g = grid2(L0.g.Nx + 2 * Ng, L0.g.Ny + 2 * Ng, L0.g.Nz + 2 * Ng, kBlk);
        CFDNN_LAUNCH(k_copy_out, g, kBlk, 0, st_, L0.g, Ng, L0.u.p, p, nullsp ? scalars_.p + 9 : (const double*)nullptr);
    }
    vcycle_count = cycles_used;
    if (cur_ev_) cur_ev_->cycles = cycles_used;
    return cycles_used;
}

double Multigrid::time_level0_smoother(int reps, cudaStream_t st) {
    if (levels_.empty() || reps < 1) return 0.0;
    st_ = st;
    Level& L = *levels_[0];
    const double w[2] = {0.6, 0.9};
    double* a = L.u.p;
    double* b = L.tmp.p;
    if (!sweeps_tb(0, a, b, w, 2)) return 0.0;          // warm-up launch, also the eligibility check
    cudaEvent_t e0, e1;
    CFDNN_CUDA(cudaEventCreate(&e0));
    CFDNN_CUDA(cudaEventCreate(&e1));
    CFDNN_CUDA(cudaEventRecord(e0, st));
    for (int r = 0; r < reps; ++r) { std::swap(a, b); sweeps_tb(0, a, b, w, 2); }
    CFDNN_CUDA(cudaEventRecord(e1, st));
    CFDNN_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    check_launch("time_level0_smoother");
    return (double)ms / reps;
}

void Multigrid::collect_timing() {
    double last = 0.0;
    for (size_t e = 0; e < evused_; ++e) {
        float ms = 0.f;
        if (cudaEventSynchronize(evpool_[e].b) == cudaSuccess &&
            cudaEventElapsedTime(&ms, evpool_[e].a, evpool_[e].b) == cudaSuccess) {
            acc_vcycle_ms += ms;
            acc_vcycles += evpool_[e].cycles;
            last = ms;
        }
    }
    if (evused_ > 0) vcycle_ms_total = last;
    evused_ = 0;
}

int Multigrid::solve_host(const double* rhs, double* p, const cfdnn_poisson_cfg& cfg) {
    const size_t n = (size_t)(mesh_.Nx + 2 * mesh_.Ng) * (mesh_.Ny + 2 * mesh_.Ng) * (mesh_.Nz + 2 * mesh_.Ng);
    if (h_rhs_.n != n) { h_rhs_.alloc(n); h_p_.alloc(n); }
    cudaStream_t st = own_stream();
    CFDNN_CUDA(cudaMemcpyAsync(h_rhs_.p, rhs, n * sizeof(double), cudaMemcpyHostToDevice, st));
    CFDNN_CUDA(cudaMemcpyAsync(h_p_.p, p, n * sizeof(double), cudaMemcpyHostToDevice, st));
    const int cyc = solve_dev(h_rhs_.p, h_p_.p, cfg, st);
    CFDNN_CUDA(cudaMemcpyAsync(p, h_p_.p, n * sizeof(double), cudaMemcpyDeviceToHost, st));
    CFDNN_CUDA(cudaStreamSynchronize(st));
    check_launch("mg solve");
    collect_timing();
    return cyc;
}

cudaStream_t Multigrid::own_stream() {
    if (!own_st_) CFDNN_CUDA(cudaStreamCreateWithFlags(&own_st_, cudaStreamNonBlocking));
    return own_st_;
}

}  // namespace cfdnn
