// solver.cu — host orchestration of the fractional-step path (see solver.cuh).
#include "solver_kernels.cuh"
#include "momentum_kernels.cuh"
#include "mg_kernels.cuh"
#include "comm.cuh"
#include "nn.cuh"
#include "les.cuh"
#include "sst.cuh"
#include <cmath>
#include <algorithm>
#include <cstring>
#include <cstdlib>

namespace cfdnn {

static const dim3 kB(64, 4, 1);
constexpr int MOM_BX = 32, MOM_BY = 8;      // tile of the z-marching momentum kernel (momentum_kernels.cuh)
static inline dim3 grid3(int nx, int ny, int nz) {
    return dim3((nx + kB.x - 1) / kB.x, (ny + kB.y - 1) / kB.y, nz);
}

Geom make_geom(const cfdnn_mesh_desc& m) {
    Geom g{};
    const int Ng = m.Ng, Nx = m.Nx, Ny = m.Ny;
    g.Nx = Nx; g.Ny = Ny; g.Nz = m.Nz; g.Ng = Ng;
    g.us = Nx + 1 + 2 * Ng; g.vs = Nx + 2 * Ng; g.ws = Nx + 2 * Ng; g.cs = Nx + 2 * Ng;
    g.up = (long long)g.us * (Ny + 2 * Ng);
    g.vp = (long long)g.vs * (Ny + 1 + 2 * Ng);
    g.wp = (long long)g.ws * (Ny + 2 * Ng);
    g.cp = (long long)g.cs * (Ny + 2 * Ng);
    g.dx = make_div_host(m.dx); g.dy = make_div_host(m.dy); g.dz = make_div_host(m.dz);
    g.dx_2 = make_div_host(2.0 * m.dx); g.dy_2 = make_div_host(2.0 * m.dy); g.dz_2 = make_div_host(2.0 * m.dz);
    g.dx2 = make_div_host(m.dx * m.dx); g.dy2 = make_div_host(m.dy * m.dy); g.dz2 = make_div_host(m.dz * m.dz);
    g.y_stretched = m.y_stretched;
    return g;
}

// ---------------------------------------------------------------- construction
Solver::Solver(const cfdnn_mesh_desc& m, const cfdnn_config& c) : mesh_(m), cfg_(c) {
    require_device();
    if (m.Nx < 1 || m.Ny < 1 || m.Nz < 1) throw std::invalid_argument("Solver: bad mesh size");
    if (m.Ng != 1) throw std::invalid_argument("Solver: only Nghost == 1 (O2 schemes) is supported");
    if (c.convective_scheme != CFDNN_CONV_CENTRAL && c.convective_scheme != CFDNN_CONV_SKEW)
        throw std::invalid_argument("Solver: convective_scheme must be central or skew");
    if (c.time_integrator < CFDNN_TI_EULER || c.time_integrator > CFDNN_TI_RK3)
        throw std::invalid_argument("Solver: bad time_integrator");
    // 2-D meshes (reference Mesh::is2D()) run as one periodic z-plane of the 3-D kernels: with w = 0
    // and identical z-neighbour planes every z term of the 3-D operators is an exact 0.0 added last,
    // which leaves the 2-D kernels' arithmetic (include/solver_kernels.hpp:341-426, 1282-1374, ...)
    // bit for bit.  The data plane is k = Ng on the device and k = 0 on the host (fields.hpp:18-19).
    is2d_ = m.Nz == 1;
    if (is2d_ && !(mesh_.dz > 0.0)) mesh_.dz = 1.0;
    const int Ng = m.Ng, Nx = m.Nx, Ny = m.Ny, Nz = m.Nz;
    g_ = make_geom(mesh_);
    su_ = (size_t)g_.up * (Nz + 2 * Ng);
    sv_ = (size_t)g_.vp * (Nz + 2 * Ng);
    sw_ = (size_t)g_.wp * (Nz + 1 + 2 * Ng);
    sc_ = (size_t)g_.cp * (Nz + 2 * Ng);
    g_.y_stretched = m.y_stretched;
    dy_min_ = m.dy;
    if (m.y_stretched) {
        if (!m.dyv || !m.dyc) throw std::invalid_argument("Solver: stretched mesh needs dyv/dyc");
        const int nv = Ny + 2 * Ng, nc = Ny + 2 * Ng + 1;
        h_dyv_.assign(m.dyv, m.dyv + nv);
        h_dyc_.assign(m.dyc, m.dyc + nc);
        std::vector<double> r(nv), rc(nc);
        for (int j = 0; j < nv; ++j) r[j] = 1.0 / m.dyv[j];
        for (int j = 0; j < nc; ++j) rc[j] = 1.0 / m.dyc[j];
        dyv_.alloc(nv); rdyv_.alloc(nv); dyc_.alloc(nc); rdyc_.alloc(nc);
        CFDNN_CUDA(cudaMemcpy(dyv_.p, m.dyv, nv * sizeof(double), cudaMemcpyHostToDevice));
        CFDNN_CUDA(cudaMemcpy(rdyv_.p, r.data(), nv * sizeof(double), cudaMemcpyHostToDevice));
        CFDNN_CUDA(cudaMemcpy(dyc_.p, m.dyc, nc * sizeof(double), cudaMemcpyHostToDevice));
        CFDNN_CUDA(cudaMemcpy(rdyc_.p, rc.data(), nc * sizeof(double), cudaMemcpyHostToDevice));
        g_.dyv = dyv_.p; g_.rdyv = rdyv_.p; g_.dyc = dyc_.p; g_.rdyc = rdyc_.p;
        dy_min_ = *std::min_element(h_dyv_.begin(), h_dyv_.end());   // src/solver.cpp:3672-3675
    }
    {
        std::vector<double> yc(Ny + 2 * Ng);
        for (int j = 0; j < Ny + 2 * Ng; ++j) yc[j] = m.yc ? m.yc[j] : m.y_lo + (j - Ng + 0.5) * m.dy;   // src/mesh.cpp:66-68
        yc_.alloc(yc.size(), false);
        CFDNN_CUDA(cudaMemcpy(yc_.p, yc.data(), yc.size() * sizeof(double), cudaMemcpyHostToDevice));
    }
    if (m.yc) h_yc_.assign(m.yc, m.yc + (Ny + 2 * Ng));
    if (m.yf) h_yf_.assign(m.yf, m.yf + (Ny + 2 * Ng + 1));
    mesh_.dyv = mesh_.dyc = mesh_.yLap_aS = mesh_.yLap_aN = mesh_.yLap_aP = mesh_.yc = mesh_.yf = nullptr;  // host pointers not kept

    vel_ = alloc_vel(); star_ = alloc_vel(); old_ = alloc_vel(); rk_ = alloc_vel();
    conv_ = alloc_vel(); diff_ = alloc_vel();
    p_ = alloc_cell(); pcorr_ = alloc_cell(); div_ = alloc_cell(); rhs_ = alloc_cell();
    nut_ = alloc_cell(); nueff_ = alloc_cell(); k_ = alloc_cell(); omega_ = alloc_cell();
    for (int t = 0; t < 3; ++t) tau_[t] = alloc_cell();

    const dim3 gmax = grid3(Nx + 1, Ny + 1, Nz + 1);
    npartial_ = (size_t)gmax.x * gmax.y * gmax.z;
    partial_.alloc(5 * std::max<size_t>(npartial_, 148 * 8) + 16);   // per-block partials (k_cfl_maxima: 5 rows of up to 1184)
    scalars_.alloc(32);
    CFDNN_CUDA(cudaMallocHost(&h_scalars_, 32 * sizeof(double)));
    CFDNN_CUDA(cudaStreamCreateWithFlags(&st_, cudaStreamNonBlocking));
    CFDNN_CUDA(cudaEventCreate(&ev0_));
    CFDNN_CUDA(cudaEventCreate(&ev1_));

    {
        // A y-stretched slab with the FFT solver requested: the semi-coarsening multigrid has no slab variant and is never
        // called there (comm_init / set_velocity_bc enforce it); it is built on the local box so the handle stays complete.
        cfdnn_mesh_desc mm = m;
        if (c.poisson_solver == CFDNN_POISSON_FFT && m.y_stretched && m.Nz_global > m.Nz) { mm.Nz_global = m.Nz; mm.k_offset = 0; }
        mg_ = std::make_unique<Multigrid>(mm);
    }
    if (is2d_) mg_->set_2d_plane(Ng, /*dup=*/true);
    if (c.poisson_solver == CFDNN_POISSON_FFT) {                 // selection: src/solver.cpp:432-447
        if (is2d_) throw std::invalid_argument("Solver: poisson_solver = FFT needs a 3-D mesh (the 2-D variant FFT2D is not built)");
        fft_ = std::make_unique<FftPoisson>(m);
    } else if (c.poisson_solver != CFDNN_POISSON_MG) {
        throw std::invalid_argument("Solver: poisson_solver must be CFDNN_POISSON_MG or CFDNN_POISSON_FFT");
    }
    if (const char* e = std::getenv("CFDNN_MOMENTUM_Z")) mom_z_ = std::atoi(e) != 0;          // 0: one-thread-per-point kernel
    if (const char* e = std::getenv("CFDNN_MOMENTUM_EXACT")) mom_exact_ = std::atoi(e) != 0;  // 1: exact divisions everywhere
    {
        using S = MomShape<MOM_BX, MOM_BY>;
#define CFDNN_MOMZ_ATTR(SK, EX, XZ_) CFDNN_CUDA(cudaFuncSetAttribute(k_momentum_z<SK, EX, XZ_, MOM_BX, MOM_BY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::SMEM))
        CFDNN_MOMZ_ATTR(true, true, false); CFDNN_MOMZ_ATTR(true, false, false); CFDNN_MOMZ_ATTR(false, true, false); CFDNN_MOMZ_ATTR(false, false, false);
        CFDNN_MOMZ_ATTR(true, true, true); CFDNN_MOMZ_ATTR(true, false, true); CFDNN_MOMZ_ATTR(false, true, true); CFDNN_MOMZ_ATTR(false, false, true);
#undef CFDNN_MOMZ_ATTR
    }
    dt_ = c.dt;
    set_velocity_bc(vbc_);
}

Solver::~Solver() {
    for (auto& e : phpool_) { if (e.a) cudaEventDestroy(e.a); if (e.b) cudaEventDestroy(e.b); }
    if (h_scalars_) cudaFreeHost(h_scalars_);
    if (ev0_) cudaEventDestroy(ev0_);
    if (ev1_) cudaEventDestroy(ev1_);
    if (st_) cudaStreamDestroy(st_);
}

Vel Solver::alloc_vel() {
    Vel v;
    store_.emplace_back(); store_.back().alloc(su_); v.u = store_.back().p;
    store_.emplace_back(); store_.back().alloc(sv_); v.v = store_.back().p;
    store_.emplace_back(); store_.back().alloc(sw_); v.w = store_.back().p;
    return v;
}
double* Solver::alloc_cell() {
    store_.emplace_back(); store_.back().alloc(sc_);
    return store_.back().p;
}

// set_velocity_bc (src/solver.cpp:650-697): validates periodic pairs and maps to Poisson BCs
void Solver::set_velocity_bc(const int bc[6]) {
    for (int d = 0; d < 6; ++d)
        if (bc[d] < 0 || bc[d] > 3) throw std::invalid_argument("set_velocity_bc: bad VelocityBC type");
    const char* axis[3] = {"x", "y", "z"};
    for (int a = 0; a < 3; ++a)
        if ((bc[2 * a] == CFDNN_VBC_PERIODIC) != (bc[2 * a + 1] == CFDNN_VBC_PERIODIC))
            throw std::invalid_argument(std::string("Periodic BC in ") + axis[a] + " requires both " + axis[a] +
                                        "_lo and " + axis[a] + "_hi to be Periodic");
    for (int d = 0; d < 6; ++d) vbc_[d] = bc[d];
    if (is2d_) vbc_[4] = vbc_[5] = CFDNN_VBC_PERIODIC;          // the embedding plane is periodic in z
    int pbc[6];
    for (int d = 0; d < 6; ++d) pbc[d] = vbc_[d] == CFDNN_VBC_PERIODIC ? CFDNN_PBC_PERIODIC : CFDNN_PBC_NEUMANN;
    mg_->set_bc(pbc);
    // FFT was requested: it stays selected while x and z are periodic and y is walled, otherwise the multigrid takes
    // over, as the reference does when the BCs change under a selected FFT solver (src/solver.cpp:727-739)
    fft_active_ = false;
    if (fft_) {
        const bool ok = pbc[0] == CFDNN_PBC_PERIODIC && pbc[1] == CFDNN_PBC_PERIODIC && pbc[4] == CFDNN_PBC_PERIODIC &&
                        pbc[5] == CFDNN_PBC_PERIODIC && pbc[2] == CFDNN_PBC_NEUMANN && pbc[3] == CFDNN_PBC_NEUMANN;
        if (ok) { fft_->set_bc(pbc); fft_active_ = true; }
    }
    if (multi() && g_.y_stretched && !fft_active_)
        throw std::invalid_argument("set_velocity_bc: on z-slabs a y-stretched mesh needs the FFT Poisson solver (periodic x/z, "
                                    "walls y); the semi-coarsening multigrid has no slab variant");
    update_links();
}

void Solver::comm_init(Comm* c) {
    if (c && c->nranks > 1) {
        // The slab layout is block distribution with EQUAL slabs (reference Decomposition::nz_for_rank with
        // Nz_global % nprocs == 0, include/decomposition.hpp:81-92): the coarse-level offsets (rank * Nz / 2),
        // the in-place all-gather of the agglomerated level and the 2:1 vertex-aligned z coarsening all rely on it.
        // A remainder (ranks owning Nz_global / P + 1 planes) is rejected here instead of giving wrong coarse
        // right-hand sides.
        const int NzG = mesh_.Nz_global, Nz = g_.Nz;
        if (NzG <= Nz)
            throw std::invalid_argument("comm_init: mesh.Nz_global must be the global plane count (> local Nz) in a multi-rank run");
        if ((long long)Nz * c->nranks != NzG)
            throw std::invalid_argument("comm_init: unequal z-slabs are not supported (need Nz_global == Nz_local * nranks; "
                                        "Nz_global = " + std::to_string(NzG) + ", Nz_local = " + std::to_string(Nz) +
                                        ", nranks = " + std::to_string(c->nranks) + ")");
        if (mesh_.k_offset != c->rank * Nz)
            throw std::invalid_argument("comm_init: mesh.k_offset must be rank * Nz_local (" + std::to_string(c->rank * Nz) +
                                        "), got " + std::to_string(mesh_.k_offset));
    }
    comm_ = c;
    // the semi-coarsening (y-stretched) multigrid has no slab variant; with the FFT solver selected and active it is never
    // called, so a stretched channel can still be decomposed
    const bool mg_unused = c && c->nranks > 1 && g_.y_stretched && fft_ && fft_active_;
    if (!mg_unused) mg_->set_comm(c);
    if (fft_) fft_->set_comm(c);                     // the reference's FFT_MPI role: slabs + all-to-all transposes
    update_links();
}
bool Solver::multi() const { return comm_ && comm_->nranks > 1; }
void Solver::update_links() {
    lo_int_ = hi_int_ = lo_wrap_ = hi_wrap_ = false;
    if (!multi()) return;
    const bool pz = vbc_[4] == CFDNN_VBC_PERIODIC && vbc_[5] == CFDNN_VBC_PERIODIC;
    lo_int_ = comm_->rank > 0;
    hi_int_ = comm_->rank < comm_->nranks - 1;
    lo_wrap_ = pz && comm_->rank == 0;
    hi_wrap_ = pz && comm_->rank == comm_->nranks - 1;
    if (seam_scratch_.n != (size_t)g_.wp) seam_scratch_.alloc((size_t)g_.wp);
}

// ---------------------------------------------------------------- phase timers
Solver::PhaseEv* Solver::ph_begin(int phase) {
    if (phused_ >= 16384) return nullptr;
    if (phused_ == phpool_.size()) {
        PhaseEv e;
        CFDNN_CUDA(cudaEventCreate(&e.a));
        CFDNN_CUDA(cudaEventCreate(&e.b));
        phpool_.push_back(e);
    }
    PhaseEv* e = &phpool_[phused_++];
    e->phase = phase;
    CFDNN_CUDA(cudaEventRecord(e->a, st_));
    return e;
}
void Solver::ph_end(PhaseEv* e) {
    if (e) CFDNN_CUDA(cudaEventRecord(e->b, st_));
}

// ---------------------------------------------------------------- field transfer
double* Solver::field_ptr(int field, size_t* n) const {
    switch (field) {
        case CFDNN_F_U: *n = su_; return vel_.u;
        case CFDNN_F_V: *n = sv_; return vel_.v;
        case CFDNN_F_W: *n = sw_; return vel_.w;
        case CFDNN_F_P: *n = sc_; return p_;
        case CFDNN_F_PCORR: *n = sc_; return pcorr_;
        case CFDNN_F_DIV: *n = sc_; return div_;
        case CFDNN_F_RHS: *n = sc_; return rhs_;
        case CFDNN_F_NUT: *n = sc_; return nut_;
        case CFDNN_F_K: *n = sc_; return k_;
        case CFDNN_F_OMEGA: *n = sc_; return omega_;
        case CFDNN_F_NUEFF: *n = sc_; return nueff_;
        case CFDNN_F_TAU_XX: *n = sc_; return tau_[0];
        case CFDNN_F_TAU_XY: *n = sc_; return tau_[1];
        case CFDNN_F_TAU_YY: *n = sc_; return tau_[2];
    }
    throw std::invalid_argument("unknown field id");
}
double* Solver::view_ptr(int field, size_t* n) const {
    switch (field) {
        case CFDNN_F_U_STAR: *n = su_; return star_.u;
        case CFDNN_F_V_STAR: *n = sv_; return star_.v;
        case CFDNN_F_W_STAR: *n = sw_; return star_.w;
        case CFDNN_F_U_OLD: *n = su_; return old_.u;
        case CFDNN_F_V_OLD: *n = sv_; return old_.v;
        case CFDNN_F_W_OLD: *n = sw_; return old_.w;
        case CFDNN_F_CONV_U: *n = su_; return conv_.u;
        case CFDNN_F_CONV_V: *n = sv_; return conv_.v;
        case CFDNN_F_CONV_W: *n = sw_; return conv_.w;
        case CFDNN_F_DIFF_U: *n = su_; return diff_.u;
        case CFDNN_F_DIFF_V: *n = sv_; return diff_.v;
        case CFDNN_F_DIFF_W: *n = sw_; return diff_.w;
    }
    return field_ptr(field, n);
}
size_t Solver::field_size(int field) const { size_t n; field_ptr(field, &n); return n; }
void Solver::upload(int field, const double* host, size_t n) {
    size_t want; double* d = field_ptr(field, &want);
    if (n != want) throw std::invalid_argument("upload: size mismatch for field " + std::to_string(field));
    if (is2d_) {     // host plane 0 holds the data: replicate it into every device plane
        const size_t pl = plane_elems(field), np = n / pl;
        for (size_t k = 0; k < np; ++k)
            CFDNN_CUDA(cudaMemcpyAsync(d + k * pl, host, pl * sizeof(double), cudaMemcpyHostToDevice, st_));
    } else {
        CFDNN_CUDA(cudaMemcpyAsync(d, host, n * sizeof(double), cudaMemcpyHostToDevice, st_));
    }
    CFDNN_CUDA(cudaStreamSynchronize(st_));
}
size_t Solver::plane_elems(int field) const {
    switch (field) {
        case CFDNN_F_U: case CFDNN_F_U_STAR: case CFDNN_F_U_OLD: case CFDNN_F_CONV_U: case CFDNN_F_DIFF_U: return (size_t)g_.up;
        case CFDNN_F_V: case CFDNN_F_V_STAR: case CFDNN_F_V_OLD: case CFDNN_F_CONV_V: case CFDNN_F_DIFF_V: return (size_t)g_.vp;
        case CFDNN_F_W: case CFDNN_F_W_STAR: case CFDNN_F_W_OLD: case CFDNN_F_CONV_W: case CFDNN_F_DIFF_W: return (size_t)g_.wp;
        default: return (size_t)g_.cp;
    }
}
void Solver::download(int field, double* host, size_t n) {          // the work fields of the device view can be read too
    size_t want; double* d = view_ptr(field, &want);
    if (n != want) throw std::invalid_argument("download: size mismatch for field " + std::to_string(field));
    if (is2d_) {     // device plane Ng -> host plane 0; the reference never touches the other planes (zeros)
        const size_t pl = plane_elems(field);
        std::fill(host + pl, host + n, 0.0);
        CFDNN_CUDA(cudaMemcpyAsync(host, d + (size_t)g_.Ng * pl, pl * sizeof(double), cudaMemcpyDeviceToHost, st_));
    } else {
        CFDNN_CUDA(cudaMemcpyAsync(host, d, n * sizeof(double), cudaMemcpyDeviceToHost, st_));
    }
    CFDNN_CUDA(cudaStreamSynchronize(st_));
    check_launch("download");
}
void Solver::initialize() {
    apply_velocity_bc();
    CFDNN_CUDA(cudaStreamSynchronize(st_));
    check_launch("initialize");
}

// ---------------------------------------------------------------- operators
void Solver::apply_velocity_bc() {
    const int Nx = g_.Nx, Ny = g_.Ny, Nz = g_.Nz, Ng = g_.Ng;
    const int xlo = vbc_[0], xhi = vbc_[1], ylo = vbc_[2], yhi = vbc_[3], zlo = vbc_[4], zhi = vbc_[5];
    // src/solver_operators.cpp:70-81, 283-288
    if (xlo != CFDNN_VBC_PERIODIC && xlo != CFDNN_VBC_NOSLIP && xlo != CFDNN_VBC_INFLOW)
        throw Error(CFDNN_ERR_RUNTIME, "Unsupported velocity BC type for x_lo");
    if (xhi != CFDNN_VBC_PERIODIC && xhi != CFDNN_VBC_NOSLIP && xhi != CFDNN_VBC_OUTFLOW)
        throw Error(CFDNN_ERR_RUNTIME, "Unsupported velocity BC type for x_hi");
    if (ylo != CFDNN_VBC_PERIODIC && ylo != CFDNN_VBC_NOSLIP)
        throw Error(CFDNN_ERR_RUNTIME, "Unsupported velocity BC type for y_lo (only Periodic and NoSlip are implemented)");
    if (yhi != CFDNN_VBC_PERIODIC && yhi != CFDNN_VBC_NOSLIP)
        throw Error(CFDNN_ERR_RUNTIME, "Unsupported velocity BC type for y_hi (only Periodic and NoSlip are implemented)");
    // z sides that are slab links carry no physical BC here: the halo exchange below is their z pass
    const bool lo_link = lo_int_ || lo_wrap_, hi_link = hi_int_ || hi_wrap_;
    if (!lo_link && zlo != CFDNN_VBC_PERIODIC && zlo != CFDNN_VBC_NOSLIP)
        throw Error(CFDNN_ERR_RUNTIME, "Unsupported velocity BC type for z_lo (only Periodic and NoSlip are implemented)");
    if (!hi_link && zhi != CFDNN_VBC_PERIODIC && zhi != CFDNN_VBC_NOSLIP)
        throw Error(CFDNN_ERR_RUNTIME, "Unsupported velocity BC type for z_hi (only Periodic and NoSlip are implemented)");
    if (multi() && ((zlo == CFDNN_VBC_PERIODIC) != (zhi == CFDNN_VBC_PERIODIC)))
        throw std::invalid_argument("Periodic BC in z requires both z_lo and z_hi to be Periodic");
    BcFlags bc{xlo, xhi, ylo, yhi, lo_link || (multi() && zlo == CFDNN_VBC_PERIODIC) ? -1 : zlo,
               hi_link || (multi() && zhi == CFDNN_VBC_PERIODIC) ? -1 : zhi};
    const int Pz = Nz + 2 * Ng;   // all planes of u, v
    auto pass = [&](int id, int na, int nb) {
        dim3 gr((na + kB.x - 1) / kB.x, (nb + kB.y - 1) / kB.y, 1);
        CFDNN_LAUNCH(k_velocity_bc, gr, kB, 0, st_, g_, vel_, bc, id, na, nb);
    };
    pass(BC_U_X, Ny + 2 * Ng, Pz);
    pass(BC_U_Y, Nx + 1 + 2 * Ng, Pz);
    pass(BC_V_X, Ny + 1 + 2 * Ng, Pz);
    pass(BC_V_Y, Nx + 2 * Ng, Pz);
    if (xlo == CFDNN_VBC_PERIODIC && xhi == CFDNN_VBC_PERIODIC) pass(BC_U_X2, Ny + 2 * Ng, Pz);
    if (ylo == CFDNN_VBC_PERIODIC && yhi == CFDNN_VBC_PERIODIC) pass(BC_V_Y2, Nx + 2 * Ng, Pz);
    if (xhi == CFDNN_VBC_OUTFLOW) pass(BC_OUTFLOW_UV, Ny + 1 + 2 * Ng, Pz);
    if (xlo == CFDNN_VBC_INFLOW) pass(BC_INFLOW_UV, Ny + 1 + 2 * Ng, Pz);
    pass(BC_U_Z, Nx + 1 + 2 * Ng, Ny + 2 * Ng);
    pass(BC_V_Z, Nx + 2 * Ng, Ny + 1 + 2 * Ng);
    pass(BC_W_Z, Nx + 2 * Ng, Ny + 2 * Ng);
    if (multi()) halo_vel(vel_, /*bc_site=*/true);
    const int Pw = Nz + 1 + 2 * Ng;
    pass(BC_W_X, Ny + 2 * Ng, Pw);
    if (xlo == CFDNN_VBC_INFLOW) pass(BC_W_X_INFLOW, Ny + 2 * Ng, Pw);
    if (xhi == CFDNN_VBC_OUTFLOW) pass(BC_W_X_OUTFLOW, Ny + 2 * Ng, Pw);
    pass(BC_W_Y, Nx + 2 * Ng, Pw);
}

void Solver::compute_convective_term() {
    const dim3 gr = grid3(g_.Nx + 1, g_.Ny + 1, g_.Nz + 1);
    if (cfg_.convective_scheme == CFDNN_CONV_SKEW)
        CFDNN_LAUNCH((k_convective<true>), gr, kB, 0, st_, g_, vel_.u, vel_.v, vel_.w, conv_.u, conv_.v, conv_.w);
    else
        CFDNN_LAUNCH((k_convective<false>), gr, kB, 0, st_, g_, vel_.u, vel_.v, vel_.w, conv_.u, conv_.v, conv_.w);
}
void Solver::compute_diffusive_term() {
    const dim3 gr = grid3(g_.Nx + 1, g_.Ny + 1, g_.Nz + 1);
    CFDNN_LAUNCH(k_diffusive, gr, kB, 0, st_, g_, vel_.u, vel_.v, vel_.w, nueff_, diff_.u, diff_.v, diff_.w);
}
// periodic seam of the duplicated faces of `out` (see k_seam)
void Solver::seam(const Vel& out, bool seam_average) {
    const int px = vbc_[0] == CFDNN_VBC_PERIODIC && vbc_[1] == CFDNN_VBC_PERIODIC;
    const int py = vbc_[2] == CFDNN_VBC_PERIODIC && vbc_[3] == CFDNN_VBC_PERIODIC;
    const int pzg = vbc_[4] == CFDNN_VBC_PERIODIC && vbc_[5] == CFDNN_VBC_PERIODIC;
    const int pz = pzg && !multi();
    if (px || py || pz) {
        const int na = std::max(g_.Nx, g_.Ny), nb = std::max(g_.Ny, g_.Nz);
        dim3 gs((na + kB.x - 1) / kB.x, (nb + kB.y - 1) / kB.y, 3);
        if (seam_average) CFDNN_LAUNCH((k_seam<true>), gs, kB, 0, st_, g_, out, px, py, pz);
        else CFDNN_LAUNCH((k_seam<false>), gs, kB, 0, st_, g_, out, px, py, pz);
    }
    if (pzg && multi()) seam_z(out, seam_average);
}
// out = in + dt * (-conv + diff + f) with conv/diff evaluated on vel_ (the reference's operators
// always read velocity_), fused into one pass.  When `out` IS vel_ (RK stages 2-3) the
// result goes to a scratch buffer that also receives vel_'s ghost entries, and the two are swapped.
void Solver::momentum(const Vel& in, Vel& out, double dt, bool seam_average) {
    const int Ng = g_.Ng;
    const dim3 gr = grid3(g_.Nx + 1 + 2 * Ng, g_.Ny + 1 + 2 * Ng, g_.Nz + 1 + 2 * Ng);
    const bool alias = out.u == vel_.u;
    if (is2d_) {
        // the operators read vel_'s z-neighbour planes: make them copies of the data plane (x/y
        // ghosts included, stale or not — exactly what the reference's 2-D kernels would read)
        const size_t up = (size_t)g_.up, vp = (size_t)g_.vp;
        for (int k = 0; k <= 2; k += 2) {
            CFDNN_CUDA(cudaMemcpyAsync(vel_.u + k * up, vel_.u + up, up * sizeof(double), cudaMemcpyDeviceToDevice, st_));
            CFDNN_CUDA(cudaMemcpyAsync(vel_.v + k * vp, vel_.v + vp, vp * sizeof(double), cudaMemcpyDeviceToDevice, st_));
        }
    }
    const Vel src_in = in;                       // copies of the pointer triples (in may be vel_ itself)
    const Vel dst = alias ? conv_ : out;
    if (mom_z_ || implicit_y()) {   // (the xz-only explicit term of implicit y-diffusion exists in this kernel only)
        // z-marching tile kernel (momentum_kernels.cuh).  Exact divisions + unfused products wherever the result
        // must carry the x86 build's bits: y-stretched meshes (their Poisson solver amplifies a last-bit change)
        // and the 2-D embedding; reciprocal multiplies + FMA on uniform 3-D meshes.
        constexpr int BX = MOM_BX, BY = MOM_BY;
        using S = MomShape<BX, BY>;
        MomCoef c{};
        c.dx = g_.dx; c.dy = g_.dy; c.dz = g_.dz; c.dx_2 = g_.dx_2; c.dy_2 = g_.dy_2; c.dz_2 = g_.dz_2;
        c.dx2 = g_.dx2; c.dy2 = g_.dy2; c.dz2 = g_.dz2;
        c.h_rdx = 0.5 / g_.dx.d; c.h_rdy = 0.5 / g_.dy.d; c.h_rdz = 0.5 / g_.dz.d;
        c.h_r2dx = 0.5 / g_.dx_2.d; c.h_r2dy = 0.5 / g_.dy_2.d; c.h_r2dz = 0.5 / g_.dz_2.d;
        c.dt = dt; c.fx = fx_; c.fy = fy_; c.fz = fz_;
        const int ex = g_.Nx + 2 * Ng, ey = g_.Ny + 2 * Ng, ez = g_.Nz + 2 * Ng;
        const int tiles = ((ex + BX) / BX) * ((ey + BY) / BY);
        int zc = 64;                              // planes per CTA: 2 extra plane loads per chunk
        while (zc > 8 && (long long)tiles * ((ez + zc) / zc) < 4LL * 148) zc /= 2;
        const dim3 grid((ex + BX) / BX, (ey + BY) / BY, (ez + zc) / zc);
        const bool exact = mom_exact_ || g_.y_stretched || is2d_;
        const bool skew = cfg_.convective_scheme == CFDNN_CONV_SKEW;
#define CFDNN_MOMZ(SK, EX, XZ_)                                                                                      \
        CFDNN_LAUNCH_COOP((k_momentum_z<SK, EX, XZ_, BX, BY>), grid, BX * BY, S::SMEM, st_, g_, c, vel_, nueff_, src_in, dst, \
                          alias ? 1 : 0, zc)
        const bool xz = implicit_y();
        if (skew) {
            if (exact) { if (xz) CFDNN_MOMZ(true, true, true); else CFDNN_MOMZ(true, true, false); }
            else { if (xz) CFDNN_MOMZ(true, false, true); else CFDNN_MOMZ(true, false, false); }
        } else {
            if (exact) { if (xz) CFDNN_MOMZ(false, true, true); else CFDNN_MOMZ(false, true, false); }
            else { if (xz) CFDNN_MOMZ(false, false, true); else CFDNN_MOMZ(false, false, false); }
        }
#undef CFDNN_MOMZ
    } else if (cfg_.convective_scheme == CFDNN_CONV_SKEW)
        CFDNN_LAUNCH((k_momentum<true>), gr, kB, 0, st_, g_, vel_, nueff_, src_in, dst, dt, fx_, fy_, fz_, alias ? 1 : 0);
    else
        CFDNN_LAUNCH((k_momentum<false>), gr, kB, 0, st_, g_, vel_, nueff_, src_in, dst, dt, fx_, fy_, fz_, alias ? 1 : 0);
    if (alias) std::swap(vel_, conv_);           // `out` refers to vel_: it now names the new buffer
    seam(out, seam_average);
}
void Solver::blend(const Vel& a, const Vel& b, const Vel& c, double a1, double a2) {
    const dim3 gr = grid3(g_.Nx + 1, g_.Ny + 1, g_.Nz + 1);
    CFDNN_LAUNCH(k_blend, gr, kB, 0, st_, g_, a, b, c, a1, a2);
}
void Solver::compute_divergence(const Vel& v, bool uniform_dy) {
    const dim3 gr = grid3(g_.Nx, g_.Ny, g_.Nz);
    Geom g = g_;
    if (uniform_dy) g.y_stretched = 0;
    CFDNN_LAUNCH_COOP(k_divergence, gr, kB, 0, st_, g, v, div_, partial_.p);
}
// mean divergence + rhs (src/solver_time.cpp:812-866; src/solver.cpp:1920-2078)
void Solver::build_rhs(double dt) {
    const dim3 gr = grid3(g_.Nx, g_.Ny, g_.Nz);
    const int nb = gr.x * gr.y * gr.z;
    const bool dist = comm_ && comm_->nranks > 1;
    const int NzG = mesh_.Nz_global > 0 ? mesh_.Nz_global : g_.Nz;
    const double count = (double)g_.Nx * g_.Ny * NzG;
    // Stretched meshes under the multigrid: the reference's loop-order sum, bit for bit (the semi-coarsening solver
    // amplifies a last-bit change of the right-hand side).  The FFT solver is direct and removes the mean itself, so the
    // serial sum (25 ms at 192x96x192, 3 per step) is not paid there.
    if (g_.y_stretched && !dist && !fft_active_) {
        CFDNN_LAUNCH_COOP(k_seq_sum_cells, 1, 256, 0, st_, g_, div_, scalars_.p + 0, count);   // bit-exact mean
    } else {
        CFDNN_LAUNCH_COOP(k_final_sum, 1, 1024, 0, st_, partial_.p, nb, scalars_.p + 0, dist ? 0.0 : count);
        if (dist) {
            comm_->allreduce_sum(scalars_.p + 0, 1, st_);
            CFDNN_LAUNCH_COOP(k_final_sum, 1, 32, 0, st_, scalars_.p + 0, 1, scalars_.p + 0, count);
        }
    }
    CFDNN_LAUNCH(k_build_rhs, gr, kB, 0, st_, g_, div_, scalars_.p + 1, 1.0 / dt, rhs_);
}
void Solver::correct_velocity() {
    const dim3 gr = grid3(g_.Nx + 1, g_.Ny + 1, g_.Nz + 1);
    CFDNN_LAUNCH(k_correct, gr, kB, 0, st_, g_, star_, vel_, pcorr_, p_, dt_);
    seam(vel_, true);
}
// RK path: velocity_star_ <- velocity_ (interior) and the correction in one pass
void Solver::copy_correct_velocity() {
    const dim3 gr = grid3(g_.Nx + 1, g_.Ny + 1, g_.Nz + 1);
    CFDNN_LAUNCH(k_copy_correct, gr, kB, 0, st_, g_, star_, vel_, pcorr_, p_, dt_);
    seam(vel_, true);
}

int Solver::poisson_solve(const cfdnn_poisson_cfg& pc) {
    if (fft_active_) {                               // dispatch on the selected solver, src/solver_time.cpp:875-884
        PhaseEv* pe = ph_begin(CFDNN_PHASE_POISSON);
        const int it = fft_->solve_dev(rhs_, pcorr_, st_);
        ph_end(pe);
        cycle_log.push_back(it);
        return it;
    }
    PhaseEv* pe = ph_begin(CFDNN_PHASE_POISSON);
    const int cyc = mg_->solve_dev(rhs_, pcorr_, pc, st_);
    ph_end(pe);
    cycle_log.push_back(cyc);
    return cyc;
}

// ---------------------------------------------------------------- z-slab exchanges
void Solver::halo_vel(const Vel& v, bool bc_site) {
    if (!multi()) return;
    const bool lo = lo_int_ || (bc_site && lo_wrap_), hi = hi_int_ || (bc_site && hi_wrap_);
    if (!lo && !hi) return;
    const int Ng = g_.Ng, Nz = g_.Nz;
    comm_->exchange_planes(v.u, (size_t)g_.up, Nz, Ng, lo, hi, st_);
    comm_->exchange_planes(v.v, (size_t)g_.vp, Nz, Ng, lo, hi, st_);
    // w: my faces Ng..Ng+Nz-1 are owned, face Ng+Nz duplicates the upper neighbour's face Ng.
    // down: faces Ng, Ng+1 -> the lower rank's Ng+Nz (far <- near, BC_W_Z) and Ng+Nz+1 ghost;
    // up  : face Ng+Nz-1   -> the upper rank's Ng-1 ghost.
    const size_t wp = (size_t)g_.wp;
    comm_->exchange_ptrs(v.w + (size_t)Ng * wp, 2 * wp, v.w + (size_t)(Ng - 1) * wp, wp,
                         v.w + (size_t)(Ng + Nz - 1) * wp, wp, v.w + (size_t)(Ng + Nz) * wp, 2 * wp, lo, hi, st_);
}
void Solver::halo_cell(double* c, bool bc_site) {
    if (!multi()) return;
    const bool lo = lo_int_ || (bc_site && lo_wrap_), hi = hi_int_ || (bc_site && hi_wrap_);
    if (!lo && !hi) return;
    comm_->exchange_planes(c, (size_t)g_.cp, g_.Nz, g_.Ng, lo, hi, st_);
}
// Periodic seam of w between the end ranks: the duplicated faces are rank 0's Ng ("near") and the
// last rank's Ng+Nz ("far").  average = false: far <- near (euler_substep, src/solver_time.cpp:518-559);
// true: both <- 0.5*(near + far) (correct_velocity src/solver_operators.cpp:1469-1485, Euler predictor).
void Solver::seam_z(const Vel& v, bool average) {
    if (!lo_wrap_ && !hi_wrap_) return;
    const size_t wp = (size_t)g_.wp;
    double* near_face = v.w + (size_t)g_.Ng * wp;
    double* far_face = v.w + (size_t)(g_.Ng + g_.Nz) * wp;
    if (!average) {
        comm_->exchange_ptrs(near_face, wp, nullptr, 0, nullptr, 0, far_face, wp, lo_wrap_, hi_wrap_, st_);
        return;
    }
    comm_->exchange_ptrs(near_face, wp, seam_scratch_.p, wp, far_face, wp, seam_scratch_.p, wp, lo_wrap_, hi_wrap_, st_);
    const int blocks = (int)((wp + 255) / 256);
    if (lo_wrap_) CFDNN_LAUNCH(k_avg_plane, blocks, 256, 0, st_, near_face, near_face, seam_scratch_.p, wp);
    if (hi_wrap_) CFDNN_LAUNCH(k_avg_plane, blocks, 256, 0, st_, far_face, seam_scratch_.p, far_face, wp);
}

// ---------------------------------------------------------------- integrators
// euler_substep (src/solver_time.cpp:269-661): operators always read velocity_
void Solver::euler_substep(const Vel& in, Vel& out, double dt) {
    halo_vel(vel_, /*bc_site=*/false);
    PhaseEv* e = ph_begin(CFDNN_PHASE_MOMENTUM);
    momentum(in, out, dt, /*seam_average=*/false);
    ph_end(e);
    if (implicit_y()) implicit_y_diffusion_step(out, dt);           // solver_time.cpp:657-660
}

bool Solver::implicit_y() const {                                    // solver_time.cpp:668-670
    return cfg_.implicit_y_diffusion && !(vbc_[2] == CFDNN_VBC_PERIODIC && vbc_[3] == CFDNN_VBC_PERIODIC);
}

// implicit_y_diffusion_step (src/solver_time.cpp:667-735): Thomas solve per y-line of u, v, w (2-D: u, v on the data
// plane).  c' goes to the diff_ buffers, which the fused momentum kernel leaves unused.
void Solver::implicit_y_diffusion_step(const Vel& v, double dt) {
    const int Ng = g_.Ng, Nx = g_.Nx, Nz = g_.Nz;
    const double dt_inv_dy2 = dt / (mesh_.dy * mesh_.dy);
    const dim3 blk(128);
    auto grid = [&](int ni, int nk) { return dim3((ni + 127) / 128, nk); };
    const int kc = is2d_ ? 1 : Nz;                  // cell planes: the data plane k = Ng of the 2-D embedding only
    PhaseEv* pe = ph_begin(CFDNN_PHASE_IMPLICIT_Y);
    if (g_.y_stretched) {
        CFDNN_LAUNCH((k_thomas_y<0, true>), grid(Nx + 1, kc), blk, 0, st_, g_, v.u, diff_.u, nueff_, dt, dt_inv_dy2, Ng, kc);
        if (g_.Ny > 1) CFDNN_LAUNCH((k_thomas_y<1, true>), grid(Nx, kc), blk, 0, st_, g_, v.v, diff_.v, nueff_, dt, dt_inv_dy2, Ng, kc);
        if (!is2d_) CFDNN_LAUNCH((k_thomas_y<2, true>), grid(Nx, Nz + 1), blk, 0, st_, g_, v.w, diff_.w, nueff_, dt, dt_inv_dy2, Ng, Nz + 1);
    } else {
        CFDNN_LAUNCH((k_thomas_y<0, false>), grid(Nx + 1, kc), blk, 0, st_, g_, v.u, diff_.u, nueff_, dt, dt_inv_dy2, Ng, kc);
        if (g_.Ny > 1) CFDNN_LAUNCH((k_thomas_y<1, false>), grid(Nx, kc), blk, 0, st_, g_, v.v, diff_.v, nueff_, dt, dt_inv_dy2, Ng, kc);
        if (!is2d_) CFDNN_LAUNCH((k_thomas_y<2, false>), grid(Nx, Nz + 1), blk, 0, st_, g_, v.w, diff_.w, nueff_, dt, dt_inv_dy2, Ng, Nz + 1);
    }
    ph_end(pe);
}

// project_velocity (src/solver_time.cpp:736-1211); which == &vel_ means "no swap"
void Solver::project_velocity(Vel* which, double dt) {
    const bool need_swap = which != &vel_;
    if (need_swap) std::swap(vel_, *which);
    struct SwapBack {                                // an exception mid-projection must not leave the buffers swapped
        Vel* a; Vel* b; bool on;
        ~SwapBack() { if (on) std::swap(*a, *b); }
    } guard{&vel_, which, need_swap};
    apply_velocity_bc();
    // the RK path's inline 2-D divergence divides by the UNIFORM dy even on a y-stretched mesh
    // (src/solver_time.cpp:789-806; the 3-D branch and the Euler path use dyv[j]) — reproduced as is
    compute_divergence(vel_, /*uniform_dy=*/is2d_);
    build_rhs(dt);
    cfdnn_poisson_cfg pc;
    cfdnn_poisson_cfg_default(&pc);
    pc.max_vcycles = cfg_.poisson_max_vcycles;     // :868-871 — everything else stays at defaults
    pc.tol_rhs = cfg_.poisson_tol_rhs;
    pc.fixed_cycles = cfg_.poisson_fixed_cycles;
    poisson_solve(pc);
    copy_correct_velocity();                       // velocity_ -> velocity_star_ (interior) :972-995, then correct
    apply_velocity_bc();
}

void Solver::ssprk3_step(double dt) {              // src/solver_time.cpp:1303-1513
    // u^n -> velocity_rk_ was done together with velocity_old_ <- velocity_ in step_device (k_copy2)
    euler_substep(vel_, star_, dt);
    project_velocity(&star_, dt);
    euler_substep(star_, vel_, dt);
    blend(rk_, vel_, star_, 0.75, 0.25);           // blend_to: star = 0.75 rk + 0.25 velocity_
    project_velocity(&star_, dt);
    euler_substep(star_, vel_, dt);
    blend(rk_, vel_, vel_, 1.0 / 3.0, 2.0 / 3.0);
    project_velocity(&vel_, dt);
    apply_velocity_bc();
}
void Solver::ssprk2_step(double dt) {              // :1213-1301
    euler_substep(vel_, star_, dt);
    project_velocity(&star_, dt);
    euler_substep(star_, vel_, dt);
    blend(rk_, vel_, vel_, 0.5, 0.5);
    project_velocity(&vel_, dt);
}

// Euler path of step() (src/solver.cpp:1505-2507)
void Solver::euler_step() {
    halo_vel(vel_, /*bc_site=*/false);
    PhaseEv* pe = ph_begin(CFDNN_PHASE_MOMENTUM);
    momentum(vel_, star_, dt_, /*seam_average=*/true);         // conv + diff + predictor :1505-1722
    ph_end(pe);
    trip_force(star_, dt_);                                      // :1724-1810
    const bool px = vbc_[0] == CFDNN_VBC_PERIODIC && vbc_[1] == CFDNN_VBC_PERIODIC;
    const bool py = vbc_[2] == CFDNN_VBC_PERIODIC && vbc_[3] == CFDNN_VBC_PERIODIC;
    const bool pz = vbc_[4] == CFDNN_VBC_PERIODIC && vbc_[5] == CFDNN_VBC_PERIODIC;
    if (!px || !py || !pz) {                                    // :1814-1845
        std::swap(vel_, star_);
        apply_velocity_bc();
        std::swap(vel_, star_);
    }
    if (implicit_y()) implicit_y_diffusion_step(star_, dt_);     // :1849-1853
    compute_divergence(star_);
    build_rhs(dt_);
    // iter_ is only advanced by solve_steady*, so a step()-driven run zeroes p' every step
    // (:2006-2012, 2081-2090)
    CFDNN_CUDA(cudaMemsetAsync(pcorr_, 0, sc_ * sizeof(double), st_));
    cfdnn_poisson_cfg pc;                                       // :2179-2205
    cfdnn_poisson_cfg_default(&pc);
    pc.max_vcycles = cfg_.poisson_max_vcycles;
    pc.tol_abs = cfg_.poisson_tol_abs;
    pc.tol_rhs = cfg_.poisson_tol_rhs;
    pc.tol_rel = cfg_.poisson_tol_rel;
    pc.check_interval = cfg_.poisson_check_interval;
    pc.use_l2_norm = cfg_.poisson_use_l2_norm;
    pc.linf_safety_factor = cfg_.poisson_linf_safety;
    pc.fixed_cycles = cfg_.poisson_fixed_cycles;
    pc.nu1 = cfg_.poisson_nu1;
    pc.nu2 = cfg_.poisson_nu2;
    pc.chebyshev_degree = cfg_.poisson_chebyshev_degree;
    if (pc.fixed_cycles <= 0) {
        // effective legacy tolerance needs rms(rhs) (:2135-2176, 2207-2209)
        const dim3 gr = grid3(g_.Nx, g_.Ny, g_.Nz);
        const int nb = gr.x * gr.y * gr.z;
        CFDNN_LAUNCH_COOP(k_sumsq_cells, gr, kB, 0, st_, g_, rhs_, partial_.p);
        CFDNN_LAUNCH_COOP(k_final_sum, 1, 1024, 0, st_, partial_.p, nb, scalars_.p + 2, 0.0);
        if (comm_ && comm_->nranks > 1) comm_->allreduce_sum(scalars_.p + 2, 1, st_);
        CFDNN_CUDA(cudaMemcpyAsync(h_scalars_ + 2, scalars_.p + 2, sizeof(double), cudaMemcpyDeviceToHost, st_));
        CFDNN_CUDA(cudaStreamSynchronize(st_));
        const int NzG = mesh_.Nz_global > 0 ? mesh_.Nz_global : g_.Nz;
        const double cnt = std::max((double)g_.Nx * g_.Ny * NzG, 1.0);
        const double rhs_rms = std::sqrt(h_scalars_[2] / cnt);
        pc.tol = std::max(cfg_.poisson_tol * std::max(rhs_rms, 1e-12), cfg_.poisson_abs_tol_floor);
    }
    poisson_solve(pc);
    correct_velocity();                                         // u = u* - dt grad p' :2443
    apply_velocity_bc();                                        // :2503-2507
}

void Solver::build_nu_eff() {
    const dim3 gr = grid3(g_.Nx + 2 * g_.Ng, g_.Ny + 2 * g_.Ng, g_.Nz + 2 * g_.Ng);
    CFDNN_LAUNCH(k_nu_eff, gr, kB, 0, st_, g_, cfg_.nu, nut_, has_model() ? 1 : 0, nueff_);
    if (has_model()) halo_cell(nueff_, /*bc_site=*/false);
}

// ---------------------------------------------------------------- trip forcing, velocity filter
void Solver::set_trip_forcing(const cfdnn_trip_desc* d) {
    if (!d) { trip_on_ = false; return; }
    if (!d->env_x || !d->g_y) throw std::invalid_argument("set_trip_forcing: env_x and g_y are required");
    if (!is2d_ && !d->Fz_v) throw std::invalid_argument("set_trip_forcing: a 3-D mesh needs Fz_v");
    const int Ng = g_.Ng;
    auto up = [](DevBuf<double>& b, const double* h, size_t n) {
        b.alloc(n, false);
        CFDNN_CUDA(cudaMemcpy(b.p, h, n * sizeof(double), cudaMemcpyHostToDevice));
    };
    up(trip_env_x_, d->env_x, g_.Nx + 2 * Ng);
    up(trip_g_y_, d->g_y, g_.Ny + 2 * Ng + 1);
    if (!is2d_) up(trip_fz_v_, d->Fz_v, g_.Nz + 2 * Ng);
    trip_w_ = !is2d_ && d->Fz_w != nullptr;                     // force_w = config.trip_force_w && is_3d :1730
    if (trip_w_) up(trip_fz_w_, d->Fz_w, g_.Nz + 2 * Ng + 1);
    trip_A_ = d->amplitude; trip_duration_ = d->duration; trip_ramp_start_ = d->ramp_off_start; trip_w_scale_ = d->w_scale;
    trip_on_ = true;
}

void Solver::trip_force(Vel& star, double dt) {
    if (!trip_on_ || !(time_ < trip_duration_)) return;         // is_trip_active(), include/solver.hpp:721
    double ramp = 1.0;                                          // get_trip_time_ramp :939-950
    if (time_ > trip_ramp_start_) {
        const double frac = (time_ - trip_ramp_start_) / (trip_duration_ - trip_ramp_start_);
        ramp = 0.5 * (1.0 + std::cos(M_PI * frac));
    }
    const double A = trip_A_ * ramp;
    const int nk = is2d_ ? 1 : g_.Nz;
    const dim3 blk(64, 4, 1);
    CFDNN_LAUNCH(k_trip_v, dim3((g_.Nx + 63) / 64, (g_.Ny + 1 + 3) / 4, nk), blk, 0, st_, g_, star.v, trip_env_x_.p, trip_g_y_.p,
                 is2d_ ? nullptr : trip_fz_v_.p, A, dt, g_.Ng, nk);
    if (trip_w_)
        CFDNN_LAUNCH(k_trip_w, dim3((g_.Nx + 63) / 64, (g_.Ny + 3) / 4, g_.Nz + 1), blk, 0, st_, g_, star.w, trip_env_x_.p,
                     trip_g_y_.p, trip_fz_w_.p, A * trip_w_scale_, dt);
}

void Solver::apply_velocity_filter(double strength) {           // src/solver.cpp:4065-4162
    if (strength <= 0.0) return;
    if (is2d_) return;                                           // the reference filters 3-D meshes only (:4090)
    if (multi()) throw std::invalid_argument("apply_velocity_filter: single-GPU only");
    const double alpha = strength * 0.25, alpha_y = alpha * 0.5;
    CFDNN_CUDA(cudaMemcpyAsync(conv_.u, vel_.u, su_ * sizeof(double), cudaMemcpyDeviceToDevice, st_));
    CFDNN_CUDA(cudaMemcpyAsync(conv_.v, vel_.v, sv_ * sizeof(double), cudaMemcpyDeviceToDevice, st_));
    CFDNN_CUDA(cudaMemcpyAsync(conv_.w, vel_.w, sw_ * sizeof(double), cudaMemcpyDeviceToDevice, st_));
    const dim3 blk(64, 4, 1);
    auto grid = [&](int nx, int ny, int nz) { return dim3((nx + 63) / 64, (ny + 3) / 4, nz); };
    CFDNN_LAUNCH((k_velocity_filter<0>), grid(g_.Nx + 1, g_.Ny, g_.Nz), blk, 0, st_, g_, vel_.u, conv_.u, alpha, alpha_y);
    CFDNN_LAUNCH((k_velocity_filter<1>), grid(g_.Nx, g_.Ny + 1, g_.Nz), blk, 0, st_, g_, vel_.v, conv_.v, alpha, alpha_y);
    CFDNN_LAUNCH((k_velocity_filter<2>), grid(g_.Nx, g_.Ny, g_.Nz + 1), blk, 0, st_, g_, vel_.w, conv_.w, alpha, alpha_y);
    CFDNN_CUDA(cudaStreamSynchronize(st_));
    check_launch("apply_velocity_filter");
}

void Solver::set_les_model(int model, double constant) {
    if (model == CFDNN_LES_NONE) { les_model_ = 0; return; }
    (void)les_default_constant(model);               // validates the id
    if (!les_met_) {
        les_mesh_ = mesh_;
        les_mesh_.yc = h_yc_.empty() ? nullptr : h_yc_.data();
        les_mesh_.yf = h_yf_.empty() ? nullptr : h_yf_.data();
        if (is2d_) les_mesh_.Nz = 1;
        les_met_ = std::make_unique<LesMetrics>(les_mesh_);
    }
    les_model_ = model;
    les_c_ = constant;
}

void Solver::set_sst_model(int kind) {
    sst_kind_ = kind;
    if (kind == CFDNN_RANS_NONE) { sst_.reset(); return; }
    if (!is2d_)
        throw std::invalid_argument("set_sst_model: the reference's SST k-omega transport is two-dimensional (k, omega, nu_t "
                                    "indexed (i, j)); on a 3-D mesh bridge the host model instead");
    cfdnn_mesh_desc m = mesh_;
    m.yc = h_yc_.empty() ? nullptr : h_yc_.data();
    m.dyc = h_dyc_.empty() ? nullptr : h_dyc_.data();
    m.dyv = h_dyv_.empty() ? nullptr : h_dyv_.data();
    sst_ = std::make_unique<SstModel>(m, cfg_.nu, kind);
}

void Solver::set_mixing_length_params(double kappa, double A_plus, double delta) {
    if (!sst_) throw std::invalid_argument("set_mixing_length_params: attach CFDNN_RANS_MIXING_LENGTH first");
    sst_->ml_kappa = kappa; sst_->ml_A_plus = A_plus; sst_->ml_delta = delta;
}

void Solver::set_gep_params(int variant, double model_nu) {
    if (!sst_) throw std::invalid_argument("set_gep_params: attach CFDNN_RANS_GEP first");
    if (variant < 0 || variant > 2) throw std::invalid_argument("set_gep_params: variant must be 0, 1 or 2");
    sst_->gep_variant = variant;
    sst_->gep_nu = model_nu > 0.0 ? model_nu : 0.001;
}

void Solver::set_pope_constants(double C1, double C2) {
    if (!sst_) throw std::invalid_argument("set_pope_constants: attach CFDNN_RANS_EARSM_POPE first");
    sst_->pope_C1 = C1; sst_->pope_C2 = C2;
}

void Solver::update_turbulence() {
    if (sst_) {
        // transport first, with the previous step's nu_t, then the closure (src/solver.cpp:1255-1306)
        sst_->advance(g_, vel_, k_, omega_, nut_, dt_, g_.Ng, st_);
        sst_->closure(g_, vel_, k_, omega_, nut_, tau_[0], tau_[1], tau_[2], g_.Ng, st_);
        return;
    }
    if (les_model_) {                                // LESModel::update: nu_t = nu_sgs (src/turbulence_les.cpp:292-335)
        les_update_device(g_, les_mesh_, *les_met_, les_model_, les_c_, vel_, nut_, g_.Ng, is2d_ ? 1 : g_.Nz, is2d_, st_);
        return;
    }
    if (!nn_) return;
    nn_->update_device(g_, yc_.p, mesh_.y_lo, mesh_.y_hi, vel_, k_, omega_, nut_, tau_[0], tau_[1], tau_[2], st_,
                       /*data_plane=*/g_.Ng);
}

void Solver::check_nan(int step) {                              // check_for_nan_inf :3195-3350
    if (!cfg_.turb_guard_enabled) return;
    if (cfg_.turb_guard_interval <= 0 || step % cfg_.turb_guard_interval != 0) return;
    int* flag = reinterpret_cast<int*>(scalars_.p + 16);
    CFDNN_CUDA(cudaMemsetAsync(flag, 0, sizeof(int), st_));
    const int blocks = 148 * 8;
    CFDNN_LAUNCH(k_nonfinite, blocks, 256, 0, st_, vel_.u, su_, flag);
    CFDNN_LAUNCH(k_nonfinite, blocks, 256, 0, st_, vel_.v, sv_, flag);
    CFDNN_LAUNCH(k_nonfinite, blocks, 256, 0, st_, vel_.w, sw_, flag);
    CFDNN_LAUNCH(k_nonfinite, blocks, 256, 0, st_, p_, sc_, flag);
    if (has_model()) CFDNN_LAUNCH(k_nonfinite, blocks, 256, 0, st_, nut_, sc_, flag);
    CFDNN_CUDA(cudaMemcpyAsync(h_scalars_ + 16, flag, sizeof(int), cudaMemcpyDeviceToHost, st_));
    CFDNN_CUDA(cudaStreamSynchronize(st_));
    if (*reinterpret_cast<int*>(h_scalars_ + 16) != 0)
        throw Error(CFDNN_ERR_RUNTIME, "NaN/Inf detected in solution fields");
}

// step() (src/solver.cpp:1200-2786)
void Solver::step_device(bool want_residual) {
    if (cfg_.time_integrator == CFDNN_TI_EULER) {
        blend(vel_, old_, old_, 1.0, 0.0);                      // velocity_old_ <- velocity_ :1210-1237
    } else {                                                    // ... and u^n -> velocity_rk_ (solver_time.cpp:1313-1330)
        const dim3 gr = grid3(g_.Nx + 1, g_.Ny + 1, g_.Nz + 1);
        CFDNN_LAUNCH(k_copy2, gr, kB, 0, st_, g_, vel_, old_, rk_);
    }
    PhaseEv* pe = ph_begin(CFDNN_PHASE_CLOSURE);
    update_turbulence();                                        // :1284-1324
    ph_end(pe);
    build_nu_eff();                                             // :1326-1410
    if (cfg_.adaptive_dt && has_model()) {                      // diffusion-CFL re-check :1416-1491
        const int nb = std::min(148 * 8, g_.Ny * g_.Nz);
        CFDNN_LAUNCH_COOP(k_cfl_maxima, nb, 256, 0, st_, g_, vel_, nut_, cfg_.nu, 1, partial_.p, nb);
        CFDNN_LAUNCH_COOP(k_final_max_rows, 1, 1024, 0, st_, partial_.p, nb, 5, scalars_.p + 4);
        if (comm_ && comm_->nranks > 1) comm_->allreduce_max(scalars_.p + 4, 5, st_);
        CFDNN_CUDA(cudaMemcpyAsync(h_scalars_ + 4, scalars_.p + 4, 5 * sizeof(double), cudaMemcpyDeviceToHost, st_));
        CFDNN_CUDA(cudaStreamSynchronize(st_));
        const double nu_eff_max = std::max(cfg_.nu, h_scalars_[8]);
        // :1475-1486 (the 2-D branch takes min(dx, dy); implicit y-diffusion without a transport model drops dy)
        double dx_min = is2d_ ? std::min(mesh_.dx, dy_min_) : std::min({mesh_.dx, dy_min_, mesh_.dz});
        if (cfg_.implicit_y_diffusion && !transport_model()) dx_min = is2d_ ? mesh_.dx : std::min(mesh_.dx, mesh_.dz);
        const double dt_diff_limit = 0.25 * dx_min * dx_min / nu_eff_max;
        if (dt_ > dt_diff_limit) dt_ = std::max(cfg_.dt_safety * dt_diff_limit, 1e-15);
    }
    if (cfg_.time_integrator == CFDNN_TI_RK3) ssprk3_step(dt_);
    else if (cfg_.time_integrator == CFDNN_TI_RK2) ssprk2_step(dt_);
    else euler_step();
    if (want_residual) {                                        // :2659-2775
        const dim3 gr = grid3(g_.Nx + 1, g_.Ny + 1, g_.Nz + 1);
        const int nb = gr.x * gr.y * gr.z;
        CFDNN_LAUNCH_COOP(k_max_change, gr, kB, 0, st_, g_, vel_, old_, partial_.p);
        CFDNN_LAUNCH_COOP(k_final_max, 1, 1024, 0, st_, partial_.p, nb, scalars_.p + 3);
        if (comm_ && comm_->nranks > 1) comm_->allreduce_max(scalars_.p + 3, 1, st_);
        CFDNN_CUDA(cudaMemcpyAsync(h_scalars_ + 3, scalars_.p + 3, sizeof(double), cudaMemcpyDeviceToHost, st_));
    }
    check_nan(step_count_);
    ++step_count_;
    time_ += dt_;
}

double Solver::step() {
    double r = 0.0;
    run(1, &r);
    return r;
}

void Solver::run(int nsteps, double* last_residual) {
    if (nsteps <= 0) return;
    mg_->acc_vcycle_ms = 0.0; mg_->acc_vcycles = 0;
    phused_ = 0;
    CFDNN_CUDA(cudaEventRecord(ev0_, st_));
    for (int s = 0; s < nsteps; ++s) step_device(s == nsteps - 1);
    CFDNN_CUDA(cudaEventRecord(ev1_, st_));
    CFDNN_CUDA(cudaStreamSynchronize(st_));
    check_launch("step");
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev0_, ev1_);
    last_step_ms = ms;
    mg_->collect_timing();
    last_vcycle_ms = mg_->acc_vcycle_ms;
    last_vcycles = mg_->acc_vcycles;
    for (int q = 0; q < CFDNN_PHASE_COUNT; ++q) last_phase_ms[q] = 0.0;
    for (size_t e = 0; e < phused_; ++e) {
        float pm = 0.f;
        if (cudaEventElapsedTime(&pm, phpool_[e].a, phpool_[e].b) == cudaSuccess) last_phase_ms[phpool_[e].phase] += pm;
    }
    if (last_residual) *last_residual = h_scalars_[3];
}

double Solver::compute_adaptive_dt() {                          // src/solver.cpp:3369-3729
    const int nb = std::min(148 * 8, g_.Ny * g_.Nz);
    CFDNN_LAUNCH_COOP(k_cfl_maxima, nb, 256, 0, st_, g_, vel_, nut_, cfg_.nu, has_model() ? 1 : 0, partial_.p, nb);
    CFDNN_LAUNCH_COOP(k_final_max_rows, 1, 1024, 0, st_, partial_.p, nb, 5, scalars_.p + 4);
    CFDNN_CUDA(cudaMemcpyAsync(h_scalars_ + 4, scalars_.p + 4, 5 * sizeof(double), cudaMemcpyDeviceToHost, st_));
    CFDNN_CUDA(cudaStreamSynchronize(st_));
    check_launch("compute_adaptive_dt");
    const double u_abs_max = std::max(1e-10, h_scalars_[4]);
    const double w_abs_max = std::max(1e-10, h_scalars_[6]);
    const double v_dy_ratio_max = std::max(1e-10, h_scalars_[7]);
    const double nu_eff_max = std::max(cfg_.nu, h_scalars_[8]);
    const double cfl_xz = cfg_.CFL_xz > 0 ? cfg_.CFL_xz : cfg_.CFL_max;
    const double dt_cfl_x = cfl_xz * mesh_.dx / u_abs_max;
    const double dt_cfl_y = cfg_.CFL_max / v_dy_ratio_max;
    const double dt_cfl_z = is2d_ ? 1e10 : cfl_xz * mesh_.dz / w_abs_max;          // :3688
    const double dx_min_full = is2d_ ? std::min(mesh_.dx, dy_min_) : std::min({mesh_.dx, dy_min_, mesh_.dz});
    const double dx_min = cfg_.implicit_y_diffusion ? (is2d_ ? mesh_.dx : std::min(mesh_.dx, mesh_.dz)) : dx_min_full;   // :3697-3706
    double dt_diff = 0.25 * dx_min * dx_min / nu_eff_max;
    if (cfg_.implicit_y_diffusion && transport_model())                                    // :3709-3712
        dt_diff = std::min(dt_diff, 0.25 * dx_min_full * dx_min_full / nu_eff_max);
    double dt_new = cfg_.dt_safety * std::min({dt_cfl_x, dt_cfl_y, dt_cfl_z, dt_diff});
    if (comm_ && comm_->nranks > 1) {
        CFDNN_CUDA(cudaMemcpyAsync(scalars_.p + 10, &dt_new, sizeof(double), cudaMemcpyHostToDevice, st_));
        comm_->allreduce_min(scalars_.p + 10, 1, st_);
        CFDNN_CUDA(cudaMemcpyAsync(&dt_new, scalars_.p + 10, sizeof(double), cudaMemcpyDeviceToHost, st_));
        CFDNN_CUDA(cudaStreamSynchronize(st_));
    }
    return dt_new;
}

}  // namespace cfdnn
