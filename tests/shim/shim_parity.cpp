// tests/shim/shim_parity.cpp — TEST: the reference's RANSSolver / MultigridPoissonSolver and the
// B200 shim classes (include/cfdnn_b200_shim.hpp) side by side in one program, same Mesh / Config /
// VelocityBC / initial VectorField.  Prints one JSON line; exit 0 when every gate holds.
// Built against /root/reference/include + oracle/_ref/libnn_cfd_core.a (tests/shim/build.sh).
#include "cfdnn_b200_shim.hpp"
#include "poisson_solver_multigrid.hpp"
#include "turbulence_nn_tbnn.hpp"
#include "turbulence_les.hpp"
#include "turbulence_transport.hpp"
#include "turbulence_baseline.hpp"
#include "turbulence_earsm.hpp"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <sstream>
using namespace nncfd;

static double rel_l2(const std::vector<double>& a, const std::vector<double>& b) {
    double num = 0, den = 0;
    for (size_t i = 0; i < a.size(); ++i) { const double d = a[i] - b[i]; num += d * d; den += b[i] * b[i]; }
    return den > 0 ? std::sqrt(num / den) : std::sqrt(num);
}

// the data plane of a 2-D field only (host plane 0; the reference's fill() also writes the two unused planes)
static double rel_l2_plane0(const std::vector<double>& a, const std::vector<double>& b, size_t n) {
    return rel_l2(std::vector<double>(a.begin(), a.begin() + n), std::vector<double>(b.begin(), b.begin() + n));
}

// ---- closures through the TurbulenceModel plugin interface (include/turbulence_model.hpp:81-165) on a 2-D
// channel (the reference's closures are 2-D models): (a) reference solver + reference TBNN = the oracle;
// (b) B200 solver + TurbulenceNNB200 handed over as unique_ptr<TurbulenceModel> (device-resident closure);
// (c) REFERENCE solver + TurbulenceNNB200 (the plugin used from the reference's own step(), host update());
// (d) B200 solver + the reference's own TBNN model (bridged host model); (e) update() through get_device_view().
static int closure_section(const std::string& wdir, std::string& json) {
    Mesh mesh;
    mesh.init_uniform(32, 48, 0.0, 4 * M_PI, -1.0, 1.0, 1);
    Config cfg;
    cfg.nu = 1.0 / 180.0; cfg.dt = 0.002; cfg.adaptive_dt = false; cfg.verbose = false;
    cfg.time_integrator = TimeIntegrator::Euler; cfg.convective_scheme = ConvectiveScheme::Skew;
    cfg.poisson_solver = PoissonSolverType::MG; cfg.poisson_fixed_cycles = 8;
    cfg.turb_model = TurbulenceModelType::NNTBNN;
    VelocityBC bc;
    bc.x_lo = bc.x_hi = VelocityBC::Periodic; bc.y_lo = bc.y_hi = VelocityBC::NoSlip;
    bc.z_lo = bc.z_hi = VelocityBC::Periodic;
    VectorField v0(mesh);
    for (int j = mesh.j_begin(); j < mesh.j_end(); ++j)
        for (int i = mesh.i_begin(); i <= mesh.i_end(); ++i) {
            const double y = mesh.y(j);
            v0.u(i, j) = (1.0 - y * y) + 0.05 * std::cos(0.5 * M_PI * y) * std::sin(2.0 * mesh.xf[i]);
        }
    for (int j = mesh.j_begin(); j <= mesh.j_end(); ++j)
        for (int i = mesh.i_begin(); i < mesh.i_end(); ++i)
            v0.v(i, j) = 0.05 * std::sin(M_PI * (mesh.yf[j] + 1.0) * 0.5) * std::cos(2.0 * mesh.x(i));
    const int nsteps = 3;
    auto make_b200_model = [&] {
        auto m = std::make_unique<b200::TurbulenceNNB200>(CFDNN_TURB_NN_TBNN);
        m->load(wdir, wdir);
        return m;
    };
    // (a)
    RANSSolver ref(mesh, cfg);
    ref.set_velocity_bc(bc); ref.set_body_force(1.0, 0.0);
    { auto m = create_turbulence_model(TurbulenceModelType::NNTBNN, wdir, wdir); m->set_nu(cfg.nu); ref.set_turbulence_model(std::move(m)); }
    ref.initialize(v0);
    // (b)
    b200::RANSSolverB200 b(mesh, cfg);
    b.set_velocity_bc(bc); b.set_body_force(1.0, 0.0);
    { std::unique_ptr<TurbulenceModel> m = make_b200_model(); b.set_turbulence_model(std::move(m)); }
    b.initialize(v0);
    // (c)
    RANSSolver refp(mesh, cfg);
    refp.set_velocity_bc(bc); refp.set_body_force(1.0, 0.0);
    { std::unique_ptr<TurbulenceModel> m = make_b200_model(); m->set_nu(cfg.nu); refp.set_turbulence_model(std::move(m)); }
    refp.initialize(v0);
    // (d)
    b200::RANSSolverB200 bh(mesh, cfg);
    bh.set_velocity_bc(bc); bh.set_body_force(1.0, 0.0);
    { auto m = create_turbulence_model(TurbulenceModelType::NNTBNN, wdir, wdir); bh.set_turbulence_model(std::move(m)); }
    bh.initialize(v0);
    double r_ref = 0, r_b = 0, r_p = 0, r_h = 0;
    for (int s = 0; s < nsteps; ++s) { r_ref = ref.step(); r_b = b.step(); r_p = refp.step(); r_h = bh.step(); }
    b.sync_from_gpu(); bh.sync_from_gpu();
    const double e_b_u = rel_l2(b.velocity().u_data(), ref.velocity().u_data());
    const double e_b_nut = rel_l2(b.nu_t().data(), ref.nu_t().data());
    const double e_p_u = rel_l2(refp.velocity().u_data(), ref.velocity().u_data());
    const double e_p_nut = rel_l2(refp.nu_t().data(), ref.nu_t().data());
    const double e_h_u = rel_l2(bh.velocity().u_data(), ref.velocity().u_data());
    const double e_h_nut = rel_l2(bh.nu_t().data(), ref.nu_t().data());
    // (e) device view: the closure updates nu_t of solver b in place on the device; must equal a host update()
    double e_view = 1.0;
    bool name_ok = false, views_ok = false;
    {
        auto m = make_b200_model();
        m->set_nu(cfg.nu);
        name_ok = m->name() == "NNTBNN" && m->provides_reynolds_stresses() && !m->uses_transport_equations() &&
                  m->is_gpu_ready();
        ScalarField k(mesh), om(mesh), nut_host(mesh), dummy(mesh);
        m->update(mesh, b.velocity(), k, om, nut_host, nullptr, nullptr);            // host path on the synced velocity
        const TurbulenceDeviceView view = b.get_device_view();
        const SolverDeviceView sv = b.get_solver_view();
        views_ok = view.u_face && view.v_face && view.nu_t && view.Nx == 32 && view.Ny == 48 && sv.is_valid() &&
                   sv.u_stride == b.velocity().u_stride() && sv.cell_stride == mesh.total_Nx();
        m->update(mesh, b.velocity(), k, om, dummy, nullptr, &view);                 // device path, in place
        b.sync_solution_from_gpu();
        e_view = rel_l2(b.nu_t().data(), nut_host.data());
    }
    const bool ok = e_b_u <= 1e-9 && e_b_nut <= 1e-6 && e_p_u <= 1e-9 && e_p_nut <= 1e-6 && e_h_u <= 1e-9 &&
                    e_h_nut <= 1e-6 && e_view <= 1e-12 && name_ok && views_ok &&
                    std::fabs(r_ref - r_b) <= 1e-7 * std::fabs(r_ref) && std::fabs(r_ref - r_p) <= 1e-7 * std::fabs(r_ref) &&
                    std::fabs(r_ref - r_h) <= 1e-7 * std::fabs(r_ref);
    char buf[1024];
    std::snprintf(buf, sizeof buf,
                  "\"closure\": {\"b200_solver_b200_model\": {\"u\": %.3e, \"nu_t\": %.3e}, "
                  "\"ref_solver_b200_model\": {\"u\": %.3e, \"nu_t\": %.3e}, "
                  "\"b200_solver_ref_model_bridged\": {\"u\": %.3e, \"nu_t\": %.3e}, \"device_view_vs_host_update\": %.3e, "
                  "\"interface_ok\": %s, \"views_ok\": %s, \"ok\": %s}",
                  e_b_u, e_b_nut, e_p_u, e_p_nut, e_h_u, e_h_nut, e_view, name_ok ? "true" : "false",
                  views_ok ? "true" : "false", ok ? "true" : "false");
    json = buf;
    return ok ? 0 : 1;
}

int main(int argc, char** argv) {
    const int N = argc > 1 ? std::atoi(argv[1]) : 16;
    const int nsteps = argc > 2 ? std::atoi(argv[2]) : 2;
    const std::string wdir = argc > 3 ? argv[3] : "";
    std::ostringstream sink;                       // the reference is chatty on stdout/stderr
    std::streambuf* oc = std::cout.rdbuf(sink.rdbuf());
    std::streambuf* oe = std::cerr.rdbuf(sink.rdbuf());
    Mesh mesh;
    mesh.init_uniform(N, N, N, 0, 2 * M_PI, 0, 2 * M_PI, 0, 2 * M_PI, 1);
    Config cfg;
    cfg.nu = 0.01; cfg.dt = 0.01; cfg.adaptive_dt = false; cfg.verbose = false;
    cfg.time_integrator = TimeIntegrator::RK3; cfg.convective_scheme = ConvectiveScheme::Skew;
    cfg.poisson_solver = PoissonSolverType::MG; cfg.poisson_fixed_cycles = 0;
    cfg.poisson_tol_rhs = 1e-8; cfg.poisson_max_vcycles = 50;
    VelocityBC bc;
    bc.x_lo = bc.x_hi = bc.y_lo = bc.y_hi = bc.z_lo = bc.z_hi = VelocityBC::Periodic;
    VectorField v0(mesh);
    for (int k = mesh.k_begin(); k < mesh.k_end(); ++k)
        for (int j = mesh.j_begin(); j < mesh.j_end(); ++j)
            for (int i = mesh.i_begin(); i <= mesh.i_end(); ++i)
                v0.u(i, j, k) = std::sin(mesh.xf[i]) * std::cos(mesh.y(j)) * std::cos(mesh.z(k));
    for (int k = mesh.k_begin(); k < mesh.k_end(); ++k)
        for (int j = mesh.j_begin(); j <= mesh.j_end(); ++j)
            for (int i = mesh.i_begin(); i < mesh.i_end(); ++i)
                v0.v(i, j, k) = -std::cos(mesh.x(i)) * std::sin(mesh.yf[j]) * std::cos(mesh.z(k));

    int rc = 0;
    std::string closure_json = "\"closure\": null";
    double r_ref = 0, r_b = 0, eu = 0, ep = 0, emg = 0, les_u = 1, les_nut = 1, sst_u = 1, sst_k = 1, sst_nut = 1;
    int cyc_ref = 0, cyc_b = 0, mg_ref = 0, mg_b = 0;
    bool threw_invalid = false;
    try {
        RANSSolver ref(mesh, cfg);
        ref.set_velocity_bc(bc);
        ref.initialize(v0);
        b200::RANSSolverB200 b(mesh, cfg);
        b.set_velocity_bc(bc);
        b.initialize(v0);
        for (int s = 0; s < nsteps; ++s) { r_ref = ref.step(); r_b = b.step(); }
        b.sync_from_gpu();
        eu = rel_l2(b.velocity().u_data(), ref.velocity().u_data());
        ep = rel_l2(b.pressure().data(), ref.pressure().data());
        cyc_ref = ref.poisson_stats().cycles; cyc_b = b.poisson_stats().cycles;

        // MultigridPoissonSolver drop-in: same rhs, same config
        ScalarField rhs(mesh), p1(mesh), p2(mesh);
        double mean = 0;
        for (int k = mesh.k_begin(); k < mesh.k_end(); ++k)
            for (int j = mesh.j_begin(); j < mesh.j_end(); ++j)
                for (int i = mesh.i_begin(); i < mesh.i_end(); ++i) {
                    rhs(i, j, k) = std::sin(2 * mesh.x(i)) * std::cos(mesh.y(j)) + 0.3 * std::sin(mesh.z(k));
                    mean += rhs(i, j, k);
                }
        (void)mean;
        PoissonConfig pc;
        pc.max_vcycles = 30; pc.tol_rhs = 1e-9; pc.tol_rel = 0; pc.tol = 0;
        MultigridPoissonSolver mref(mesh);
        mref.set_bc(PoissonBC::Periodic, PoissonBC::Periodic, PoissonBC::Neumann, PoissonBC::Neumann,
                    PoissonBC::Periodic, PoissonBC::Periodic);
        b200::MultigridPoissonSolverB200 mb(mesh);
        mb.set_bc(PoissonBC::Periodic, PoissonBC::Periodic, PoissonBC::Neumann, PoissonBC::Neumann,
                  PoissonBC::Periodic, PoissonBC::Periodic);
        mg_ref = mref.solve(rhs, p1, pc);
        mg_b = mb.solve(rhs, p2, pc);
        emg = rel_l2(p2.data(), p1.data());
        if (mref.converged() != mb.converged()) rc = 1;

        // error mapping: asymmetric periodic pair -> std::invalid_argument (src/solver.cpp:653-661)
        VelocityBC bad = bc;
        bad.x_hi = VelocityBC::NoSlip;
        try { b.set_velocity_bc(bad); } catch (const std::invalid_argument&) { threw_invalid = true; }
        // initialize_uniform (include/solver.hpp:180) and the trivial single-process decomposition (:93)
        {
            b200::RANSSolverB200 bu(mesh, cfg);
            bu.set_velocity_bc(bc);
            Decomposition single(mesh.Nz);
            bu.set_decomposition(&single);
            bu.initialize_uniform(0.3, -0.2);
            RANSSolver ru(mesh, cfg);
            ru.set_velocity_bc(bc);
            ru.initialize_uniform(0.3, -0.2);
            bu.step(); ru.step();
            bu.sync_from_gpu();
            if (rel_l2(bu.velocity().u_data(), ru.velocity().u_data()) > 1e-12 || bu.decomposition() != &single) rc = 1;
        }
        // the reference's own LES model object handed to both solvers: device kernel under the B200 solver
        {
            Config lc = cfg;
            lc.poisson_fixed_cycles = 4;
            RANSSolver rl(mesh, lc);
            rl.set_velocity_bc(bc);
            rl.set_turbulence_model(std::make_unique<WALEModel>());
            rl.initialize(v0);
            b200::RANSSolverB200 bl(mesh, lc);
            bl.set_velocity_bc(bc);
            bl.set_turbulence_model(std::make_unique<WALEModel>());
            bl.initialize(v0);
            rl.step(); bl.step(); rl.step(); bl.step();
            bl.sync_from_gpu();
            les_u = rel_l2(bl.velocity().u_data(), rl.velocity().u_data());
            les_nut = rel_l2(bl.nu_t().data(), rl.nu_t().data());
            if (!(les_u <= 1e-10 && les_nut <= 1e-6)) rc = 1;
            // DynamicSmagorinskyModel: the reference's CPU path evaluates it as static Smagorinsky (Cs = 0.17); same kernel here
            RANSSolver rd(mesh, lc);
            rd.set_velocity_bc(bc);
            rd.set_turbulence_model(std::make_unique<DynamicSmagorinskyModel>());
            rd.initialize(v0);
            b200::RANSSolverB200 bd(mesh, lc);
            bd.set_velocity_bc(bc);
            bd.set_turbulence_model(std::make_unique<DynamicSmagorinskyModel>());
            bd.initialize(v0);
            rd.step(); bd.step();
            bd.sync_from_gpu();
            if (!(rel_l2(bd.velocity().u_data(), rd.velocity().u_data()) <= 1e-10 &&
                  rel_l2(bd.nu_t().data(), rd.nu_t().data()) <= 1e-6)) rc = 1;
        }
        // the reference's SST k-omega transport model on a 2-D channel: native wavefront sweep under the B200 solver
        {
            Mesh m2;
            m2.init_uniform(32, 24, 0.0, 4 * M_PI, -1.0, 1.0);
            Config sc = cfg;
            sc.Nx = 32; sc.Ny = 24; sc.Nz = 1;
            sc.nu = 1.0 / 180.0; sc.dt = 0.002;
            sc.poisson_fixed_cycles = 6;
            VelocityBC b2;
            b2.x_lo = b2.x_hi = VelocityBC::Periodic;
            b2.y_lo = b2.y_hi = VelocityBC::NoSlip;
            VectorField w0(m2);
            for (int j = m2.j_begin(); j < m2.j_end(); ++j)
                for (int i = m2.i_begin(); i <= m2.i_end(); ++i)
                    w0.u(i, j) = (1.0 - m2.y(j) * m2.y(j)) * (1.0 + 0.05 * std::sin(m2.xf[i]));
            RANSSolver rs(m2, sc);
            rs.set_velocity_bc(b2);
            rs.set_body_force(1.0, 0.0);
            rs.set_turbulence_model(std::make_unique<SSTKOmegaTransport>());
            rs.initialize(w0);
            b200::RANSSolverB200 bs(m2, sc);
            bs.set_velocity_bc(b2);
            bs.set_body_force(1.0, 0.0);
            bs.set_turbulence_model(std::make_unique<SSTKOmegaTransport>());
            bs.initialize(w0);
            for (int s = 0; s < 3; ++s) { rs.step(); bs.step(); }
            bs.sync_from_gpu();
            sst_u = rel_l2(bs.velocity().u_data(), rs.velocity().u_data());
            const size_t pl = (size_t)m2.total_Nx() * m2.total_Ny();
            sst_k = std::max(rel_l2_plane0(bs.k().data(), rs.k().data(), pl), rel_l2_plane0(bs.omega().data(), rs.omega().data(), pl));
            sst_nut = rel_l2(bs.nu_t().data(), rs.nu_t().data());
            if (!(sst_u <= 1e-10 && sst_k <= 1e-10 && sst_nut <= 1e-9)) rc = 1;
            // the mixing-length baseline (turb_model = baseline), same route
            {
                RANSSolver rm(m2, sc);
                rm.set_velocity_bc(b2);
                rm.set_body_force(1.0, 0.0);
                rm.set_turbulence_model(std::make_unique<MixingLengthModel>());
                rm.initialize(w0);
                b200::RANSSolverB200 bm(m2, sc);
                bm.set_velocity_bc(b2);
                bm.set_body_force(1.0, 0.0);
                bm.set_turbulence_model(std::make_unique<MixingLengthModel>());
                bm.initialize(w0);
                for (int s = 0; s < 3; ++s) { rm.step(); bm.step(); }
                bm.sync_from_gpu();
                if (!(rel_l2(bm.velocity().u_data(), rm.velocity().u_data()) <= 1e-10 &&
                      rel_l2(bm.nu_t().data(), rm.nu_t().data()) <= 1e-9)) rc = 1;
            }
            // SST + Wallin-Johansson EARSM, same route (the reference's SSTWithEARSM object)
            {
                RANSSolver re(m2, sc);
                re.set_velocity_bc(b2);
                re.set_body_force(1.0, 0.0);
                re.set_turbulence_model(std::make_unique<SSTWithEARSM>(EARSMType::WallinJohansson2000));
                re.initialize(w0);
                b200::RANSSolverB200 be(m2, sc);
                be.set_velocity_bc(b2);
                be.set_body_force(1.0, 0.0);
                be.set_turbulence_model(std::make_unique<SSTWithEARSM>(EARSMType::WallinJohansson2000));
                be.initialize(w0);
                for (int s = 0; s < 2; ++s) { re.step(); be.step(); }
                be.sync_from_gpu();
                if (!(rel_l2(be.velocity().u_data(), re.velocity().u_data()) <= 1e-10 &&
                      rel_l2(be.nu_t().data(), re.nu_t().data()) <= 1e-9)) rc = 1;
            }
            // the standard k-omega model, same route
            RANSSolver rk(m2, sc);
            rk.set_velocity_bc(b2);
            rk.set_body_force(1.0, 0.0);
            rk.set_turbulence_model(std::make_unique<KOmegaTransport>());
            rk.initialize(w0);
            b200::RANSSolverB200 bk(m2, sc);
            bk.set_velocity_bc(b2);
            bk.set_body_force(1.0, 0.0);
            bk.set_turbulence_model(std::make_unique<KOmegaTransport>());
            bk.initialize(w0);
            for (int s = 0; s < 2; ++s) { rk.step(); bk.step(); }
            bk.sync_from_gpu();
            if (!(rel_l2(bk.velocity().u_data(), rk.velocity().u_data()) <= 1e-10 &&
                  rel_l2_plane0(bk.k().data(), rk.k().data(), pl) <= 1e-10 &&
                  rel_l2(bk.nu_t().data(), rk.nu_t().data()) <= 1e-9)) rc = 1;
        }
        if (!wdir.empty() && closure_section(wdir, closure_json) != 0) rc = 1;
    } catch (const std::exception& e) {
        std::cout.rdbuf(oc); std::cerr.rdbuf(oe);
        std::printf("{\"error\": \"%s\"}\n", e.what());
        return 2;
    }
    std::cout.rdbuf(oc); std::cerr.rdbuf(oe);
    if (!(eu <= 1e-10 && ep <= 1e-10 && emg <= 1e-10 && cyc_ref == cyc_b && mg_ref == mg_b && threw_invalid &&
          std::fabs(r_ref - r_b) <= 1e-9 * std::fabs(r_ref)))
        rc = 1;
    std::printf("{\"N\": %d, \"steps\": %d, \"residual_ref\": %.17g, \"residual_b200\": %.17g, \"u_rel_l2\": %.3e, "
                "\"p_rel_l2\": %.3e, \"cycles_ref\": %d, \"cycles_b200\": %d, \"mg_cycles_ref\": %d, "
                "\"mg_cycles_b200\": %d, \"mg_p_rel_l2\": %.3e, \"invalid_argument\": %s, \"les_wale\": {\"u\": %.3e, \"nu_t\": %.3e}, \"sst\": {\"u\": %.3e, \"k\": %.3e, \"nu_t\": %.3e}, %s, \"ok\": %s}\n",
                N, nsteps, r_ref, r_b, eu, ep, cyc_ref, cyc_b, mg_ref, mg_b, emg, threw_invalid ? "true" : "false",
                les_u, les_nut, sst_u, sst_k, sst_nut, closure_json.c_str(), rc == 0 ? "true" : "false");
    return rc;
}
