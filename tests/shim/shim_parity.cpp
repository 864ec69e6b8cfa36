// tests/shim/shim_parity.cpp — TEST: the reference's RANSSolver / MultigridPoissonSolver and the
// B200 shim classes (include/cfdnn_b200_shim.hpp) side by side in one program, same Mesh / Config /
// VelocityBC / initial VectorField.  Prints one JSON line; exit 0 when every gate holds.
// Built against /root/reference/include + oracle/_ref/libnn_cfd_core.a (tests/shim/build.sh).
#include "cfdnn_b200_shim.hpp"
#include "poisson_solver_multigrid.hpp"
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

int main(int argc, char** argv) {
    const int N = argc > 1 ? std::atoi(argv[1]) : 16;
    const int nsteps = argc > 2 ? std::atoi(argv[2]) : 2;
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
    double r_ref = 0, r_b = 0, eu = 0, ep = 0, emg = 0;
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
                "\"mg_cycles_b200\": %d, \"mg_p_rel_l2\": %.3e, \"invalid_argument\": %s, \"ok\": %s}\n",
                N, nsteps, r_ref, r_b, eu, ep, cyc_ref, cyc_b, mg_ref, mg_b, emg, threw_invalid ? "true" : "false",
                rc == 0 ? "true" : "false");
    return rc;
}
