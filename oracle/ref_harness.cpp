// oracle/ref_harness.cpp — TEST INFRASTRUCTURE (not product code).
//
// A driver written for this repo that links against the UNMODIFIED reference
// (comp-physics/cfd-nn CPU build, compiled by oracle/Makefile into oracle/_ref/)
// and exercises exactly the public entry points the B200 backend replaces:
//
//   mode=step   RANSSolver::step()                       (reference include/solver.hpp:208)
//   mode=mg     MultigridPoissonSolver::solve()          (reference include/poisson_solver_multigrid.hpp:46)
//   mode=nn     TurbulenceModel::update() for nn_tbnn / nn_mlp (reference include/turbulence_model.hpp:96)
//   mode=bench  step() timed the way tests/bench_mg.cpp:66-78 does (3 warm-up steps)
//
// It only reads/writes raw little-endian float64 arrays in the reference's own
// ghost-inclusive layout (include/fields.hpp:118-136), so the Python tests can
// feed identical inputs to the CUDA path and compare outputs.
//
// usage: ref_harness key=value ...   (see parse below)

#include "turbulence_gep.hpp"
#include "turbulence_baseline.hpp"
#include "solver.hpp"
#include "config.hpp"
#include "mesh.hpp"
#include "fields.hpp"
#include "poisson_solver_multigrid.hpp"
#include "turbulence_model.hpp"
#include "turbulence_nn_tbnn.hpp"
#include "turbulence_nn_mlp.hpp"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <map>
#include <sstream>
#include <string>
#include <vector>

using namespace nncfd;

namespace {

std::map<std::string, std::string> g_args;

std::string arg_s(const std::string& k, const std::string& def) {
    auto it = g_args.find(k);
    return it == g_args.end() ? def : it->second;
}
int arg_i(const std::string& k, int def) {
    auto it = g_args.find(k);
    return it == g_args.end() ? def : std::atoi(it->second.c_str());
}
double arg_d(const std::string& k, double def) {
    auto it = g_args.find(k);
    return it == g_args.end() ? def : std::strtod(it->second.c_str(), nullptr);
}

void write_f64(const std::string& path, const std::vector<double>& v) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) { std::perror(path.c_str()); std::exit(2); }
    std::fwrite(v.data(), sizeof(double), v.size(), f);
    std::fclose(f);
}
bool read_f64(const std::string& path, std::vector<double>& v) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    size_t n = std::fread(v.data(), sizeof(double), v.size(), f);
    std::fclose(f);
    if (n != v.size()) {
        std::fprintf(stderr, "size mismatch reading %s: got %zu want %zu\n", path.c_str(), n, v.size());
        std::exit(2);
    }
    return true;
}

Mesh make_mesh() {
    const int Nx = arg_i("Nx", 16), Ny = arg_i("Ny", 16), Nz = arg_i("Nz", 16);
    const int ng = arg_i("ng", 1);
    const double two_pi = 2.0 * M_PI;
    const double x0 = arg_d("x0", 0.0), x1 = arg_d("x1", two_pi);
    const double y0 = arg_d("y0", 0.0), y1 = arg_d("y1", two_pi);
    const double z0 = arg_d("z0", 0.0), z1 = arg_d("z1", two_pi);
    const double beta = arg_d("stretch_beta", 0.0);
    Mesh m;
    if (beta > 0.0) {
        if (Nz > 1) m.init_stretched_y(Nx, Ny, Nz, x0, x1, y0, y1, z0, z1, Mesh::tanh_stretching(beta), ng);
        else        m.init_stretched_y(Nx, Ny, x0, x1, y0, y1, Mesh::tanh_stretching(beta), ng);
    } else {
        if (Nz > 1) m.init_uniform(Nx, Ny, Nz, x0, x1, y0, y1, z0, z1, ng);
        else        m.init_uniform(Nx, Ny, x0, x1, y0, y1, ng);
    }
    return m;
}

VelocityBC::Type bc_from_char(char c) {
    switch (c) {
        case 'P': return VelocityBC::Periodic;
        case 'W': return VelocityBC::NoSlip;
        case 'I': return VelocityBC::Inflow;
        case 'O': return VelocityBC::Outflow;
    }
    std::fprintf(stderr, "bad velocity bc char %c\n", c);
    std::exit(2);
}
PoissonBC pbc_from_char(char c) {
    switch (c) {
        case 'P': return PoissonBC::Periodic;
        case 'N': return PoissonBC::Neumann;
        case 'D': return PoissonBC::Dirichlet;
    }
    std::fprintf(stderr, "bad poisson bc char %c\n", c);
    std::exit(2);
}

Config make_config(const Mesh& mesh) {
    Config c;
    c.Nx = mesh.Nx; c.Ny = mesh.Ny; c.Nz = mesh.Nz;
    c.nu = arg_d("nu", 0.01);
    c.dt = arg_d("dt", 0.01);
    c.adaptive_dt = arg_i("adaptive_dt", 0) != 0;
    c.CFL_max = arg_d("CFL_max", c.CFL_max);
    c.CFL_xz = arg_d("CFL_xz", c.CFL_xz);
    c.dt_safety = arg_d("dt_safety", c.dt_safety);
    c.verbose = false;
    c.postprocess = false;
    c.write_fields = false;
    const std::string ti = arg_s("ti", "rk3");
    c.time_integrator = ti == "rk3" ? TimeIntegrator::RK3 : ti == "rk2" ? TimeIntegrator::RK2 : TimeIntegrator::Euler;
    const std::string sch = arg_s("scheme", "skew");
    c.convective_scheme = sch == "skew" ? ConvectiveScheme::Skew
                        : sch == "upwind" ? ConvectiveScheme::Upwind : ConvectiveScheme::Central;
    c.poisson_solver = PoissonSolverType::MG;
    c.poisson_fixed_cycles = arg_i("fixed", 8);
    c.poisson_tol_rhs = arg_d("tol_rhs", c.poisson_tol_rhs);
    c.poisson_max_vcycles = arg_i("maxv", c.poisson_max_vcycles);
    c.turb_guard_enabled = arg_i("guard", 1) != 0;
    const std::string model = arg_s("model", "none");
    if (model == "nn_tbnn") c.turb_model = TurbulenceModelType::NNTBNN;
    else if (model == "nn_mlp") c.turb_model = TurbulenceModelType::NNMLP;
    else if (model == "sst") c.turb_model = TurbulenceModelType::SSTKOmega;
    else if (model == "komega") c.turb_model = TurbulenceModelType::KOmega;
    else if (model == "baseline") c.turb_model = TurbulenceModelType::Baseline;
    else if (model == "gep") c.turb_model = TurbulenceModelType::GEP;
    else if (model == "earsm_wj") c.turb_model = TurbulenceModelType::EARSM_WJ;
    else if (model == "earsm_gs") c.turb_model = TurbulenceModelType::EARSM_GS;
    else if (model == "earsm_pope") c.turb_model = TurbulenceModelType::EARSM_Pope;
    else if (model == "smagorinsky") c.turb_model = TurbulenceModelType::Smagorinsky;
    else if (model == "wale") c.turb_model = TurbulenceModelType::WALE;
    else if (model == "vreman") c.turb_model = TurbulenceModelType::Vreman;
    else if (model == "sigma") c.turb_model = TurbulenceModelType::Sigma;
    else c.turb_model = TurbulenceModelType::None;
    c.nn_weights_path = arg_s("nn_weights", "");
    c.nn_scaling_path = arg_s("nn_scaling", c.nn_weights_path);
    c.nu_t_max = arg_d("nu_t_max", c.nu_t_max);
    c.implicit_y_diffusion = arg_i("implicit_y", 0) != 0;
    c.trip_enabled = arg_i("trip", 0) != 0;                 // trip-region forcing (Euler path; needs call_initialize=1)
    c.trip_amplitude = arg_d("trip_amp", c.trip_amplitude);
    c.trip_duration = arg_d("trip_duration", c.trip_duration);
    c.trip_ramp_off_start = arg_d("trip_ramp", c.trip_ramp_off_start);
    c.trip_n_modes_z = arg_i("trip_modes", c.trip_n_modes_z);
    c.trip_force_w = arg_i("trip_force_w", c.trip_force_w ? 1 : 0) != 0;
    c.trip_w_scale = arg_d("trip_w_scale", c.trip_w_scale);
    c.trip_x_start = arg_d("trip_x0", c.trip_x_start);
    c.trip_x_end = arg_d("trip_x1", c.trip_x_end);
    return c;
}

void init_velocity(const Mesh& mesh, RANSSolver& solver) {
    const std::string init = arg_s("init", "tgv");
    VectorField& vel = solver.velocity();
    const int Ng = mesh.Nghost;
    if (init == "tgv") {
        // Taylor-Green vortex on faces, as reference app/main_taylor_green_3d.cpp:253-276
        const double V0 = arg_d("V0", 1.0);
        for (int k = mesh.k_begin(); k < mesh.k_end(); ++k)
            for (int j = mesh.j_begin(); j < mesh.j_end(); ++j)
                for (int i = mesh.i_begin(); i <= mesh.i_end(); ++i)
                    vel.u(i, j, k) = V0 * std::sin(mesh.xf[i]) * std::cos(mesh.y(j)) * std::cos(mesh.z(k));
        for (int k = mesh.k_begin(); k < mesh.k_end(); ++k)
            for (int j = mesh.j_begin(); j <= mesh.j_end(); ++j)
                for (int i = mesh.i_begin(); i < mesh.i_end(); ++i)
                    vel.v(i, j, k) = -V0 * std::cos(mesh.x(i)) * std::sin(mesh.yf[j]) * std::cos(mesh.z(k));
    } else if (init == "shear") {
        // wall-bounded start, as reference tests/bench_mg.cpp:59-64 (index-based loops incl. ghosts)
        for (int k = 0; k <= mesh.Nz; ++k)
            for (int j = 0; j <= mesh.Ny; ++j)
                for (int i = 0; i <= mesh.Nx + 1; ++i)
                    vel.u(i, j, k) = 0.01 * std::sin(j * 3.14159 / mesh.Ny);
    } else if (init.rfind("file:", 0) == 0) {
        const std::string pre = init.substr(5);
        read_f64(pre + "_u.f64", vel.u_data());
        read_f64(pre + "_v.f64", vel.v_data());
        if (!mesh.is2D()) read_f64(pre + "_w.f64", vel.w_data());
    } else {
        std::fprintf(stderr, "unknown init %s\n", init.c_str());
        std::exit(2);
    }
    (void)Ng;
}

void dump_mesh(const Mesh& mesh, const std::string& pre) {
    write_f64(pre + "_xc.f64", mesh.xc); write_f64(pre + "_xf.f64", mesh.xf);
    write_f64(pre + "_yc.f64", mesh.yc); write_f64(pre + "_yf.f64", mesh.yf);
    write_f64(pre + "_zc.f64", mesh.zc); write_f64(pre + "_zf.f64", mesh.zf);
    if (mesh.is_y_stretched()) {
        write_f64(pre + "_dyv.f64", mesh.dyv); write_f64(pre + "_dyc.f64", mesh.dyc);
        write_f64(pre + "_aS.f64", mesh.yLap_aS); write_f64(pre + "_aN.f64", mesh.yLap_aN);
        write_f64(pre + "_aP.f64", mesh.yLap_aP);
    }
}

void dump_state(const Mesh& mesh, const RANSSolver& s, const std::string& pre) {
    write_f64(pre + "_u.f64", s.velocity().u_data());
    write_f64(pre + "_v.f64", s.velocity().v_data());
    if (!mesh.is2D()) write_f64(pre + "_w.f64", s.velocity().w_data());
    write_f64(pre + "_p.f64", s.pressure().data());
    write_f64(pre + "_pcorr.f64", s.pressure_correction().data());
    write_f64(pre + "_div.f64", s.div_velocity().data());
    write_f64(pre + "_rhs.f64", s.rhs_poisson().data());
    write_f64(pre + "_nut.f64", s.nu_t().data());
    write_f64(pre + "_k.f64", s.k().data());
    write_f64(pre + "_omega.f64", s.omega().data());
}

// parse "[Poisson] solve #n vcycles=k" lines the reference prints with POISSON_STATS set
// (reference src/solver_time.cpp:926-935)
std::vector<int> parse_cycles(const std::string& log) {
    std::vector<int> out;
    std::istringstream is(log);
    std::string line;
    while (std::getline(is, line)) {
        auto p = line.find("vcycles=");
        if (line.rfind("[Poisson] solve", 0) == 0 && p != std::string::npos)
            out.push_back(std::atoi(line.c_str() + p + 8));
    }
    return out;
}

int run_step(bool bench) {
    Mesh mesh = make_mesh();
    Config cfg = make_config(mesh);
    const std::string out = arg_s("out", "");
    const int nsteps = arg_i("nsteps", 1);
    const int warmup = arg_i("warmup", bench ? 3 : 0);

    setenv("POISSON_STATS", "1", 1);
    std::ostringstream cerr_cap;
    std::streambuf* old_cerr = std::cerr.rdbuf(cerr_cap.rdbuf());
    std::ostringstream cout_cap;
    std::streambuf* old_cout = std::cout.rdbuf(cout_cap.rdbuf());

    RANSSolver solver(mesh, cfg);
    const std::string bc = arg_s("bc", "PPPPPP");
    VelocityBC vbc;
    vbc.x_lo = bc_from_char(bc[0]); vbc.x_hi = bc_from_char(bc[1]);
    vbc.y_lo = bc_from_char(bc[2]); vbc.y_hi = bc_from_char(bc[3]);
    vbc.z_lo = bc_from_char(bc[4]); vbc.z_hi = bc_from_char(bc[5]);
    solver.set_velocity_bc(vbc);
    solver.set_body_force(arg_d("fx", 0.0), arg_d("fy", 0.0), arg_d("fz", 0.0));

    if (cfg.turb_model != TurbulenceModelType::None) {
        auto model = create_turbulence_model(cfg.turb_model, cfg.nn_weights_path, cfg.nn_scaling_path,
                                             cfg.pope_C1, cfg.pope_C2);
        if (model) {
            if (auto* gp = dynamic_cast<TurbulenceGEP*>(model.get()))
                gp->set_variant(static_cast<TurbulenceGEP::Variant>(arg_i("gep_variant", 2)));
            model->set_nu(cfg.nu);
            solver.set_turbulence_model(std::move(model));
        }
    }

    if (arg_s("model", "none") == "algebraic_komega") {      // not reachable through the reference's factory
        auto model = std::make_unique<AlgebraicKOmegaModel>();
        model->set_nu(cfg.nu);
        solver.set_turbulence_model(std::move(model));
    }

    init_velocity(mesh, solver);
    if (arg_i("call_initialize", 0)) {
        VectorField copy = solver.velocity();
        solver.initialize(copy);
    }
    if (!out.empty() && !bench) {
        dump_mesh(mesh, out + "_mesh");
        dump_state(mesh, solver, out + "_init");
    }

    std::vector<double> residuals, dts;
    for (int s = 0; s < warmup; ++s) solver.step();

    const std::string dump_at = "," + arg_s("dump_at", "") + ",";
    auto t0 = std::chrono::high_resolution_clock::now();
    for (int s = 1; s <= nsteps; ++s) {
        if (cfg.adaptive_dt) {
            double dt = solver.compute_adaptive_dt();
            solver.set_dt(dt);
            dts.push_back(dt);
        }
        residuals.push_back(solver.step());
        if (!bench && !out.empty() && dump_at.find("," + std::to_string(s) + ",") != std::string::npos)
            dump_state(mesh, solver, out + "_s" + std::to_string(s));
    }
    auto t1 = std::chrono::high_resolution_clock::now();
    const double secs = std::chrono::duration<double>(t1 - t0).count();

    std::cerr.rdbuf(old_cerr);
    std::cout.rdbuf(old_cout);
    if (!bench && !out.empty()) dump_state(mesh, solver, out + "_final");

    std::vector<int> cyc = parse_cycles(cerr_cap.str());
    std::ostringstream js;
    js.precision(17);
    js << "{\"mode\":\"" << (bench ? "bench" : "step") << "\",\"Nx\":" << mesh.Nx << ",\"Ny\":" << mesh.Ny
       << ",\"Nz\":" << mesh.Nz << ",\"Ng\":" << mesh.Nghost << ",\"dx\":" << mesh.dx << ",\"dy\":" << mesh.dy
       << ",\"dz\":" << mesh.dz << ",\"nsteps\":" << nsteps << ",\"warmup\":" << warmup
       << ",\"seconds\":" << secs << ",\"ms_per_step\":" << 1e3 * secs / nsteps
       << ",\"mcell_steps_per_s\":" << (double)mesh.Nx * mesh.Ny * mesh.Nz * nsteps / secs / 1e6
       << ",\"last_cycles\":" << solver.poisson_stats().cycles << ",\"residuals\":[";
    for (size_t i = 0; i < residuals.size(); ++i) js << (i ? "," : "") << residuals[i];
    js << "],\"dts\":[";
    for (size_t i = 0; i < dts.size(); ++i) js << (i ? "," : "") << dts[i];
    js << "],\"cycles\":[";
    for (size_t i = 0; i < cyc.size(); ++i) js << (i ? "," : "") << cyc[i];
    js << "]}";
    std::cout << js.str() << std::endl;
    if (!out.empty()) { std::ofstream f(out + "_meta.json"); f << js.str() << "\n"; }
    return 0;
}

int run_mg() {
    Mesh mesh = make_mesh();
    const std::string out = arg_s("out", "");
    std::ostringstream cout_cap;
    std::streambuf* old_cout = std::cout.rdbuf(cout_cap.rdbuf());

    MultigridPoissonSolver mg(mesh);
    const std::string bc = arg_s("bc", "PPPPPP");
    if (mesh.is2D())
        mg.set_bc(pbc_from_char(bc[0]), pbc_from_char(bc[1]), pbc_from_char(bc[2]), pbc_from_char(bc[3]));
    else
        mg.set_bc(pbc_from_char(bc[0]), pbc_from_char(bc[1]), pbc_from_char(bc[2]), pbc_from_char(bc[3]),
                  pbc_from_char(bc[4]), pbc_from_char(bc[5]));

    PoissonConfig pc;
    pc.fixed_cycles = arg_i("fixed", 0);
    pc.max_vcycles = arg_i("maxv", 20);
    pc.tol_rhs = arg_d("tol_rhs", pc.tol_rhs);
    pc.tol_rel = arg_d("tol_rel", pc.tol_rel);
    pc.tol_abs = arg_d("tol_abs", pc.tol_abs);
    pc.tol = arg_d("tol", pc.tol);
    pc.check_interval = arg_i("check_interval", pc.check_interval);
    pc.nu1 = arg_i("nu1", pc.nu1);
    pc.nu2 = arg_i("nu2", pc.nu2);
    pc.chebyshev_degree = arg_i("degree", pc.chebyshev_degree);
    pc.use_l2_norm = arg_i("use_l2", 1) != 0;

    ScalarField rhs(mesh), p(mesh);
    if (!read_f64(arg_s("rhs", ""), rhs.data())) { std::fprintf(stderr, "mg: need rhs=<file>\n"); return 2; }
    const std::string p0 = arg_s("p0", "");
    if (!p0.empty()) read_f64(p0, p.data());

    const int repeat = arg_i("repeat", 1);
    int cycles = 0;
    auto t0 = std::chrono::high_resolution_clock::now();
    for (int r = 0; r < repeat; ++r) cycles = mg.solve(rhs, p, pc);
    auto t1 = std::chrono::high_resolution_clock::now();
    std::cout.rdbuf(old_cout);

    if (!out.empty()) write_f64(out + "_p.f64", p.data());
    if (!out.empty() && mesh.is_y_stretched()) dump_mesh(mesh, out + "_mesh");
    std::ostringstream js;
    js.precision(17);
    js << "{\"mode\":\"mg\",\"cycles\":" << cycles << ",\"residual\":" << mg.residual()
       << ",\"residual_l2\":" << mg.residual_l2() << ",\"b_inf\":" << mg.rhs_norm()
       << ",\"b_l2\":" << mg.rhs_norm_l2() << ",\"r0\":" << mg.initial_residual()
       << ",\"r0_l2\":" << mg.initial_residual_l2() << ",\"converged\":" << (mg.converged() ? 1 : 0)
       << ",\"fixed_mode\":" << (mg.used_fixed_cycles() ? 1 : 0)
       << ",\"seconds\":" << std::chrono::duration<double>(t1 - t0).count() << "}";
    std::cout << js.str() << std::endl;
    if (!out.empty()) { std::ofstream f(out + "_meta.json"); f << js.str() << "\n"; }
    return 0;
}

// TurbulenceModel::update on given velocity (+k, omega) files; 2-D reference semantics
// (reference src/turbulence_nn_tbnn.cpp:640-871, src/turbulence_nn_mlp.cpp:140-292).
int run_nn() {
    Mesh mesh = make_mesh();
    const std::string out = arg_s("out", "");
    std::ostringstream cout_cap;
    std::streambuf* old_cout = std::cout.rdbuf(cout_cap.rdbuf());

    const std::string model = arg_s("model", "nn_tbnn");
    const std::string wdir = arg_s("nn_weights", "");
    auto type = model == "nn_mlp" ? TurbulenceModelType::NNMLP : TurbulenceModelType::NNTBNN;
    auto tm = create_turbulence_model(type, wdir, arg_s("nn_scaling", wdir), 0.1, 0.1);
    tm->set_nu(arg_d("nu", 1.0 / 180.0));
    if (g_args.count("delta")) tm->set_delta(arg_d("delta", 1.0));
    if (auto* mlp = dynamic_cast<TurbulenceNNMLP*>(tm.get())) {
        if (g_args.count("nu_t_max")) mlp->set_nu_t_max(arg_d("nu_t_max", 1.0));
    }

    VectorField vel(mesh);
    const std::string pre = arg_s("in", "");
    read_f64(pre + "_u.f64", vel.u_data());
    read_f64(pre + "_v.f64", vel.v_data());
    if (!mesh.is2D()) read_f64(pre + "_w.f64", vel.w_data());
    ScalarField k(mesh), omega(mesh), nu_t(mesh);
    { std::vector<double> tmp(k.data().size());
      if (read_f64(pre + "_k.f64", tmp)) k.data() = tmp;
      if (read_f64(pre + "_omega.f64", tmp)) omega.data() = tmp; }
    TensorField tau(mesh);

    const int repeat = arg_i("repeat", 1);
    auto t0 = std::chrono::high_resolution_clock::now();
    for (int r = 0; r < repeat; ++r)
        tm->update(mesh, vel, k, omega, nu_t, tm->provides_reynolds_stresses() ? &tau : nullptr, nullptr);
    auto t1 = std::chrono::high_resolution_clock::now();
    std::cout.rdbuf(old_cout);

    if (!out.empty()) {
        write_f64(out + "_nut.f64", nu_t.data());
        write_f64(out + "_tau_xx.f64", tau.xx_data());
        write_f64(out + "_tau_xy.f64", tau.xy_data());
        write_f64(out + "_tau_yy.f64", tau.yy_data());
    }
    const double secs = std::chrono::duration<double>(t1 - t0).count();
    std::cout.precision(17);
    std::cout << "{\"mode\":\"nn\",\"model\":\"" << tm->name() << "\",\"seconds\":" << secs
              << ",\"cells\":" << (long)mesh.Nx * mesh.Ny << ",\"repeat\":" << repeat << "}" << std::endl;
    return 0;
}

// LESModel::update (CPU path: GradientComputer + compute_nu_sgs_cell) on a given velocity; 2-D or 3-D mesh
// (reference src/turbulence_les.cpp:292-335, src/velocity_gradient.cpp:20-119)
int run_les() {
    Mesh mesh = make_mesh();
    const std::string out = arg_s("out", "");
    const std::string model = arg_s("model", "smagorinsky");
    TurbulenceModelType type = model == "wale" ? TurbulenceModelType::WALE
                             : model == "vreman" ? TurbulenceModelType::Vreman
                             : model == "sigma" ? TurbulenceModelType::Sigma : TurbulenceModelType::Smagorinsky;
    auto tm = create_turbulence_model(type, "", "", 0.1, 0.1);
    VectorField vel(mesh);
    const std::string pre = arg_s("in", "");
    read_f64(pre + "_u.f64", vel.u_data());
    read_f64(pre + "_v.f64", vel.v_data());
    if (!mesh.is2D()) read_f64(pre + "_w.f64", vel.w_data());
    ScalarField k(mesh), omega(mesh), nu_t(mesh);
    tm->update(mesh, vel, k, omega, nu_t, nullptr, nullptr);
    if (!out.empty()) {
        write_f64(out + "_nut.f64", nu_t.data());
        dump_mesh(mesh, out + "_mesh");
    }
    std::cout << "{\"mode\":\"les\",\"model\":\"" << tm->name() << "\"}" << std::endl;
    return 0;
}

} // namespace

int main(int argc, char** argv) {
    for (int i = 1; i < argc; ++i) {
        std::string a(argv[i]);
        auto eq = a.find('=');
        if (eq == std::string::npos) { std::fprintf(stderr, "bad arg %s (want key=value)\n", argv[i]); return 2; }
        g_args[a.substr(0, eq)] = a.substr(eq + 1);
    }
    const std::string mode = arg_s("mode", "step");
    try {
        if (mode == "step") return run_step(false);
        if (mode == "bench") return run_step(true);
        if (mode == "mg") return run_mg();
        if (mode == "nn") return run_nn();
        if (mode == "les") return run_les();
        if (mode == "mesh") {                      // dump the reference's mesh arrays (y metrics of stretched meshes)
            Mesh mesh = make_mesh();
            dump_mesh(mesh, arg_s("out", "mesh") + "_mesh");
            std::printf("{\"mode\":\"mesh\",\"stretched\":%d}\n", mesh.is_y_stretched() ? 1 : 0);
            return 0;
        }
    } catch (const std::exception& e) {
        std::fprintf(stderr, "ref_harness: exception: %s\n", e.what());
        return 3;
    }
    std::fprintf(stderr, "unknown mode %s\n", mode.c_str());
    return 2;
}
