// sst.cuh — SST k-omega (+ SST closure) and standard k-omega (+ Boussinesq closure) transport models on 2-D meshes
// (SURVEY §8(f)2); see sst.cu.
#pragma once
#include "common.cuh"
#include "solver.cuh"

namespace cfdnn {

struct SstParams {                  // SSTConstants, include/turbulence_transport.hpp:32-58 (defaults), + nu
    double nu;
    double sigma_k1 = 0.85, sigma_omega1 = 0.5, beta1 = 0.075, alpha1 = 5.0 / 9.0;
    double sigma_k2 = 1.0, sigma_omega2 = 0.856, beta2 = 0.0828, alpha2 = 0.44;
    double beta_star = 0.09, a1 = 0.31;
    double CD_omega_min = 1e-10;
    double k_min = 1e-10, omega_min = 1e-10, k_max = 100.0, omega_max = 1e8;
};

class SstModel {
public:
    SstModel(const cfdnn_mesh_desc& m, double nu, int kind);      // kind: CFDNN_RANS_SST ... CFDNN_RANS_ALGEBRAIC_KOMEGA
    // SSTKOmegaTransport::advance_turbulence (CPU path, src/turbulence_transport.cpp:768-865): one in-place sweep over
    // the cells in (j, i) order + wall BCs of k and omega.  Fields are planes of the solver's cell layout.
    void advance(const Geom& g, const Vel& vel, double* k, double* omega, const double* nu_t, double dt, int plane,
                 cudaStream_t st);
    // SSTKOmegaTransport::update -> SSTClosure::compute_nu_t (:259-313)
    // (the EARSM kinds also write the Reynolds stresses when txx / txy / tyy are given)
    void closure(const Geom& g, const Vel& vel, const double* k, const double* omega, double* nu_t, double* txx, double* txy,
                 double* tyy, int plane, cudaStream_t st);
    SstParams prm;
    double pope_C1 = 0.1, pope_C2 = 0.1;                           // PopeQuadraticEARSM defaults (include/turbulence_earsm.hpp:255)
    double gep_nu = 0.001;                                         // TurbulenceGEP's own nu_ (see cfdnn_solver_set_gep_params)
    int gep_variant = 2;                                           // TurbulenceGEP::Variant (0 channel, 1 periodic hill, 2 simple = default)
    double ml_kappa = 0.41, ml_A_plus = 26.0, ml_delta = 1.0;     // MixingLengthModel, include/turbulence_baseline.hpp:43-45

private:
    int kind_;
    double dx_, dy_, dy_wall_;
    DevBuf<double> utau_;           // friction velocity of the algebraic models (device scalar)
    DevBuf<double> wall_;           // wall distance per row, Ny + 2 Ng (Mesh::wall_distance, src/mesh.cpp:6-12)
    DevBuf<double> dyc_;            // centre spacings of a stretched mesh (closure gradients), else empty
};

}  // namespace cfdnn
