// nn.cuh — per-cell neural-network turbulence closures (TBNN and scalar-nu_t MLP).
//
// Stands in for TurbulenceNNTBNN / TurbulenceNNMLP (reference include/turbulence_nn_tbnn.hpp:35-56,
// include/turbulence_nn_mlp.hpp, src/turbulence_nn_tbnn.cpp:640-871, src/turbulence_nn_mlp.cpp:140-292),
// the FP64 MLP of src/nn_core.cpp:36-204 and the features / tensor basis of src/features.cpp:67-326.
#pragma once
#include "common.cuh"
#include "solver.cuh"
#include <vector>

namespace cfdnn {

class NNClosure {
public:
    NNClosure(int kind, int n_layers, const int* dims, const int* act, const double* const* W,
              const double* const* b, const double* means, const double* stds);
    ~NNClosure();
    void set_params(double nu, double delta, double u_ref, double k_min, double nu_t_max);

    // TurbulenceModel::update on host arrays (2-D mesh: reference semantics; 3-D: every interior plane)
    void update_host(const cfdnn_mesh_desc& m, const double* u, const double* v, const double* k,
                     const double* omega, double* nu_t, double* txx, double* txy, double* tyy);
    // device-resident twin used by Solver::step
    void update_device(const Geom& g, const cfdnn_mesh_desc& m, const Vel& vel, const double* k,
                       const double* omega, double* nu_t, double* txx, double* txy, double* tyy,
                       cudaStream_t st);
    // MLP::forward on a batch: x[n][in] -> y[n][out]
    void forward_host(const double* x, double* y, size_t n);
    double last_ms = 0.0;
    int kind() const { return kind_; }

private:
    int kind_;
    std::vector<int> dims_, act_;
    DevBuf<double> W_, b_, means_, stds_;
    std::vector<size_t> woff_, boff_;
    bool has_scaling_ = false;
    double nu_ = 0.001, delta_ = 1.0, u_ref_ = 1.0, k_min_ = 1e-10, nu_t_max_ = 1.0;
    cudaStream_t st_ = nullptr;
    cudaEvent_t ev0_ = nullptr, ev1_ = nullptr;
    DevBuf<double> scratch_[12];
    DevBuf<double> yc_, dyc_;
};

}  // namespace cfdnn
