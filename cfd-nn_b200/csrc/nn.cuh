// nn.cuh — per-cell neural-network turbulence closures (TBNN and scalar-nu_t MLP).
//
// Stands in for TurbulenceNNTBNN / TurbulenceNNMLP (reference include/turbulence_nn_tbnn.hpp:35-56,
// include/turbulence_nn_mlp.hpp, src/turbulence_nn_tbnn.cpp:640-871, src/turbulence_nn_mlp.cpp:140-292),
// the FP64 MLP of src/nn_core.cpp:36-204 and the features / tensor basis of src/features.cpp:67-326.
// CPU-path semantics: k/omega estimated algebraically when k is not provided,
// uniform-spacing gradients for the features, stretched-aware gradients for the tensor basis.
// 3-D meshes: the 2-D update is applied to every interior z-plane.
#pragma once
#include "common.cuh"
#include "solver.cuh"
#include <vector>

namespace cfdnn {

constexpr int NN_MAX_LAYERS = 8;
constexpr int NN_MAX_WIDTH = 64;

struct NNDev {                       // by-value kernel argument
    int n_layers;
    int dims[NN_MAX_LAYERS + 1];
    int act[NN_MAX_LAYERS];          // 0 linear, 1 tanh, 2 relu
    int woff[NN_MAX_LAYERS], boff[NN_MAX_LAYERS];
    const double *W, *b;             // row-major (out,in) per layer, concatenated
    const double *means, *stds;      // null when the model has no input scaling
    int n_scale;
};

struct NNFrag {                      // by-value kernel argument: weights in DMMA fragment order (nn.cu)
    int n_layers;
    int nti[NN_MAX_LAYERS], nto[NN_MAX_LAYERS];   // 8-wide K / N tiles per layer
    int act[NN_MAX_LAYERS];
    int woff[NN_MAX_LAYERS];         // double2 offset of layer l: [nto][nti][32 lanes] double2
    int boff[NN_MAX_LAYERS];         // double offset of the zero-padded bias (nto*8 entries)
    int w_total, b_total;            // double2 / double counts
    const double2* W;
    const double* b;
    const double *means, *stds;
    int n_in, n_out;
};

struct NNParams { double nu, delta, u_ref, k_min, nu_t_max; };

class NNClosure {
public:
    NNClosure(int kind, int n_layers, const int* dims, const int* act, const double* const* W,
              const double* const* b, const double* means, const double* stds);
    ~NNClosure();
    void set_params(double nu, double delta, double u_ref, double k_min, double nu_t_max);

    // TurbulenceModel::update on host arrays (2-D mesh: reference semantics; 3-D: every interior plane)
    void update_host(const cfdnn_mesh_desc& m, const double* u, const double* v, const double* k,
                     const double* omega, double* nu_t, double* txx, double* txy, double* tyy);
    // device-resident twin used by Solver::step.  yc: device array of Ny+2Ng cell-centre y.
    void update_device(const Geom& g, const double* yc, double y_lo, double y_hi, const Vel& vel,
                       const double* k, const double* omega, double* nu_t, double* txx, double* txy,
                       double* tyy, cudaStream_t st, int data_plane = -1);
    // TurbulenceModel::update with a device view: every pointer is a device twin in the reference layout
    void update_dev_ptrs(const cfdnn_mesh_desc& m, const double* u, const double* v, const double* k,
                         const double* omega, double* nu_t, double* txx, double* txy, double* tyy);
    // MLP::forward on a batch: x[n][in] -> y[n][out]
    void forward_host(const double* x, double* y, size_t n);
    double last_ms = 0.0;
    int kind() const { return kind_; }
    int in_dim() const { return net_.dims[0]; }
    int out_dim() const { return net_.dims[net_.n_layers]; }
    bool uses_tensor_cores() const { return mma_ok_; }

private:
    int kind_;
    NNDev net_{};
    NNParams prm_{0.001, 1.0, 1.0, 1e-10, 1.0};
    DevBuf<double> W_, b_, means_, stds_;
    DevBuf<int> kflag_;
    // tensor-core (DMMA) path: pre-permuted weight fragments
    NNFrag frag_{};
    DevBuf<double2> Wf_;
    DevBuf<double> bf_;
    size_t mma_smem_ = 0;
    bool mma_ok_ = false;
    int mmah_ = 0;                   // > 0: hidden width / 8 of the shape-specialised kernel (k_nn_update_mmah)
    size_t mmah_smem_ = 0;
    void build_fragments(const std::vector<double>& hw, const std::vector<double>& hb);
    cudaStream_t st_ = nullptr;
    cudaEvent_t ev0_ = nullptr, ev1_ = nullptr;
};

}  // namespace cfdnn
