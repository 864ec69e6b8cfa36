// nn.cu — NN closures (see nn.cuh).  Placeholder bodies until the kernels land.
#include "nn.cuh"
namespace cfdnn {
NNClosure::NNClosure(int kind, int, const int*, const int*, const double* const*, const double* const*,
                     const double*, const double*) : kind_(kind) {
    throw Error(CFDNN_ERR_RUNTIME, "NN closure not built yet");
}
NNClosure::~NNClosure() {}
void NNClosure::set_params(double, double, double, double, double) {}
void NNClosure::update_host(const cfdnn_mesh_desc&, const double*, const double*, const double*, const double*,
                            double*, double*, double*, double*) {}
void NNClosure::update_device(const Geom&, const cfdnn_mesh_desc&, const Vel&, const double*, const double*,
                              double*, double*, double*, double*, cudaStream_t) {}
void NNClosure::forward_host(const double*, double*, size_t) {}
}  // namespace cfdnn
