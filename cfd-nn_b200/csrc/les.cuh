// les.cuh — static LES subgrid-scale closures (Smagorinsky, WALE, Vreman, Sigma); see les.cu.
#pragma once
#include "common.cuh"
#include "solver.cuh"

namespace cfdnn {

double les_default_constant(int model);

struct LesMetrics {                 // device copies of the y arrays the gradient tensor and the filter width need
    explicit LesMetrics(const cfdnn_mesh_desc& m);
    DevBuf<double> yc, yf, delta;   // Ny+2Ng centres, Ny+2Ng+1 faces, filter width per row
};

// nu_t = nu_sgs on `nplanes` z-planes starting at plane0 (device arrays, reference layouts)
void les_update_device(const Geom& g, const cfdnn_mesh_desc& m, const LesMetrics& met, int model, double C, const Vel& vel,
                       double* nu_t, int plane0, int nplanes, bool is2d, cudaStream_t st);
// LESModel::update on host arrays
void les_update_host(const cfdnn_mesh_desc& m, int model, double C, const double* u, const double* v, const double* w,
                     double* nu_t);

}  // namespace cfdnn
