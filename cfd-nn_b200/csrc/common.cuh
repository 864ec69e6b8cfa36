// common.cuh — shared host/device helpers for libcfdnn_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <atomic>

#include "../../include/cfdnn_b200.h"

namespace cfdnn {

// ---------------------------------------------------------------- errors
struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

void set_last_error(const std::string& m);
extern std::atomic<unsigned long long> g_launches;

#define CFDNN_CUDA(call)                                                                     \
    do {                                                                                     \
        cudaError_t e__ = (call);                                                            \
        if (e__ != cudaSuccess) {                                                            \
            int c__ = (e__ == cudaErrorNoDevice || e__ == cudaErrorInsufficientDriver ||     \
                       e__ == cudaErrorNoKernelImageForDevice)                               \
                          ? CFDNN_ERR_NO_DEVICE : CFDNN_ERR_CUDA;                            \
            throw ::cfdnn::Error(c__, std::string(#call) + ": " + cudaGetErrorString(e__) + \
                                          " (" + __FILE__ + ":" + std::to_string(__LINE__) + ")"); \
        }                                                                                    \
    } while (0)

// Count every kernel we launch (bench.py reports it as gpu_launches).
// CFDNN_LAUNCH_COOP marks kernels whose threads cooperate (__syncthreads / warp shuffles); on
// the GPU the two are the same launch, the distinction only matters to the host-side test
// double in tests/devsim (which is never part of this library).
#ifndef CFDNN_DEVSIM
#define CFDNN_LAUNCH(kernel, grid, block, smem, stream, ...)                                 \
    do {                                                                                     \
        kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);                          \
        ::cfdnn::g_launches.fetch_add(1, std::memory_order_relaxed);                         \
    } while (0)
#define CFDNN_LAUNCH_COOP CFDNN_LAUNCH
#define CFDNN_DYN_SMEM(T, name) extern __shared__ T name[]
#else
#define CFDNN_LAUNCH(kernel, grid, block, smem, stream, ...)                                 \
    do {                                                                                     \
        ::devsim::launch_plain(dim3(grid), dim3(block), (smem), [&] { kernel(__VA_ARGS__); }); \
        ::cfdnn::g_launches.fetch_add(1, std::memory_order_relaxed);                         \
    } while (0)
#define CFDNN_LAUNCH_COOP(kernel, grid, block, smem, stream, ...)                            \
    do {                                                                                     \
        ::devsim::launch_coop(dim3(grid), dim3(block), (smem), [&] { kernel(__VA_ARGS__); }); \
        ::cfdnn::g_launches.fetch_add(1, std::memory_order_relaxed);                         \
    } while (0)
#endif

inline void check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess)
        throw Error(e == cudaErrorNoKernelImageForDevice ? CFDNN_ERR_NO_DEVICE : CFDNN_ERR_CUDA,
                    std::string(what) + ": " + cudaGetErrorString(e));
}

template <typename F>
int guarded(F&& f) {
    try {
        f();
        return CFDNN_OK;
    } catch (const Error& e) {
        set_last_error(e.what());
        return e.code;
    } catch (const std::invalid_argument& e) {
        set_last_error(e.what());
        return CFDNN_ERR_INVALID;
    } catch (const std::exception& e) {
        set_last_error(e.what());
        return CFDNN_ERR_RUNTIME;
    }
}

void require_device();

// ---------------------------------------------------------------- device memory
template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    DevBuf(DevBuf&& o) noexcept : p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
    DevBuf& operator=(DevBuf&& o) noexcept {
        if (this != &o) { release(); p = o.p; n = o.n; o.p = nullptr; o.n = 0; }
        return *this;
    }
    ~DevBuf() { release(); }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
    void alloc(size_t count, bool zero = true) {
        release();
        n = count;
        if (count == 0) return;
        CFDNN_CUDA(cudaMalloc(&p, count * sizeof(T)));
        if (zero) CFDNN_CUDA(cudaMemset(p, 0, count * sizeof(T)));
    }
};

// ---------------------------------------------------------------- device math
struct Div {       // a divisor and its reciprocal, passed by value to kernels
    double d, r;
};
#if defined(__CUDACC__) || defined(CFDNN_DEVSIM)
// Correctly rounded x/d from a precomputed rd = RN(1/d) (Markstein correction step).
// The FP64 stencils of the reference are written with true divisions (e.g.
// src/poisson_solver_multigrid.cpp:976-979); this reproduces them with 3 FMA-pipe
// instructions instead of the ~15-instruction IEEE division sequence.  Files are compiled
// with -fmad=false so only the explicit fma() calls here are fused.
__device__ __forceinline__ double fdiv(double x, double d, double rd) {
    double q = x * rd;
    double e = fma(-q, d, x);
    return fma(e, rd, q);
}

__device__ __forceinline__ double operator/(double x, const Div& v) { return fdiv(x, v.d, v.r); }

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_down_sync(0xffffffffu, v, o));
    return v;
}
// Block reductions (deterministic: fixed shuffle tree, fixed warp order). Result valid in thread 0.
__device__ __forceinline__ double block_sum(double v, double* sh /* >=32 */) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    if (w == 0) {
        v = (lane < nw) ? sh[lane] : 0.0;
        v = warp_sum(v);
    }
    return v;
}
__device__ __forceinline__ double block_max(double v, double* sh) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int nw = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    if (w == 0) {
        v = (lane < nw) ? sh[lane] : 0.0;
        v = warp_max(v);
    }
    return v;
}
#endif

inline Div make_div_host(double d) { return Div{d, 1.0 / d}; }

}  // namespace cfdnn
