// tests/devsim/cuda_runtime.h — TEST INFRASTRUCTURE (never part of libcfdnn_b200.so).
//
// A host-side test double of the small slice of the CUDA runtime that cfd-nn_b200/csrc uses,
// so the *host logic* of the backend (V-cycle recursion, RK3 buffer choreography, BC pass
// order, convergence checks, slab bookkeeping) and the arithmetic of the plain kernels can be
// exercised by `pytest -m "not gpu"` in a container without a GPU, against the oracle.
//
// It is selected only by compiling the .cu sources as C++ with -DCFDNN_DEVSIM and
// -Itests/devsim (this header then shadows <cuda_runtime.h>); tests/devsim/build.sh writes
// the result to tests/devsim/libcfdnn_devsim.so.  The product library is built by nvcc for
// sm_100a only and has no CPU path; nothing under cfd-nn_b200/ loads the devsim library.
//
// Execution model: a kernel launch runs every (block, thread) sequentially on the calling
// host thread.  Plain launches call the kernel body in nested loops.  Cooperative launches
// (kernels that use __syncthreads / warp shuffles; CFDNN_LAUNCH_COOP) run each block's
// threads as ucontext fibers with block- and warp-level barriers.
#pragma once
#ifndef CFDNN_DEVSIM
#error "tests/devsim/cuda_runtime.h is only for -DCFDNN_DEVSIM test builds"
#endif

#include <ucontext.h>
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline __attribute__((always_inline))
#define __restrict__ __restrict
#define __launch_bounds__(...)
#define __shared__ static
#define __constant__ static
#define __align__(n) __attribute__((aligned(n)))

struct dim3 {
    unsigned x, y, z;
    dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {}
};
struct uint3 { unsigned x, y, z; };
struct double2 { double x, y; };

typedef int cudaError_t;
enum { cudaSuccess = 0, cudaErrorNoDevice = 100, cudaErrorInsufficientDriver = 35,
       cudaErrorNoKernelImageForDevice = 209, cudaErrorInvalidValue = 1 };
typedef struct devsim_stream* cudaStream_t;
struct devsim_event { std::chrono::steady_clock::time_point t; };
typedef devsim_event* cudaEvent_t;
typedef void* cudaGraph_t;
typedef void* cudaGraphExec_t;
enum cudaMemcpyKind { cudaMemcpyHostToHost, cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost,
                      cudaMemcpyDeviceToDevice, cudaMemcpyDefault };
enum { cudaStreamNonBlocking = 1, cudaStreamCaptureModeThreadLocal = 1, cudaEventDisableTiming = 2 };

namespace devsim {
extern uint3 g_tid, g_bid;
extern dim3 g_bdim, g_gdim;
extern void* g_dyn_smem;
extern bool g_in_coop;
void sync_block();
void sync_warp();
int lane_id();
uint64_t* warp_slot(int lane);
void launch_coop(dim3 grid, dim3 block, size_t smem, const std::function<void()>& body);
template <typename F>
inline void launch_plain(dim3 grid, dim3 block, size_t smem, F&& body) {
    static std::vector<char> dyn;
    if (dyn.size() < smem) dyn.resize(smem);
    g_dyn_smem = dyn.data();
    g_gdim = grid; g_bdim = block; g_in_coop = false;
    for (unsigned bz = 0; bz < grid.z; ++bz)
     for (unsigned by = 0; by < grid.y; ++by)
      for (unsigned bx = 0; bx < grid.x; ++bx) {
        g_bid = uint3{bx, by, bz};
        for (unsigned tz = 0; tz < block.z; ++tz)
         for (unsigned ty = 0; ty < block.y; ++ty)
          for (unsigned tx = 0; tx < block.x; ++tx) {
            g_tid = uint3{tx, ty, tz};
            body();
          }
      }
}
}  // namespace devsim

#define threadIdx (::devsim::g_tid)
#define blockIdx (::devsim::g_bid)
#define blockDim (::devsim::g_bdim)
#define gridDim (::devsim::g_gdim)

inline void __syncthreads() { devsim::sync_block(); }
inline void __syncwarp(unsigned = 0xffffffffu) { devsim::sync_warp(); }
template <typename T>
inline T devsim_shfl(T v, int src) {
    static_assert(sizeof(T) <= 8, "shuffle payload");
    const int lane = devsim::lane_id();
    uint64_t raw = 0;
    std::memcpy(&raw, &v, sizeof(T));
    *devsim::warp_slot(lane) = raw;
    devsim::sync_warp();
    T r = v;
    if (src >= 0 && src < 32) {
        uint64_t o = *devsim::warp_slot(src);
        std::memcpy(&r, &o, sizeof(T));
    }
    devsim::sync_warp();
    return r;
}
template <typename T> inline T __shfl_down_sync(unsigned, T v, int d) { int s = devsim::lane_id() + d; return devsim_shfl(v, s < 32 ? s : -1); }
template <typename T> inline T __shfl_up_sync(unsigned, T v, int d) { int s = devsim::lane_id() - d; return devsim_shfl(v, s >= 0 ? s : -1); }
template <typename T> inline T __shfl_xor_sync(unsigned, T v, int m) { return devsim_shfl(v, devsim::lane_id() ^ m); }
template <typename T> inline T __shfl_sync(unsigned, T v, int s) { return devsim_shfl(v, s & 31); }
template <typename T> inline T __ldg(const T* p) { return *p; }
inline double __dmul_rn(double a, double b) { return a * b; }
inline double __dadd_rn(double a, double b) { return a + b; }
inline double __fma_rn(double a, double b, double c) { return std::fma(a, b, c); }
inline double __drcp_rn(double a) { return 1.0 / a; }
using std::fma; using std::fmax; using std::fmin; using std::fabs; using std::min; using std::max;
inline bool devsim_isnan(double x) { return x != x; }

// ---- runtime entry points (synchronous; "device memory" is host memory) ----
inline cudaError_t cudaMalloc(void** p, size_t n) { *p = std::malloc(n ? n : 1); return *p ? cudaSuccess : 2; }
template <typename T> inline cudaError_t cudaMalloc(T** p, size_t n) { return cudaMalloc((void**)p, n); }
inline cudaError_t cudaFree(void* p) { std::free(p); return cudaSuccess; }
inline cudaError_t cudaMallocHost(void** p, size_t n) { return cudaMalloc(p, n); }
template <typename T> inline cudaError_t cudaMallocHost(T** p, size_t n) { return cudaMalloc((void**)p, n); }
inline cudaError_t cudaFreeHost(void* p) { std::free(p); return cudaSuccess; }
inline cudaError_t cudaMemset(void* p, int v, size_t n) { std::memset(p, v, n); return cudaSuccess; }
inline cudaError_t cudaMemsetAsync(void* p, int v, size_t n, cudaStream_t = nullptr) { std::memset(p, v, n); return cudaSuccess; }
inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { std::memmove(d, s, n); return cudaSuccess; }
inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t = nullptr) { std::memmove(d, s, n); return cudaSuccess; }
inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned) { *s = (cudaStream_t)std::malloc(1); return cudaSuccess; }
inline cudaError_t cudaStreamCreate(cudaStream_t* s) { return cudaStreamCreateWithFlags(s, 0); }
inline cudaError_t cudaStreamDestroy(cudaStream_t s) { std::free(s); return cudaSuccess; }
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned = 0) { return cudaSuccess; }
inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
inline int __clz(int x) { return x ? __builtin_clz((unsigned)x) : 32; }
inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = new devsim_event(); return cudaSuccess; }
inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, unsigned) { return cudaEventCreate(e); }
inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return cudaSuccess; }
inline cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t = nullptr) { e->t = std::chrono::steady_clock::now(); return cudaSuccess; }
inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t a, cudaEvent_t b) {
    *ms = std::chrono::duration<float, std::milli>(b->t - a->t).count(); return cudaSuccess; }
inline cudaError_t cudaGetLastError() { return cudaSuccess; }
inline cudaError_t cudaPeekAtLastError() { return cudaSuccess; }
inline const char* cudaGetErrorString(cudaError_t) { return "devsim error"; }
inline cudaError_t cudaGetDeviceCount(int* n) { *n = 1; return cudaSuccess; }
inline cudaError_t cudaSetDevice(int) { return cudaSuccess; }
inline cudaError_t cudaGetDevice(int* d) { *d = 0; return cudaSuccess; }
template <typename T>
inline cudaError_t cudaGetSymbolAddress(void** p, T& symbol) { *p = (void*)&symbol; return cudaSuccess; }
inline cudaError_t cudaGraphExecDestroy(cudaGraphExec_t) { return cudaSuccess; }
inline cudaError_t cudaGraphDestroy(cudaGraph_t) { return cudaSuccess; }
template <typename K> inline cudaError_t cudaFuncSetAttribute(K, int, int) { return cudaSuccess; }
enum { cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };

#define CFDNN_DYN_SMEM(T, name) T* name = reinterpret_cast<T*>(::devsim::g_dyn_smem)
// CUDA graphs are not modelled: the backend is told so (Multigrid::use_graph_ = false under CFDNN_DEVSIM)
inline cudaError_t cudaStreamBeginCapture(cudaStream_t, int) { return cudaErrorInvalidValue; }
inline cudaError_t cudaStreamEndCapture(cudaStream_t, cudaGraph_t*) { return cudaErrorInvalidValue; }
inline cudaError_t cudaGraphInstantiate(cudaGraphExec_t*, cudaGraph_t, unsigned long long) { return cudaErrorInvalidValue; }
inline cudaError_t cudaGraphLaunch(cudaGraphExec_t, cudaStream_t) { return cudaErrorInvalidValue; }
