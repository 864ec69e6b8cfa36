// comm.cuh — z-slab communicator: halo planes, scalar all-reduces, coarse-level gather.
//
// Replaces the reference's Decomposition + HaloExchange pair (include/decomposition.hpp:81-92,
// src/decomposition.cpp:60-113, src/halo_exchange.cpp:93-192: MPI Isend/Irecv of Ng z-planes,
// host-staged on GPU builds) with NCCL point-to-point / collectives issued on the solver's
// stream over NVLink.  z is the slowest index, so a halo plane is one contiguous run of
// doubles: no pack/unpack kernels (reference src/cuda_kernels/halo_pack.cu:149-173).
//
// libnccl is dlopen()ed on first use so that the single-GPU path has no NCCL dependency and,
// inside a torch process, the already-loaded bundled libnccl.so.2 is the one that is used.
#pragma once
#include "common.cuh"
#include <cstddef>

namespace cfdnn {

struct Comm {
    int rank = 0, nranks = 1;
    virtual ~Comm() {}
    // field: level/field array whose z-planes are `plane` doubles; interior planes live at
    // plane indices [ng, ng+nz).  Sends the first/last ng interior planes to the lower/upper
    // neighbour's opposite ghost planes (ring when the end ranks pass lo/hi = true).
    void exchange_planes(double* field, size_t plane, int nz, int ng, bool lo, bool hi, cudaStream_t st) {
        const size_t cnt = plane * (size_t)ng;
        exchange_ptrs(field + cnt, cnt, field, cnt, field + (size_t)nz * plane, cnt,
                      field + (size_t)(nz + ng) * plane, cnt, lo, hi, st);
    }
    // general form: send `n_send_lo` doubles to the lower neighbour and receive `n_recv_lo` from it
    // (if lo), same with the upper neighbour (if hi).  A count of 0 skips that transfer; the two
    // ends of a link must agree on the counts.  Posting order is send-up, recv-lo, send-lo,
    // recv-up so that the pairs also match when both neighbours are the same rank (P = 2 ring).
    virtual void exchange_ptrs(const double* send_lo, size_t n_send_lo, double* recv_lo, size_t n_recv_lo,
                               const double* send_hi, size_t n_send_hi, double* recv_hi, size_t n_recv_hi,
                               bool lo, bool hi, cudaStream_t st) = 0;
    virtual void allreduce_sum(double* d, int n, cudaStream_t st) = 0;
    virtual void allreduce_max(double* d, int n, cudaStream_t st) = 0;
    virtual void allreduce_min(double* d, int n, cudaStream_t st) = 0;
    // in place: rank r's `count` doubles already sit at base + r*count
    virtual void allgather_planes(double* base, size_t count, cudaStream_t st) = 0;
    // block q of `send` (count doubles at send + q*count) goes to rank q, which stores it as block `rank` of its
    // `recv`; send and recv must not overlap (the FFT Poisson solver's transposes)
    virtual void alltoall(const double* send, double* recv, size_t count, cudaStream_t st) = 0;
    // true when the ranks are GPUs of one node driven by NCCL (device buffers, CUDA IPC possible); false for the host
    // callback transport of the CPU test double
    virtual bool device_transport() const { return false; }
};

// NCCL transport (product).  id128 = ncclUniqueId bytes made by cfdnn_comm_unique_id on rank 0.
Comm* make_nccl_comm(int rank, int nranks, const void* id128);
void nccl_unique_id(void* id128);

// Host-callback transport: used by the CPU test double only (tests/devsim, gloo world_size 2).
struct cfdnn_comm_callbacks_t {
    void* ctx;
    // exchange with two peers at once; a peer < 0 means "no transfer on that side"
    void (*sendrecv)(void* ctx, int lo_peer, const double* send_lo, size_t n_send_lo, double* recv_lo,
                     size_t n_recv_lo, int hi_peer, const double* send_hi, size_t n_send_hi, double* recv_hi,
                     size_t n_recv_hi);
    void (*allreduce)(void* ctx, double* d, int n, int op /*0 sum 1 max 2 min*/);
    void (*allgather)(void* ctx, double* base, size_t count);
};
Comm* make_callback_comm(int rank, int nranks, const cfdnn_comm_callbacks_t& cb);

}  // namespace cfdnn
