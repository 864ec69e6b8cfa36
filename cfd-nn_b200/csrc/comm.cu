// comm.cu — NCCL-backed slab communicator (see comm.cuh).
#include "comm.cuh"
#include <vector>
#include <dlfcn.h>
#include <cstring>
#include <mutex>

namespace cfdnn {

#ifndef CFDNN_DEVSIM
namespace {
// Minimal NCCL surface, bound at run time (ABI of nccl.h 2.x).
typedef struct ncclComm* ncclComm_t;
struct ncclUniqueId { char internal[128]; };
enum { ncclSuccess = 0 };
enum { ncclFloat64 = 8 };
enum { ncclSum = 0, ncclProd = 1, ncclMax = 2, ncclMin = 3 };

struct Nccl {
    void* lib = nullptr;
    int (*GetUniqueId)(ncclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
Nccl& nccl() {
    static Nccl n;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* nm : names) {
            n.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
            if (n.lib) break;
        }
        if (!n.lib) return;
#define BIND(f) n.f = reinterpret_cast<decltype(n.f)>(dlsym(n.lib, "nccl" #f))
        BIND(GetUniqueId); BIND(CommInitRank); BIND(CommDestroy); BIND(GroupStart); BIND(GroupEnd);
        BIND(Send); BIND(Recv); BIND(AllReduce); BIND(AllGather); BIND(GetErrorString);
#undef BIND
    });
    if (!n.lib || !n.CommInitRank || !n.Send || !n.AllReduce)
        throw Error(CFDNN_ERR_RUNTIME, "libnccl.so.2 could not be loaded: multi-GPU slabs need NCCL");
    return n;
}
#define NCCL_OK(call)                                                                          \
    do {                                                                                       \
        int r__ = (call);                                                                      \
        if (r__ != ncclSuccess)                                                                \
            throw Error(CFDNN_ERR_RUNTIME, std::string(#call) + ": " +                         \
                        (nccl().GetErrorString ? nccl().GetErrorString(r__) : "nccl error")); \
    } while (0)

struct NcclComm final : Comm {
    ncclComm_t comm = nullptr;
    ~NcclComm() override { if (comm) nccl().CommDestroy(comm); }
    void exchange_ptrs(const double* send_lo, size_t n_send_lo, double* recv_lo, size_t n_recv_lo,
                       const double* send_hi, size_t n_send_hi, double* recv_hi, size_t n_recv_hi,
                       bool lo, bool hi, cudaStream_t st) override {
        if (!lo && !hi) return;
        Nccl& n = nccl();
        const int dn = (rank + nranks - 1) % nranks, up = (rank + 1) % nranks;
        NCCL_OK(n.GroupStart());
        if (hi && n_send_hi) NCCL_OK(n.Send(send_hi, n_send_hi, ncclFloat64, up, comm, st));
        if (lo && n_recv_lo) NCCL_OK(n.Recv(recv_lo, n_recv_lo, ncclFloat64, dn, comm, st));
        if (lo && n_send_lo) NCCL_OK(n.Send(send_lo, n_send_lo, ncclFloat64, dn, comm, st));
        if (hi && n_recv_hi) NCCL_OK(n.Recv(recv_hi, n_recv_hi, ncclFloat64, up, comm, st));
        NCCL_OK(n.GroupEnd());
    }
    void allreduce(double* d, int cnt, int op, cudaStream_t st) {
        NCCL_OK(nccl().AllReduce(d, d, (size_t)cnt, ncclFloat64, op, comm, st));
    }
    void allreduce_sum(double* d, int n, cudaStream_t st) override { allreduce(d, n, ncclSum, st); }
    void allreduce_max(double* d, int n, cudaStream_t st) override { allreduce(d, n, ncclMax, st); }
    void allreduce_min(double* d, int n, cudaStream_t st) override { allreduce(d, n, ncclMin, st); }
    void allgather_planes(double* base, size_t count, cudaStream_t st) override {
        NCCL_OK(nccl().AllGather(base + (size_t)rank * count, base, count, ncclFloat64, comm, st));
    }
    bool device_transport() const override { return true; }
    void alltoall(const double* send, double* recv, size_t count, cudaStream_t st) override {
        Nccl& n = nccl();
        CFDNN_CUDA(cudaMemcpyAsync(recv + (size_t)rank * count, send + (size_t)rank * count, count * sizeof(double),
                                   cudaMemcpyDeviceToDevice, st));
        NCCL_OK(n.GroupStart());
        for (int d = 1; d < nranks; ++d) {           // pair (rank + d, rank - d): every rank posts the same pattern
            const int to = (rank + d) % nranks, from = (rank + nranks - d) % nranks;
            NCCL_OK(n.Send(send + (size_t)to * count, count, ncclFloat64, to, comm, st));
            NCCL_OK(n.Recv(recv + (size_t)from * count, count, ncclFloat64, from, comm, st));
        }
        NCCL_OK(n.GroupEnd());
    }
};
}  // namespace

void nccl_unique_id(void* id128) {
    ncclUniqueId id;
    NCCL_OK(nccl().GetUniqueId(&id));
    std::memcpy(id128, &id, sizeof(id));
}

Comm* make_nccl_comm(int rank, int nranks, const void* id128) {
    if (nranks < 1 || rank < 0 || rank >= nranks) throw std::invalid_argument("comm_init: bad rank/nranks");
    auto* c = new NcclComm();
    c->rank = rank; c->nranks = nranks;
    if (nranks > 1) {
        ncclUniqueId id;
        std::memcpy(&id, id128, sizeof(id));
        int r = nccl().CommInitRank(&c->comm, nranks, id, rank);
        if (r != ncclSuccess) { delete c; throw Error(CFDNN_ERR_RUNTIME, "ncclCommInitRank failed"); }
    }
    return c;
}
#else   // ---- CFDNN_DEVSIM: no NCCL in the CPU test double
void nccl_unique_id(void* id128) { std::memset(id128, 0, 128); }
Comm* make_nccl_comm(int, int, const void*) {
    throw Error(CFDNN_ERR_RUNTIME, "devsim build has no NCCL transport");
}
#endif

#ifdef CFDNN_DEVSIM
// Host-callback transport: exists in the CPU test double only.
namespace {
struct CallbackComm final : Comm {
    cfdnn_comm_callbacks_t cb;
    void exchange_ptrs(const double* send_lo, size_t n_send_lo, double* recv_lo, size_t n_recv_lo,
                       const double* send_hi, size_t n_send_hi, double* recv_hi, size_t n_recv_hi,
                       bool lo, bool hi, cudaStream_t) override {
        if (!lo && !hi) return;
        const int dn = (rank + nranks - 1) % nranks, up = (rank + 1) % nranks;
        cb.sendrecv(cb.ctx, lo ? dn : -1, send_lo, n_send_lo, recv_lo, n_recv_lo,
                    hi ? up : -1, send_hi, n_send_hi, recv_hi, n_recv_hi);
    }
    void allreduce_sum(double* d, int n, cudaStream_t) override { cb.allreduce(cb.ctx, d, n, 0); }
    void allreduce_max(double* d, int n, cudaStream_t) override { cb.allreduce(cb.ctx, d, n, 1); }
    void allreduce_min(double* d, int n, cudaStream_t) override { cb.allreduce(cb.ctx, d, n, 2); }
    void allgather_planes(double* base, size_t count, cudaStream_t) override { cb.allgather(cb.ctx, base, count); }
    void alltoall(const double* send, double* recv, size_t count, cudaStream_t) override {
        // test transport: everybody gathers everybody's send buffer and keeps its own blocks
        const size_t per = count * (size_t)nranks;
        std::vector<double> all(per * (size_t)nranks);
        std::memcpy(all.data() + per * (size_t)rank, send, per * sizeof(double));
        cb.allgather(cb.ctx, all.data(), per);
        for (int s = 0; s < nranks; ++s)
            std::memcpy(recv + (size_t)s * count, all.data() + per * (size_t)s + count * (size_t)rank, count * sizeof(double));
    }
};
}  // namespace
Comm* make_callback_comm(int rank, int nranks, const cfdnn_comm_callbacks_t& cb) {
    auto* c = new CallbackComm();
    c->rank = rank; c->nranks = nranks; c->cb = cb;
    return c;
}
#endif

}  // namespace cfdnn
