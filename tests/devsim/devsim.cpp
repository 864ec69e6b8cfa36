// tests/devsim/devsim.cpp — fiber scheduler of the CUDA test double (see cuda_runtime.h).
// TEST INFRASTRUCTURE; not linked into the product library.
#include "cuda_runtime.h"

namespace devsim {

uint3 g_tid{0, 0, 0}, g_bid{0, 0, 0};
dim3 g_bdim, g_gdim;
void* g_dyn_smem = nullptr;
bool g_in_coop = false;

namespace {
constexpr size_t kStack = 256 * 1024;

// Context switch between the scheduler and a fiber.  glibc's swapcontext makes a sigprocmask system call on every
// switch (minutes of system time over the CPU suite); on x86-64 a switch only has to save the callee-saved registers
// and the stack pointer, so it is done by hand there.  Elsewhere ucontext is used as before.
#if defined(__x86_64__)
#define DEVSIM_ASM_SWITCH 1
struct Ctx { void* sp = nullptr; };
extern "C" __attribute__((naked, noinline)) void devsim_switch(void** /*save_sp: rdi*/, void* /*to_sp: rsi*/) {
    __asm__ volatile(
        "pushq %rbp\n\t pushq %rbx\n\t pushq %r12\n\t pushq %r13\n\t pushq %r14\n\t pushq %r15\n\t"
        "movq %rsp, (%rdi)\n\t"
        "movq %rsi, %rsp\n\t"
        "popq %r15\n\t popq %r14\n\t popq %r13\n\t popq %r12\n\t popq %rbx\n\t popq %rbp\n\t"
        "ret\n\t");
}
inline void ctx_switch(Ctx& from, Ctx& to) { devsim_switch(&from.sp, to.sp); }
// a fresh fiber: six zeroed callee-saved registers, then the entry point as the return address; the stack pointer at
// entry is 8 modulo 16 as after a call
inline void ctx_make(Ctx& c, char* stack, size_t size, void (*entry)()) {
    uintptr_t top = (reinterpret_cast<uintptr_t>(stack) + size) & ~uintptr_t(15);
    void** sp = reinterpret_cast<void**>(top - 64);
    for (int i = 0; i < 6; ++i) sp[i] = nullptr;
    sp[6] = reinterpret_cast<void*>(entry);
    sp[7] = nullptr;                                  // the entry never returns
    c.sp = sp;
}
#else
#define DEVSIM_ASM_SWITCH 0
struct Ctx { ucontext_t uc; };
inline void ctx_switch(Ctx& from, Ctx& to) { swapcontext(&from.uc, &to.uc); }
inline void ctx_make(Ctx& c, char* stack, size_t size, void (*entry)()) {
    getcontext(&c.uc);
    c.uc.uc_stack.ss_sp = stack;
    c.uc.uc_stack.ss_size = size;
    c.uc.uc_link = nullptr;
    makecontext(&c.uc, entry, 0);
}
#endif

struct Fiber {
    Ctx ctx;
    char* stack = nullptr;
    bool done = true;
    int wait = 0;          // 0 runnable, 1 block barrier, 2 warp barrier
    unsigned gen = 0;
    uint3 tid;
};
std::vector<Fiber> fibers;
Ctx main_ctx;
int cur = -1, n_threads = 0, alive = 0;
unsigned block_gen = 0;
int block_arrived = 0;
struct WarpState { unsigned gen = 0; int arrived = 0, alive = 0; uint64_t slot[32]; };
std::vector<WarpState> warps;
const std::function<void()>* cur_body = nullptr;

void release_block_if_ready() {
    if (alive > 0 && block_arrived >= alive) { block_arrived = 0; ++block_gen; }
}
void release_warp_if_ready(WarpState& w) {
    if (w.alive > 0 && w.arrived >= w.alive) { w.arrived = 0; ++w.gen; }
}
void trampoline() {
    (*cur_body)();
    Fiber& f = fibers[cur];
    f.done = true;
    --alive;
    WarpState& w = warps[cur >> 5];
    --w.alive;
    release_block_if_ready();
    release_warp_if_ready(w);
    ctx_switch(f.ctx, main_ctx);
    std::abort();                                     // a finished fiber is never resumed
}
void yield() {
    Fiber& f = fibers[cur];
    ctx_switch(f.ctx, main_ctx);
    g_tid = f.tid;
}
}  // namespace

int lane_id() {
    if (!g_in_coop) { std::fprintf(stderr, "devsim: warp op in a plain launch (use CFDNN_LAUNCH_COOP)\n"); std::abort(); }
    return cur & 31;
}
uint64_t* warp_slot(int lane) { return &warps[cur >> 5].slot[lane]; }

void sync_block() {
    if (!g_in_coop) { std::fprintf(stderr, "devsim: __syncthreads in a plain launch (use CFDNN_LAUNCH_COOP)\n"); std::abort(); }
    Fiber& f = fibers[cur];
    const unsigned my = block_gen;
    ++block_arrived;
    release_block_if_ready();
    if (block_gen != my) return;
    f.wait = 1; f.gen = my;
    yield();
}
void sync_warp() {
    if (!g_in_coop) { std::fprintf(stderr, "devsim: warp sync in a plain launch\n"); std::abort(); }
    Fiber& f = fibers[cur];
    WarpState& w = warps[cur >> 5];
    const unsigned my = w.gen;
    ++w.arrived;
    release_warp_if_ready(w);
    if (w.gen != my) return;
    f.wait = 2; f.gen = my;
    yield();
}

void launch_coop(dim3 grid, dim3 block, size_t smem, const std::function<void()>& body) {
    static std::vector<char> dyn;
    if (dyn.size() < smem) dyn.resize(smem);
    g_dyn_smem = dyn.data();
    g_gdim = grid; g_bdim = block;
    n_threads = (int)(block.x * block.y * block.z);
    if ((int)fibers.size() < n_threads) {
        size_t old = fibers.size();
        fibers.resize(n_threads);
        for (size_t t = old; t < fibers.size(); ++t) fibers[t].stack = (char*)std::malloc(kStack);
    }
    warps.assign((n_threads + 31) / 32, WarpState());
    cur_body = &body;
    g_in_coop = true;
    for (unsigned bz = 0; bz < grid.z; ++bz)
     for (unsigned by = 0; by < grid.y; ++by)
      for (unsigned bx = 0; bx < grid.x; ++bx) {
        g_bid = uint3{bx, by, bz};
        alive = n_threads; block_arrived = 0;
        for (auto& w : warps) { w.arrived = 0; w.alive = 0; }
        int t = 0;
        for (unsigned tz = 0; tz < block.z; ++tz)
         for (unsigned ty = 0; ty < block.y; ++ty)
          for (unsigned tx = 0; tx < block.x; ++tx, ++t) {
            Fiber& f = fibers[t];
            f.done = false; f.wait = 0; f.tid = uint3{tx, ty, tz};
            ctx_make(f.ctx, f.stack, kStack, trampoline);
            ++warps[t >> 5].alive;
          }
        while (alive > 0) {
            bool progressed = false;
            for (int i = 0; i < n_threads; ++i) {
                Fiber& f = fibers[i];
                if (f.done) continue;
                if (f.wait == 1 && block_gen == f.gen) continue;
                if (f.wait == 2 && warps[i >> 5].gen == f.gen) continue;
                f.wait = 0;
                cur = i;
                g_tid = f.tid;
                ctx_switch(main_ctx, f.ctx);
                progressed = true;
            }
            if (!progressed) { std::fprintf(stderr, "devsim: deadlock in cooperative kernel\n"); std::abort(); }
        }
      }
    g_in_coop = false;
    cur = -1;
}

}  // namespace devsim
