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
struct Fiber {
    ucontext_t ctx;
    char* stack = nullptr;
    bool done = true;
    int wait = 0;          // 0 runnable, 1 block barrier, 2 warp barrier
    unsigned gen = 0;
    uint3 tid;
};
std::vector<Fiber> fibers;
ucontext_t main_ctx;
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
    swapcontext(&f.ctx, &main_ctx);
}
void yield() {
    Fiber& f = fibers[cur];
    swapcontext(&f.ctx, &main_ctx);
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
            getcontext(&f.ctx);
            f.ctx.uc_stack.ss_sp = f.stack;
            f.ctx.uc_stack.ss_size = kStack;
            f.ctx.uc_link = nullptr;
            makecontext(&f.ctx, (void (*)())trampoline, 0);
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
                swapcontext(&main_ctx, &f.ctx);
                progressed = true;
            }
            if (!progressed) { std::fprintf(stderr, "devsim: deadlock in cooperative kernel\n"); std::abort(); }
        }
      }
    g_in_coop = false;
    cur = -1;
}

}  // namespace devsim
