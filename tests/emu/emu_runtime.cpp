// tests/emu/emu_runtime.cpp — TEST INFRASTRUCTURE ONLY: the fiber scheduler behind tests/emu/shim/cuda_runtime.h.
//
// launch() runs the blocks of a grid one after another.  Inside a block every CUDA thread is a ucontext fiber; a
// fiber runs until it reaches a warp or block barrier (every *_sync intrinsic is one), then the next fiber runs.
// A barrier releases when all lanes that exist in the warp (all threads of the block) have arrived — the semantics
// of converged full-mask code, which is what the render kernels promise.  A lane that exits, or waits at a different
// barrier, while its warp waits is a bug in the kernel: the scheduler detects that no fiber can advance and aborts
// with a message instead of hanging.
// rtemu_set_schedule_seed(s != 0) shuffles the order in which the blocks of a grid run and, at every sweep, the order in
// which the runnable fibers of a block are resumed: atomics land in another order, so a result that silently depends
// on thread order (it must not: on the device there is none) changes with the seed and a test can see it.
#include <ucontext.h>
#include <sys/mman.h>

#include <cstdio>
#include <utility>
#include <mutex>
#include <vector>

#include "shim/cuda_runtime.h"

namespace rtemu {

uint3 g_threadIdx, g_blockIdx;
dim3 g_blockDim, g_gridDim;
static char g_sharedAnchor;
uintptr_t g_sharedBase = (uintptr_t)&g_sharedAnchor;
unsigned long long g_kernelLaunches = 0;
unsigned long long g_scheduleSeed = 0;
unsigned long long g_shuffles = 0;

namespace {

const size_t kStackBytes = 256 * 1024;

struct Fiber {
    ucontext_t ctx;
    char* stack = nullptr;
    bool done = false;
    unsigned parity = 0;       // exchange buffer this lane uses for its current collective
};

struct BlockState {
    std::vector<Fiber> fibers;
    ucontext_t sched;
    unsigned nThreads = 0;
    int current = 0;
    std::vector<int> warpArrived;
    std::vector<unsigned> warpGen;
    int blockArrived = 0;
    unsigned blockGen = 0;
    std::vector<uint64_t> xbuf[2];
    unsigned long long progress = 0;
    const std::function<void()>* body = nullptr;
};

BlockState* g_blk = nullptr;
std::mutex g_launchMutex;
std::vector<char*> g_stackPool;

struct Shuffler {                       // xorshift64*; Fisher-Yates
    unsigned long long s;
    unsigned long long next() { s ^= s >> 12; s ^= s << 25; s ^= s >> 27; return s * 2685821657736338717ULL; }
    void shuffle(std::vector<unsigned>& v) {
        for (size_t i = v.size(); i > 1; i--) std::swap(v[i - 1], v[next() % i]);
    }
};

char* takeStack() {
    if (!g_stackPool.empty()) { char* s = g_stackPool.back(); g_stackPool.pop_back(); return s; }
    void* p = mmap(nullptr, kStackBytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
    if (p == MAP_FAILED) { std::fprintf(stderr, "rtemu: cannot map a fiber stack\n"); std::abort(); }
    return (char*)p;
}

void yieldToScheduler() {
    BlockState& b = *g_blk;
    swapcontext(&b.fibers[b.current].ctx, &b.sched);
}

void fiberEntry() {
    BlockState& b = *g_blk;
    (*b.body)();
    b.fibers[b.current].done = true;
    b.progress++;
    // returning resumes uc_link = the scheduler
}

}  // namespace

int warpLanes() {
    const unsigned w = g_threadIdx.x >> 5;
    const unsigned left = g_blk->nThreads - w * 32;
    return (int)(left < 32 ? left : 32);
}

void nextExchange() { g_blk->fibers[g_blk->current].parity ^= 1u; }

uint64_t* exchangeSlot(int lane) {
    BlockState& b = *g_blk;
    return &b.xbuf[b.fibers[b.current].parity][(g_threadIdx.x & ~31u) + (unsigned)lane];
}

void warpBarrier() {
    BlockState& b = *g_blk;
    const unsigned w = g_threadIdx.x >> 5;
    const unsigned gen = b.warpGen[w];
    b.progress++;
    if (++b.warpArrived[w] == warpLanes()) {
        b.warpArrived[w] = 0;
        b.warpGen[w]++;
        return;
    }
    while (b.warpGen[w] == gen) yieldToScheduler();
}

void blockBarrier() {
    BlockState& b = *g_blk;
    const unsigned gen = b.blockGen;
    b.progress++;
    if (++b.blockArrived == (int)b.nThreads) {
        b.blockArrived = 0;
        b.blockGen++;
        return;
    }
    while (b.blockGen == gen) yieldToScheduler();
}

void launch(dim3 grid, dim3 block, const std::function<void()>& body) {
    std::lock_guard<std::mutex> lock(g_launchMutex);
    g_kernelLaunches++;
    g_gridDim = grid;
    g_blockDim = block;
    const unsigned nThreads = block.x * block.y * block.z;
    Shuffler rng{(g_scheduleSeed + g_kernelLaunches) * 0x9E3779B97F4A7C15ULL | 1ULL};
    std::vector<unsigned> blockOrder(grid.x), threadOrder(nThreads);
    for (unsigned i = 0; i < grid.x; i++) blockOrder[i] = i;
    for (unsigned i = 0; i < nThreads; i++) threadOrder[i] = i;
    if (g_scheduleSeed) rng.shuffle(blockOrder);
    for (unsigned bo = 0; bo < grid.x; bo++) {
        const unsigned bi = blockOrder[bo];
        BlockState b;
        b.nThreads = nThreads;
        b.fibers.resize(nThreads);
        b.warpArrived.assign((nThreads + 31) / 32, 0);
        b.warpGen.assign((nThreads + 31) / 32, 0);
        b.xbuf[0].assign((nThreads + 31) / 32 * 32, 0);
        b.xbuf[1].assign((nThreads + 31) / 32 * 32, 0);
        b.body = &body;
        g_blk = &b;
        g_blockIdx.x = bi; g_blockIdx.y = g_blockIdx.z = 0;
        for (unsigned t = 0; t < nThreads; t++) {
            Fiber& f = b.fibers[t];
            f.stack = takeStack();
            getcontext(&f.ctx);
            f.ctx.uc_stack.ss_sp = f.stack;
            f.ctx.uc_stack.ss_size = kStackBytes;
            f.ctx.uc_link = &b.sched;
            makecontext(&f.ctx, (void (*)())fiberEntry, 0);
        }
        for (;;) {
            bool alive = false;
            const unsigned long long before = b.progress;
            if (g_scheduleSeed) { rng.shuffle(threadOrder); g_shuffles++; }
            for (unsigned to = 0; to < nThreads; to++) {
                const unsigned t = threadOrder[to];
                if (b.fibers[t].done) continue;
                alive = true;
                b.current = (int)t;
                g_threadIdx.x = t; g_threadIdx.y = g_threadIdx.z = 0;
                swapcontext(&b.sched, &b.fibers[t].ctx);
            }
            if (!alive) break;
            if (b.progress == before) {
                std::fprintf(stderr, "rtemu: deadlock in block %u — lanes of a warp wait at different barriers or some exited "
                                     "(divergent use of a full-mask *_sync intrinsic)\n", bi);
                std::abort();
            }
        }
        for (Fiber& f : b.fibers) g_stackPool.push_back(f.stack);
        g_blk = nullptr;
    }
}

}  // namespace rtemu

extern "C" unsigned long long rtemu_kernel_launches(void) { return rtemu::g_kernelLaunches; }
extern "C" unsigned long long rtemu_schedule_shuffles(void) { return rtemu::g_shuffles; }
extern "C" void rtemu_set_schedule_seed(unsigned long long seed) { rtemu::g_scheduleSeed = seed; }
