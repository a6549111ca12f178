// cuemu.cpp -- fiber scheduler behind tools/cuemu/cuemu.h (development aid only, see header).
#include "cuemu.h"
#include <stdio.h>
#include <time.h>
#include <ucontext.h>
#include <map>
#include <vector>

cuemu_dim3 threadIdx, blockIdx, blockDim, gridDim;
unsigned char *cuemu_dyn_smem_ptr = nullptr;

namespace {
const size_t kStack = 192 * 1024;
struct Fiber { ucontext_t ctx; unsigned char *stack; bool done; dim3 tid; };
std::vector<Fiber> fibers;
std::vector<unsigned char *> stack_pool;
ucontext_t main_ctx;
int cur = -1, nthreads = 0, nalive = 0;
const std::function<void()> *g_body = nullptr;

// CTA barrier
int bar_arrived = 0; unsigned bar_gen = 0; int bar_acc[2] = {0, 0};
// per-warp rendezvous
struct WarpSync { int arrived; unsigned gen; uint64_t slot[2][32]; unsigned ballot[2]; };
// one rendezvous state per (warp, participant mask): groups of lanes with different masks may be
// waiting at different rendezvous at the same time
std::vector<std::map<unsigned, WarpSync>> warps;

int linear_tid(const dim3 &t) { return (int)(t.x + blockDim.x * (t.y + blockDim.y * t.z)); }

void fiber_entry()
{
    (*g_body)();
    fibers[cur].done = true;
    nalive--;
    // a finished thread counts as arrived for any barrier still pending
    swapcontext(&fibers[cur].ctx, &main_ctx);
}

void yield_fiber() { swapcontext(&fibers[cur].ctx, &main_ctx); }

int alive_in_warp(int w)
{
    int n = 0;
    for (int l = 0; l < 32; ++l) { int t = w * 32 + l; if (t < nthreads && !fibers[t].done) n++; }
    return n;
}
} // namespace

void cuemu_launch(dim3 grid, dim3 block, size_t dyn_smem, const std::function<void()> &body)
{
    nthreads = (int)(block.x * block.y * block.z);
    gridDim = grid; blockDim = block;
    g_body = &body;
    while ((int)stack_pool.size() < nthreads) stack_pool.push_back((unsigned char *)malloc(kStack));
    unsigned char *smem = (unsigned char *)calloc(dyn_smem + 64, 1);
    cuemu_dyn_smem_ptr = smem;
    for (unsigned bz = 0; bz < grid.z; ++bz)
        for (unsigned by = 0; by < grid.y; ++by)
            for (unsigned bx = 0; bx < grid.x; ++bx) {
                blockIdx = dim3(bx, by, bz);
                fibers.assign((size_t)nthreads, Fiber());
                warps.assign((size_t)(nthreads + 31) / 32, std::map<unsigned, WarpSync>());
                bar_arrived = 0; nalive = nthreads;
                for (int t = 0; t < nthreads; ++t) {
                    Fiber &f = fibers[t];
                    f.done = false; f.stack = stack_pool[t];
                    f.tid = dim3(t % block.x, (t / block.x) % block.y, t / (block.x * block.y));
                    getcontext(&f.ctx);
                    f.ctx.uc_stack.ss_sp = f.stack; f.ctx.uc_stack.ss_size = kStack; f.ctx.uc_link = &main_ctx;
                    makecontext(&f.ctx, fiber_entry, 0);
                }
                while (nalive > 0) {
                    for (int t = 0; t < nthreads; ++t) {
                        if (fibers[t].done) continue;
                        cur = t; threadIdx = fibers[t].tid;
                        swapcontext(&main_ctx, &fibers[t].ctx);
                    }
                }
            }
    free(smem);
    cuemu_dyn_smem_ptr = nullptr;
}

static int cta_barrier(int pred, int mode)
{
    unsigned g = bar_gen;
    if (bar_arrived == 0) bar_acc[g & 1] = (mode == 2) ? 1 : 0;
    if (mode == 0 || mode == 1) bar_acc[g & 1] += pred ? 1 : 0;
    if (mode == 2) bar_acc[g & 1] &= pred ? 1 : 0;
    bar_arrived++;
    for (;;) {
        if (bar_gen != g) break;
        if (bar_arrived >= nalive) { bar_gen++; bar_arrived = 0; break; }
        yield_fiber();
    }
    return bar_acc[g & 1];
}

void __syncthreads() { cta_barrier(0, 0); }
int __syncthreads_or(int pred) { return cta_barrier(pred, 0) != 0; }
int __syncthreads_count(int pred) { return cta_barrier(pred, 1); }
int __syncthreads_and(int pred) { return cta_barrier(pred, 2); }

// warp rendezvous among the lanes named in mask; returns the state and the generation used
static WarpSync &warp_rendezvous(unsigned mask, uint64_t v, int pred, unsigned &g)
{
    int t = linear_tid(threadIdx), w = t / 32, lane = t % 32;
    WarpSync &ws = warps[w][mask];
    g = ws.gen;
    if (ws.arrived == 0) ws.ballot[g & 1] = 0;
    ws.slot[g & 1][lane] = v;
    if (pred) ws.ballot[g & 1] |= 1u << lane;
    ws.arrived++;
    for (;;) {
        int need = 0;
        for (int l = 0; l < 32; ++l) { int tt = w * 32 + l; if ((mask >> l & 1) && tt < nthreads && !fibers[tt].done) need++; }
        if (ws.gen != g) break;
        if (ws.arrived >= need) { ws.gen++; ws.arrived = 0; break; }
        yield_fiber();
    }
    return ws;
}

void __syncwarp(unsigned mask) { unsigned g; warp_rendezvous(mask, 0, 0, g); }
unsigned __ballot_sync(unsigned mask, int pred)
{
    unsigned g;
    WarpSync &ws = warp_rendezvous(mask, 0, pred, g);
    return ws.ballot[g & 1] & mask;
}
int __any_sync(unsigned mask, int pred) { return __ballot_sync(mask, pred) != 0; }
int __all_sync(unsigned mask, int pred) { return __ballot_sync(mask, !pred) == 0; }
uint64_t cuemu_shfl_partial(unsigned mask, uint64_t v, int src)
{
    unsigned g;
    WarpSync &ws = warp_rendezvous(mask, v, 0, g);
    return ws.slot[g & 1][src & 31];
}
uint64_t cuemu_shfl(uint64_t v, int src) { return cuemu_shfl_partial(0xffffffffu, v, src); }

static double now_ms() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; }
cudaError_t cudaEventCreate(cudaEvent_t *e) { *e = (cudaEvent_t)calloc(1, sizeof(cuemu_event)); return 0; }
cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t) { e->t = now_ms(); return 0; }
cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t a, cudaEvent_t b) { *ms = (float)(b->t - a->t); return 0; }
