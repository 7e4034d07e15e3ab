/*
 * cuda_on_host.h -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Lets the *device code* of the reference's three CUDA extensions
 *   imaginaire/third_party/resample2d/src/resample2d_kernel.cu:15-203
 *   imaginaire/third_party/channelnorm/src/channelnorm_kernel.cu:18-96
 *   imaginaire/third_party/correlation/src/correlation_cuda_kernel.cu:16-333
 * compile with g++ and run on host cores, so that the reference's own kernel bodies (not a
 * restatement) produce the golden vectors for those ops.  oracle/ref_cuda/build_ref.py reads the
 * sources where they lie under /root/reference, drops their host launchers (the `<<< >>>` syntax and
 * the at::Tensor plumbing are not C++), and compiles the rest against this header into oracle/_ref/.
 *
 * Execution model: blocks run one after another.  Threads of a block are cooperative fibers
 * (ucontext): a fiber runs until it reaches a block-wide or warp-wide synchronisation point and then
 * hands over to the next one, so one pass over all fibers advances every thread to the same barrier.
 * That is exact for code whose barriers are reached by every thread of the block, which holds for
 * these kernels (their early returns depend on block indices only).  Kernels without any
 * synchronisation can be run with fibers = false: plain nested loops.
 *   atomicAdd        -> `+=` (threads never run concurrently; accumulation order = thread order)
 *   __shfl_down_sync -> exchange through a 32-entry array with two hand-overs (all lanes of the warp
 *                       take part, as the full mask says)
 *   __shared__       -> static storage (one block is resident at a time)
 */
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <ucontext.h>
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __shared__ static
#define __launch_bounds__(...)

struct long4 { long x, y, z, w; };
static inline long4 make_long4(long x, long y, long z, long w) { return long4{x, y, z, w}; }
struct uint3_ { unsigned x, y, z; };
struct dim3 {
    unsigned x, y, z;
    dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
typedef void* cudaStream_t;

namespace cuda_on_host {
inline uint3_ g_blockIdx, g_threadIdx;
inline dim3 g_blockDim, g_gridDim;
inline ucontext_t g_sched;
inline ucontext_t* g_current = nullptr;
inline float g_xchg_f[1024];

inline void yield_to_scheduler() {
    if (!g_current) { std::fprintf(stderr, "cuda_on_host: synchronisation outside a fiber launch\n"); std::abort(); }
    swapcontext(g_current, &g_sched);
}

struct Fiber {
    ucontext_t ctx;
    std::vector<char> stack;
    bool done = false;
    const std::function<void()>* body = nullptr;
};
inline Fiber* g_starting = nullptr;
inline void fiber_entry() {
    Fiber* f = g_starting;
    (*f->body)();
    f->done = true;
    swapcontext(&f->ctx, &g_sched);
}

// Runs `kernel` for every thread of every block of the launch.
inline void launch(dim3 grid, dim3 block, bool fibers, const std::function<void()>& kernel) {
    g_gridDim = grid;
    g_blockDim = block;
    const unsigned nthreads = block.x * block.y * block.z;
    std::vector<Fiber> fb(fibers ? nthreads : 0);
    for (auto& f : fb) f.stack.resize(256 * 1024);
    for (unsigned bz = 0; bz < grid.z; ++bz)
        for (unsigned by = 0; by < grid.y; ++by)
            for (unsigned bx = 0; bx < grid.x; ++bx) {
                g_blockIdx = uint3_{bx, by, bz};
                if (!fibers) {
                    for (unsigned t = 0; t < nthreads; ++t) {
                        g_threadIdx = uint3_{t % block.x, (t / block.x) % block.y, t / (block.x * block.y)};
                        kernel();
                    }
                    continue;
                }
                for (unsigned t = 0; t < nthreads; ++t) {
                    Fiber& f = fb[t];
                    f.done = false;
                    f.body = &kernel;
                    getcontext(&f.ctx);
                    f.ctx.uc_stack.ss_sp = f.stack.data();
                    f.ctx.uc_stack.ss_size = f.stack.size();
                    f.ctx.uc_link = nullptr;
                    makecontext(&f.ctx, fiber_entry, 0);
                }
                bool any = true;
                while (any) {       // one sweep = every live thread runs up to its next synchronisation point
                    any = false;
                    for (unsigned t = 0; t < nthreads; ++t) {
                        Fiber& f = fb[t];
                        if (f.done) continue;
                        g_threadIdx = uint3_{t % block.x, (t / block.x) % block.y, t / (block.x * block.y)};
                        g_current = &f.ctx;
                        g_starting = &f;
                        swapcontext(&g_sched, &f.ctx);
                        g_current = nullptr;
                        any = any || !f.done;
                    }
                }
            }
}
}  // namespace cuda_on_host

#define blockIdx (cuda_on_host::g_blockIdx)
#define threadIdx (cuda_on_host::g_threadIdx)
#define blockDim (cuda_on_host::g_blockDim)
#define gridDim (cuda_on_host::g_gridDim)
static const int warpSize = 32;

static inline void __syncthreads() { cuda_on_host::yield_to_scheduler(); }
static inline void __syncwarp(unsigned = 0xffffffffu) { cuda_on_host::yield_to_scheduler(); }
static inline float __shfl_down_sync(unsigned, float val, int offset) {
    using namespace cuda_on_host;
    const unsigned t = g_threadIdx.x;
    const unsigned lane = t % 32;
    g_xchg_f[t] = val;
    yield_to_scheduler();
    const float r = (lane + offset < 32 && t + offset < g_blockDim.x) ? g_xchg_f[t + offset] : val;
    yield_to_scheduler();
    return r;
}
// not inlined: an atomic is its own instruction on the GPU, its addend is never fused with the product that feeds it
__attribute__((noinline)) static float atomicAdd(float* addr, float v) { const float old = *addr; *addr = old + v; return old; }
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }
using std::floor;
using std::sqrt;

namespace at { struct Half {}; struct Tensor; }
