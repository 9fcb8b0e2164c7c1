// cuda_emu.cpp — TEST HARNESS ONLY (see cuda_emu.h).  Fiber scheduler + trivial runtime shim.
#include "cuda_emu.h"

#include <chrono>
#include <cstdio>
#include <ucontext.h>
#include <vector>

namespace emu {

dim3 threadIdx, blockIdx, blockDim, gridDim;
unsigned char *dyn_smem = nullptr;

namespace {
struct Fiber {
    ucontext_t ctx;
    char *stack = nullptr;
    bool done = false;
    dim3 tid;
};
constexpr size_t STACK_BYTES = 96 * 1024;
ucontext_t g_sched;
std::vector<Fiber> g_fibers;
int g_current = -1;
const std::function<void()> *g_body = nullptr;
std::vector<double> g_xbuf;

void fiber_entry()
{
    (*g_body)();
    g_fibers[g_current].done = true;
    swapcontext(&g_fibers[g_current].ctx, &g_sched);
}
} // namespace

void yield() { swapcontext(&g_fibers[g_current].ctx, &g_sched); }

namespace {
struct Bar { unsigned arrived = 0; unsigned long gen = 0; };
Bar g_bars[16 + 64]; // 0..15: bar.sync ids (0 = __syncthreads), 16..: per-warp barriers
unsigned g_nthreads = 0;
void barrier(Bar &b, unsigned count)
{
    unsigned long gen = b.gen;
    if (++b.arrived >= count) {
        b.arrived = 0;
        ++b.gen;
        return;
    }
    while (b.gen == gen) yield();
}
unsigned lin_tid() { return threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z); }
unsigned warp_size_of(unsigned warp) { unsigned lo = warp * 32; return g_nthreads - lo < 32 ? g_nthreads - lo : 32; }
} // namespace

void syncthreads() { barrier(g_bars[0], g_nthreads); }
void named_barrier(int id, unsigned count) { barrier(g_bars[id & 15], count); }
void syncwarp()
{
    unsigned w = lin_tid() >> 5;
    barrier(g_bars[16 + (w & 63)], warp_size_of(w));
}

double shfl_down(double v, unsigned delta)
{
    unsigned lin = lin_tid();
    g_xbuf[lin] = v;
    syncwarp();
    unsigned lane = lin & 31u, src = lin + delta;
    double r = (lane + delta < 32u && src < g_nthreads) ? g_xbuf[src] : v;
    syncwarp();
    return r;
}

double shfl_idx(double v, unsigned src_lane)
{
    unsigned lin = lin_tid();
    g_xbuf[lin] = v;
    syncwarp();
    unsigned src = (lin & ~31u) + (src_lane & 31u);
    double r = src < g_nthreads ? g_xbuf[src] : v;
    syncwarp();
    return r;
}

void launch(dim3 grid, dim3 block, const std::function<void()> &body, size_t dyn_smem_bytes)
{
    const unsigned nthreads = block.x * block.y * block.z;
    g_nthreads = nthreads;
    static std::vector<unsigned char> smem_buf;
    if (smem_buf.size() < dyn_smem_bytes + 256) smem_buf.resize(dyn_smem_bytes + 256);
    dyn_smem = (unsigned char *)(((uintptr_t)smem_buf.data() + 127) & ~(uintptr_t)127);
    for (auto &b : g_bars) b = Bar();
    if (g_fibers.size() < nthreads) {
        size_t old = g_fibers.size();
        g_fibers.resize(nthreads);
        for (size_t i = old; i < nthreads; ++i) g_fibers[i].stack = (char *)malloc(STACK_BYTES);
    }
    if (g_xbuf.size() < nthreads) g_xbuf.resize(nthreads);
    g_body = &body;
    gridDim = grid;
    blockDim = block;
    for (unsigned bz = 0; bz < grid.z; ++bz)
        for (unsigned by = 0; by < grid.y; ++by)
            for (unsigned bx = 0; bx < grid.x; ++bx) {
                blockIdx = dim3(bx, by, bz);
                unsigned t = 0;
                for (unsigned tz = 0; tz < block.z; ++tz)
                    for (unsigned ty = 0; ty < block.y; ++ty)
                        for (unsigned tx = 0; tx < block.x; ++tx, ++t) {
                            Fiber &f = g_fibers[t];
                            f.done = false;
                            f.tid = dim3(tx, ty, tz);
                            getcontext(&f.ctx);
                            f.ctx.uc_stack.ss_sp = f.stack;
                            f.ctx.uc_stack.ss_size = STACK_BYTES;
                            f.ctx.uc_link = nullptr;
                            makecontext(&f.ctx, fiber_entry, 0);
                        }
                unsigned alive = nthreads;
                while (alive) {
                    for (unsigned i = 0; i < nthreads; ++i) {
                        Fiber &f = g_fibers[i];
                        if (f.done) continue;
                        g_current = (int)i;
                        threadIdx = f.tid;
                        swapcontext(&g_sched, &f.ctx);
                        if (f.done) --alive;
                    }
                }
            }
    g_body = nullptr;
}

} // namespace emu

// ---- runtime shim ---------------------------------------------------------------------------------
struct emu_stream { int dummy; };
struct emu_event { std::chrono::steady_clock::time_point t; };

static size_t g_allocated = 0;
static size_t mem_total()
{
    const char *e = getenv("PHB_EMU_MEM_BYTES");
    return e ? (size_t)strtoull(e, nullptr, 10) : ((size_t)2 << 30);
}

cudaError_t cudaMalloc(void **p, size_t bytes)
{
    if (g_allocated + bytes > mem_total()) return cudaErrorMemoryAllocation;
    size_t total = ((bytes + 255) / 256) * 256 + 256;
    char *raw = (char *)aligned_alloc(256, total);
    if (!raw) return cudaErrorMemoryAllocation;
    *(size_t *)raw = bytes;
    memset(raw + 256, 0xA5, total - 256); // poison: uninitialised reads show up as garbage
    g_allocated += bytes;
    *p = raw + 256;
    return cudaSuccess;
}
cudaError_t cudaFree(void *p)
{
    if (!p) return cudaSuccess;
    char *raw = (char *)p - 256;
    g_allocated -= *(size_t *)raw;
    free(raw);
    return cudaSuccess;
}
cudaError_t cudaHostAlloc(void **p, size_t bytes, unsigned) { *p = malloc(bytes); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
cudaError_t cudaFreeHost(void *p) { free(p); return cudaSuccess; }
cudaError_t cudaMemcpyAsync(void *dst, const void *src, size_t n, cudaMemcpyKind, cudaStream_t) { memcpy(dst, src, n); return cudaSuccess; }
cudaError_t cudaMemcpy2DAsync(void *dst, size_t dpitch, const void *src, size_t spitch, size_t width, size_t height,
                              cudaMemcpyKind, cudaStream_t)
{
    for (size_t r = 0; r < height; ++r) memcpy((char *)dst + r * dpitch, (const char *)src + r * spitch, width);
    return cudaSuccess;
}
cudaError_t cudaMemsetAsync(void *p, int v, size_t n, cudaStream_t) { memset(p, v, n); return cudaSuccess; }
cudaError_t cudaMemGetInfo(size_t *free_b, size_t *total_b)
{
    *total_b = mem_total();
    *free_b = *total_b > g_allocated ? *total_b - g_allocated : 0;
    return cudaSuccess;
}
cudaError_t cudaGetDeviceCount(int *n)
{
    const char *e = getenv("PHB_EMU_DEVICES");
    *n = e ? atoi(e) : 4;
    return cudaSuccess;
}
cudaError_t cudaSetDevice(int) { return cudaSuccess; }
cudaError_t cudaDeviceGetAttribute(int *v, int attr, int)
{
    if (attr == cudaDevAttrMultiProcessorCount) *v = 3;            // few "SMs": several site blocks per CTA
    else if (attr == cudaDevAttrMaxSharedMemoryPerBlockOptin) {
        const char *e = getenv("PHB_EMU_SMEM_BYTES");              // small values force global stack levels
        *v = e ? atoi(e) : 232448;
    } else *v = 10;
    return cudaSuccess;
}
cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, unsigned) { *s = new emu_stream(); return cudaSuccess; }
cudaError_t cudaStreamDestroy(cudaStream_t s) { delete s; return cudaSuccess; }
cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
cudaError_t cudaEventCreate(cudaEvent_t *e) { *e = new emu_event(); return cudaSuccess; }
cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return cudaSuccess; }
cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned) { return cudaSuccess; } // streams are synchronous here
cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t) { e->t = std::chrono::steady_clock::now(); return cudaSuccess; }
cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t a, cudaEvent_t b)
{
    *ms = std::chrono::duration<float, std::milli>(b->t - a->t).count();
    return cudaSuccess;
}
cudaError_t cudaGetLastError() { return cudaSuccess; }
const char *cudaGetErrorString(cudaError_t e) { return e == cudaSuccess ? "success" : (e == cudaErrorMemoryAllocation ? "out of memory" : "error"); }
