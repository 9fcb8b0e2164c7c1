// cuda_emu.h — TEST HARNESS ONLY.  A minimal CUDA-runtime + kernel-execution shim so that
// phylo-rs_b200/csrc/*.cu can be compiled with g++ and driven by `pytest -m "not gpu"` in a
// container that has no GPU.  Every CUDA thread of a block is a ucontext fiber; __syncthreads()
// yields to a round-robin scheduler, which gives barrier semantics; warp shuffles go through a
// block-wide exchange buffer.  Blocks run one after another on the calling OS thread.
//
// This is NOT a product code path: the emulation library lives under tests/emu/, is never loaded
// by the phylo_rs_b200 package, and exists to check scheduler / tiling / indexing logic (also under
// AddressSanitizer) before GPU time is spent.  Numerical parity claims are made only by the
// `-m gpu` tests, which call the real sm_100a library.
#pragma once
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline __attribute__((always_inline))
#define __launch_bounds__(...)
#define __shared__ static
#define __align__(n) __attribute__((aligned(n)))

struct dim3 {
    unsigned x, y, z;
    dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct alignas(16) double2 { double x, y; };
struct uchar2 { unsigned char x, y; };
struct alignas(16) int4 { int x, y, z, w; };
struct alignas(16) uint4 { unsigned int x, y, z, w; };
struct alignas(8) int2 { int x, y; };
struct uchar4 { unsigned char x, y, z, w; };
struct uint2 { unsigned int x, y; };
static inline double2 make_double2(double x, double y) { double2 r; r.x = x; r.y = y; return r; }
static inline int2 make_int2(int x, int y) { int2 r; r.x = x; r.y = y; return r; }

namespace emu {
extern dim3 threadIdx, blockIdx, blockDim, gridDim;
extern unsigned char *dyn_smem; // dynamic shared memory of the running block
void launch(dim3 grid, dim3 block, const std::function<void()> &body, size_t dyn_smem_bytes = 0);
void yield();                                  // let the other fibers of the block run
void syncthreads();                            // counting barrier over all live threads of the block
void named_barrier(int id, unsigned count);    // bar.sync id, count
void syncwarp();
double shfl_down(double v, unsigned delta);    // warp-level
double shfl_idx(double v, unsigned src_lane);  // warp-level
// mbarrier / bulk-copy stand-ins (product: PTX mbarrier.* and cp.async.bulk, see csrc/sweep.cuh)
struct Mbar { int count, pending; long long tx; int phase; };
inline void mbar_check(Mbar *b) { if (b->pending == 0 && b->tx == 0) { b->phase ^= 1; b->pending = b->count; } }
} // namespace emu
using emu::blockDim;
using emu::blockIdx;
using emu::gridDim;
using emu::threadIdx;

static inline void __syncthreads() { emu::syncthreads(); }
static inline void __syncwarp(unsigned = 0xffffffffu) { emu::syncwarp(); }
static inline double __shfl_sync(unsigned, double v, int src) { return emu::shfl_idx(v, (unsigned)src); }
static inline int __ffs(int v) { return __builtin_ffs(v); }
static inline int __double2hiint(double d) { long long v; memcpy(&v, &d, 8); return (int)(v >> 32); }
static inline double __shfl_down_sync(unsigned, double v, unsigned delta) { return emu::shfl_down(v, delta); }
template <typename T> static inline T __ldg(const T *p) { return *p; }
static inline int __popc(unsigned v) { return __builtin_popcount(v); }
// round-to-nearest product / sum that the device compiler may not contract (the harness is built with -ffp-contract=off)
static inline double __dmul_rn(double a, double b) { return a * b; }
static inline double __dadd_rn(double a, double b) { return a + b; }
static inline double __longlong_as_double(long long v) { double d; memcpy(&d, &v, 8); return d; }
static inline int atomicAdd(int *p, int v) { int o = *p; *p += v; return o; }
static inline unsigned atomicOr(unsigned *p, unsigned v) { unsigned o = *p; *p |= v; return o; }
static inline int atomicAnd(int *p, int v) { int o = *p; *p &= v; return o; }
using std::isfinite;
using std::max;
using std::min;

// ---- runtime API subset -------------------------------------------------------------------------
typedef int cudaError_t;
enum { cudaSuccess = 0, cudaErrorMemoryAllocation = 2, cudaErrorInvalidValue = 1, cudaErrorInvalidSymbol = 13, cudaErrorInvalidDevice = 101 };
typedef struct emu_stream *cudaStream_t;
typedef struct emu_event *cudaEvent_t;
enum cudaMemcpyKind { cudaMemcpyHostToDevice = 1, cudaMemcpyDeviceToHost = 2, cudaMemcpyDeviceToDevice = 3 };
enum { cudaStreamNonBlocking = 1, cudaHostAllocDefault = 0, cudaDevAttrComputeCapabilityMajor = 75,
       cudaDevAttrMultiProcessorCount = 16, cudaDevAttrMaxSharedMemoryPerBlockOptin = 97 };

cudaError_t cudaMalloc(void **p, size_t bytes);
// emulated device memory is host memory: a __device__ array's address is the array itself
template <typename T>
inline cudaError_t cudaGetSymbolAddress(void **p, const T &symbol)
{
    *p = const_cast<void *>(static_cast<const void *>(&symbol));
    return cudaSuccess;
}
cudaError_t cudaFree(void *p);
cudaError_t cudaHostAlloc(void **p, size_t bytes, unsigned flags);
cudaError_t cudaFreeHost(void *p);
cudaError_t cudaMemcpyAsync(void *dst, const void *src, size_t n, cudaMemcpyKind, cudaStream_t);
cudaError_t cudaMemcpy2DAsync(void *dst, size_t dpitch, const void *src, size_t spitch, size_t width, size_t height,
                              cudaMemcpyKind, cudaStream_t);
cudaError_t cudaMemsetAsync(void *p, int v, size_t n, cudaStream_t);
cudaError_t cudaMemGetInfo(size_t *free_b, size_t *total_b);
cudaError_t cudaGetDeviceCount(int *n);
cudaError_t cudaSetDevice(int d);
cudaError_t cudaDeviceGetAttribute(int *v, int attr, int dev);
cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, unsigned flags);
cudaError_t cudaStreamDestroy(cudaStream_t s);
cudaError_t cudaStreamSynchronize(cudaStream_t s);
cudaError_t cudaEventCreate(cudaEvent_t *e);
cudaError_t cudaEventDestroy(cudaEvent_t e);
cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t s);
cudaError_t cudaStreamWaitEvent(cudaStream_t s, cudaEvent_t e, unsigned flags = 0);
inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; } // the emulation executes in issue order
cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t a, cudaEvent_t b);
cudaError_t cudaGetLastError();
const char *cudaGetErrorString(cudaError_t e);
