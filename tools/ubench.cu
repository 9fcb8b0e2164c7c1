// tools/ubench.cu — sm_100a micro-benchmarks behind the design numbers in DESIGN.md §4.1 (not product code):
//   * DFMA throughput per SM as a function of resident warps and independent chains per thread
//   * DFMA dependent-issue latency
//   * LDS.128 warp-broadcast and 4-row gather throughput (the P(t) block / tip-table access patterns)
//   * the op-loop mix (8 broadcast LDS.128 + 64 DFMA + 16 DMUL per iteration) at 2, 2.5 and 3 warps per scheduler
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench tools/ubench.cu ; run under gpurun.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

template <int ILP>
__global__ void dfma_kernel(double *out, int iters, double b, double c)
{
    double a[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) a[i] = threadIdx.x * 1e-3 + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) a[i] = a[i] * b + c;
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// broadcast: every lane reads the same 16 bytes; gather: lane reads row (lane & 3) of 32-byte rows
template <int MODE>
__global__ void lds_kernel(double *out, int iters)
{
    __shared__ double2 buf[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) buf[i] = make_double2(i, 1.0);
    __syncthreads();
    double2 acc = make_double2(0, 0);
    int base = MODE == 0 ? 0 : 2 * (threadIdx.x & 3);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            double2 v = buf[(base + 8 * k + (it & 1)) & 1023];
            acc.x += v.x; // keep the loads alive without FP64 pressure mattering: 2 DADD per LDS.128
            acc.y += v.y;
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc.x + acc.y;
}

// gather patterns of the tip-table lookups: every lane picks one of NROWS rows (pseudo-random, fixed per lane and k);
// W = 8: rows of 8 bytes read with LDS.64, W = 16: rows of 16 bytes read with LDS.128.  Reports SM clocks per warp
// instruction, i.e. the cost of delivering W bytes to each lane.
template <int W, int NROWS>
__global__ void gather_kernel(double *out, int iters)
{
    __shared__ double2 buf[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) buf[i] = make_double2(i, 1.0);
    __syncthreads();
    unsigned off[16];
    unsigned h = threadIdx.x * 2654435761u + 12345u;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        h = h * 1664525u + 1013904223u;
        off[k] = ((h >> 16) % NROWS) * W + k * 16 * ((NROWS * W + 15) / 16);
    }
    double acc = 0, acc2 = 0;
    const char *base = reinterpret_cast<const char *>(buf);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const unsigned o = (off[k] + (it & 1) * 4096u) & 16383u;
            if (W == 8) {
                acc += *reinterpret_cast<const double *>(base + o);
            } else {
                double2 v = *reinterpret_cast<const double2 *>(base + o);
                acc += v.x;
                acc2 += v.y;
            }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc + acc2;
}

// the op-loop mix: per iteration 8 broadcast LDS.128 (a 4x4 matrix), 4 sites x (16 FMA + 4 MUL)
template <int R>
__global__ void mix_kernel(double *out, int iters)
{
    __shared__ double2 P[8 * 64];
    for (int i = threadIdx.x; i < 8 * 64; i += blockDim.x) P[i] = make_double2(0.25 + 1e-9 * i, 0.25);
    __syncthreads();
    double v[R][4];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int i = 0; i < 4; ++i) v[r][i] = 1.0 + 0.01 * i + threadIdx.x * 1e-6;
    for (int it = 0; it < iters; ++it) {
        const double2 *q = P + 8 * (it & 63);
        double p[16];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            double2 t = q[i];
            p[2 * i] = t.x;
            p[2 * i + 1] = t.y;
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            double y[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) y[i] = p[4 * i] * v[r][0] + p[4 * i + 1] * v[r][1] + p[4 * i + 2] * v[r][2] + p[4 * i + 3] * v[r][3];
#pragma unroll
            for (int i = 0; i < 4; ++i) v[r][i] = y[i] * 1.0000001;
        }
    }
    double s = 0;
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int i = 0; i < 4; ++i) s += v[r][i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static float time_ms(F launch)
{
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    launch();
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(a));
    launch();
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, a, b));
    CK(cudaGetLastError());
    return ms;
}

int main()
{
    int n_sm = 0, khz = 0;
    CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, 0));
    CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0));
    double *out;
    CK(cudaMalloc(&out, sizeof(double) * n_sm * 2048));
    printf("{\"sms\": %d, \"clock_khz\": %d}\n", n_sm, khz);
    const int iters = 20000;
    const int warps[] = {1, 2, 4, 8, 10, 12, 16, 24, 32};
    for (int w : warps) {
        float m1 = time_ms([&] { dfma_kernel<1><<<n_sm, 32 * w>>>(out, iters, 1.0000001, 1e-9); });
        float m4 = time_ms([&] { dfma_kernel<4><<<n_sm, 32 * w>>>(out, iters, 1.0000001, 1e-9); });
        float m8 = time_ms([&] { dfma_kernel<8><<<n_sm, 32 * w>>>(out, iters, 1.0000001, 1e-9); });
        float m16 = time_ms([&] { dfma_kernel<16><<<n_sm, 32 * w>>>(out, iters, 1.0000001, 1e-9); });
        auto tf = [&](float ms, int ilp) { return 2.0 * iters * ilp * 32.0 * w * n_sm / (ms * 1e-3) / 1e12; };
        // clocks per warp-level DFMA per SM sub-partition (4 schedulers) at the nominal clock
        auto cpi = [&](float ms, int ilp) { return ms * 1e-3 * khz * 1e3 / ((double)iters * ilp * w / 4.0); };
        printf("{\"bench\": \"dfma\", \"warps_per_sm\": %d, \"tflops_ilp1\": %.2f, \"tflops_ilp4\": %.2f, \"tflops_ilp8\": %.2f, \"tflops_ilp16\": %.2f, "
               "\"clk_per_warp_dfma_per_smsp_ilp16\": %.2f, \"dep_latency_clk_ilp1_1warp\": %.1f}\n",
               w, tf(m1, 1), tf(m4, 4), tf(m8, 8), tf(m16, 16), cpi(m16, 16), w == 1 ? m1 * 1e-3 * khz * 1e3 / iters : 0.0);
    }
    for (int w : {4, 8, 12, 16}) {
        float mb = time_ms([&] { lds_kernel<0><<<n_sm, 32 * w>>>(out, iters / 4); });
        float mg = time_ms([&] { lds_kernel<1><<<n_sm, 32 * w>>>(out, iters / 4); });
        auto clk = [&](float ms) { return ms * 1e-3 * khz * 1e3 / ((double)(iters / 4) * 16 * w); }; // SM clocks per warp LDS.128
        printf("{\"bench\": \"lds128\", \"warps_per_sm\": %d, \"clk_per_warp_lds_broadcast\": %.2f, \"clk_per_warp_lds_gather4rows\": %.2f}\n", w, clk(mb), clk(mg));
    }
    for (int w : {4, 12, 16}) {
        auto clk = [&](float ms) { return ms * 1e-3 * khz * 1e3 / ((double)(iters / 4) * 16 * w); };
        float a = time_ms([&] { gather_kernel<8, 5><<<n_sm, 32 * w>>>(out, iters / 4); });
        float b = time_ms([&] { gather_kernel<16, 5><<<n_sm, 32 * w>>>(out, iters / 4); });
        float c = time_ms([&] { gather_kernel<8, 25><<<n_sm, 32 * w>>>(out, iters / 4); });
        float d = time_ms([&] { gather_kernel<16, 25><<<n_sm, 32 * w>>>(out, iters / 4); });
        printf("{\"bench\": \"gather\", \"warps_per_sm\": %d, \"clk_lds64_5rows\": %.2f, \"clk_lds128_5rows\": %.2f, \"clk_lds64_25rows\": %.2f, "
               "\"clk_lds128_25rows\": %.2f}\n", w, clk(a), clk(b), clk(c), clk(d));
    }
    for (int w : {4, 8, 10, 12, 16}) {
        float m4 = time_ms([&] { mix_kernel<4><<<n_sm, 32 * w>>>(out, iters); });
        float m8 = time_ms([&] { mix_kernel<8><<<n_sm, 32 * w>>>(out, iters); });
        auto tf = [&](float ms, int r) { return 2.0 * iters * r * 20.0 * 32.0 * w * n_sm / (ms * 1e-3) / 1e12; };
        printf("{\"bench\": \"mix\", \"warps_per_sm\": %d, \"tflops_r4\": %.2f, \"tflops_r8\": %.2f, \"clk_per_iter_r4\": %.1f, \"clk_per_iter_r8\": %.1f}\n", w,
               tf(m4, 4), tf(m8, 8), m4 * 1e-3 * khz * 1e3 / iters, m8 * 1e-3 * khz * 1e3 / iters);
    }
    return 0;
}
