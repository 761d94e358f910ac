// Dev tool: L1 data-pipe cost of vector loads whose 32 lanes read D distinct chunks (lanes in contiguous groups, the chunk
// of group q at q * 288 B -- a warp whose rays sit in D different field cells reading one entry of each cell's record):
// LDG.128, LDG.256 (sm_100: ld.global.v4.f64) and LDS.128.  Cycles per warp-level load and per 16 bytes per lane.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o lsu_groups tools/lsu_groups_microbench.cu
#include <cstdio>
#include <cuda_runtime.h>

// MODE 0: LDG.128, 1: LDG.256, 2: LDS.128
template <int MODE, int D>
__global__ void loads(const double2 *g, double *out, int iters)
{
    __shared__ double2 sm[2048];
    for (int k = threadIdx.x; k < 2048; k += blockDim.x) sm[k] = g[k];
    __syncthreads();
    double acc = 0;
    const int grp = ((threadIdx.x & 31) * D) >> 5;
    unsigned idx = blockIdx.x & 7;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            const unsigned j = (((idx + u) & 31) << (MODE == 1 ? 1 : 0)) + grp * 18 * 3; // records 3 apart: 864 B
            if (MODE == 0) { const double2 v = __ldg(g + j); acc += v.x + v.y; }
            if (MODE == 2) { const double2 v = sm[j & 2047]; acc += v.x + v.y; }
            if (MODE == 1) {
                double a, b, c, d;
                asm volatile("ld.global.nc.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(g + j));
                acc += (a + b) + (c + d);
            }
        }
        idx = (idx * 5 + 1) & 31;
    }
    if (acc == 12345.678) out[0] = acc;
}

template <class F> static float time_ms(F launch)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(); cudaDeviceSynchronize();
    cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    return ms;
}

template <int MODE, int D> static void run(const double2 *g, double *o, int sms, double clk)
{
    const int iters = 4096, threads = 512;
    const float t = time_ms([&] { loads<MODE, D><<<sms, threads>>>(g, o, iters); });
    const double cyc = t * 1e-3 * clk / (iters * 16.0 * (threads / 32));
    const char *nm[3] = {"LDG.128", "LDG.256", "LDS.128"};
    printf("%s  D=%2d: %.2f SM-cycles per warp-level load, %.2f per 16 B per lane\n", nm[MODE], D, cyc, MODE == 1 ? cyc / 2 : cyc);
}

int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const double clk = p.clockRate * 1e3; const int sms = p.multiProcessorCount;
    double2 *g; double *o;
    cudaMalloc(&g, 8192 * sizeof(double2)); cudaMemset(g, 0, 8192 * sizeof(double2)); cudaMalloc(&o, 8);
    run<0, 1>(g, o, sms, clk); run<0, 2>(g, o, sms, clk); run<0, 4>(g, o, sms, clk); run<0, 8>(g, o, sms, clk);
    run<0, 16>(g, o, sms, clk); run<0, 32>(g, o, sms, clk);
    run<1, 1>(g, o, sms, clk); run<1, 2>(g, o, sms, clk); run<1, 4>(g, o, sms, clk); run<1, 8>(g, o, sms, clk);
    run<1, 16>(g, o, sms, clk); run<1, 32>(g, o, sms, clk);
    run<2, 1>(g, o, sms, clk); run<2, 4>(g, o, sms, clk); run<2, 16>(g, o, sms, clk); run<2, 32>(g, o, sms, clk);
    return 0;
}
