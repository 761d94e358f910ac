// Dev tool: L1 data-pipe cost of warp-uniform (broadcast) loads on B200: LDG.128 / LDG.64 from global (L1 hit),
// LDS.128 / LDS.64 from shared, generic LD.128 of a shared address.  The trace kernel reads 288 B of per-cell
// coefficients per step; ncu shows its LDG.128s cost 4 data-pipe wavefronts each even when all 32 lanes read the
// same 16 bytes.  nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o lsu_microbench tools/lsu_microbench.cu
#include <cstdio>
#include <cuda_runtime.h>

// MODE 0: LDG.128 uniform address; 1: LDG.64 uniform; 2: LDS.128 uniform; 3: LDS.64 uniform; 4: generic LD.128 (shared);
// 5: LDG.128 distinct 16-B per lane (coalesced 512 B); 6: LDS.128 distinct per lane
template <int MODE>
__global__ void loads(const double2 *g, double *out, int iters)
{
    __shared__ double2 sm[1024];
    for (int k = threadIdx.x; k < 1024; k += blockDim.x) sm[k] = g[k];
    __syncthreads();
    double acc = 0;
    const int lane_off = (MODE >= 5) ? (threadIdx.x & 31) : 0;
    unsigned idx = blockIdx.x & 7;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            const unsigned j = ((idx + u * 37) & 511) + lane_off;
            if (MODE == 0 || MODE == 5) { const double2 v = __ldg(g + j); acc += v.x + v.y; }
            if (MODE == 1) { acc += __ldg(&g[j].x); }
            if (MODE == 2 || MODE == 6) { const double2 v = sm[j]; acc += v.x + v.y; }
            if (MODE == 3) { acc += sm[j].x; }
            if (MODE == 4) {
                const double2 *p = sm + j; // generic pointer into the shared window
                double2 v;
                asm volatile("ld.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
                acc += v.x + v.y;
            }
        }
        idx = (idx * 5 + 1) & 511;
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

int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const double clk = p.clockRate * 1e3; const int sms = p.multiProcessorCount;
    double2 *g; double *o;
    cudaMalloc(&g, 2048 * sizeof(double2)); cudaMemset(g, 0, 2048 * sizeof(double2)); cudaMalloc(&o, 8);
    const int iters = 4096, threads = 512; // 16 warps per SM
    const char *nm[7] = {"LDG.128 uniform address", "LDG.64 uniform address", "LDS.128 uniform address", "LDS.64 uniform address",
                         "generic LD.128 (shared) uniform", "LDG.128 distinct per lane", "LDS.128 distinct per lane"};
    float t[7];
    t[0] = time_ms([&] { loads<0><<<sms, threads>>>(g, o, iters); });
    t[1] = time_ms([&] { loads<1><<<sms, threads>>>(g, o, iters); });
    t[2] = time_ms([&] { loads<2><<<sms, threads>>>(g, o, iters); });
    t[3] = time_ms([&] { loads<3><<<sms, threads>>>(g, o, iters); });
    t[4] = time_ms([&] { loads<4><<<sms, threads>>>(g, o, iters); });
    t[5] = time_ms([&] { loads<5><<<sms, threads>>>(g, o, iters); });
    t[6] = time_ms([&] { loads<6><<<sms, threads>>>(g, o, iters); });
    for (int k = 0; k < 7; ++k)
        printf("%-34s %.2f SM-cycles per warp-level load instruction\n", nm[k], t[k] * 1e-3 * clk / (iters * 16.0 * (threads / 32)));
    return 0;
}
