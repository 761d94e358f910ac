// Dev tool: FP64 pipe characteristics of the B200 SM that shape the trace kernel's design.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fp64_microbench tools/fp64_microbench.cu
// Prints (1) dependent-DFMA latency, (2) DFMA throughput per SM sub-partition vs warps x ILP,
// (3) relative cost of DMUL / DADD / DSETP / F2I / MUFU.RCP64H, (4) whether integer work issued
// between DFMAs is free (co-issue with the 2-cycle FP64 pipe occupancy).
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP>
__global__ void dfma_chain(double *out, int iters, double a, double b)
{
    double v[ILP];
#pragma unroll
    for (int k = 0; k < ILP; ++k) v[k] = threadIdx.x + k;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u)
#pragma unroll
            for (int k = 0; k < ILP; ++k) v[k] = fma(v[k], a, b);
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < ILP; ++k) s += v[k];
    if (s == 12345.678) out[0] = s;
}

// OP: 0 dfma, 1 dmul, 2 dadd, 3 dsetp(+sel), 4 f2i, 5 mufu.rcp64h, 6 dfma + 1 integer op, 7 dfma + 2 integer ops
template <int OP>
__global__ void op_mix(double *out, int iters, double a, double b)
{
    double v0 = threadIdx.x, v1 = v0 + 1, v2 = v0 + 2, v3 = v0 + 3;
    int n0 = threadIdx.x, n1 = 1, n2 = 2, n3 = 3;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            if (OP == 0) { v0 = fma(v0, a, b); v1 = fma(v1, a, b); v2 = fma(v2, a, b); v3 = fma(v3, a, b); }
            if (OP == 1) { v0 = v0 * a; v1 = v1 * a; v2 = v2 * a; v3 = v3 * a; }
            if (OP == 2) { v0 = v0 + b; v1 = v1 + b; v2 = v2 + b; v3 = v3 + b; }
            if (OP == 3) { n0 += v0 > a; n1 += v1 > b; n2 += v2 > a; n3 += v3 > b; v0 += 0; }
            if (OP == 4) { n0 += __double2int_rd(v0); n1 += __double2int_rd(v1); n2 += __double2int_rd(v2); n3 += __double2int_rd(v3); }
            if (OP == 5) {
                asm volatile("rcp.approx.ftz.f64 %0, %0;" : "+d"(v0)); asm volatile("rcp.approx.ftz.f64 %0, %0;" : "+d"(v1));
                asm volatile("rcp.approx.ftz.f64 %0, %0;" : "+d"(v2)); asm volatile("rcp.approx.ftz.f64 %0, %0;" : "+d"(v3));
            }
            if (OP == 6 || OP == 7) {
                v0 = fma(v0, a, b); n0 = n0 * 3 + n1; v1 = fma(v1, a, b); n1 = n1 ^ (n0 >> 3);
                v2 = fma(v2, a, b); n2 = n2 * 5 + n3; v3 = fma(v3, a, b); n3 = n3 ^ (n2 >> 5);
                if (OP == 7) { n0 += n2 & 7; n1 -= n3 | 1; n2 += n1 & 3; n3 -= n0 | 2; }
            }
        }
    }
    const double s = v0 + v1 + v2 + v3 + (double)(n0 + n1 + n2 + n3);
    if (s == 12345.678) out[0] = s;
}

// Operand-source variants (register-file read bandwidth hypothesis): ILP independent chains per warp with
//  SRC 0: v = fma(v, a, b), a and b the same registers for every chain (operand reuse cache can serve them)
//  SRC 1: v = fma(v, c[A], c[B]) with constant-bank operands (no register read for them)
//  SRC 2: v_k = fma(v_k, a_k, b_k) with distinct registers per chain (three fresh 64-bit register reads each)
__constant__ double kc[2] = {0.999999, 1e-9};
template <int ILP, int SRC>
__global__ void dfma_src(double *out, int iters, double a, double b)
{
    double v[ILP], aa[ILP], bb[ILP];
#pragma unroll
    for (int k = 0; k < ILP; ++k) { v[k] = threadIdx.x + k; aa[k] = a + 1e-9 * (k + threadIdx.x); bb[k] = b * (k + 1 + threadIdx.x); }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u)
#pragma unroll
            for (int k = 0; k < ILP; ++k) {
                if (SRC == 0) v[k] = fma(v[k], a, b);
                if (SRC == 1) v[k] = fma(v[k], kc[0], kc[1]);
                if (SRC == 2) v[k] = fma(v[k], aa[k], bb[k]);
            }
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < ILP; ++k) s += v[k];
    if (s == 12345.678) out[0] = s;
}

template <class F> static float time_ms(F launch)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch();
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms;
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const double clk = p.clockRate * 1e3; // Hz (max SM clock)
    const int sms = p.multiProcessorCount;
    double *d;
    cudaMalloc(&d, 8);
    const int iters = 2048;
    printf("device %s, %d SMs, %.0f MHz\n", p.name, sms, clk / 1e6);
    // (1) latency: one warp per SM, ILP 1
    {
        float ms = time_ms([&] { dfma_chain<1><<<sms, 32>>>(d, iters, 0.999999, 1e-9); });
        printf("dependent DFMA latency: %.2f cycles\n", ms * 1e-3 * clk / (iters * 16.0));
    }
    // (2) throughput per sub-partition: warps per SMSP x ILP (blocks of 128 threads = 1 warp per SMSP)
    printf("DFMA issue rate per SM sub-partition (warp-instr / cycle; pipe peak = 0.5):\n warps/SMSP");
    for (int ilp = 1; ilp <= 4; ++ilp) printf("   ILP%d", ilp);
    printf("\n");
    for (int w = 1; w <= 8; ++w) {
        printf("   %d      ", w);
        float ms[4];
        ms[0] = time_ms([&] { dfma_chain<1><<<sms, 128 * w>>>(d, iters, 0.999999, 1e-9); });
        ms[1] = time_ms([&] { dfma_chain<2><<<sms, 128 * w>>>(d, iters, 0.999999, 1e-9); });
        ms[2] = time_ms([&] { dfma_chain<3><<<sms, 128 * w>>>(d, iters, 0.999999, 1e-9); });
        ms[3] = time_ms([&] { dfma_chain<4><<<sms, 128 * w>>>(d, iters, 0.999999, 1e-9); });
        for (int k = 0; k < 4; ++k) printf("  %.3f", (double)iters * 16 * (k + 1) * w / (ms[k] * 1e-3 * clk));
        printf("\n");
    }
    // (2b) operand sources, 5 warps/SMSP (the trace kernel's occupancy)
    {
        const int w = 5;
        const char *nm[3] = {"shared a,b registers", "constant-bank a,b", "distinct a_k,b_k registers"};
        float ms[3][3];
        ms[0][0] = time_ms([&] { dfma_src<1, 0><<<sms, 128 * w>>>(d, iters, 0.999999, 1e-9); });
        ms[0][1] = time_ms([&] { dfma_src<2, 0><<<sms, 128 * w>>>(d, iters, 0.999999, 1e-9); });
        ms[0][2] = time_ms([&] { dfma_src<4, 0><<<sms, 128 * w>>>(d, iters, 0.999999, 1e-9); });
        ms[1][0] = time_ms([&] { dfma_src<1, 1><<<sms, 128 * w>>>(d, iters, 0.999999, 1e-9); });
        ms[1][1] = time_ms([&] { dfma_src<2, 1><<<sms, 128 * w>>>(d, iters, 0.999999, 1e-9); });
        ms[1][2] = time_ms([&] { dfma_src<4, 1><<<sms, 128 * w>>>(d, iters, 0.999999, 1e-9); });
        ms[2][0] = time_ms([&] { dfma_src<1, 2><<<sms, 128 * w>>>(d, iters, 0.999999, 1e-9); });
        ms[2][1] = time_ms([&] { dfma_src<2, 2><<<sms, 128 * w>>>(d, iters, 0.999999, 1e-9); });
        ms[2][2] = time_ms([&] { dfma_src<4, 2><<<sms, 128 * w>>>(d, iters, 0.999999, 1e-9); });
        printf("DFMA rate per SMSP by operand source at 5 warps/SMSP (ILP 1, 2, 4):\n");
        const int il[3] = {1, 2, 4};
        for (int r = 0; r < 3; ++r) {
            printf("  %-28s", nm[r]);
            for (int k = 0; k < 3; ++k) printf("  %.3f", (double)iters * 16 * il[k] * w / (ms[r][k] * 1e-3 * clk));
            printf("\n");
        }
    }
    // (3)/(4) op mix at 4 warps/SMSP x ILP 4 (saturating): cycles per warp-instruction group of 4
    const char *names[8] = {"DFMA", "DMUL", "DADD", "DSETP+int add", "F2I.F64 + int add", "MUFU.RCP64H", "DFMA + 1 int op", "DFMA + 2 int ops"};
    float t[8];
    t[0] = time_ms([&] { op_mix<0><<<sms, 512>>>(d, iters, 0.999999, 1e-9); });
    t[1] = time_ms([&] { op_mix<1><<<sms, 512>>>(d, iters, 0.999999, 1e-9); });
    t[2] = time_ms([&] { op_mix<2><<<sms, 512>>>(d, iters, 0.999999, 1e-9); });
    t[3] = time_ms([&] { op_mix<3><<<sms, 512>>>(d, iters, 0.999999, 1e-9); });
    t[4] = time_ms([&] { op_mix<4><<<sms, 512>>>(d, iters, 0.999999, 1e-9); });
    t[5] = time_ms([&] { op_mix<5><<<sms, 512>>>(d, iters, 0.999999, 1e-9); });
    t[6] = time_ms([&] { op_mix<6><<<sms, 512>>>(d, iters, 0.999999, 1e-9); });
    t[7] = time_ms([&] { op_mix<7><<<sms, 512>>>(d, iters, 0.999999, 1e-9); });
    for (int k = 0; k < 8; ++k)
        printf("%-20s %.2f SMSP-cycles per warp-instruction of this kind (4 warps/SMSP)\n", names[k],
               t[k] * 1e-3 * clk / (iters * 16.0 * 4 * 4));
    return 0;
}
