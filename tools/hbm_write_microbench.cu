// Dev tool: what HBM3e sustains for the trajectory write-back pattern of trace_kernel (TRAJ mode) on B200.
//   mode 0: plain fill, every thread streams consecutive doubles (the write-only roofline);
//   mode 1: the s_ray pattern -- a warp writes one 256-B segment per (row, column), segments of one warp are R*8 bytes
//           apart, st.global.cs.f64 per lane (what store_row does);
//   mode 2: the same segments written with cp.async.bulk shared->global (one 256-B bulk copy per segment, issued by
//           6 lanes of the warp at once) -- the staged path.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/bin/hbm_write_microbench tools/hbm_write_microbench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void fill_kernel(double *p, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) __stcs(p + i, (double)i);
}

// R rays, rows rows: thread r writes s[(row*6+col)*R + r] for all rows, cols
__global__ void __launch_bounds__(128) traj_kernel(double *s, int64_t R, int rows)
{
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    double *p = s + r;
    double v = (double)r;
    for (int i = 0; i < rows; ++i) {
#pragma unroll
        for (int c = 0; c < 6; ++c) __stcs(p + c * R, v + c);
        p += 6 * R;
        v += 1.0;
    }
}

// staged: each warp parks K rows x 6 columns x 32 lanes in shared memory, then 6*K lanes issue one 256-B bulk copy each
template <int K>
__global__ void __launch_bounds__(128) traj_bulk_kernel(double *s, int64_t R, int rows)
{
    __shared__ __align__(128) double buf[4][2][K * 6][32]; // per warp, double-buffered
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t r0 = r - lane;
    if (r0 >= R) return;
    double v = (double)r;
    int phase = 0;
    for (int i0 = 0; i0 < rows; i0 += K) {
        // the bulk copies that read this half two flushes ago must have finished reading shared memory
        asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        __syncwarp();
        const int kk = min(K, rows - i0);
        for (int k = 0; k < kk; ++k) {
#pragma unroll
            for (int c = 0; c < 6; ++c) buf[w][phase][k * 6 + c][lane] = v + c;
            v += 1.0;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane < kk * 6) {
            double *g = s + ((int64_t)i0 * 6 + lane) * R + r0;
            const unsigned sa = (unsigned)__cvta_generic_to_shared(&buf[w][phase][lane][0]);
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], 256;" ::"l"(g), "r"(sa) : "memory");
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        phase ^= 1;
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

int main()
{
    const int64_t R = 1 << 20;
    const int rows = 910;
    const size_t n = (size_t)R * 6 * rows;
    double *s;
    if (cudaMalloc(&s, n * 8) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto time = [&](const char *name, auto launch) {
        float best = 1e30f;
        for (int rep = 0; rep < 4; ++rep) {
            cudaEventRecord(e0);
            launch();
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (rep > 0 && ms < best) best = ms;
        }
        cudaError_t e = cudaGetLastError();
        printf("%-34s %8.3f ms  %7.1f GB/s  %s\n", name, best, n * 8 / (best * 1e-3) / 1e9, e == cudaSuccess ? "" : cudaGetErrorString(e));
    };
    time("fill (148*16 blocks x 256)", [&] { fill_kernel<<<148 * 16, 256>>>(s, n); });
    time("fill (148*64 blocks x 256)", [&] { fill_kernel<<<148 * 64, 256>>>(s, n); });
    time("traj st.cs.f64 per lane", [&] { traj_kernel<<<(unsigned)(R / 128), 128>>>(s, R, rows); });
    time("traj bulk K=1", [&] { traj_bulk_kernel<1><<<(unsigned)(R / 128), 128>>>(s, R, rows); });
    time("traj bulk K=2", [&] { traj_bulk_kernel<2><<<(unsigned)(R / 128), 128>>>(s, R, rows); });
    time("traj bulk K=4", [&] { traj_bulk_kernel<4><<<(unsigned)(R / 128), 128>>>(s, R, rows); });
    // verify the bulk pattern wrote what the per-lane pattern writes
    traj_kernel<<<(unsigned)(R / 128), 128>>>(s, R, 8);
    double *h0 = (double *)malloc(48 * R * 8), *h1 = (double *)malloc(48 * R * 8);
    cudaMemcpy(h0, s, 48 * R * 8, cudaMemcpyDeviceToHost);
    cudaMemset(s, 0, 48 * R * 8);
    traj_bulk_kernel<4><<<(unsigned)(R / 128), 128>>>(s, R, 8);
    cudaMemcpy(h1, s, 48 * R * 8, cudaMemcpyDeviceToHost);
    size_t bad = 0;
    for (size_t i = 0; i < (size_t)48 * R; ++i) bad += h0[i] != h1[i];
    printf("bulk vs per-lane mismatches: %zu (%s)\n", bad, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
