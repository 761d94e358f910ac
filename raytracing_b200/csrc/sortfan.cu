// sortfan.cu -- RTB_F_SORT_FAN: rtb_trace of an unsorted fan.
//
// The reference traces its rays one by one in the order they come (RT:807 `for k in range(ray_count)`); here 32
// neighbouring rays share a warp, and a warp whose rays start in different places or directions fetches 32 different
// cell records per step and idles on its longest ray: the shuffled 2^20-ray vert fan takes 73.5 ms against 15.3 ms in
// launch-angle order (profiles/r2_time_r2_sort_fan_full.txt).  With the flag the library orders the rays by
// (x0, y0, theta0) -- three stable radix sorts of order-preserving keys, least significant first -- traces the
// ordered fan and scatters every result column back to the caller's ray index.  Rays are independent, so the results
// are those of the unsorted call, bit for bit (tests/test_gpu_round2.py).  FINAL-type outputs only: a trajectory keeps
// the caller's order by construction of its rows.
#include "ctx.cuh"

#include <cub/device/device_radix_sort.cuh>

namespace {

// doubles order like these unsigned keys (negative values: all bits flipped, others: sign bit set)
__device__ __forceinline__ unsigned long long order_key(double v)
{
    const long long b = __double_as_longlong(v);
    return (unsigned long long)(b ^ ((b >> 63) | (long long)0x8000000000000000ull));
}

__global__ void sort_key_kernel(const double *v, const unsigned *perm, int64_t R, unsigned long long *keys, unsigned *iota)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    keys[i] = order_key(v[perm ? perm[i] : i]);
    if (iota) iota[i] = (unsigned)i;
}

__global__ void gather_fan_kernel(const unsigned *perm, int64_t R, const double *theta0, const double *x0, const double *y0,
                                  const double *steps, const int64_t *max_sizes, double *s_theta, double *s_x, double *s_y,
                                  double *s_steps, int64_t *s_msz)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    const unsigned k = perm[i];
    s_theta[i] = theta0[k]; s_x[i] = x0[k]; s_y[i] = y0[k];
    if (steps) s_steps[i] = steps[k];
    if (max_sizes) s_msz[i] = max_sizes[k];
}

// dst[c][perm[i]] = src[c][i] for the `rows` columns-of-R of a result array
__global__ void scatter_rows_kernel(const double *src, double *dst, const unsigned *perm, int64_t R, int rows)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    const unsigned k = perm[i];
    for (int c = 0; c < rows; ++c) dst[(int64_t)c * R + k] = src[(int64_t)c * R + i];
}

} // namespace

int rtb_trace_sorted(rtb_ctx *ctx, const rtb_params *p, const double *theta0, const double *x0, const double *y0, int64_t R,
                     double *d_ray, double *errors, double *last_row, double *px_stats, rtb_stats *stats)
{
    if (R >= ((int64_t)1 << 31)) return fail(ctx, RTB_ERR_ARG, "rtb_trace: RTB_F_SORT_FAN takes fewer than 2^31 rays");
    const bool host = p->ptr_kind == RTB_PTR_HOST;
    if (ctx->gather_world > 0 && !host)
        return fail(ctx, RTB_ERR_ARG, "rtb_trace: RTB_F_SORT_FAN with a fused gather (the gather index is the ray's position in the launch)");
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    Staging sg(ctx, host);
    auto bail = [&](int code) { cudaStreamSynchronize(st); return code; };
    const double *d_th = sg.in(theta0, R), *d_x0 = sg.in(x0, R), *d_y0 = sg.in(y0, R);
    const double *d_steps = sg.in(p->steps, R);
    const int64_t *d_msz = sg.in(p->max_sizes, R);
    unsigned long long *keys = (unsigned long long *)sg.alloc(sizeof(unsigned long long) * R);
    unsigned long long *keys2 = (unsigned long long *)sg.alloc(sizeof(unsigned long long) * R);
    unsigned *pa = (unsigned *)sg.alloc(sizeof(unsigned) * R), *pb = (unsigned *)sg.alloc(sizeof(unsigned) * R);
    size_t tmp_bytes = 0;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, keys2, pa, pb, (int)R, 0, 64, st));
    void *tmp = sg.alloc(tmp_bytes);
    double *s_th = (double *)sg.alloc(sizeof(double) * R), *s_x = (double *)sg.alloc(sizeof(double) * R);
    double *s_y = (double *)sg.alloc(sizeof(double) * R);
    double *s_steps = p->steps ? (double *)sg.alloc(sizeof(double) * R) : nullptr;
    int64_t *s_msz = p->max_sizes ? (int64_t *)sg.alloc(sizeof(int64_t) * R) : nullptr;
    double *t_dray = (double *)sg.alloc(sizeof(double) * 3 * R);
    double *t_err = errors ? (double *)sg.alloc(sizeof(double) * R) : nullptr;
    double *t_last = last_row ? (double *)sg.alloc(sizeof(double) * 6 * R) : nullptr;
    double *t_px = px_stats ? (double *)sg.alloc(sizeof(double) * 3 * R) : nullptr;
    if (sg.err != cudaSuccess) return bail(fail(ctx, sg.err == cudaErrorMemoryAllocation ? RTB_ERR_NOMEM : RTB_ERR_CUDA,
                                                 "rtb_trace: buffers of RTB_F_SORT_FAN", sg.err));
    const unsigned g = (unsigned)((R + 255) / 256);
    // least significant key first; every pass is stable
    sort_key_kernel<<<g, 256, 0, st>>>(d_th, nullptr, R, keys, pa);
    CK(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys, keys2, pa, pb, (int)R, 0, 64, st));
    sort_key_kernel<<<g, 256, 0, st>>>(d_y0, pb, R, keys, nullptr);
    CK(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys, keys2, pb, pa, (int)R, 0, 64, st));
    sort_key_kernel<<<g, 256, 0, st>>>(d_x0, pa, R, keys, nullptr);
    CK(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys, keys2, pa, pb, (int)R, 0, 64, st));
    const unsigned *perm = pb; // launch position i holds the caller's ray perm[i]
    gather_fan_kernel<<<g, 256, 0, st>>>(perm, R, d_th, d_x0, d_y0, d_steps, d_msz, s_th, s_x, s_y, s_steps, s_msz);
    CK(cudaGetLastError());

    rtb_params pp = *p;
    pp.flags &= ~RTB_F_SORT_FAN;
    pp.ptr_kind = RTB_PTR_DEVICE;
    pp.steps = s_steps; pp.max_sizes = s_msz;
    int rc = rtb_trace(ctx, &pp, s_th, s_x, s_y, R, nullptr, t_dray, t_err, t_last, t_px, stats);
    if (rc) return bail(rc);

    double *o_dray = sg.out(d_ray, (size_t)3 * R), *o_err = sg.out(errors, (size_t)R);
    double *o_last = sg.out(last_row, (size_t)6 * R), *o_px = sg.out(px_stats, (size_t)3 * R);
    if (sg.err != cudaSuccess) return bail(fail(ctx, RTB_ERR_NOMEM, "rtb_trace: result buffers of RTB_F_SORT_FAN", sg.err));
    scatter_rows_kernel<<<g, 256, 0, st>>>(t_dray, o_dray, perm, R, 3);
    if (errors) scatter_rows_kernel<<<g, 256, 0, st>>>(t_err, o_err, perm, R, 1);
    if (last_row) scatter_rows_kernel<<<g, 256, 0, st>>>(t_last, o_last, perm, R, 6);
    if (px_stats) scatter_rows_kernel<<<g, 256, 0, st>>>(t_px, o_px, perm, R, 3);
    CK(cudaGetLastError());
    sg.copy_back();
    CK(cudaStreamSynchronize(st));
    if (sg.err != cudaSuccess) return fail(ctx, RTB_ERR_CUDA, "rtb_trace: copy (RTB_F_SORT_FAN)", sg.err);
    if (stats) {
        stats->launches += 8 + (errors ? 1 : 0) + (last_row ? 1 : 0) + (px_stats ? 1 : 0); // keys x3, sorts x3, gather, scatters
        stats->h2d_bytes = sg.h2d; stats->d2h_bytes = sg.d2h;
    }
    return RTB_OK;
}
