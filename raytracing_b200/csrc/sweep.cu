// sweep.cu -- rtb_sweep: the DELTA_S sweep of RT_bench.py:1303-1318 (executor.map(search_delta, ...) over the
// candidate steps) as batched traces of candidates x rays, reduced on the device to what the selection rules
// (find_index, RT:1323-1373) consume: per candidate the sum / max / count of a per-ray metric.
//   interface  : Snell error of every ray (RT:896-919)          -> np.mean(errors), np.max(errors)   RT:954
//   fisheye    : closure % of the fan's first ray (RT:956)
//   vert/aniso : coefficient of variation of every ray's non-zero p_x entries (RT:1354-1360)
// The batched fan is never materialised: the trace kernel indexes the R-ray fan modulo R and the candidate arrays by
// r / R (TraceArgs.cand_R).  Candidates are processed in groups of at most 2^26 rays.
#include "ctx.cuh"

namespace {

constexpr int RED_T = 256;

// One block per candidate; every thread accumulates rays tid, tid + 256, ... in order, then a fixed tree: the result
// does not depend on scheduling (bitwise reproducible run to run and across shardings of the candidate axis).
__global__ void __launch_bounds__(RED_T) sweep_reduce_kernel(int metric, int64_t R, int64_t skip_lo, int64_t skip_hi,
                                                             const double *errors, const double *d_ray, const double *last_row,
                                                             const double *px_stats, const int64_t *max_sizes, int64_t n,
                                                             double *sum, double *mx, int64_t *count)
{
    __shared__ double ssum[RED_T], smax[RED_T];
    __shared__ long long scnt[RED_T];
    const int64_t c = blockIdx.x;
    const int64_t base = c * R;
    double s = 0.0, m = -INFINITY;
    long long cnt = 0;
    if (metric == RTB_SWEEP_CLOSURE) {
        if (threadIdx.x == 0 && n > 0) {
            // rays[-1, 0:2, 0]: row max_size-1 of the first ray, zeros unless the ray ran to the last row (RT:802)
            const bool done = d_ray[2 * n + base] == (double)(max_sizes[c] - 1);
            const double x = done ? last_row[base] : 0.0, y = done ? last_row[n + base] : 0.0;
            const double dx = 1.0 - x, dy = 0.0 - y;
            s = m = 100.0 * sqrt(dx * dx + dy * dy) / (2.0 * RTB_PI);
            cnt = 1;
        }
    } else {
        for (int64_t k = skip_lo + threadIdx.x; k < R - skip_hi; k += RED_T) {
            double v;
            if (metric == RTB_SWEEP_SNELL) v = errors[base + k];
            else v = 100.0 * px_stats[2 * n + base + k] / px_stats[n + base + k]; // 100 std / mean, RT:1359
            s += v;
            m = fmax(m, v); // like np.max a NaN metric is not filtered: fmax drops it, so flag it through the sum
            ++cnt;
        }
    }
    ssum[threadIdx.x] = s; smax[threadIdx.x] = m; scnt[threadIdx.x] = cnt;
    __syncthreads();
    for (int o = RED_T / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) {
            ssum[threadIdx.x] += ssum[threadIdx.x + o];
            smax[threadIdx.x] = fmax(smax[threadIdx.x], smax[threadIdx.x + o]);
            scnt[threadIdx.x] += scnt[threadIdx.x + o];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        sum[c] = ssum[0];
        mx[c] = (ssum[0] != ssum[0]) ? ssum[0] : smax[0]; // a NaN among the metrics shows in both outputs
        count[c] = scnt[0];
    }
}

} // namespace

extern "C" int rtb_sweep(rtb_ctx *ctx, const rtb_params *p, const double *theta0, const double *x0, const double *y0,
                         int64_t R, const rtb_sweep_spec *spec, double *sum, double *mx, int64_t *count, rtb_stats *stats)
{
    if (!ctx) return RTB_ERR_ARG;
    if (!ctx->has_field) return fail(ctx, RTB_ERR_NOFIELD, "rtb_sweep: no field set (rtb_set_field)");
    if (!p || !spec || !theta0 || !x0 || !y0 || !sum || !mx || !count || !spec->steps || !spec->max_sizes)
        return fail(ctx, RTB_ERR_ARG, "rtb_sweep: null argument");
    if (p->method < 1 || p->method > 11) return fail(ctx, RTB_ERR_ARG, "rtb_sweep: method must be 1..11");
    if (spec->metric < RTB_SWEEP_SNELL || spec->metric > RTB_SWEEP_PX_CV) return fail(ctx, RTB_ERR_ARG, "rtb_sweep: unknown metric");
    if (R < 0 || spec->C < 0 || spec->skip_lo < 0 || spec->skip_hi < 0) return fail(ctx, RTB_ERR_ARG, "rtb_sweep: negative size");
    if (stats) std::memset(stats, 0, sizeof(*stats));
    const int64_t C = spec->C;
    if (C == 0) return RTB_OK;
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const bool host = p->ptr_kind == RTB_PTR_HOST;
    const int metric = spec->metric;

    // warp-uniform loop bound of the kernels = the largest candidate
    std::vector<int64_t> hm(C);
    if (host) std::memcpy(hm.data(), spec->max_sizes, sizeof(int64_t) * C);
    else {
        CK(cudaMemcpyAsync(hm.data(), spec->max_sizes, sizeof(int64_t) * C, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
    }
    int64_t Mmax = 1;
    for (int64_t v : hm) {
        if (v < 1 || v > 0x7fffffff) return fail(ctx, RTB_ERR_ARG, "rtb_sweep: bad max_sizes entry");
        Mmax = std::max(Mmax, v);
    }

    CK(cudaEventRecord(ctx->ev[0], st));
    CK(cudaMemsetAsync(ctx->d_total, 0, 2 * sizeof(unsigned long long), st));
    Staging sg(ctx, host);
    auto bail = [&](int code) { cudaStreamSynchronize(st); return code; };
    const double *d_th = sg.in(theta0, R), *d_x0 = sg.in(x0, R), *d_y0 = sg.in(y0, R);
    const double *d_steps = sg.in(spec->steps, C);
    const int64_t *d_msz = sg.in(spec->max_sizes, C);
    double *d_sum = sg.out(sum, C), *d_mx = sg.out(mx, C);
    int64_t *d_cnt = sg.out(count, C);
    const int64_t group = R > 0 ? std::max<int64_t>(1, std::min<int64_t>(C, ((int64_t)1 << 26) / std::max<int64_t>(R, 1))) : C;
    const int64_t nmax = group * R;
    double *d_ray = nullptr, *d_err = nullptr, *d_last = nullptr, *d_px = nullptr;
    if (R > 0) {
        d_ray = (double *)sg.alloc(sizeof(double) * 3 * nmax);
        if (metric == RTB_SWEEP_SNELL) d_err = (double *)sg.alloc(sizeof(double) * nmax);
        if (metric == RTB_SWEEP_CLOSURE) d_last = (double *)sg.alloc(sizeof(double) * 6 * nmax);
        if (metric == RTB_SWEEP_PX_CV) d_px = (double *)sg.alloc(sizeof(double) * 3 * nmax);
    }
    if (sg.err != cudaSuccess) return bail(fail(ctx, sg.err == cudaErrorMemoryAllocation ? RTB_ERR_NOMEM : RTB_ERR_CUDA,
                                                 "rtb_sweep: staging", sg.err));
    rtb_params pp = *p;
    pp.max_size = Mmax; pp.steps = nullptr; pp.max_sizes = nullptr; pp.step = 1.0;
    if (metric == RTB_SWEEP_SNELL) pp.flags |= RTB_F_INTERFACE;
    int64_t launches = 0;
    CK(cudaEventRecord(ctx->ev[1], st));
    for (int64_t c0 = 0; c0 < C; c0 += group) {
        const int64_t ng = std::min(group, C - c0), n = ng * R;
        if (n > 0) {
            TraceArgs a = rtb_base_args(ctx, &pp, 1);
            a.theta0 = d_th; a.x0 = d_x0; a.y0 = d_y0;
            a.steps = d_steps + c0; a.max_sizes = d_msz + c0;
            a.cand_R = R; a.R = n;
            a.lane_shift = pick_lane_shift(n);
            a.stage = n >= (1 << 16);
            a.d_ray = d_ray; a.last_row = d_last; a.px_stats = d_px;
            const size_t mk = sg.mark(); // the Snell metric's checkpoints and capture rows live for one group
            int rc = rtb_trace_core(ctx, st, &pp, a, d_err, metric == RTB_SWEEP_SNELL, metric == RTB_SWEEP_SNELL, false, 1,
                                    sg, launches);
            if (rc) return bail(rc);
            sg.release(mk);
        }
        // n == 0 (empty fan): count 0, sum 0, max -inf, like np.mean/np.max of nothing would warn about
        sweep_reduce_kernel<<<(unsigned)ng, RED_T, 0, st>>>(metric, R, spec->skip_lo, spec->skip_hi, d_err, d_ray, d_last,
                                                            d_px, d_msz + c0, n, d_sum + c0, d_mx + c0, d_cnt + c0);
        CK(cudaGetLastError());
        ++launches;
    }
    CK(cudaEventRecord(ctx->ev[2], st));
    sg.copy_back();
    unsigned long long counters[2] = {0, 0};
    CK(cudaMemcpyAsync(counters, ctx->d_total, sizeof(counters), cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(ctx->ev[3], st));
    CK(cudaStreamSynchronize(st));
    if (sg.err != cudaSuccess) return fail(ctx, RTB_ERR_CUDA, "rtb_sweep: copy", sg.err);
    if (stats) {
        float ms = 0;
        cudaEventElapsedTime(&ms, ctx->ev[1], ctx->ev[2]); stats->kernel_ms = ms;
        cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[3]); stats->total_ms = ms;
        stats->ray_steps = (int64_t)counters[0];
        stats->nonfinite_rays = (int64_t)counters[1];
        stats->launches = launches;
        stats->h2d_bytes = sg.h2d; stats->d2h_bytes = sg.d2h;
    }
    return RTB_OK;
}
