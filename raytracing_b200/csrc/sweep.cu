// sweep.cu -- rtb_sweep: the DELTA_S sweep of RT_bench.py:1303-1318 (executor.map(search_delta, ...) over the
// candidate steps) as batched traces of candidates x rays, reduced on the device to what the selection rules
// (find_index, RT:1323-1373) consume: per candidate the sum / max / count of a per-ray metric.
//   interface  : Snell error of every ray (RT:896-919)          -> np.mean(errors), np.max(errors)   RT:954
//   fisheye    : closure % of the fan's first ray (RT:956)
//   vert/aniso : coefficient of variation of every ray's non-zero p_x entries (RT:1354-1360)
// The batched fan is never materialised: the trace kernel indexes the R-ray fan modulo R and the candidate arrays by
// r / R (TraceArgs.cand_R).  Candidates are processed in groups of at most 2^26 rays.
#include "ctx.cuh"

#include <mutex>

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

// ---------------------------------------------------------------------------------------------------------------
// rtb_sweep_noalloc: the same sweep for callers that must not allocate (CUDA graph capture, serving loops, harnesses
// that forbid cudaMalloc).  No counterpart in the reference.
//
// rtb_sweep parks the per-ray results of the trace kernel (d_ray, last rows / p_x statistics) in pool buffers obtained
// with cudaMalloc and hands them to sweep_reduce_kernel; here the same kernels use one fixed array in device memory
// instead: RTB_NOALLOC_MAX_RAYS rays x 9 doubles = 9 MiB, resident on every GPU that loads the library until the
// process exits.  Every call on a device shares it (calls that overlap on the GPU race), and it keeps the previous
// call's contents, so the part a call uses is cleared first.
// ---------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) double g_sweep_scratch[(size_t)RTB_NOALLOC_MAX_RAYS * 9];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
double *g_sweep_scratch_addr[NOALLOC_MAX_DEVICES]; // device address of g_sweep_scratch per device (host code may not
std::mutex g_sweep_scratch_mutex;                  // pass the symbol itself: its name is a host-side stand-in)
} // namespace

extern "C" int rtb_sweep_noalloc(rtb_ctx *ctx, const rtb_params *p, const double *theta0, const double *x0,
                                 const double *y0, int64_t R, const rtb_sweep_spec *spec, double *sum, double *mx,
                                 int64_t *count, rtb_stats *stats)
{
    if (!ctx) return RTB_ERR_ARG;
    if (!ctx->has_field) return fail(ctx, RTB_ERR_NOFIELD, "rtb_sweep_noalloc: no field set (rtb_set_field)");
    if (!p || !spec || !theta0 || !x0 || !y0 || !sum || !mx || !count || !spec->steps || !spec->max_sizes)
        return fail(ctx, RTB_ERR_ARG, "rtb_sweep_noalloc: null argument");
    if (p->ptr_kind != RTB_PTR_DEVICE)
        return fail(ctx, RTB_ERR_ARG, "rtb_sweep_noalloc: device pointers only (RTB_PTR_DEVICE): host arrays would need staging buffers");
    if (p->method < 1 || p->method > 11) return fail(ctx, RTB_ERR_ARG, "rtb_sweep_noalloc: method must be 1..11");
    if (spec->metric != RTB_SWEEP_CLOSURE && spec->metric != RTB_SWEEP_PX_CV)
        return fail(ctx, RTB_ERR_ARG, "rtb_sweep_noalloc: metric must be RTB_SWEEP_CLOSURE or RTB_SWEEP_PX_CV (the Snell metric "
                                      "re-integrates from checkpoint buffers sized by the call: use rtb_sweep)");
    if (R < 0 || spec->C < 0 || spec->skip_lo < 0 || spec->skip_hi < 0) return fail(ctx, RTB_ERR_ARG, "rtb_sweep_noalloc: negative size");
    const int64_t C = spec->C;
    if (C > RTB_NOALLOC_MAX_CANDIDATES)
        return fail(ctx, RTB_ERR_ARG, "rtb_sweep_noalloc: more than RTB_NOALLOC_MAX_CANDIDATES (256) candidates");
    if (R > RTB_NOALLOC_MAX_RAYS || C * R > RTB_NOALLOC_MAX_RAYS)
        return fail(ctx, RTB_ERR_ARG, "rtb_sweep_noalloc: C * R exceeds RTB_NOALLOC_MAX_RAYS (131072 rays in all): split the "
                                      "candidates over several calls or use rtb_sweep");
    if (ctx->device < 0 || ctx->device >= NOALLOC_MAX_DEVICES) return fail(ctx, RTB_ERR_ARG, "rtb_sweep_noalloc: device index");
    if (stats) std::memset(stats, 0, sizeof(*stats));
    if (C == 0) return RTB_OK;
    CK(cudaSetDevice(ctx->device));
    double *scratch;
    { // once per device, with that device current
        std::lock_guard<std::mutex> lock(g_sweep_scratch_mutex);
        if (!g_sweep_scratch_addr[ctx->device]) {
            double *ptr = nullptr;
            CK(cudaGetSymbolAddress((void **)&ptr, g_sweep_scratch));
            g_sweep_scratch_addr[ctx->device] = ptr;
        }
        scratch = g_sweep_scratch_addr[ctx->device];
    }
    cudaStream_t st = ctx->stream;
    const int metric = spec->metric;

    // warp-uniform loop bound of the kernel = the largest candidate (read back into a fixed host array)
    int64_t hm[RTB_NOALLOC_MAX_CANDIDATES];
    CK(cudaMemcpyAsync(hm, spec->max_sizes, sizeof(int64_t) * C, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    int64_t Mmax = 1;
    for (int64_t c = 0; c < C; ++c) {
        if (hm[c] < 1 || hm[c] > 0x7fffffff) return fail(ctx, RTB_ERR_ARG, "rtb_sweep_noalloc: bad max_sizes entry");
        Mmax = std::max(Mmax, hm[c]);
    }

    const int64_t n = C * R;
    CK(cudaEventRecord(ctx->ev[0], st));
    CK(cudaMemsetAsync(ctx->d_total, 0, 2 * sizeof(unsigned long long), st));
    // the array keeps the last call's contents: rows a ray does not reach must read as zeros (np.zeros RT:802)
    if (n > 0) CK(cudaMemsetAsync(scratch, 0, sizeof(double) * 9 * n, st));
    double *d_ray = scratch, *d_last = nullptr, *d_px = nullptr; // (3, n) | (6, n) or (3, n)
    if (metric == RTB_SWEEP_CLOSURE) d_last = scratch + 3 * n;
    else d_px = scratch + 3 * n;
    rtb_params pp = *p;
    pp.max_size = Mmax; pp.steps = nullptr; pp.max_sizes = nullptr; pp.step = 1.0;
    int64_t launches = 0;
    CK(cudaEventRecord(ctx->ev[1], st));
    if (n > 0) {
        TraceArgs a = rtb_base_args(ctx, &pp, 1);
        a.theta0 = theta0; a.x0 = x0; a.y0 = y0;
        a.steps = spec->steps; a.max_sizes = spec->max_sizes;
        a.cand_R = R; a.R = n;
        a.lane_shift = pick_lane_shift(n);
        a.stage = n >= (1 << 16);
        a.d_ray = d_ray; a.last_row = d_last; a.px_stats = d_px;
        const int rc = rtb_launch(ctx, st, &pp, a, false);
        if (rc) { cudaStreamSynchronize(st); return rc; }
        ++launches;
    }
    sweep_reduce_kernel<<<(unsigned)C, RED_T, 0, st>>>(metric, R, spec->skip_lo, spec->skip_hi, nullptr, d_ray, d_last, d_px,
                                                       spec->max_sizes, n, sum, mx, count);
    CK(cudaGetLastError());
    ++launches;
    CK(cudaEventRecord(ctx->ev[2], st));
    unsigned long long counters[2] = {0, 0};
    CK(cudaMemcpyAsync(counters, ctx->d_total, sizeof(counters), cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(ctx->ev[3], st));
    CK(cudaStreamSynchronize(st));
    if (stats) {
        float ms = 0;
        cudaEventElapsedTime(&ms, ctx->ev[1], ctx->ev[2]); stats->kernel_ms = ms;
        cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[3]); stats->total_ms = ms;
        stats->ray_steps = (int64_t)counters[0];
        stats->nonfinite_rays = (int64_t)counters[1];
        stats->launches = launches;
    }
    return RTB_OK;
}
