// shoot.cu -- batched two-point shooting search (BASELINE.json configs[4]).
//
// The reference has no two-point search (README.md:6 "shooting method" = a forward fan through
// trazar), so this is DEFINED here from reference parts and is parity-unpinned:
//   coarse : one fan of A launch angles per DELTA_S candidate is integrated by the trace kernel
//            (trazar RT_bench.py:766-948); for every target point the angle whose exit point is
//            closest is the bracket centre (first minimum, like np.argmin);
//   refine : the launch angle is refined by the reference's own golden-section routine
//            (golden RT_bench.py:175-199: both probes re-evaluated every iteration, stop when
//            |c-d| <= GOLD_TOL, result (a+b)/2), one full re-trace of two rays per probe pair,
//            all (candidate, target) pairs batched into one launch per iteration.
// "Exit point" of a ray = where its last segment crosses the box of constants() (linear
// interpolation between the last row inside and the first row outside, RT_bench.py:878), with the
// travel time s_ray[.,4] interpolated the same way.  miss^2 = squared distance exit point - target.
#include "ctx.cuh"

#ifndef RTB_SHOOT_GROUPS_DEFAULT
#define RTB_SHOOT_GROUPS_DEFAULT 4
#endif

namespace {

struct FanBuffers { // device pointers of one batched launch
    double *theta, *x0, *y0, *steps;
    int64_t *max_sizes;
    double *d_ray, *exit_out;
};

// Launch the integrator on prepared device buffers (FINAL mode + exit points), no staging, no sync.
int launch_fan(rtb_ctx *ctx, const rtb_params *p, const FanBuffers &b, int64_t n, int64_t max_size, cudaStream_t st = nullptr,
               bool count_steps = true)
{
    rtb_params pp = *p;
    pp.max_size = max_size;
    TraceArgs a = rtb_base_args(ctx, &pp, 1);
    a.tail_zero = 0;
    a.theta0 = b.theta; a.x0 = b.x0; a.y0 = b.y0;
    a.steps = b.steps; a.max_sizes = b.max_sizes;
    a.R = n;
    a.d_ray = b.d_ray;
    a.exit_out = b.exit_out;
    a.nonfinite = nullptr;
    if (!count_steps) a.total_steps = nullptr; // the refinement counts the steps of the probes its passes consume
    // small batches: rays spread over the warps of all SMs (pick_lane_shift), 32-thread blocks; large ones run on
    // the ray queue (candidates of different length share warps at the candidate boundaries and, in the refinement,
    // converged pairs take no steps at all: the warps compact)
    a.lane_shift = pick_lane_shift(n);
    a.stage = n >= (1 << 16);
    return rtb_launch(ctx, st ? st : ctx->stream, &pp, a, false);
}

// coarse fan: ray (c, j) = angle j with candidate c's step / max_size
__global__ void coarse_fill_kernel(const double *angles, int64_t A, const double *steps, const int64_t *max_sizes,
                                   int64_t C, double sx, double sy, FanBuffers b)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= A * C) return;
    const int64_t c = i / A, j = i - c * A;
    b.theta[i] = angles[j];
    b.x0[i] = sx; b.y0[i] = sy;
    b.steps[i] = steps[c];
    b.max_sizes[i] = max_sizes[c];
}

// For each (candidate c, target t): first j minimising |exit(c,j) - target t|^2.
// grid (ceil(Tn/128), C); exit points of candidate c are staged through shared memory in tiles.
#define SHOOT_TILE 1024
__global__ void __launch_bounds__(128) bracket_kernel(const double *exit_out, int64_t A, int64_t C, const double *tx,
                                                      const double *ty, int64_t Tn, int64_t index0, double *miss2,
                                                      int64_t *index)
{
    __shared__ double sx[SHOOT_TILE], sy[SHOOT_TILE];
    const int64_t c = blockIdx.y;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const double px = t < Tn ? tx[t] : 0.0, py = t < Tn ? ty[t] : 0.0;
    const double *ex = exit_out + c * A, *ey = exit_out + A * C + c * A;
    double best = INFINITY;
    int64_t bj = 0;
    for (int64_t j0 = 0; j0 < A; j0 += SHOOT_TILE) {
        const int m = (int)min((int64_t)SHOOT_TILE, A - j0);
        for (int k = threadIdx.x; k < m; k += blockDim.x) { sx[k] = ex[j0 + k]; sy[k] = ey[j0 + k]; }
        __syncthreads();
        for (int k = 0; k < m; ++k) {
            const double dx = sx[k] - px, dy = sy[k] - py;
            const double d2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)); // no contraction: numpy's value
            if (d2 < best) { best = d2; bj = j0 + k; }
        }
        __syncthreads();
    }
    if (t < Tn) {
        miss2[c * Tn + t] = best;
        index[c * Tn + t] = index0 + bj;
    }
}

// golden RT:175-199, state per (candidate, target) pair i: bracket [a,b]; probes c = theta[i], d = theta[n+i].
__device__ __forceinline__ void golden_probes(double a, double b, double &c, double &d)
{
    c = __dsub_rn(b, __dmul_rn(__dsub_rn(b, a), RTB_GOLD_RATIO));
    d = __dadd_rn(a, __dmul_rn(__dsub_rn(b, a), RTB_GOLD_RATIO));
}

// Probe rays of one pair in a group's fan: P = 2 (c, d of the current bracket) or, when the passes are speculated two at
// a time, P = 6: c, d, then the probes of the NEXT pass for both outcomes of this one -- c1, d1 of [a, d] (b = d, "left")
// and of [c, b] (a = c).  Ray j of pair i sits at slot i*P + j (probe_slot).  Both outcomes are traced before either is known, so one
// launch carries two passes of golden's loop (RT:188-194); the probes are formed by the same expressions the
// one-pass-per-launch form uses after its decision, hence the same bits.
//
// Slot order: the probes of one pair are neighbours in the fan.  c and d are 0.24 of the bracket apart and the bracket
// shrinks by 0.618 per pass, so from the third pass on a warp covers the launch angles of 16 targets instead of 32 and
// its lanes sit in about half as many field cells -- the refinement is bound by the L1 data pipe (one wavefront per
// distinct cell record per load, profiles/r2_ncu_shoot_refine.csv), not by FP64.
__device__ __forceinline__ int64_t probe_slot(int P, int j, int64_t i) { return i * P + j; }

__device__ __forceinline__ void write_probes(const FanBuffers &f, int64_t n, int64_t i, int P, double a, double b, int64_t M,
                                             bool &active)
{
    double c, d;
    golden_probes(a, b, c, d);
    active = fabs(c - d) > RTB_GOLD_TOL;           // while abs(c - d) > tol
    f.theta[probe_slot(P, 0, i)] = c; f.theta[probe_slot(P, 1, i)] = d;
    const int64_t m0 = active ? M : 1;             // converged pairs stop tracing (max_size 1 = no step)
    f.max_sizes[probe_slot(P, 0, i)] = m0; f.max_sizes[probe_slot(P, 1, i)] = m0;
    if (P == 6) {
        double cl, dl, cr, dr;
        golden_probes(a, d, cl, dl);               // after b = d
        golden_probes(c, b, cr, dr);               // after a = c
        f.theta[i * 6 + 2] = cl; f.theta[i * 6 + 3] = dl; f.theta[i * 6 + 4] = cr; f.theta[i * 6 + 5] = dr;
        const int64_t ml = (active && fabs(cl - dl) > RTB_GOLD_TOL) ? M : 1, mr = (active && fabs(cr - dr) > RTB_GOLD_TOL) ? M : 1;
        f.max_sizes[i * 6 + 2] = ml; f.max_sizes[i * 6 + 3] = ml; f.max_sizes[i * 6 + 4] = mr; f.max_sizes[i * 6 + 5] = mr;
    }
}

// Pairs [i0, i0 + n) of the (C, Tn) problem; `f` holds this group's P*n probe rays.  active[0] counts the pairs that
// still iterate, active[1] takes the widest |c - d| of the group (positive doubles order like their bit patterns).
__global__ void refine_init_kernel(const double *lo, const double *hi, const double *steps, const int64_t *max_sizes,
                                   int64_t Tn, int64_t i0, int64_t n, int P, double sx, double sy, double *a, double *b,
                                   FanBuffers f, unsigned long long *active)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t c = (i0 + i) / Tn;
    lo += i0; hi += i0; a += i0; b += i0;
    a[i] = lo[i]; b[i] = hi[i];
    for (int j = 0; j < P; ++j) { f.x0[i * P + j] = sx; f.y0[i * P + j] = sy; f.steps[i * P + j] = steps[c]; }
    bool act;
    write_probes(f, n, i, P, lo[i], hi[i], max_sizes[c], act);
    if (act) {
        atomicAdd(active, 1ull);
        atomicMax(active + 1, (unsigned long long)__double_as_longlong(fabs(f.theta[probe_slot(P, 0, i)] - f.theta[probe_slot(P, 1, i)])));
    }
}

// squared miss distance of probe ray j of pair i (un-fused multiply/add: numpy's value)
__device__ __forceinline__ double miss2_of(const FanBuffers &f, int64_t n, int P, int j, int64_t i, double px, double py)
{
    const int64_t k = probe_slot(P, j, i);
    const double dx = f.exit_out[k] - px, dy = f.exit_out[(int64_t)P * n + k] - py; // exit_out is (3, P n)
    return __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
}

// One launch's worth of golden passes for every pair of the group: RT:188-194 once (P = 2) or twice (P = 6).
// `useful` gets the ray-steps of the probes the passes actually consumed (the speculated branch not taken is work done,
// not work counted).
__global__ void refine_update_kernel(const double *tx, const double *ty, const int64_t *max_sizes, int64_t Tn, int64_t i0,
                                     int64_t n, int P, double *a, double *b, FanBuffers f, unsigned long long *active,
                                     unsigned long long *useful)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long used = 0;
    if (i < n) {
        const double pc = f.theta[probe_slot(P, 0, i)], pd = f.theta[probe_slot(P, 1, i)];
        if (fabs(pc - pd) > RTB_GOLD_TOL) { // while abs(c - d) > tol
            const int64_t t = (i0 + i) % Tn;
            const double px = tx[t], py = ty[t];
            double va = a[i0 + i], vb = b[i0 + i];
            const double *steps_of = f.d_ray + 2 * (int64_t)P * n; // d_ray is (3, P n): row 2 = steps taken
            const bool left = miss2_of(f, n, P, 0, i, px, py) < miss2_of(f, n, P, 1, i, px, py); // if func(c) < func(d)
            if (left) vb = pd; else va = pc;
            used += (unsigned long long)(steps_of[probe_slot(P, 0, i)] + steps_of[probe_slot(P, 1, i)]);
            if (P == 6) {
                const int j = left ? 2 : 4;
                const double c1 = f.theta[probe_slot(P, j, i)], d1 = f.theta[probe_slot(P, j + 1, i)];
                if (fabs(c1 - d1) > RTB_GOLD_TOL) {
                    if (miss2_of(f, n, P, j, i, px, py) < miss2_of(f, n, P, j + 1, i, px, py)) vb = d1; else va = c1;
                    used += (unsigned long long)(steps_of[probe_slot(P, j, i)] + steps_of[probe_slot(P, j + 1, i)]);
                }
            }
            a[i0 + i] = va; b[i0 + i] = vb;
            bool act;
            write_probes(f, n, i, P, va, vb, max_sizes[(i0 + i) / Tn], act);
            if (act) atomicAdd(active, 1ull);
        }
    }
    used = __reduce_add_sync(0xffffffffu, (unsigned)used) + 0ull; // < 2^32 per warp: 32 lanes x 4 rays x max_size < 2^31
    if ((threadIdx.x & 31) == 0 && used) atomicAdd(useful, used);
}

__global__ void refine_final_fill_kernel(const double *a, const double *b, const double *steps, const int64_t *max_sizes,
                                         int64_t Tn, int64_t n, double sx, double sy, double *theta_out, FanBuffers f)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double th = (b[i] + a[i]) / 2.0; // return (b + a) / 2  RT:199
    theta_out[i] = th;
    f.theta[i] = th;
    f.x0[i] = sx; f.y0[i] = sy;
    f.steps[i] = steps[i / Tn];
    f.max_sizes[i] = max_sizes[i / Tn];
}

__global__ void refine_result_kernel(const double *tx, const double *ty, int64_t Tn, int64_t n, FanBuffers f,
                                     double *result)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t t = i % Tn;
    // the final launch has n rays: exit_out and d_ray are (3, n)
    const double ex = f.exit_out[i], ey = f.exit_out[n + i], eT = f.exit_out[2 * n + i];
    const double dx = ex - tx[t], dy = ey - ty[t];
    result[i] = sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
    result[n + i] = eT;
    result[2 * n + i] = ex;
    result[3 * n + i] = ey;
    result[4 * n + i] = f.d_ray[2 * n + i]; // steps taken (integer stored as double, like d_ray[2] RT:890)
}

} // namespace

static int check_problem(rtb_ctx *ctx, const rtb_params *p, const rtb_shoot_problem *q, const char *who)
{
    if (!ctx) return RTB_ERR_ARG;
    if (!ctx->has_field) return fail(ctx, RTB_ERR_NOFIELD, "rtb_shoot: no field set (rtb_set_field)");
    if (!p || !q || !q->steps || !q->max_sizes || !q->target_x || !q->target_y) return fail(ctx, RTB_ERR_ARG, who);
    if (p->method < 1 || p->method > 11) return fail(ctx, RTB_ERR_ARG, "rtb_shoot: method must be 1..11");
    if (q->C < 0 || q->Tn < 0) return fail(ctx, RTB_ERR_ARG, "rtb_shoot: negative size");
    return RTB_OK;
}

// max over the candidates' max_sizes (host or device array) = warp-uniform loop bound of the kernel
static int max_rows(rtb_ctx *ctx, const int64_t *max_sizes, int64_t C, bool host, int64_t *out)
{
    std::vector<int64_t> h(C);
    if (host) std::memcpy(h.data(), max_sizes, sizeof(int64_t) * C);
    else {
        CK(cudaMemcpyAsync(h.data(), max_sizes, sizeof(int64_t) * C, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
    }
    int64_t m = 1;
    for (int64_t v : h) {
        if (v < 1 || v > 0x7fffffff) return fail(ctx, RTB_ERR_ARG, "rtb_shoot: bad max_sizes entry");
        m = std::max(m, v);
    }
    *out = m;
    return RTB_OK;
}

static FanBuffers alloc_fan(Staging &sg, int64_t n)
{
    FanBuffers b;
    b.theta = (double *)sg.alloc(sizeof(double) * n);
    b.x0 = (double *)sg.alloc(sizeof(double) * n);
    b.y0 = (double *)sg.alloc(sizeof(double) * n);
    b.steps = (double *)sg.alloc(sizeof(double) * n);
    b.max_sizes = (int64_t *)sg.alloc(sizeof(int64_t) * n);
    b.d_ray = (double *)sg.alloc(sizeof(double) * 3 * n);
    b.exit_out = (double *)sg.alloc(sizeof(double) * 3 * n);
    return b;
}

static void finish_stats(rtb_ctx *ctx, rtb_stats *stats, const Staging &sg, unsigned long long total, int64_t launches)
{
    if (!stats) return;
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev[1], ctx->ev[2]); stats->kernel_ms = ms;
    cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[3]); stats->total_ms = ms;
    stats->ray_steps = (int64_t)total;
    stats->launches = launches;
    stats->h2d_bytes = sg.h2d; stats->d2h_bytes = sg.d2h;
}

extern "C" int rtb_shoot_coarse(rtb_ctx *ctx, const rtb_params *p, const rtb_shoot_problem *q, const double *angles,
                                int64_t A, int64_t index0, double *exit_xyT, double *miss2, int64_t *index,
                                rtb_stats *stats)
{
    int rc = check_problem(ctx, p, q, "rtb_shoot_coarse: null argument");
    if (rc) return rc;
    if (!angles || !miss2 || !index || A < 0) return fail(ctx, RTB_ERR_ARG, "rtb_shoot_coarse: null argument");
    if (stats) std::memset(stats, 0, sizeof(*stats));
    const int64_t C = q->C, Tn = q->Tn, n = C * A;
    if (n == 0 || Tn == 0) return RTB_OK;
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const bool host = p->ptr_kind == RTB_PTR_HOST;
    int64_t M = 1;
    if ((rc = max_rows(ctx, q->max_sizes, C, host, &M))) return rc;
    CK(cudaEventRecord(ctx->ev[0], st));
    Staging sg(ctx, host);
    const double *d_ang = sg.in(angles, A), *d_steps = sg.in(q->steps, C);
    const int64_t *d_msz = sg.in(q->max_sizes, C);
    const double *d_tx = sg.in(q->target_x, Tn), *d_ty = sg.in(q->target_y, Tn);
    double *d_miss = sg.out(miss2, C * Tn);
    int64_t *d_idx = sg.out(index, C * Tn);
    FanBuffers fb = alloc_fan(sg, n);
    if (exit_xyT) fb.exit_out = sg.out(exit_xyT, 3 * n);
    if (sg.err != cudaSuccess || !fb.exit_out) return fail(ctx, RTB_ERR_NOMEM, "rtb_shoot_coarse: staging", sg.err);
    CK(cudaMemsetAsync(ctx->d_total, 0, sizeof(unsigned long long), st));
    CK(cudaEventRecord(ctx->ev[1], st));
    coarse_fill_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d_ang, A, d_steps, d_msz, C, q->src_x, q->src_y, fb);
    if ((rc = launch_fan(ctx, p, fb, n, M))) return rc;
    bracket_kernel<<<dim3((unsigned)((Tn + 127) / 128), (unsigned)C), 128, 0, st>>>(fb.exit_out, A, C, d_tx, d_ty, Tn,
                                                                                    index0, d_miss, d_idx);
    CK(cudaGetLastError());
    CK(cudaEventRecord(ctx->ev[2], st));
    sg.copy_back();
    unsigned long long total = 0;
    CK(cudaMemcpyAsync(&total, ctx->d_total, sizeof(total), cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(ctx->ev[3], st));
    CK(cudaStreamSynchronize(st));
    if (sg.err != cudaSuccess) return fail(ctx, RTB_ERR_CUDA, "rtb_shoot_coarse: copy", sg.err);
    finish_stats(ctx, stats, sg, total, 3);
    return RTB_OK;
}

extern "C" int rtb_shoot_refine(rtb_ctx *ctx, const rtb_params *p, const rtb_shoot_problem *q, const double *lo,
                                const double *hi, double *theta, double *result, int64_t *iterations,
                                rtb_stats *stats)
{
    int rc = check_problem(ctx, p, q, "rtb_shoot_refine: null argument");
    if (rc) return rc;
    if (!lo || !hi || !theta || !result) return fail(ctx, RTB_ERR_ARG, "rtb_shoot_refine: null argument");
    if (stats) std::memset(stats, 0, sizeof(*stats));
    if (iterations) *iterations = 0;
    const int64_t C = q->C, Tn = q->Tn, n = C * Tn;
    if (n == 0) return RTB_OK;
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const bool host = p->ptr_kind == RTB_PTR_HOST;
    int64_t M = 1;
    if ((rc = max_rows(ctx, q->max_sizes, C, host, &M))) return rc;
    CK(cudaEventRecord(ctx->ev[0], st));
    Staging sg(ctx, host);
    const double *d_lo = sg.in(lo, n), *d_hi = sg.in(hi, n), *d_steps = sg.in(q->steps, C);
    const int64_t *d_msz = sg.in(q->max_sizes, C);
    const double *d_tx = sg.in(q->target_x, Tn), *d_ty = sg.in(q->target_y, Tn);
    double *d_theta = sg.out(theta, n), *d_res = sg.out(result, 5 * n);
    double *d_a = (double *)sg.alloc(sizeof(double) * n), *d_b = (double *)sg.alloc(sizeof(double) * n);

    // The pairs are independent, and every golden iteration of a pair is one re-trace of its two probe rays -- a
    // launch that lasts as long as its longest ray however few rays it holds.  So the pairs are cut into up to four
    // groups (contiguous in the candidate-major order: candidates of similar length) that iterate on their own
    // streams: while one group's launch drains its long rays the others fill the SMs.  Nothing comes back to the host
    // between iterations: golden's trip count follows from the bracket width alone (the loop test RT:188 is on
    // |c - d|, which shrinks by GOLD_RATIO per pass whatever f does), so each group's count is fixed after its first
    // pass over the brackets and one counter read at the very end confirms that no pair is left iterating.
    const int Gwant = ctx->shoot_groups > 0 ? ctx->shoot_groups : RTB_SHOOT_GROUPS_DEFAULT;
    const int G = (int)std::min<int64_t>(Gwant, std::max<int64_t>(1, n / 4096));
    for (int g = 0; g + 1 < G; ++g) if (!ctx->shoot_st[g]) CK(cudaStreamCreateWithFlags(&ctx->shoot_st[g], cudaStreamNonBlocking));
    for (int g = 0; g < G; ++g) if (!ctx->shoot_ev[g]) CK(cudaEventCreateWithFlags(&ctx->shoot_ev[g], cudaEventDisableTiming));
    unsigned long long *d_active = (unsigned long long *)sg.alloc(sizeof(unsigned long long) * 2 * G);
    // Two passes per launch (6 probe rays per pair instead of 2, see write_probes): 1.5x the ray-steps for half the
    // launches, for problems whose passes are bound by their longest ray rather than by throughput.  Opt-in
    // (RTB_OPT_SHOOT_SPECULATE 2): measured slower on the configs[4] problem at every size tried -- even one GPU's share
    // of it at N=8 (200 candidates x 512 targets) is throughput-bound: 32.7 ms against 48.2 ms speculated.
    const bool spec = ctx->shoot_spec == 2;
    const int P = spec ? 6 : 2;
    struct Group { int64_t i0, n; FanBuffers fb; cudaStream_t st; int iters; };
    std::vector<Group> gr(G);
    bool ok = sg.err == cudaSuccess && d_a && d_b && d_active;
    for (int g = 0; g < G && ok; ++g) {
        gr[g].i0 = n * g / G;
        gr[g].n = n * (g + 1) / G - gr[g].i0;
        gr[g].fb = alloc_fan(sg, P * gr[g].n);
        gr[g].st = g == 0 ? st : ctx->shoot_st[g - 1];
        gr[g].iters = 0;
        ok = sg.err == cudaSuccess && gr[g].fb.exit_out;
    }
    if (!ok) return fail(ctx, RTB_ERR_NOMEM, "rtb_shoot_refine: staging", sg.err);
    auto bail = [&](int code) {
        cudaStreamSynchronize(st);
        for (auto &ps : ctx->shoot_st) if (ps) cudaStreamSynchronize(ps);
        return code;
    };
    CK(cudaMemsetAsync(ctx->d_total, 0, 2 * sizeof(unsigned long long), st));
    CK(cudaMemsetAsync(d_active, 0, sizeof(unsigned long long) * 2 * G, st));
    CK(cudaEventRecord(ctx->ev[1], st));
    int64_t launches = 0;
    for (int g = 0; g < G; ++g) {
        const Group &k = gr[g];
        refine_init_kernel<<<(unsigned)((k.n + 255) / 256), 256, 0, st>>>(d_lo, d_hi, d_steps, d_msz, Tn, k.i0, k.n, P, q->src_x,
                                                                       q->src_y, d_a, d_b, k.fb, d_active + 2 * g);
        ++launches;
    }
    std::vector<unsigned long long> h_active(2 * G);
    CK(cudaMemcpyAsync(h_active.data(), d_active, sizeof(unsigned long long) * 2 * G, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    int64_t iters = 0;
    for (int g = 0; g < G; ++g) {
        if (h_active[2 * g] == 0) continue;
        double w;
        std::memcpy(&w, &h_active[2 * g + 1], sizeof(w)); // widest |c - d| of the group
        // passes until |c - d| * GOLD_RATIO^k <= GOLD_TOL, + 1: the rounding of the bracket arithmetic can delay the
        // last pass of a pair by one (a pair that has converged takes no steps, so the spare pass costs nothing)
        int k = 0;
        while (w > RTB_GOLD_TOL && k < 4096) { w *= RTB_GOLD_RATIO; ++k; }
        gr[g].iters = spec ? (k + 2) / 2 : k + 1; // launches: two passes each when speculating
        iters = std::max<int64_t>(iters, k);
    }
    // fork: the side streams start after the brackets are initialised
    CK(cudaEventRecord(ctx->shoot_ev[0], st));
    for (int g = 1; g < G; ++g) CK(cudaStreamWaitEvent(gr[g].st, ctx->shoot_ev[0], 0));
    CK(cudaMemsetAsync(d_active, 0, sizeof(unsigned long long) * 2 * G, st));
    int max_it = 0;
    for (auto &k : gr) max_it = std::max(max_it, k.iters);
    // golden's while loop (RT:188): every pass re-traces both probes of every still-active pair; the groups' passes
    // are enqueued round-robin so that the device always has several independent launches to interleave
    for (int it = 0; it < max_it; ++it) {
        for (int g = 0; g < G; ++g) {
            const Group &k = gr[g];
            if (it >= k.iters) continue;
            if ((rc = launch_fan(ctx, p, k.fb, P * k.n, M, k.st, false))) return bail(rc);
            refine_update_kernel<<<(unsigned)((k.n + 255) / 256), 256, 0, k.st>>>(d_tx, d_ty, d_msz, Tn, k.i0, k.n, P, d_a, d_b, k.fb,
                                                                               d_active + 2 * g + (it == k.iters - 1 ? 0 : 1), // [2g]: pairs left after the last pass; [2g+1]: sink
                                                                               ctx->d_total);
            launches += 2;
        }
    }
    // join
    for (int g = 1; g < G; ++g) {
        CK(cudaEventRecord(ctx->shoot_ev[g], gr[g].st));
        CK(cudaStreamWaitEvent(st, ctx->shoot_ev[g], 0));
    }
    CK(cudaMemcpyAsync(h_active.data(), d_active, sizeof(unsigned long long) * 2 * G, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    for (int g = 0; g < G; ++g)
        if (h_active[2 * g] != 0)
            return fail(ctx, RTB_ERR_ARG, "rtb_shoot_refine: golden did not converge in the predicted number of passes");
    const unsigned gn = (unsigned)((n + 255) / 256);
    FanBuffers fb = alloc_fan(sg, n); // the final re-trace of theta* = (a+b)/2, all pairs in one launch
    if (sg.err != cudaSuccess || !fb.exit_out) return fail(ctx, RTB_ERR_NOMEM, "rtb_shoot_refine: staging", sg.err);
    refine_final_fill_kernel<<<gn, 256, 0, st>>>(d_a, d_b, d_steps, d_msz, Tn, n, q->src_x, q->src_y, d_theta, fb);
    if ((rc = launch_fan(ctx, p, fb, n, M))) return bail(rc);
    refine_result_kernel<<<gn, 256, 0, st>>>(d_tx, d_ty, Tn, n, fb, d_res);
    launches += 3;
    CK(cudaGetLastError());
    CK(cudaEventRecord(ctx->ev[2], st));
    sg.copy_back();
    unsigned long long total = 0;
    CK(cudaMemcpyAsync(&total, ctx->d_total, sizeof(total), cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(ctx->ev[3], st));
    CK(cudaStreamSynchronize(st));
    if (sg.err != cudaSuccess) return fail(ctx, RTB_ERR_CUDA, "rtb_shoot_refine: copy", sg.err);
    if (iterations) *iterations = iters;
    finish_stats(ctx, stats, sg, total, launches);
    return RTB_OK;
}
