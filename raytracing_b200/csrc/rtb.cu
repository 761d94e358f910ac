// rtb.cu -- C ABI (include/rtb.h) of the B200 ray-stepping engine: context, field upload,
// staging, kernel dispatch, Snell metric, FP64 peak probe.  No CPU fallback anywhere: every
// entry point needs a CUDA device and reports failures through the return code.
#include "ctx.cuh"
#include "cell_tables.cuh"

typedef CUresult (*cuTensorMapEncodeTiled_t)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                             const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                             CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                             CUtensorMapFloatOOBfill);

extern "C" int rtb_abi_version(void) { return RTB_ABI_VERSION; }

// Hash of the CUDA sources this library was compiled from (csrc/Makefile writes build_id.h before compiling):
// __graft_entry__.build()/smoke() and tests/test_cabi.py compare it with the hash of the sources in the tree, so a
// stale librtb.so cannot pass for the current code.
#include "build_id.h"
extern "C" const char *rtb_build_id(void) { return RTB_BUILD_ID; }
// Digest of the trace-kernel sources alone (tools/build_id.py KERNEL_SOURCES): the key of profiles/ncu_latest.json.
extern "C" const char *rtb_kernel_id(void) { return RTB_KERNEL_ID; }

extern "C" const char *rtb_last_error(rtb_ctx *ctx) { return ctx ? ctx->err.c_str() : "null context"; }

extern "C" int rtb_create(int device, rtb_ctx **out)
{
    if (!out) return RTB_ERR_ARG;
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev) return RTB_ERR_CUDA;
    rtb_ctx *ctx = new rtb_ctx();
    ctx->device = device;
    if (cudaSetDevice(device) != cudaSuccess ||
        cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete ctx;
        return RTB_ERR_CUDA;
    }
    ctx->own_stream = true;
    for (auto &ev : ctx->ev) cudaEventCreate(&ev);
    if (cudaMalloc(&ctx->d_total, sizeof(unsigned long long) * RTB_COUNTER_SLOTS) != cudaSuccess) {
        delete ctx;
        return RTB_ERR_CUDA;
    }
    cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
    *out = ctx;
    return RTB_OK;
}

void rtb_free_field(rtb_ctx *ctx)
{
    for (void *p : ctx->field_allocs) cudaFree(p);
    ctx->field_allocs.clear();
    ctx->has_field = false;
}

extern "C" void rtb_destroy(rtb_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    rtb_free_field(ctx);
    for (auto &b : ctx->pool) cudaFree(b.p);
    if (ctx->d_total) cudaFree(ctx->d_total);
    for (auto &ev : ctx->ev) if (ev) cudaEventDestroy(ev);
    if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    for (auto &ps : ctx->pipe) if (ps) cudaStreamDestroy(ps);
    for (auto &pe : ctx->pipe_ev) if (pe) cudaEventDestroy(pe);
    for (auto &bb : ctx->bounce) if (bb) cudaFreeHost(bb);
    for (auto &be : ctx->bounce_ev) if (be) cudaEventDestroy(be);
    for (auto &ps : ctx->shoot_st) if (ps) cudaStreamDestroy(ps);
    for (auto &pe : ctx->shoot_ev) if (pe) cudaEventDestroy(pe);
    delete ctx;
}

extern "C" int rtb_set_option(rtb_ctx *ctx, int32_t option, int32_t value)
{
    if (!ctx) return RTB_ERR_ARG;
    if (option == RTB_OPT_QUEUE && value >= 0 && value <= 2) { ctx->queue_mode = value; return RTB_OK; }
    if (option == RTB_OPT_TRAJ_STORE && value >= 0 && value <= 1) { ctx->traj_mode = value; return RTB_OK; }
    if (option == RTB_OPT_SHOOT_SPECULATE && value >= 0 && value <= 2) { ctx->shoot_spec = value; return RTB_OK; }
    if (option == RTB_OPT_SHOOT_GROUPS && value >= 0 && value <= rtb_ctx::kShootStreams + 1) { ctx->shoot_groups = value; return RTB_OK; }
    return fail(ctx, RTB_ERR_ARG, "rtb_set_option: unknown option or value");
}

extern "C" int rtb_set_stream(rtb_ctx *ctx, void *cuda_stream)
{
    if (!ctx) return RTB_ERR_ARG;
    cudaSetDevice(ctx->device);
    if (ctx->own_stream && ctx->stream) { cudaStreamDestroy(ctx->stream); ctx->own_stream = false; }
    if (cuda_stream) {
        ctx->stream = (cudaStream_t)cuda_stream;
    } else {
        CK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
        ctx->own_stream = true;
    }
    return RTB_OK;
}

// Knot-vector metadata for the device interval guess and the uniform-basis fast path.
int rtb_make_axis(rtb_ctx *ctx, const double *t, int n, int k, DevAxis *ax)
{
    if (!t || n < 2 * (k + 1)) return fail(ctx, RTB_ERR_ARG, "spline: too few knots");
    for (int i = 0; i + 1 < n; ++i)
        if (t[i + 1] < t[i]) return fail(ctx, RTB_ERR_ARG, "spline: knots not sorted");
    std::vector<double> d;
    for (int i = k; i < n - k - 1; ++i) d.push_back(t[i + 1] - t[i]);
    std::vector<double> ds(d);
    std::nth_element(ds.begin(), ds.begin() + ds.size() / 2, ds.end());
    const double h = ds[ds.size() / 2];
    if (!(h > 0)) return fail(ctx, RTB_ERR_ARG, "spline: degenerate knots");
    // longest run [A,B] of intervals t[j+1]-t[j] equal to h within 1e-12 relative
    int bestA = 0, bestB = -1, A = -1;
    for (int j = 0; j < n - 1; ++j) {
        const bool u = std::fabs((t[j + 1] - t[j]) - h) <= 1e-12 * h;
        if (u && A < 0) A = j;
        if ((!u || j == n - 2) && A >= 0) {
            const int B = u ? j : j - 1;
            if (B - A > bestB - bestA) { bestA = A; bestB = B; }
            A = -1;
        }
    }
    double *dt = nullptr;
    CK(cudaMalloc(&dt, sizeof(double) * n));
    ctx->field_allocs.push_back(dt);
    CK(cudaMemcpy(dt, t, sizeof(double) * n, cudaMemcpyHostToDevice));
    ax->t = dt; ax->n = n; ax->k = k;
    ax->tb = t[k]; ax->te = t[n - k - 1];
    ax->inv_h = 1.0 / h;
    if (bestB >= bestA) {
        ax->inv_h = (double)(bestB + 1 - bestA) / (t[bestB + 1] - t[bestA]); // spacing over the whole run
        ax->t0 = t[bestA]; ax->off = bestA;
        ax->lo = bestA + k - 1; ax->hi = bestB - k + 1;
    } else {
        ax->t0 = t[k]; ax->off = k; ax->lo = 1; ax->hi = 0;
    }
    ax->c0 = (double)ax->off - ax->t0 * ax->inv_h - 0.5;
    // the default evaluator takes the cell fraction from t0 + j/inv_h instead of the stored knots: allowed only if
    // they agree to 1e-12 of a cell over the uniform run (linspace knots: ~2e-13), else the run is disabled
    for (int l = ax->lo; l <= ax->hi + 1 && ax->hi >= ax->lo; ++l) {
        const double dev = (t[l] - ax->t0) * ax->inv_h - (double)(l - ax->off);
        if (std::fabs(dev) > 1e-12) { ax->lo = 1; ax->hi = 0; break; }
    }
    return RTB_OK;
}

// Second half of building a field, shared by rtb_set_field and rtb_fit_field (fit.cu): detects the region where one
// interval / fraction per physical axis serves all three splines, expands the per-cell tables and publishes `f`.
// tzu/tzv/tgu/tgv are HOST copies of the knots whose device copies the axes of `f` already hold.
int rtb_finish_field(rtb_ctx *ctx, DevField &f, const double *tzu, int nzu, const double *tzv, int nzv,
                     const double *tgu, const double *tgv)
{
    // fast path: bilinear knots == cubic knots shifted by one over the uniform cubic range
    auto shared_axis = [](const double *tz, int nz, const DevAxis &az, const double *tg, const DevAxis &ag, int &lo, int &hi) {
        lo = ag.lo; hi = ag.hi;
        if (hi < lo) return false;
        if (lo - 1 < az.lo) lo = az.lo + 1;
        if (hi - 1 > az.hi) hi = az.hi + 1;
        if (hi < lo || hi + 1 >= nz + 1) return false;
        for (int l = lo; l <= hi; ++l)
            if (tz[l - 1] != tg[l] || tz[l] != tg[l + 1]) return false;
        return true;
    };
    f.fast_ok = shared_axis(tzu, nzu, f.zu, tgu, f.gu_, f.flo_u, f.fhi_u) &&
                shared_axis(tzv, nzv, f.zv, tgv, f.gv_, f.flo_v, f.fhi_v);
    // per-cell polynomial tables of the fast region (cell_tables.cuh); without them the general evaluator is used
    f.pg = nullptr; f.pstride = 0;
    if (f.fast_ok) {
        const size_t ncell = (size_t)(f.fhi_u - f.flo_u + 1) * (size_t)(f.fhi_v - f.flo_v + 1);
        double2 *pg = nullptr;
        if (cudaMalloc(&pg, ncell * RTB_CELL_REC * sizeof(double2)) == cudaSuccess) ctx->field_allocs.push_back(pg);
        if (pg) {
            f.pg = pg; f.pstride = f.fhi_v - f.flo_v + 1;
            cell_table_kernel<<<(unsigned)((ncell + 127) / 128), 128, 0, ctx->stream>>>(f, pg);
            CK(cudaGetLastError());
            CK(cudaStreamSynchronize(ctx->stream));
        } else {
            (void)cudaGetLastError(); // out of memory for the tables: correct but slow general path
            f.fast_ok = 0;
        }
    }
    ctx->f = f;
    CK(cudaMalloc(&ctx->f_dev, sizeof(DevField)));
    ctx->field_allocs.push_back(ctx->f_dev);
    CK(cudaMemcpy(ctx->f_dev, &f, sizeof(DevField), cudaMemcpyHostToDevice));
    ctx->has_field = true;
    return RTB_OK;
}

extern "C" int rtb_set_field(rtb_ctx *ctx, const rtb_spline *z, const rtb_spline *gu, const rtb_spline *gv)
{
    if (!ctx || !z || !gu || !gv) return fail(ctx, RTB_ERR_ARG, "rtb_set_field: null argument");
    if (z->ku != 1 || z->kv != 1 || gu->ku != 3 || gu->kv != 3 || gv->ku != 3 || gv->kv != 3)
        return fail(ctx, RTB_ERR_ARG, "rtb_set_field: expected bilinear z and bicubic grd (RT_bench.py:455-457)");
    if (gu->nu != gv->nu || gu->nv != gv->nv || std::memcmp(gu->tu, gv->tu, sizeof(double) * gu->nu) ||
        std::memcmp(gu->tv, gv->tv, sizeof(double) * gu->nv))
        return fail(ctx, RTB_ERR_ARG, "rtb_set_field: grd[0] and grd[1] must share their knots");
    CK(cudaSetDevice(ctx->device));
    rtb_free_field(ctx);
    DevField f;
    int rc;
    if ((rc = rtb_make_axis(ctx, z->tu, z->nu, 1, &f.zu))) return rc;
    if ((rc = rtb_make_axis(ctx, z->tv, z->nv, 1, &f.zv))) return rc;
    if ((rc = rtb_make_axis(ctx, gu->tu, gu->nu, 3, &f.gu_))) return rc;
    if ((rc = rtb_make_axis(ctx, gu->tv, gu->nv, 3, &f.gv_))) return rc;
    f.ncz = z->nv - 2;
    f.ncg = gu->nv - 4;
    const size_t nz = (size_t)(z->nu - 2) * f.ncz, ng = (size_t)(gu->nu - 4) * f.ncg;
    double *dz = nullptr;
    double2 *dg = nullptr;
    CK(cudaMalloc(&dz, nz * sizeof(double)));
    ctx->field_allocs.push_back(dz);
    CK(cudaMalloc(&dg, ng * sizeof(double2)));
    ctx->field_allocs.push_back(dg);
    CK(cudaMemcpy(dz, z->c, nz * sizeof(double), cudaMemcpyHostToDevice));
    std::vector<double2> inter(ng);
    for (size_t i = 0; i < ng; ++i) inter[i] = make_double2(gu->c[i], gv->c[i]);
    CK(cudaMemcpy(dg, inter.data(), ng * sizeof(double2), cudaMemcpyHostToDevice));
    f.cz = dz;
    f.cg = dg;
    return rtb_finish_field(ctx, f, z->tu, z->nu, z->tv, z->nv, gu->tu, gu->tv);
}

template <bool EXACT>
__global__ void field_eval_kernel(DevField f, const DevField *fd, const double *x, const double *y, int64_t n, double *on,
                                  double *ogx, double *ogy)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double vn, gx, gy;
    if (EXACT) n_gradient<true>(f, x[i], y[i], vn, gx, gy);
    else n_gradient_fast(f, fd, x[i], y[i], vn, gx, gy);
    on[i] = vn; ogx[i] = gx; ogy[i] = gy;
}

extern "C" int rtb_field_eval(rtb_ctx *ctx, int32_t ptr_kind, uint32_t flags, const double *x, const double *y,
                              int64_t n, double *out_n, double *out_gx, double *out_gy)
{
    if (!ctx) return RTB_ERR_ARG;
    if (!ctx->has_field) return fail(ctx, RTB_ERR_NOFIELD, "rtb_field_eval: no field set");
    if (n < 0 || !x || !y || !out_n || !out_gx || !out_gy) return fail(ctx, RTB_ERR_ARG, "rtb_field_eval: null argument");
    if (n == 0) return RTB_OK;
    CK(cudaSetDevice(ctx->device));
    Staging st(ctx, ptr_kind == RTB_PTR_HOST);
    const double *dx = st.in(x, n), *dy = st.in(y, n);
    double *dn = st.out(out_n, n), *dgx = st.out(out_gx, n), *dgy = st.out(out_gy, n);
    if (st.err != cudaSuccess) return fail(ctx, RTB_ERR_CUDA, "rtb_field_eval: staging", st.err);
    const unsigned grid = (unsigned)((n + 127) / 128);
    if (flags & RTB_F_EXACT_BASIS) field_eval_kernel<true><<<grid, 128, 0, ctx->stream>>>(ctx->f, ctx->f_dev, dx, dy, n, dn, dgx, dgy);
    else field_eval_kernel<false><<<grid, 128, 0, ctx->stream>>>(ctx->f, ctx->f_dev, dx, dy, n, dn, dgx, dgy);
    CK(cudaGetLastError());
    st.copy_back();
    CK(cudaStreamSynchronize(ctx->stream));
    if (st.err != cudaSuccess) return fail(ctx, RTB_ERR_CUDA, "rtb_field_eval: copy", st.err);
    return RTB_OK;
}

// ---------------------------------------------------------------------------------------
// Snell error metric RT_bench.py:896-919
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ double snell_expected(double theta0)
{
    if (theta0 < RTB_PI / 4) return 90 - 180 * theta0 / RTB_PI;
    if (theta0 == RTB_PI / 4) return 0.0;
    return 180 * asin(sqrt(2.0) * sin(RTB_PI / 2 - theta0)) / RTB_PI;
}
__device__ __forceinline__ void snell_rows(double di, int &ia, int &ib)
{
    const long long i = (long long)di;
    ib = (int)(9.5 * (double)i / 10); // int(9.5*(i)/10) RT:913
    ia = (int)((double)(9 * i) / 10); // int(9*(i)/10)
}
__global__ void snell_rows_kernel(const double *d_ray, int64_t R, int *cap_rows)
{
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    int ia, ib;
    snell_rows(d_ray[2 * R + r], ia, ib);
    cap_rows[r] = ia;
    cap_rows[R + r] = ib;
}
__global__ void snell_from_traj_kernel(const double *s_ray, const double *d_ray, const double *theta0, int64_t cand_R,
                                       int64_t R, double *errors)
{
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    int ia, ib;
    snell_rows(d_ray[2 * R + r], ia, ib);
    const double distx = s_ray[((size_t)ib * 6 + 0) * R + r] - s_ray[((size_t)ia * 6 + 0) * R + r];
    const double disty = s_ray[((size_t)ib * 6 + 1) * R + r] - s_ray[((size_t)ia * 6 + 1) * R + r];
    const double angsim = 180 * atan(fabs(distx / disty)) / RTB_PI;
    errors[r] = fabs(angsim - snell_expected(theta0[cand_R > 0 ? r % cand_R : r]));
}
__global__ void snell_from_cap_kernel(const double *cap, const double *theta0, int64_t cand_R, int64_t R, double *errors)
{
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    const double distx = cap[2 * R + r] - cap[r];
    const double disty = cap[3 * R + r] - cap[R + r];
    const double angsim = 180 * atan(fabs(distx / disty)) / RTB_PI;
    errors[r] = fabs(angsim - snell_expected(theta0[cand_R > 0 ? r % cand_R : r]));
}

// ---------------------------------------------------------------------------------------
// rtb_trace: trazar RT_bench.py:766-948
// ---------------------------------------------------------------------------------------
// Launch configuration of one integration kernel launch: which kernel (static grid / ray queue), block size, and --
// for the queue kernel -- a freshly zeroed queue head.
int rtb_launch(rtb_ctx *ctx, cudaStream_t st, const rtb_params *p, TraceArgs a, bool traj)
{
    LaunchCfg c;
    c.iso = p->gamma == 1.0;
    c.traj = traj;
    c.exact = (p->flags & RTB_F_EXACT_BASIS) != 0;
    c.block = a.lane_shift ? 32 : ctx->block;
    // Ray queue + warp compaction: FINAL-type outputs of the default kernels on fans of at least a wave; the static
    // grid keeps trajectories (coalesced rows need thread <-> ray fixed), the faithful kernels and spread small fans.
    // Opt-in (RTB_F_RAY_QUEUE per call, RTB_OPT_QUEUE 1 per context): on every fan measured so far -- angle-sorted,
    // shuffled, DELTA_S candidates mixed inside warps -- the static grid is 1-12 % faster (profiles/r2c_time_r2.txt):
    // rays that diverge in length also diverge in position, and then the step is bound by the 32 different cell records
    // a warp has to fetch, not by idle lanes; the hardware block scheduler already refills SMs block by block.
    const bool can_queue = !traj && !c.exact && a.lane_shift == 0 && a.stage && (p->flags & RTB_F_STATIC_GRID) == 0;
    c.queue = can_queue && (ctx->queue_mode == 1 || (p->flags & RTB_F_RAY_QUEUE) != 0);
    c.queue_blocks = ctx->sm_count * RTB_QUEUE_MIN_BLOCKS;
    if (c.queue) {
        const int64_t need = (a.R + RTB_BLOCK - 1) / RTB_BLOCK;
        if (need < c.queue_blocks) c.queue_blocks = (int)need;
        ctx->queue_slot = 2 + (ctx->queue_slot - 2 + 1) % (RTB_COUNTER_SLOTS - 2);
        a.queue_head = ctx->d_total + ctx->queue_slot;
        CK(cudaMemsetAsync(a.queue_head, 0, sizeof(unsigned long long), st));
    }
    // Trajectory rows through shared memory + 2-D TMA stores (traj_stage.cuh): s_ray as the matrix (rows*6, R) of
    // doubles, tiles of (6*RTB_TRAJ_ROWS) x 32 and single rows of 6 x 32.  Needs 16-byte aligned rows.
    a.traj_bulk = 0;
    const int64_t rows = traj ? ((int64_t)a.max_size + a.out_stride - 1) / a.out_stride : 0;
    if (traj && ctx->traj_mode == 1 && a.lane_shift == 0 && (a.R % 2) == 0 && ((uintptr_t)a.s_ray % 16) == 0 &&
        a.R < ((int64_t)1 << 31) && rows * 6 < ((int64_t)1 << 31)) {
        static cuTensorMapEncodeTiled_t encode = nullptr;
        if (!encode) {
            void *fn = nullptr;
            cudaDriverEntryPointQueryResult qr;
            if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) == cudaSuccess &&
                qr == cudaDriverEntryPointSuccess)
                encode = (cuTensorMapEncodeTiled_t)fn;
            (void)cudaGetLastError();
        }
        if (encode) {
            const cuuint64_t dims[2] = {(cuuint64_t)a.R, (cuuint64_t)(rows * 6)};
            const cuuint64_t strides[1] = {(cuuint64_t)a.R * sizeof(double)};
            const cuuint32_t box[2] = {32, 6 * RTB_TRAJ_ROWS}, one[2] = {32, 6}, es[2] = {1, 1};
            const CUresult r1 = encode(&a.tmaps.box, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, a.s_ray, dims, strides, box, es,
                                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                       CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            const CUresult r2 = encode(&a.tmaps.row, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, a.s_ray, dims, strides, one, es,
                                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                       CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            a.traj_bulk = (r1 == CUDA_SUCCESS && r2 == CUDA_SUCCESS) ? 1 : 0;
        }
    }
    CK(kLaunch[p->method](a, c, st));
    return RTB_OK;
}

// The kernels of one trace on stream `st`, every pointer of `a` on the device: the integration and, when asked, the
// Snell error metric (RT_bench.py:896-919).
int rtb_trace_core(rtb_ctx *ctx, cudaStream_t st, const rtb_params *p, const TraceArgs &a, double *d_err,
                   bool has_errors, bool want_err, bool traj, int stride, Staging &sg, int64_t &launches)
{
    const int64_t R = a.R;
    const bool two_pass = has_errors && want_err && !(traj && stride == 1);
    int rc;
    TraceArgs a1 = a;
    if (two_pass && !traj) { // a trajectory pass writes no checkpoints: pass 2 then integrates from the origin
        // pass 1 leaves the ray state every 2^shift steps so that pass 2 starts near its first sample row instead
        // of at the ray's origin: at most 32 checkpoints of 136 B per ray, 1 GB in all, else coarser or none
        int shift = 8;
        while (((int64_t)a.max_size >> shift) > 32) ++shift;
        while (shift < 30 && ((int64_t)a.max_size >> shift) * RTB_CKPT_FIELDS * 8 * R > ((int64_t)1 << 30)) ++shift;
        const int rows = (int)(((int64_t)a.max_size - 1) >> shift);
        if (rows > 0) {
            a1.ckpt = (double *)sg.alloc((size_t)rows * RTB_CKPT_FIELDS * 8 * R);
            if (a1.ckpt) { a1.ckpt_shift = shift; a1.ckpt_rows = rows; }
            else { (void)cudaGetLastError(); sg.err = cudaSuccess; } // no room: pass 2 integrates from the origin
        }
    }
    if ((rc = rtb_launch(ctx, st, p, a1, traj))) return rc;
    ++launches;
    const unsigned g1 = (unsigned)((R + 127) / 128);
    if (has_errors) {
        if (!want_err) {
            CK(cudaMemsetAsync(d_err, 0, sizeof(double) * R, st));
        } else if (traj && stride == 1) {
            snell_from_traj_kernel<<<g1, 128, 0, st>>>(a.s_ray, a.d_ray, a.theta0, a.cand_R, R, d_err);
            ++launches;
        } else {
            // two-pass: the sample rows int(9i/10), int(9.5i/10) depend on the final step count i,
            // so re-integrate (bit-identical arithmetic) up to the later row and capture both.
            int *cap_rows = (int *)sg.alloc(sizeof(int) * 2 * R);
            double *cap_out = (double *)sg.alloc(sizeof(double) * 4 * R);
            if (!cap_rows || !cap_out) return fail(ctx, RTB_ERR_NOMEM, "rtb_trace: capture buffers", sg.err);
            snell_rows_kernel<<<g1, 128, 0, st>>>(a.d_ray, R, cap_rows);
            TraceArgs b = a1;
            b.s_ray = nullptr; b.d_ray = nullptr; b.last_row = nullptr; b.px_stats = nullptr; b.exit_out = nullptr;
            b.total_steps = nullptr; b.nonfinite = nullptr; b.cap_rows = cap_rows; b.cap_out = cap_out;
            if ((rc = rtb_launch(ctx, st, p, b, false))) return rc;
            snell_from_cap_kernel<<<g1, 128, 0, st>>>(cap_out, a.theta0, a.cand_R, R, d_err);
            launches += 3;
        }
        CK(cudaGetLastError());
    }
    return RTB_OK;
}

// Common part of the TraceArgs of a call.
TraceArgs rtb_base_args(rtb_ctx *ctx, const rtb_params *p, int stride)
{
    TraceArgs base;
    std::memset(&base, 0, sizeof(base));
    base.f = ctx->f;
    base.f_dev = ctx->f_dev;
    base.max_size = (int)p->max_size; base.out_stride = stride;
    base.tail_zero = (p->flags & RTB_F_NO_TAIL_ZERO) ? 0 : 1;
    base.step = p->step; base.gamma = p->gamma;
    base.limx_i = p->limx_i; base.limx_s = p->limx_s; base.limy_i = p->limy_i; base.limy_s = p->limy_s;
    base.in = make_cell_box(ctx->f, p->limx_i, p->limx_s, p->limy_i, p->limy_s);
    base.total_steps = ctx->d_total;
    base.nonfinite = ctx->d_total + 1;
    base.gather_bshift = -1;
    return base;
}

extern "C" int rtb_trace(rtb_ctx *ctx, const rtb_params *p, const double *theta0, const double *x0,
                         const double *y0, int64_t R, double *s_ray, double *d_ray, double *errors,
                         double *last_row, double *px_stats, rtb_stats *stats)
{
    if (!ctx) return RTB_ERR_ARG;
    if (!ctx->has_field) return fail(ctx, RTB_ERR_NOFIELD, "rtb_trace: no field set (rtb_set_field)");
    if (!p || !theta0 || !x0 || !y0 || !d_ray) return fail(ctx, RTB_ERR_ARG, "rtb_trace: null argument");
    if (p->method < 1 || p->method > 11) return fail(ctx, RTB_ERR_ARG, "rtb_trace: method must be 1..11");
    if (R < 0 || p->max_size < 1 || p->max_size > 0x7fffffff) return fail(ctx, RTB_ERR_ARG, "rtb_trace: bad R/max_size");
    if (s_ray && (p->out_stride < 1)) return fail(ctx, RTB_ERR_ARG, "rtb_trace: out_stride must be >= 1");
    if (s_ray && (p->steps || p->max_sizes)) return fail(ctx, RTB_ERR_ARG, "rtb_trace: s_ray with per-ray steps");
    if (!(p->step > 0) && !p->steps) return fail(ctx, RTB_ERR_ARG, "rtb_trace: step must be > 0");
    if ((p->flags & RTB_F_SORT_FAN) && s_ray) return fail(ctx, RTB_ERR_ARG, "rtb_trace: RTB_F_SORT_FAN with s_ray (FINAL-type outputs only)");
    if (stats) std::memset(stats, 0, sizeof(*stats));
    if (R == 0) return RTB_OK;
    if ((p->flags & RTB_F_SORT_FAN) && R > 1) return rtb_trace_sorted(ctx, p, theta0, x0, y0, R, d_ray, errors, last_row, px_stats, stats);
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const bool host = p->ptr_kind == RTB_PTR_HOST;
    const bool traj = s_ray != nullptr;
    const bool want_err = errors && (p->flags & RTB_F_INTERFACE);
    const int stride = traj ? p->out_stride : 1;
    const int64_t rows = traj ? (p->max_size + stride - 1) / stride : 0;

    CK(cudaEventRecord(ctx->ev[0], st));
    CK(cudaMemsetAsync(ctx->d_total, 0, 2 * sizeof(unsigned long long), st));
    TraceArgs base = rtb_base_args(ctx, p, stride);
    if (ctx->gather_world > 0 && !host && d_ray) {
        if (ctx->gather_bshift < 0 && ctx->gather_off + R > ctx->gather_R)
            return fail(ctx, RTB_ERR_ARG, "rtb_trace: fan exceeds the gather window (rtb_set_gather)");
        for (int pe = 0; pe < ctx->gather_world; ++pe) base.gather[pe] = ctx->gather[pe];
        base.gather_world = ctx->gather_world; base.gather_R = ctx->gather_R; base.gather_off = ctx->gather_off;
        base.gather_bshift = ctx->gather_bshift; base.gather_nranks = ctx->gather_nranks; base.gather_rank = ctx->gather_rank;
    }
    base.stage = R >= (1 << 16); // below ~one wave of warps a launch is a latency chain: no staging detour
    int64_t launches = 0, h2d = 0, d2h = 0;
    unsigned long long counters[2] = {0, 0};
    int rc;

    // Host buffers, no trajectory, a large fan: cut the rays into chunks and run each chunk's H2D copy, kernels and
    // D2H copy on one of three side streams, so that the copies of one chunk overlap the stepping of the others
    // (rays are independent: the results are those of the single launch, bit for bit).
    const int64_t kPipeMinRays = 1 << 17;
    if (host && !traj && R >= kPipeMinRays) {
        for (auto &ps : ctx->pipe) if (!ps) CK(cudaStreamCreateWithFlags(&ps, cudaStreamNonBlocking));
        for (auto &pe : ctx->pipe_ev) if (!pe) CK(cudaEventCreateWithFlags(&pe, cudaEventDisableTiming));
        const int64_t nchunk = std::min<int64_t>(8, R / (kPipeMinRays / 2));
        const int64_t chunk = (((R + nchunk - 1) / nchunk) + 127) / 128 * 128;
        CK(cudaEventRecord(ctx->ev[1], st));
        CK(cudaEventRecord(ctx->pipe_ev[0], st));
        for (auto &ps : ctx->pipe) CK(cudaStreamWaitEvent(ps, ctx->pipe_ev[0], 0));
        // On any failure the side streams are drained before the staging buffers go back to the pool (`bail`).
        std::vector<std::unique_ptr<Staging>> sgs;
        auto bail = [&](int code) {
            for (auto &ps : ctx->pipe) cudaStreamSynchronize(ps);
            return code;
        };
        int k = 0;
        for (int64_t r0 = 0; r0 < R; r0 += chunk, ++k) {
            const int64_t rn = std::min(chunk, R - r0);
            cudaStream_t ps = ctx->pipe[k % 3];
            sgs.emplace_back(new Staging(ctx, true, ps));
            Staging &sg = *sgs.back();
            TraceArgs a = base;
            a.theta0 = sg.in(theta0 + r0, rn); a.x0 = sg.in(x0 + r0, rn); a.y0 = sg.in(y0 + r0, rn);
            a.steps = sg.in(p->steps ? p->steps + r0 : nullptr, rn);
            a.max_sizes = sg.in(p->max_sizes ? p->max_sizes + r0 : nullptr, rn);
            a.R = rn;
            a.d_ray = sg.out_cols(d_ray, 3, R, r0, rn);
            a.last_row = sg.out_cols(last_row, 6, R, r0, rn);
            a.px_stats = sg.out_cols(px_stats, 3, R, r0, rn);
            double *d_err = sg.out(errors ? errors + r0 : nullptr, (size_t)rn);
            if (sg.err != cudaSuccess)
                return bail(fail(ctx, sg.err == cudaErrorMemoryAllocation ? RTB_ERR_NOMEM : RTB_ERR_CUDA,
                                 "rtb_trace: staging", sg.err));
            if ((rc = rtb_trace_core(ctx, ps, p, a, d_err, errors != nullptr, want_err, false, stride, sg, launches)))
                return bail(rc);
            sg.copy_back();
        }
        cudaError_t je = cudaSuccess;
        for (int j = 0; j < 3 && je == cudaSuccess; ++j) {
            je = cudaEventRecord(ctx->pipe_ev[1 + j], ctx->pipe[j]);
            if (je == cudaSuccess) je = cudaStreamWaitEvent(st, ctx->pipe_ev[1 + j], 0);
        }
        if (je == cudaSuccess) je = cudaEventRecord(ctx->ev[2], st);
        if (je == cudaSuccess) je = cudaMemcpyAsync(counters, ctx->d_total, sizeof(counters), cudaMemcpyDeviceToHost, st);
        if (je == cudaSuccess) je = cudaEventRecord(ctx->ev[3], st);
        if (je == cudaSuccess) je = cudaStreamSynchronize(st);
        if (je != cudaSuccess) return bail(fail(ctx, RTB_ERR_CUDA, "rtb_trace: joining the chunk streams", je));
        for (auto &sg : sgs) {
            if (sg->err != cudaSuccess) return fail(ctx, RTB_ERR_CUDA, "rtb_trace: copy", sg->err);
            h2d += sg->h2d; d2h += sg->d2h;
        }
    } else {
        Staging sg(ctx, host);
        // an error after the first launch must not hand buffers with work in flight back to the pool
        auto bail = [&](int code) { cudaStreamSynchronize(st); return code; };
        TraceArgs a = base;
        a.theta0 = sg.in(theta0, R); a.x0 = sg.in(x0, R); a.y0 = sg.in(y0, R);
        a.steps = sg.in(p->steps, R); a.max_sizes = sg.in(p->max_sizes, R);
        a.R = R;
        a.lane_shift = pick_lane_shift(R);
        a.s_ray = sg.out(s_ray, (size_t)rows * 6 * R);
        a.d_ray = sg.out(d_ray, (size_t)3 * R);
        a.last_row = sg.out(last_row, (size_t)6 * R);
        a.px_stats = sg.out(px_stats, (size_t)3 * R);
        double *d_err = sg.out(errors, (size_t)R);
        if (sg.err != cudaSuccess) return bail(fail(ctx, sg.err == cudaErrorMemoryAllocation ? RTB_ERR_NOMEM : RTB_ERR_CUDA,
                                                     "rtb_trace: staging", sg.err));
        CK(cudaEventRecord(ctx->ev[1], st));
        if ((rc = rtb_trace_core(ctx, st, p, a, d_err, errors != nullptr, want_err, traj, stride, sg, launches))) return bail(rc);
        CK(cudaEventRecord(ctx->ev[2], st));
        sg.copy_back();
        CK(cudaMemcpyAsync(counters, ctx->d_total, sizeof(counters), cudaMemcpyDeviceToHost, st));
        CK(cudaEventRecord(ctx->ev[3], st));
        CK(cudaStreamSynchronize(st));
        if (sg.err != cudaSuccess) return fail(ctx, RTB_ERR_CUDA, "rtb_trace: copy", sg.err);
        h2d = sg.h2d; d2h = sg.d2h;
    }
    if (stats) {
        float ms = 0;
        cudaEventElapsedTime(&ms, ctx->ev[1], ctx->ev[2]); stats->kernel_ms = ms;
        cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[3]); stats->total_ms = ms;
        stats->ray_steps = (int64_t)counters[0];
        stats->nonfinite_rays = (int64_t)counters[1];
        stats->launches = launches;
        stats->h2d_bytes = h2d; stats->d2h_bytes = d2h;
    }
    return RTB_OK;
}

// ---------------------------------------------------------------------------------------
// Fused multi-GPU result gather: buffers shared between the per-GPU processes of one node through CUDA IPC, written
// by the integration kernel of every rank (trace_kernel.cuh epilogue).
// ---------------------------------------------------------------------------------------
extern "C" int rtb_gather_alloc(rtb_ctx *ctx, int64_t bytes, void **ptr, void *handle64)
{
    if (!ctx || !ptr || bytes <= 0) return fail(ctx, RTB_ERR_ARG, "rtb_gather_alloc: bad argument");
    CK(cudaSetDevice(ctx->device));
    void *d = nullptr;
    cudaError_t e = cudaMalloc(&d, (size_t)bytes);
    if (e != cudaSuccess) return fail(ctx, RTB_ERR_NOMEM, "rtb_gather_alloc", e);
    if (handle64) {
        static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
        e = cudaIpcGetMemHandle((cudaIpcMemHandle_t *)handle64, d);
        if (e != cudaSuccess) { cudaFree(d); return fail(ctx, RTB_ERR_CUDA, "cudaIpcGetMemHandle", e); }
    }
    *ptr = d;
    return RTB_OK;
}

extern "C" int rtb_gather_open(rtb_ctx *ctx, const void *handle64, void **ptr)
{
    if (!ctx || !handle64 || !ptr) return fail(ctx, RTB_ERR_ARG, "rtb_gather_open: null argument");
    CK(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle64, sizeof(h));
    CK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return RTB_OK;
}

extern "C" int rtb_gather_close(rtb_ctx *ctx, void *ptr, int32_t own)
{
    if (!ctx || !ptr) return RTB_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    if (own) CK(cudaFree(ptr));
    else CK(cudaIpcCloseMemHandle(ptr));
    return RTB_OK;
}

extern "C" int rtb_set_gather(rtb_ctx *ctx, int32_t world, double *const *bases, int64_t R_total, int64_t offset)
{
    if (!ctx) return RTB_ERR_ARG;
    if (world < 0 || world > RTB_MAX_PEERS) return fail(ctx, RTB_ERR_ARG, "rtb_set_gather: world must be 0..8");
    if (world > 0 && (!bases || R_total <= 0 || offset < 0 || offset >= R_total))
        return fail(ctx, RTB_ERR_ARG, "rtb_set_gather: bad window");
    for (int pe = 0; pe < world; ++pe)
        if (!bases[pe]) return fail(ctx, RTB_ERR_ARG, "rtb_set_gather: null peer buffer");
    for (int pe = 0; pe < RTB_MAX_PEERS; ++pe) ctx->gather[pe] = pe < world ? bases[pe] : nullptr;
    ctx->gather_world = world; ctx->gather_R = R_total; ctx->gather_off = offset;
    ctx->gather_bshift = -1; ctx->gather_nranks = 1; ctx->gather_rank = 0;
    return RTB_OK;
}

extern "C" int rtb_set_gather_cyclic(rtb_ctx *ctx, int32_t world, double *const *bases, int64_t R_total, int64_t block,
                                     int32_t nranks, int32_t rank)
{
    if (!ctx) return RTB_ERR_ARG;
    if (world < 1 || world > RTB_MAX_PEERS || !bases || R_total <= 0)
        return fail(ctx, RTB_ERR_ARG, "rtb_set_gather_cyclic: bad window");
    if (block < 1 || (block & (block - 1)) != 0 || nranks < 1 || rank < 0 || rank >= nranks)
        return fail(ctx, RTB_ERR_ARG, "rtb_set_gather_cyclic: block must be a power of two, 0 <= rank < nranks");
    for (int pe = 0; pe < world; ++pe)
        if (!bases[pe]) return fail(ctx, RTB_ERR_ARG, "rtb_set_gather_cyclic: null peer buffer");
    for (int pe = 0; pe < RTB_MAX_PEERS; ++pe) ctx->gather[pe] = pe < world ? bases[pe] : nullptr;
    ctx->gather_world = world; ctx->gather_R = R_total; ctx->gather_off = 0;
    int b = 0;
    while (((int64_t)1 << b) < block) ++b;
    ctx->gather_bshift = b; ctx->gather_nranks = nranks; ctx->gather_rank = rank;
    return RTB_OK;
}

// ---------------------------------------------------------------------------------------
// FP64 pipe peak: 8 independent DFMA chains per thread, all SMs busy.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) dfma_peak_kernel(double *out, int iters, double a, double b)
{
    double v0 = threadIdx.x, v1 = v0 + 1, v2 = v0 + 2, v3 = v0 + 3, v4 = v0 + 4, v5 = v0 + 5, v6 = v0 + 6, v7 = v0 + 7;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            v0 = fma(v0, a, b); v1 = fma(v1, a, b); v2 = fma(v2, a, b); v3 = fma(v3, a, b);
            v4 = fma(v4, a, b); v5 = fma(v5, a, b); v6 = fma(v6, a, b); v7 = fma(v7, a, b);
        }
    }
    const double s = v0 + v1 + v2 + v3 + v4 + v5 + v6 + v7;
    if (s == 12345.678) out[0] = s; // never true; keeps the chains alive
}

extern "C" int rtb_fp64_peak(rtb_ctx *ctx, double *flops_per_s)
{
    if (!ctx || !flops_per_s) return RTB_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, ctx->device));
    double *d = nullptr;
    CK(cudaMalloc(&d, 8));
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 4096;
    double best = 0;
    for (int rep = 0; rep < 4; ++rep) {
        CK(cudaEventRecord(ctx->ev[0], ctx->stream));
        dfma_peak_kernel<<<blocks, threads, 0, ctx->stream>>>(d, iters, 0.999999, 1e-9);
        CK(cudaEventRecord(ctx->ev[1], ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        float ms = 0;
        cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
        const double fl = 2.0 * 64.0 * iters * (double)blocks * threads / (ms * 1e-3);
        if (rep > 0 && fl > best) best = fl;
    }
    cudaFree(d);
    *flops_per_s = best;
    return RTB_OK;
}
