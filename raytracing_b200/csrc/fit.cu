// fit.cu -- interpolacion() (RT_bench.py:435-464) on the device: np.gradient of the sampled index field and the
// three RectBivariateSpline fits, so that a field goes from the (my, mx) samples Z to trace-ready tables without
// scipy on the path.  On the host the reference spends 0.25 s here for the interface scenario
// (two FITPACK regrid calls on 1.08 M points) against 2 ms of ray stepping.
//
//   np.gradient(Z, DELTA, edge_order=2)   -> gradient_kernel, bit-exact (same operations in the same order)
//   RectBivariateSpline(y, x, Z, kx=1..)  -> s=0 interpolation with knots on the samples: c == Z
//   RectBivariateSpline(y, x, G, kx=3..)  -> FITPACK fpregr/fpgrre with s=0: knots t = [x0 x4, x[2..m-3], x[m-1] x4]
//       (the second and second-last sample are not knots) and coefficients solving  A_u C A_v^T = G, A = collocation
//       matrix of the cubic B-splines at the samples.  FITPACK triangularises A with Givens rotations; here the
//       O(m) part (collocation + banded LU, bandwidths 2/2, no pivoting: B-spline collocation matrices are totally
//       positive) runs on the host and the O(my*mx) part -- two sweeps of banded substitutions over all right-hand
//       sides, both gradient components at once as double2 -- on the device.  Same linear system, different
//       elimination order: coefficients agree with scipy's to a few ulp of max|G| (tests/test_gpu_fit.py).
#include "ctx.cuh"

namespace {

// ---------------------------------------------------------------------------------------
// np.gradient(f, delta, edge_order=2) for a 2-D array, uniform spacing (numpy/lib/function_base.py):
//   interior  (f[i+1] - f[i-1]) / (2*delta)
//   first     a*f[0] + b*f[1] + c*f[2],  a = -1.5/delta, b = 2/delta, c = -0.5/delta
//   last      a*f[-3] + b*f[-2] + c*f[-1], a = 0.5/delta, b = -2/delta, c = 1.5/delta
// numpy evaluates every product and sum as its own ufunc call, so nothing is fused: __dmul_rn/__dadd_rn.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ double grad1(const double *f, int64_t i, int64_t m, int64_t stride, double delta)
{
    if (i == 0) {
        const double a = __ddiv_rn(-1.5, delta), b = __ddiv_rn(2.0, delta), c = __ddiv_rn(-0.5, delta);
        return __dadd_rn(__dadd_rn(__dmul_rn(a, f[0]), __dmul_rn(b, f[stride])), __dmul_rn(c, f[2 * stride]));
    }
    if (i == m - 1) {
        const double a = __ddiv_rn(0.5, delta), b = __ddiv_rn(-2.0, delta), c = __ddiv_rn(1.5, delta);
        return __dadd_rn(__dadd_rn(__dmul_rn(a, f[-2 * stride]), __dmul_rn(b, f[-stride])), __dmul_rn(c, f[0]));
    }
    return __ddiv_rn(__dsub_rn(f[stride], f[-stride]), __dmul_rn(2.0, delta));
}

// G[i*mx + j] = {d/d(axis 0), d/d(axis 1)}; optionally also as two planes g0, g1.
__global__ void gradient_kernel(const double *__restrict__ Z, int64_t my, int64_t mx, double delta, double2 *G,
                                double *g0, double *g1)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= my * mx) return;
    const int64_t i = e / mx, j = e - i * mx;
    const double *f = Z + e;
    const double a0 = grad1(f, i, my, mx, delta), a1 = grad1(f, j, mx, 1, delta);
    if (G) G[e] = make_double2(a0, a1);
    if (g0) g0[e] = a0;
    if (g1) g1[e] = a1;
}

// ---------------------------------------------------------------------------------------
// Host: collocation matrix of the interpolating cubic spline and its banded LU.
// ---------------------------------------------------------------------------------------
// FITPACK fpbspl, degree 3: h[0..3] = B_{l-3..l}(x) for t[l] <= x <= t[l+1].
static void bspl3_host(const double *t, int l, double x, double *h)
{
    double hh[4];
    h[0] = 1.0;
    for (int j = 1; j <= 3; ++j) {
        for (int i = 0; i < j; ++i) hh[i] = h[i];
        h[0] = 0.0;
        for (int i = 1; i <= j; ++i) {
            const double tli = t[l + i], tlj = t[l + i - j];
            if (tli == tlj) { h[i] = 0.0; continue; }
            const double f = hh[i - 1] / (tli - tlj);
            h[i - 1] = h[i - 1] + f * (tli - x);
            h[i] = f * (x - tlj);
        }
    }
}

// Knots of the s=0 spline of degree k on samples x[0..m) (fpregr: every sample is a knot except, for k=3, the second
// and the second-last).
static std::vector<double> interp_knots(const double *x, int m, int k)
{
    std::vector<double> t;
    for (int i = 0; i <= k; ++i) t.push_back(x[0]);
    const int skip = (k + 1) / 2; // k=1: 1, k=3: 2
    for (int i = skip; i < m - skip; ++i) t.push_back(x[i]);
    for (int i = 0; i <= k; ++i) t.push_back(x[m - 1]);
    return t;
}

// Band LU of the cubic collocation matrix: fac[5*i + {0,1,2,3,4}] = L[i][i-2], L[i][i-1], 1/U[i][i], U[i][i+1],
// U[i][i+2].  Returns false on a vanishing pivot.
static bool colloc_lu(const std::vector<double> &t, const double *x, int m, std::vector<double> &fac)
{
    std::vector<double> A((size_t)m * 5, 0.0); // A[5*i + (d+2)] = A[i][i+d]
    for (int i = 0; i < m; ++i) {
        const int l = std::min(std::max(i + 2, 3), m - 1);
        double h[4];
        bspl3_host(t.data(), l, x[i], h);
        for (int q = 0; q < 4; ++q) {
            const int d = (l - 3 + q) - i;
            if (d >= -2 && d <= 2) A[(size_t)5 * i + d + 2] = h[q];
            else if (h[q] != 0.0) return false; // cannot happen for knots from interp_knots
        }
    }
    auto at = [&](int i, int j) -> double & { return A[(size_t)5 * i + (j - i) + 2]; };
    for (int k = 0; k < m; ++k) {
        const double piv = at(k, k);
        if (!(std::fabs(piv) > 1e-300)) return false;
        for (int i = k + 1; i <= std::min(k + 2, m - 1); ++i) {
            const double l = at(i, k) / piv;
            at(i, k) = l;
            if (l != 0.0)
                for (int j = k + 1; j <= std::min(k + 2, m - 1); ++j) at(i, j) -= l * at(k, j);
        }
    }
    fac.resize((size_t)m * 5);
    for (int i = 0; i < m; ++i) {
        fac[(size_t)5 * i + 0] = A[(size_t)5 * i + 0];
        fac[(size_t)5 * i + 1] = A[(size_t)5 * i + 1];
        fac[(size_t)5 * i + 2] = 1.0 / A[(size_t)5 * i + 2];
        fac[(size_t)5 * i + 3] = A[(size_t)5 * i + 3];
        fac[(size_t)5 * i + 4] = A[(size_t)5 * i + 4];
    }
    return true;
}

// ---------------------------------------------------------------------------------------
// Device: banded substitutions over all right-hand sides.  One thread owns one line of the array (a column for
// the first axis, a row for the second) and walks it twice; the loads of a chunk of CH elements are issued before
// the dependent arithmetic of the chunk so the walk is bound by the 2-FMA recurrence, not by memory latency.
// `len` elements spaced `es` apart per line, lines spaced `ls` apart.
// ---------------------------------------------------------------------------------------
constexpr int CH = 8;

__global__ void __launch_bounds__(32) band_solve_kernel(double2 *__restrict__ G, const double *__restrict__ fac, int len,
                                                        int64_t es, int64_t ls, int lines)
{
    const int line = blockIdx.x * blockDim.x + threadIdx.x;
    if (line >= lines) return;
    double2 *g = G + (int64_t)line * ls;
    // forward: y[i] = b[i] - L[i][i-1] y[i-1] - L[i][i-2] y[i-2]
    double2 y1 = make_double2(0, 0), y2 = y1;
    for (int i0 = 0; i0 < len; i0 += CH) {
        double2 b[CH];
#pragma unroll
        for (int u = 0; u < CH; ++u)
            if (i0 + u < len) b[u] = g[(int64_t)(i0 + u) * es];
#pragma unroll
        for (int u = 0; u < CH; ++u) {
            const int i = i0 + u;
            if (i < len) {
                const double l2 = __ldg(fac + 5 * i), l1 = __ldg(fac + 5 * i + 1);
                double2 y;
                y.x = fma(-l2, y2.x, fma(-l1, y1.x, b[u].x));
                y.y = fma(-l2, y2.y, fma(-l1, y1.y, b[u].y));
                g[(int64_t)i * es] = y;
                y2 = y1; y1 = y;
            }
        }
    }
    // backward: c[i] = (y[i] - U[i][i+1] c[i+1] - U[i][i+2] c[i+2]) / U[i][i]
    double2 c1 = make_double2(0, 0), c2 = c1;
    for (int i0 = len - 1; i0 >= 0; i0 -= CH) {
        double2 b[CH];
#pragma unroll
        for (int u = 0; u < CH; ++u)
            if (i0 - u >= 0) b[u] = g[(int64_t)(i0 - u) * es];
#pragma unroll
        for (int u = 0; u < CH; ++u) {
            const int i = i0 - u;
            if (i >= 0) {
                const double iu = __ldg(fac + 5 * i + 2), u1 = __ldg(fac + 5 * i + 3), u2 = __ldg(fac + 5 * i + 4);
                double2 c;
                c.x = fma(-u2, c2.x, fma(-u1, c1.x, b[u].x)) * iu;
                c.y = fma(-u2, c2.y, fma(-u1, c1.y, b[u].y)) * iu;
                g[(int64_t)i * es] = c;
                c2 = c1; c1 = c;
            }
        }
    }
}

__global__ void plane_kernel(const double2 *__restrict__ G, size_t n, int comp, double *out)
{
    const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e < n) out[e] = comp ? G[e].y : G[e].x;
}

static int upload(rtb_ctx *ctx, const std::vector<double> &v, double **out, Staging &sg)
{
    double *d = (double *)sg.alloc(v.size() * sizeof(double));
    if (!d) return fail(ctx, RTB_ERR_NOMEM, "rtb_fit_field: factor buffers", sg.err);
    CK(cudaMemcpyAsync(d, v.data(), v.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    *out = d;
    return RTB_OK;
}

} // namespace

extern "C" int rtb_gradient(rtb_ctx *ctx, int32_t ptr_kind, const double *Z, int64_t my, int64_t mx, double delta,
                            double *g0, double *g1)
{
    if (!ctx) return RTB_ERR_ARG;
    if (!Z || !g0 || !g1) return fail(ctx, RTB_ERR_ARG, "rtb_gradient: null argument");
    if (my < 3 || mx < 3) // numpy: "Shape of array too small to calculate a numerical gradient"
        return fail(ctx, RTB_ERR_ARG, "rtb_gradient: edge_order=2 needs at least 3 samples per axis");
    if (!(delta != 0.0)) return fail(ctx, RTB_ERR_ARG, "rtb_gradient: delta must be non-zero");
    CK(cudaSetDevice(ctx->device));
    Staging sg(ctx, ptr_kind == RTB_PTR_HOST);
    const size_t n = (size_t)my * mx;
    const double *dZ = sg.in(Z, n);
    double *d0 = sg.out(g0, n), *d1 = sg.out(g1, n);
    if (sg.err != cudaSuccess || !dZ || !d0 || !d1) return fail(ctx, RTB_ERR_NOMEM, "rtb_gradient: staging", sg.err);
    gradient_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(dZ, my, mx, delta, nullptr, d0, d1);
    CK(cudaGetLastError());
    sg.copy_back();
    CK(cudaStreamSynchronize(ctx->stream));
    if (sg.err != cudaSuccess) return fail(ctx, RTB_ERR_CUDA, "rtb_gradient: copy", sg.err);
    return RTB_OK;
}

extern "C" int rtb_fit_field(rtb_ctx *ctx, int32_t ptr_kind, const double *y, int64_t my, const double *x,
                             int64_t mx, const double *Z, double delta, rtb_stats *stats)
{
    if (!ctx) return RTB_ERR_ARG;
    if (!y || !x || !Z) return fail(ctx, RTB_ERR_ARG, "rtb_fit_field: null argument");
    // scipy: RectBivariateSpline needs more samples than the degree (m > k); the knot layout above needs m >= 4,
    // np.gradient(edge_order=2) m >= 3
    if (my < 4 || mx < 4 || my > (1 << 24) || mx > (1 << 24))
        return fail(ctx, RTB_ERR_ARG, "rtb_fit_field: need 4 <= samples per axis <= 2^24");
    if (!(delta != 0.0)) return fail(ctx, RTB_ERR_ARG, "rtb_fit_field: delta must be non-zero");
    if (stats) std::memset(stats, 0, sizeof(*stats));
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const bool host = ptr_kind == RTB_PTR_HOST;
    const int ny = (int)my, nx = (int)mx;
    // sample positions on the host (O(m)): knots, collocation, LU
    std::vector<double> hy(ny), hx(nx);
    if (host) {
        std::memcpy(hy.data(), y, sizeof(double) * ny);
        std::memcpy(hx.data(), x, sizeof(double) * nx);
    } else {
        CK(cudaMemcpyAsync(hy.data(), y, sizeof(double) * ny, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(hx.data(), x, sizeof(double) * nx, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
    }
    for (int i = 0; i + 1 < ny; ++i)
        if (!(hy[i + 1] > hy[i])) return fail(ctx, RTB_ERR_ARG, "rtb_fit_field: y must be strictly increasing");
    for (int i = 0; i + 1 < nx; ++i)
        if (!(hx[i + 1] > hx[i])) return fail(ctx, RTB_ERR_ARG, "rtb_fit_field: x must be strictly increasing");
    const std::vector<double> tzu = interp_knots(hy.data(), ny, 1), tzv = interp_knots(hx.data(), nx, 1),
                              tgu = interp_knots(hy.data(), ny, 3), tgv = interp_knots(hx.data(), nx, 3);
    std::vector<double> fu, fv;
    if (!colloc_lu(tgu, hy.data(), ny, fu) || !colloc_lu(tgv, hx.data(), nx, fv))
        return fail(ctx, RTB_ERR_ARG, "rtb_fit_field: singular collocation matrix");

    CK(cudaEventRecord(ctx->ev[0], st));
    rtb_free_field(ctx);
    DevField f;
    int rc;
    if ((rc = rtb_make_axis(ctx, tzu.data(), (int)tzu.size(), 1, &f.zu))) return rc;
    if ((rc = rtb_make_axis(ctx, tzv.data(), (int)tzv.size(), 1, &f.zv))) return rc;
    if ((rc = rtb_make_axis(ctx, tgu.data(), (int)tgu.size(), 3, &f.gu_))) return rc;
    if ((rc = rtb_make_axis(ctx, tgv.data(), (int)tgv.size(), 3, &f.gv_))) return rc;
    f.ncz = nx;
    f.ncg = nx;
    const size_t n = (size_t)ny * nx;
    double *dz = nullptr;
    double2 *dg = nullptr;
    CK(cudaMalloc(&dz, n * sizeof(double)));
    ctx->field_allocs.push_back(dz);
    CK(cudaMalloc(&dg, n * sizeof(double2)));
    ctx->field_allocs.push_back(dg);
    // bilinear spline: the coefficients are the samples
    CK(cudaMemcpyAsync(dz, Z, n * sizeof(double), host ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, st));
    f.cz = dz;
    f.cg = dg;
    {
        Staging sg(ctx, host);
        double *dfu = nullptr, *dfv = nullptr;
        if ((rc = upload(ctx, fu, &dfu, sg)) || (rc = upload(ctx, fv, &dfv, sg))) return rc;
        CK(cudaEventRecord(ctx->ev[1], st));
        gradient_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(dz, my, mx, delta, dg, nullptr, nullptr);
        // first axis (y): one thread per column, elements mx apart;  second axis (x): one thread per row
        band_solve_kernel<<<(nx + 31) / 32, 32, 0, st>>>(dg, dfu, ny, nx, 1, nx);
        band_solve_kernel<<<(ny + 31) / 32, 32, 0, st>>>(dg, dfv, nx, 1, nx, ny);
        CK(cudaGetLastError());
        CK(cudaEventRecord(ctx->ev[2], st));
        CK(cudaStreamSynchronize(st)); // the factor buffers return to the pool when sg goes out of scope
    }
    rc = rtb_finish_field(ctx, f, tzu.data(), (int)tzu.size(), tzv.data(), (int)tzv.size(), tgu.data(), tgv.data());
    if (rc) return rc;
    CK(cudaEventRecord(ctx->ev[3], st));
    CK(cudaStreamSynchronize(st));
    if (stats) {
        float ms = 0;
        cudaEventElapsedTime(&ms, ctx->ev[1], ctx->ev[2]); stats->kernel_ms = ms;
        cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[3]); stats->total_ms = ms;
        stats->launches = 3 + (ctx->f.pg ? 1 : 0);
        stats->h2d_bytes = host ? (int64_t)(n * sizeof(double)) : 0;
    }
    return RTB_OK;
}

extern "C" int rtb_get_field(rtb_ctx *ctx, int32_t which, int32_t *nu, int32_t *nv, double *tu, double *tv, double *c)
{
    if (!ctx) return RTB_ERR_ARG;
    if (!ctx->has_field) return fail(ctx, RTB_ERR_NOFIELD, "rtb_get_field: no field set");
    if (which < 0 || which > 2) return fail(ctx, RTB_ERR_ARG, "rtb_get_field: which must be 0 (z), 1 (grd[0]) or 2 (grd[1])");
    CK(cudaSetDevice(ctx->device));
    const DevField &f = ctx->f;
    const DevAxis &au = which ? f.gu_ : f.zu, &av = which ? f.gv_ : f.zv;
    if (nu) *nu = au.n;
    if (nv) *nv = av.n;
    if (tu) CK(cudaMemcpy(tu, au.t, sizeof(double) * au.n, cudaMemcpyDeviceToHost));
    if (tv) CK(cudaMemcpy(tv, av.t, sizeof(double) * av.n, cudaMemcpyDeviceToHost));
    if (c) {
        const size_t cnt = (size_t)(au.n - au.k - 1) * (size_t)(av.n - av.k - 1);
        if (which == 0) {
            CK(cudaMemcpy(c, f.cz, sizeof(double) * cnt, cudaMemcpyDeviceToHost));
        } else { // de-interleave {grd[0], grd[1]}
            Staging sg(ctx, true);
            double *d = sg.out(c, cnt);
            if (!d) return fail(ctx, RTB_ERR_NOMEM, "rtb_get_field: staging", sg.err);
            plane_kernel<<<(unsigned)((cnt + 255) / 256), 256, 0, ctx->stream>>>(f.cg, cnt, which - 1, d);
            CK(cudaGetLastError());
            sg.copy_back();
            CK(cudaStreamSynchronize(ctx->stream));
            if (sg.err != cudaSuccess) return fail(ctx, RTB_ERR_CUDA, "rtb_get_field: copy", sg.err);
        }
    }
    return RTB_OK;
}
