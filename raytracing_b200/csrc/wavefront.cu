// wavefront.cu -- iso-travel-time points of a traced fan: the arithmetic part of the
// reference's wavefront plot (graficar, RT_bench.py:987-1007): for every ray and every requested travel time
// tau, PchipInterpolator(t_ray, x)(tau), (t_ray, y)(tau), (t_ray, angle)(tau) over rows 0..d_ray[2] of s_ray,
// for the rays whose travel time reaches tau.  PCHIP is local: the value on [T_i, T_i+1] needs the two end
// derivatives, each from its neighbouring slopes (Fritsch-Carlson weights, scipy's three-point end rule), so one
// thread per (time, ray) does a binary search in the ray's T column and evaluates three cubic Hermite pieces with
// scipy's own coefficient formulas (CubicHermiteSpline: c0..c3 in powers of tau - T_i).
#include "ctx.cuh"

namespace {

__device__ __forceinline__ double sgn(double v) { return (v > 0.0) - (v < 0.0); }

// scipy PchipInterpolator._edge_case
__device__ __forceinline__ double pchip_edge(double h0, double h1, double m0, double m1)
{
    double d = ((2 * h0 + h1) * m0 - h0 * m1) / (h0 + h1);
    if (sgn(d) != sgn(m0)) d = 0.0;
    else if (sgn(m0) != sgn(m1) && fabs(d) > 3.0 * fabs(m0)) d = 3.0 * m0;
    return d;
}

// scipy PchipInterpolator._find_derivatives at interior knot: slopes m0 (left, width h0), m1 (right, width h1)
__device__ __forceinline__ double pchip_inner(double h0, double h1, double m0, double m1)
{
    if (sgn(m0) != sgn(m1) || m0 == 0.0 || m1 == 0.0) return 0.0;
    const double w1 = 2 * h1 + h0, w2 = h1 + 2 * h0;
    const double whmean = (w1 / m0 + w2 / m1) / (w1 + w2);
    return 1.0 / whmean;
}

struct Col { // one column of s_ray for one ray: element i at p[i * stride]
    const double *p;
    size_t stride;
    __device__ __forceinline__ double operator[](int i) const { return p[(size_t)i * stride]; }
};

// derivative of the PCHIP interpolant of (T, Y) at knot k of n knots (n >= 3)
__device__ double pchip_deriv(const Col &T, const Col &Y, int k, int n)
{
    if (k == 0) {
        const double h0 = T[1] - T[0], h1 = T[2] - T[1];
        return pchip_edge(h0, h1, (Y[1] - Y[0]) / h0, (Y[2] - Y[1]) / h1);
    }
    if (k == n - 1) {
        const double h0 = T[n - 1] - T[n - 2], h1 = T[n - 2] - T[n - 3];
        return pchip_edge(h0, h1, (Y[n - 1] - Y[n - 2]) / h0, (Y[n - 2] - Y[n - 3]) / h1);
    }
    const double h0 = T[k] - T[k - 1], h1 = T[k + 1] - T[k];
    return pchip_inner(h0, h1, (Y[k] - Y[k - 1]) / h0, (Y[k + 1] - Y[k]) / h1);
}

__device__ double pchip_eval(const Col &T, const Col &Y, int i, int n, double tau)
{
    const double dx = T[i + 1] - T[i], slope = (Y[i + 1] - Y[i]) / dx;
    if (n == 2) return Y[i] + slope * (tau - T[i]);
    const double d0 = pchip_deriv(T, Y, i, n), d1 = pchip_deriv(T, Y, i + 1, n);
    const double t = (d0 + d1 - 2 * slope) / dx; // CubicHermiteSpline.__init__
    const double c0 = t / dx, c1 = (slope - d0) / dx - t, c2 = d0, c3 = Y[i];
    const double u = tau - T[i];
    return ((c0 * u + c1) * u + c2) * u + c3;
}

// out (4, nt, R): x, y, angle, valid (1 / 0)
__global__ void wavefront_kernel(const double *s_ray, const double *d_ray, int64_t rows, int64_t R, const double *times,
                                 int nt, double *out)
{
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (r >= R) return;
    // rows 0..d_ray[2]; never past the array (a strided or truncated s_ray has fewer rows than d_ray[2] + 1: the
    // input must be the stride-1 trajectory, rtb_wavefront rejects anything else it can detect)
    const int n = (int)min((int64_t)d_ray[2 * R + r] + 1, rows);
    const size_t stride = (size_t)6 * R;
    const Col T = {s_ray + 4 * R + r, stride};
    const double tau = times[j];
    double x = 0, y = 0, ang = 0, valid = 0;
    if (n >= 2 && T[n - 1] >= tau) { // np.max(t_ray) >= travel_time, the reference's only test (RT:995; T increases
                                     // from 0, a tau below T[0] extrapolates the first interval like PchipInterpolator)
        int lo = 0, hi = n - 1; // largest i with T[i] <= tau, clamped to n-2 (PPoly interval rule)
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (T[mid] <= tau) lo = mid; else hi = mid;
        }
        const Col X = {s_ray + r, stride}, Y = {s_ray + R + r, stride}, A = {s_ray + 5 * R + r, stride};
        x = pchip_eval(T, X, lo, n, tau);
        y = pchip_eval(T, Y, lo, n, tau);
        ang = pchip_eval(T, A, lo, n, tau);
        valid = 1.0;
    }
    const size_t o = (size_t)j * R + r, plane = (size_t)nt * R;
    out[o] = x; out[plane + o] = y; out[2 * plane + o] = ang; out[3 * plane + o] = valid;
}

} // namespace

extern "C" int rtb_wavefront(rtb_ctx *ctx, int32_t ptr_kind, const double *s_ray, const double *d_ray, int64_t rows,
                             int64_t R, const double *times, int32_t nt, double *out)
{
    if (!ctx) return RTB_ERR_ARG;
    if (!s_ray || !d_ray || !times || !out || rows < 1 || R < 0 || nt < 0) return fail(ctx, RTB_ERR_ARG, "rtb_wavefront: bad argument");
    if (R == 0 || nt == 0) return RTB_OK;
    CK(cudaSetDevice(ctx->device));
    Staging sg(ctx, ptr_kind == RTB_PTR_HOST);
    const double *ds = sg.in(s_ray, (size_t)rows * 6 * R), *dd = sg.in(d_ray, (size_t)3 * R), *dt = sg.in(times, (size_t)nt);
    double *dout = sg.out(out, (size_t)4 * nt * R);
    if (sg.err != cudaSuccess) return fail(ctx, RTB_ERR_NOMEM, "rtb_wavefront: staging", sg.err);
    wavefront_kernel<<<dim3((unsigned)((R + 127) / 128), (unsigned)nt), 128, 0, ctx->stream>>>(ds, dd, rows, R, dt, nt, dout);
    CK(cudaGetLastError());
    sg.copy_back();
    CK(cudaStreamSynchronize(ctx->stream));
    if (sg.err != cudaSuccess) return fail(ctx, RTB_ERR_CUDA, "rtb_wavefront: copy", sg.err);
    return RTB_OK;
}
