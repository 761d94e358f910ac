// field.cuh -- device evaluator of the refractive-index field n(x,y) and its gradient.
//
// Replaces n_gradient (RT_bench.py:141-156), i.e. three single-point calls of
// scipy RectBivariateSpline.__call__ -> FITPACK bispev/fpbisp/fpbspl: bilinear n = z(y,x),
// bicubic dn/dx = grd[1](y,x), dn/dy = grd[0](y,x).  The first spline axis "u" is the
// physical y (RT_bench.py:455-457), the second "v" the physical x.
//
// B200 design: one thread evaluates one ray position.  The two gradient splines share
// their knot vectors (same grid, same degree), so the bicubic basis is computed once and
// their coefficients are stored interleaved as double2 {gu, gv}: one 16-byte gather
// serves both splines (16 LDG.128 instead of 32 LDG.64 per step).  Knot vectors are a few
// KB and stay in L1; coefficient tables (2-9 MB each) stay in the 126 MB L2.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "fastmath.cuh"

namespace rtb {

struct DevAxis {
    const double *t;  // knots (device)
    int n;            // number of knots
    int k;            // degree
    int off;          // interval guess: l = floor((x - t0) * inv_h) + off
    int lo, hi;       // inclusive range of l (0-based, t[l] <= x < t[l+1]) where t[l-k+1 .. l+k] is uniform
    double t0, inv_h; // origin and reciprocal nominal spacing of the interior knots
    double tb, te;    // clamp range t[k], t[n-k-1]  (fpbisp: arg clamped to [tb, te])
    double c0;        // off - t0*inv_h - 0.5: fma(x, inv_h, c0) = (interval coordinate) - 0.5 on the uniform run
};

struct DevField {
    DevAxis zu, zv;    // bilinear spline of n
    DevAxis gu_, gv_;  // shared axes of the two bicubic gradient splines
    const double *cz;  // (zu.n-2) x (zv.n-2)
    const double2 *cg; // interleaved {grd[0] = dn/dy, grd[1] = dn/dx}, (gu.n-4) x (gv.n-4)
    int ncz;           // row length of cz  = zv.n - zv.k - 1
    int ncg;           // row length of cg  = gv.n - gv.k - 1
    // Fast path (host-verified in rtb_set_field): on [flo_u, fhi_u] x [flo_v, fhi_v] (cubic interval
    // indices) all knots involved are uniform AND the bilinear knots coincide with the cubic ones
    // shifted by one (tz[l-1] == tg[l]), so one interval / one fraction per physical axis serves all
    // three splines.  fast_ok == 0 disables it (arbitrary user knots).
    int fast_ok, flo_u, fhi_u, flo_v, fhi_v;
    // Per-cell polynomial tables of the fast region (built on the device by rtb_set_field, cell_tables.cuh): for
    // the cell (lu, lv), ci = (lu - flo_u) * pstride + (lv - flo_v),
    //   pg[18*ci + 4*p + q] = coefficient of tu^p tv^q of {dn/dy, dn/dx}   (tu, tv in [-1/2, 1/2]: offsets from the
    //   pg[18*ci + 16, 17]  = {n0, d/dtu}, {d/dtv, d2/dtu dtv} of the bilinear n    cell centre in cell units)
    // i.e. the same piecewise polynomials as the B-spline form, expanded once: 33 FP64 instructions per evaluation
    // instead of 66 (two cubic bases + 32 coefficient MACs + bilinear), for 288 B of table per cell (L2 / HBM).
    const double2 *pg;
    int pstride;
};

// Cubic-interval index ranges (inclusive) inside which a position needs no further checks in the step loop:
// the default evaluator's fast region intersected with the cells that lie strictly inside the box of
// constants() (RT:878), so the common case pays neither the general-path test nor the four FP64 box compares.
constexpr int RTB_CELL_REC = 18; // double2 per cell record

struct CellBox { int lo_u, hi_u, lo_v, hi_v; };

// FITPACK fpbisp interval search, restated for the GPU: clamp, then find l (0-based) with
// t[l] <= x < t[l+1], k <= l <= n-k-2.  FITPACK walks linearly from l=k; here a
// multiply+floor guess is corrected against the real knots, which yields the same l.
__device__ __forceinline__ int find_interval(const DevAxis &a, double &x)
{
    if (x < a.tb) x = a.tb;
    if (x > a.te) x = a.te;
    const int lmin = a.k, lmax = a.n - a.k - 2;
    int l = __double2int_rd((x - a.t0) * a.inv_h) + a.off;
    l = max(lmin, min(lmax, l));
    while (l > lmin && x < __ldg(a.t + l)) --l;
    while (l < lmax && !(x < __ldg(a.t + l + 1))) ++l;
    return l;
}

// FITPACK fpbspl (Cox-de Boor recurrence with the real knots), K = degree; l 0-based.
template <int K>
__device__ __forceinline__ void basis_exact(const double *__restrict__ t, int l, double x, double *h)
{
    double hh[K + 1];
    h[0] = 1.0;
#pragma unroll
    for (int j = 1; j <= K; ++j) {
#pragma unroll
        for (int i = 0; i < j; ++i) hh[i] = h[i];
        h[0] = 0.0;
#pragma unroll
        for (int i = 1; i <= j; ++i) {
            const double tli = __ldg(t + l + i);     // 1-based t(l1+i),   l1 = l+1
            const double tlj = __ldg(t + l + i - j); // 1-based t(l1+i-j)
            if (tli == tlj) { h[i] = 0.0; continue; }
            const double f = hh[i - 1] / (tli - tlj);
            h[i - 1] = h[i - 1] + f * (tli - x);
            h[i] = f * (x - tlj);
        }
    }
}

// Closed-form uniform cubic B-spline basis at s = (x - t[l]) / h in [0,1).
__device__ __forceinline__ void basis_uniform3(double s, double *w)
{
    // 10 FP64 instructions: w3 = s^3/6, p = w0 + w3 = 1/6 - s/2 + s^2/2, w1 = 2/3 - s^2 + s^3/2, and the
    // partition of unity for w2 (weights are O(1): the subtractions cost ~1e-16 absolute)
    const double s2 = s * s, s3 = s2 * s;
    w[3] = s3 * K_SIXTH;
    const double p = fma(s2, 0.5, fma(s, -0.5, K_SIXTH));
    w[0] = p - w[3];
    w[1] = fma(s3, 0.5, K_TWOTHIRDS - s2);
    w[2] = (1.0 - p) - w[1];
}

template <bool EXACT>
__device__ __forceinline__ int basis3(const DevAxis &a, double x, double *w)
{
    const int l = find_interval(a, x);
    if (!EXACT && l >= a.lo && l <= a.hi) {
        basis_uniform3((x - __ldg(a.t + l)) * a.inv_h, w);
    } else {
        basis_exact<3>(a.t, l, x, w);
    }
    return l;
}

template <bool EXACT>
__device__ __forceinline__ int basis1(const DevAxis &a, double x, double *w)
{
    const int l = find_interval(a, x);
    if (!EXACT && l >= a.lo && l <= a.hi) {
        const double s = (x - __ldg(a.t + l)) * a.inv_h;
        w[0] = 1.0 - s;
        w[1] = s;
    } else {
        basis_exact<1>(a.t, l, x, w);
    }
    return l;
}

// n_gradient RT_bench.py:141-156: n, gx = dn/dx, gy = dn/dy at (x, y).
template <bool EXACT>
__device__ __forceinline__ void n_gradient(const DevField &f, double x, double y, double &n, double &gx,
                                           double &gy)
{
    {
        double wu[2], wv[2];
        const int lu = basis1<EXACT>(f.zu, y, wu);
        const int lv = basis1<EXACT>(f.zv, x, wv);
        const double *c = f.cz + (size_t)(lu - 1) * f.ncz + (lv - 1);
        const double r0 = fma(__ldg(c + 1), wv[1], __ldg(c) * wv[0]);
        const double r1 = fma(__ldg(c + f.ncz + 1), wv[1], __ldg(c + f.ncz) * wv[0]);
        n = fma(r1, wu[1], r0 * wu[0]);
    }
    {
        double wu[4], wv[4];
        const int lu = basis3<EXACT>(f.gu_, y, wu);
        const int lv = basis3<EXACT>(f.gv_, x, wv);
        const double2 *c = f.cg + (size_t)(lu - 3) * f.ncg + (lv - 3);
        double au = 0.0, av = 0.0;
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const double2 c0 = __ldg(c), c1 = __ldg(c + 1), c2 = __ldg(c + 2), c3 = __ldg(c + 3);
            double ru = c0.x * wv[0], rv = c0.y * wv[0];
            ru = fma(c1.x, wv[1], ru); rv = fma(c1.y, wv[1], rv);
            ru = fma(c2.x, wv[2], ru); rv = fma(c2.y, wv[2], rv);
            ru = fma(c3.x, wv[3], ru); rv = fma(c3.y, wv[3], rv);
            au = fma(ru, wu[a], au);
            av = fma(rv, wu[a], av);
            c += f.ncg;
        }
        gy = au; // grd[0] = d n / d y
        gx = av; // grd[1] = d n / d x
    }
}

// Rare general path of the default evaluator (near the grid ends / arbitrary knots): kept out of line
// so its registers do not burden the step loop.
// `fd` is the context's copy of the DevField in global memory (taking the address of the kernel
// parameter would force a 700-byte local-memory copy of the whole parameter block).
struct FieldSample { double n, gx, gy; };
static __device__ __noinline__ FieldSample n_gradient_general(const DevField *fd, double x, double y)
{
    FieldSample r; // returned by value: keeps the caller's n, gx, gy in registers
    n_gradient<false>(*fd, x, y, r.n, r.gx, r.gy);
    return r;
}

// One cell record (RTB_CELL_REC double2: 16 bicubic-pair coefficients + 4 bilinear ones) evaluated at the offsets
// (tu, tv) from the cell centre: Horner in tv for each power of tu, both gradient components at once.
template <bool SMEM>
__device__ __forceinline__ void eval_record(const double2 *c, double tu, double tv, double &n, double &gx, double &gy)
{
    auto ld = [&](int k) { return SMEM ? c[k] : __ldg(c + k); };
    {
        const double2 a = ld(16), b = ld(17); // {n0, n_u}, {n_v, n_uv}
        n = fma(tu, fma(tv, b.y, a.y), fma(tv, b.x, a.x));
    }
    double2 row[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const double2 c0 = ld(4 * p), c1 = ld(4 * p + 1), c2 = ld(4 * p + 2), c3 = ld(4 * p + 3);
        row[p].x = fma(fma(fma(c3.x, tv, c2.x), tv, c1.x), tv, c0.x);
        row[p].y = fma(fma(fma(c3.y, tv, c2.y), tv, c1.y), tv, c0.y);
    }
    gy = fma(fma(fma(row[3].x, tu, row[2].x), tu, row[1].x), tu, row[0].x);
    gx = fma(fma(fma(row[3].y, tu, row[2].y), tu, row[1].y), tu, row[0].y);
}
// Default evaluator: one interval computation per physical axis without knot comparisons (a position 1e-13 of
// a cell beyond a knot evaluates the neighbouring polynomial piece of a C2 / C0 spline: error << 1e-13), then the
// cell's expanded polynomials from the tables.  Falls back to the general routine outside the fast region.
// Returns true when the position's cell lies inside `in` (then it is also strictly inside the box).
//
// Warp staging (wbuf != nullptr, dense fans): when all 32 lanes of the warp are stepping (mask full) and sit in the
// SAME cell -- neighbouring rays of a 2^20-ray fan are ~1e-5 cells apart -- 18 lanes fetch the record once (one
// LDG.128), park it in the warp's 288 B of shared memory and every lane reads it back with broadcast LDS.128: the
// L1 data pipe moves 32 lanes x 16 B per LDG.128 whatever the addresses are (4 wavefronts, tools/lsu_microbench.cu),
// a uniform LDS.128 costs half of that.  Same table entries, same arithmetic, same results.
__device__ __forceinline__ bool n_gradient_cells(const DevField &f, const DevField *fd, const CellBox &in, double x,
                                                 double y, double &n, double &gx, double &gy, unsigned mask = 0u,
                                                 double2 *wbuf = nullptr)
{
    // Interval index and cell fraction without conversion instructions or knot loads: vh = (interval
    // coordinate) - 0.5; adding 1.5*2^52 rounds it to the nearest integer = floor(coordinate), which then sits in
    // the low word of the sum; the fraction is vh - that integer + 0.5.  (A coordinate within half an ulp of a
    // knot may land in the neighbouring cell with fraction 1 or 0: same value, the spline is continuous.)
    // The knots of the uniform run equal t0 + j/inv_h to < 3e-13 of a cell (checked in rtb_set_field).
    const double MAGIC = 6755399441055744.0;
    const double vu = fma(y, f.gu_.inv_h, f.gu_.c0), vv = fma(x, f.gv_.inv_h, f.gv_.c0);
    const double mu = vu + MAGIC, mv = vv + MAGIC;
    const int lu = __double2loint(mu), lv = __double2loint(mv);
    // |coordinate| < 2^31 <=> high word of the sum is 0x43380000 (or 0x4337ffff for negatives); NaN fails too
    const bool sane = (unsigned)(__double2hiint(mu) - 0x4337ffff) <= 1u && (unsigned)(__double2hiint(mv) - 0x4337ffff) <= 1u;
    const bool inner = sane && lu >= in.lo_u && lu <= in.hi_u && lv >= in.lo_v && lv <= in.hi_v;
#ifndef RTB_NO_STAGE
    if (wbuf != nullptr && mask == 0xffffffffu) { // warp-uniform condition
        const int cell = inner ? (lu - f.flo_u) * f.pstride + (lv - f.flo_v) : -1;
        const int cell0 = __shfl_sync(0xffffffffu, cell, 0);
        if (__all_sync(0xffffffffu, cell == cell0 && cell >= 0)) {
            const int lane = threadIdx.x & 31;
            __syncwarp(); // the previous record has been read by every lane
            if (lane < RTB_CELL_REC) wbuf[lane] = __ldg(f.pg + (size_t)RTB_CELL_REC * cell0 + lane);
            __syncwarp();
            const double tu = vu - (mu - MAGIC), tv = vv - (mv - MAGIC);
            eval_record<true>(wbuf, tu, tv, n, gx, gy);
            return true;
        }
    }
#endif
    if (!inner) { // rare: near the box / grid edge, or arbitrary knots
        if (!f.fast_ok || !sane || lu < f.flo_u || lu > f.fhi_u || lv < f.flo_v || lv > f.fhi_v) {
            const FieldSample r = n_gradient_general(fd, x, y);
            n = r.n; gx = r.gx; gy = r.gy;
            return false;
        }
    }
    const double tu = vu - (mu - MAGIC), tv = vv - (mv - MAGIC); // offsets from the cell centre, in [-1/2, 1/2]
    const size_t ci = (size_t)(lu - f.flo_u) * f.pstride + (lv - f.flo_v);
    const double2 *c = f.pg + (size_t)RTB_CELL_REC * ci;
    eval_record<false>(c, tu, tv, n, gx, gy);
    return inner;
}

// Same evaluator without a box (ray start, rtb_field_eval).
__device__ __forceinline__ void n_gradient_fast(const DevField &f, const DevField *fd, double x, double y,
                                                double &n, double &gx, double &gy)
{
    const CellBox in = {f.flo_u, f.fast_ok ? f.fhi_u : f.flo_u - 1, f.flo_v, f.fhi_v};
    n_gradient_cells(f, fd, in, x, y, n, gx, gy);
}

} // namespace rtb
