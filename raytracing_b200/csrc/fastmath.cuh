// fastmath.cuh -- branch-free FP64 building blocks for the default (non-faithful) kernels.
//
// ncu on the first faithful kernel (profiles/r1_*) showed 653 warp instructions per ray-step of
// which only ~200 were FP64 arithmetic: libdevice sincos (range-reduction slow-path check, stack
// frame), IEEE-754 division and sqrt sequences (slow-path branches) and their BSSY/BSYNC pairs made
// the kernel issue-bound.  These replacements keep < 1 ulp-level accuracy (far inside the 1e-10
// parity bar) with straight-line DFMA code so the FP64 pipe, not the issue slot, is the limiter.
#pragma once
#include <cuda_runtime.h>

namespace rtb {

// 1/x for normal x (refractive indices, curvatures, anisotropy coefficients: all O(1e-8..1e3)).
// MUFU.RCP64H seed (~20 bits) + two Newton steps -> < 1 ulp.
__device__ __forceinline__ double rcp_fast(double x)
{
    // MUFU.RCP64H seed (relative error e0 ~ 2^-20), one cubically convergent step r(1 + e + e^2), then one
    // residual correction: error ~e0^3 = 2^-60 before the final rounding (< 1 ulp), 4 DFMA like two Newton steps
    // but the last one makes the result accurate to the last bit in all but ~1e-3 of the cases
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double e = fma(-x, r, 1.0);
    r = fma(r, fma(e, e, e), r);
    e = fma(-x, r, 1.0);
    return fma(r, e, r);
}

// 1/x to < 1 ulp in 3 DFMA (no final residual step): the step loop's reciprocals of n.
__device__ __forceinline__ double rcp_fast3(double x)
{
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    const double e = fma(-x, r, 1.0);
    return fma(r, fma(e, e, e), r);
}

// 1/sqrt(x) for normal x > 0: MUFU.RSQ64H seed + two Newton steps.
__device__ __forceinline__ double rsqrt_fast(double x)
{
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    double e = fma(-x * y, y, 1.0);
    y = fma(0.5 * y, e, y);
    e = fma(-x * y, y, 1.0);
    y = fma(0.5 * y, e, y);
    return y;
}

// 1/sqrt(x) in 5 FP64 instructions: one cubically convergent step y0 (1 + e/2 + 3 e^2/8) on the seed
// (e0 ~ 2^-20 -> ~2^-60 before rounding).  For normalising direction vectors in the step loop.
__device__ __forceinline__ double rsqrt_fast3(double x)
{
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double e = fma(-x * y, y, 1.0);
    const double t = fma(e, 0.375, 0.5) * e;
    return fma(y, t, y);
}

// sqrt(x) for x >= 0 (0 -> 0), one residual correction on top of x * rsqrt(x).
__device__ __forceinline__ double sqrt_fast(double x)
{
    const double xs = fmax(x, 1e-300);
    const double y = rsqrt_fast(xs);
    double s = xs * y;
    const double r = fma(-s, s, xs);
    s = fma(r, 0.5 * y, s);
    return x > 0.0 ? s : 0.0;
}

// Polynomial coefficients live in constant memory so DFMA reads them as c[bank][offset] operands
// (as immediates each use costs two UMOV issue slots: 50 of 402 instructions per step in the first
// default kernel, profiles/r1b_*).
__constant__ double kSin[6] = {1.58969099521155010221e-10, -2.50507602534068634195e-08, 2.75573137070700676789e-06,
                               -1.98412698298579493134e-04, 8.33333333332248946124e-03, -1.66666666666666324348e-01};
__constant__ double kCos[6] = {-1.13596475577881948265e-11, 2.08757232129817482790e-09, -2.75573143513906633035e-07,
                               2.48015872894767294178e-05, -1.38888888888741095749e-03, 4.16666666666666019037e-02};
__constant__ double kPio2[4] = {6.36619772367581382433e-01, -1.57079632673412561417e+00, -6.07710050630396597660e-11,
                                -2.02226624879595063154e-21};

// Other non-"round" constants of the step loop (a 64-bit immediate costs two UMOV/IMAD.MOV issue slots at every
// use; as c[bank][offset] operands they are free).  Constants whose low word is zero (0.5, 1, 0.75 ...) encode
// as 32-bit immediates and stay literal.
__constant__ double kK[6] = {1.0 / 6.0, 2.0 / 3.0, 1.0 / 24.0, 1.0 / 120.0, -1.0 / 5040.0, -1.0 / 720.0};
#define K_SIXTH kK[0]
#define K_TWOTHIRDS kK[1]
#define K_1_24 kK[2]
#define K_1_120 kK[3]
#define K_M1_5040 kK[4]
#define K_M1_720 kK[5]

// |x| < 2^e-ish threshold tests on the high word: integer compare, keeps the FP64 pipe for arithmetic.
__device__ __forceinline__ bool abs_lt_hi(double x, int hi_word) { return (__double2hiint(x) & 0x7fffffff) < hi_word; }
#define HI_0_75 0x3fe80000   // 0.75
#define HI_2M20 0x3eb00000   // 2^-20 = 9.54e-7
#define HI_2M11 0x3f400000   // 2^-11 = 4.88e-4
#define HI_2M6 0x3f900000    // 2^-6 = 1.5625e-2
#define HI_2M8 0x3f700000    // 2^-8 = 3.90625e-3

// sin/cos kernels on |r| <= pi/4 (+ slack), minimax-quality Taylor coefficients in Horner form.
__device__ __forceinline__ double sin_kernel(double r, double r2)
{
    double p = kSin[0];
    p = fma(p, r2, kSin[1]);
    p = fma(p, r2, kSin[2]);
    p = fma(p, r2, kSin[3]);
    p = fma(p, r2, kSin[4]);
    p = fma(p, r2, kSin[5]);
    return fma(r * r2, p, r);
}
__device__ __forceinline__ double cos_kernel(double r2)
{
    double p = kCos[0];
    p = fma(p, r2, kCos[1]);
    p = fma(p, r2, kCos[2]);
    p = fma(p, r2, kCos[3]);
    p = fma(p, r2, kCos[4]);
    p = fma(p, r2, kCos[5]);
    const double h = 0.5 * r2;
    const double w = 1.0 - h;
    return w + (((1.0 - w) - h) + r2 * r2 * p);
}

// sin and cos of a small increment |k| <= SMALL_ANGLE_MAX without range reduction.
#define RTB_SMALL_ANGLE_MAX 0.75
__device__ __forceinline__ void sincos_small(double k, double &s, double &c)
{
    const double k2 = k * k;
    s = sin_kernel(k, k2);
    c = cos_kernel(k2);
}

// sqrt(x) for x > 0 strictly (step lengths): one Newton step on the rsqrt seed (40 bits) and one
// residual correction of the product (-> < 1 ulp).
__device__ __forceinline__ double sqrt_pos(double x)
{
    // y0 = MUFU.RSQ64H seed (e0 ~ 2^-20); y = y0 (1 + e/2 + 3e^2/8) with e = 1 - x y0^2 leaves ~e0^3 = 2^-60;
    // sqrt(x) = x * y, < 1 ulp.  6 FP64 instructions (two Newton steps + residual correction: 8).
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double e = fma(-x * y, y, 1.0);
    const double t = fma(e, 0.375, 0.5) * e;
    y = fma(y, t, y);
    return x * y;
}

// (sin, cos)(theta + k) from (s0, c0) = (sin, cos)(theta) by the angle-addition formulas, written as
// a correction to (s0, c0): s = s0 + (s0 (cos k - 1) + c0 sin k).  The polynomial degree adapts to
// |k| (per-step angle increments are 1e-5..1e-2; all tiers are accurate to < 1e-17 absolute).
__device__ __forceinline__ void rotate(double s0, double c0, double k, double &s, double &c)
{
    const double k2 = k * k;
    double sk, cm; // sin k, cos k - 1
    if (abs_lt_hi(k2, HI_2M20)) {          // |k| < 9.8e-4: truncation < 1e-17
        sk = fma(k * k2, -K_SIXTH, k);
        cm = k2 * fma(k2, K_1_24, -0.5);
    } else if (abs_lt_hi(k2, HI_2M11)) {   // |k| < 2.2e-2: truncation < 2e-18
        sk = fma(k * k2, fma(k2, fma(k2, K_M1_5040, K_1_120), -K_SIXTH), k);
        cm = k2 * fma(k2, fma(k2, K_M1_720, K_1_24), -0.5);
    } else {
        sk = sin_kernel(k, k2);
        double p = kCos[0];
        p = fma(p, k2, kCos[1]);
        p = fma(p, k2, kCos[2]);
        p = fma(p, k2, kCos[3]);
        p = fma(p, k2, kCos[4]);
        p = fma(p, k2, kCos[5]);
        cm = fma(k2 * k2, p, -0.5 * k2);
    }
    s = s0 + fma(s0, cm, c0 * sk);
    c = c0 + fma(c0, cm, -(s0 * sk));
}

// Branch-free rotation for |k| < 2^-8: sin k to k^5 (next term k^7/5040 < 3e-21), cos k - 1 to k^4 (next term
// k^6/720 < 5e-18).
__device__ __forceinline__ void rotate_small(double s0, double c0, double k, double &s, double &c)
{
    const double k2 = k * k;
    const double sk = fma(k * k2, fma(k2, K_1_120, -K_SIXTH), k);
    const double cm = k2 * fma(k2, K_1_24, -0.5);
    s = s0 + fma(s0, cm, c0 * sk);
    c = c0 + fma(c0, cm, -(s0 * sk));
}

static __device__ __noinline__ double2 sincos_libdevice(double x)
{
    double2 r; // (sin, cos) by value: no address-taken locals in the caller
    sincos(x, &r.x, &r.y);
    return r;
}

// sin and cos for |x| up to ~1e5 (ray angles are O(1..100)): Cody-Waite reduction by pi/2 with a
// 3-term split of pi/2, no slow path.  Larger arguments fall back to libdevice.
template <bool FALLBACK>
__device__ __forceinline__ void sincos_cw_t(double x, double &s, double &c)
{
    if (FALLBACK && fabs(x) > 1.0e5) { const double2 r = sincos_libdevice(x); s = r.x; c = r.y; return; }
    const double q = rint(x * kPio2[0]); // x * 2/pi
    double r = fma(q, kPio2[1], x);      // pi/2 hi (33 bits)
    r = fma(q, kPio2[2], r);             // pi/2 mid
    r = fma(q, kPio2[3], r);             // pi/2 lo
    const int n = __double2int_rn(q);
    const double r2 = r * r;
    const double sr = sin_kernel(r, r2);
    const double cr = cos_kernel(r2);
    double ss = (n & 1) ? cr : sr;
    double cc = (n & 1) ? sr : cr;
    s = (n & 2) ? -ss : ss;
    c = ((n + 1) & 2) ? -cc : cc;
}
__device__ __forceinline__ void sincos_cw(double x, double &s, double &c) { sincos_cw_t<true>(x, s, c); }
// Straight-line variant (no out-of-line call): for |x| > 1e5 the 3-term reduction loses accuracy gradually.
__device__ __forceinline__ void sincos_cw_inline(double x, double &s, double &c) { sincos_cw_t<false>(x, s, c); }

} // namespace rtb
