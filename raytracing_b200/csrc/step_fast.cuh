// step_fast.cuh -- the eleven step methods (RT_bench.py:469-764) in the default B200 form:
// same mathematics as step.cuh (the faithful restatement), re-associated for the FP64 pipe:
//   * 1/n carried in the ray state (one Newton reciprocal per step instead of 4 IEEE divisions);
//   * sin/cos of theta+k for the small increments of tfinal_2o / curvature_t by angle addition
//     from the carried (sin, cos) with an exact resync every RESYNC steps;
//   * theta_cost_t / finite_diff: the new unit vector is the normalised momentum / difference
//     vector itself (cos(atan2(b,a)) = a/|(a,b)|), the atan2 is evaluated only where theta is
//     stored (every kept row in trajectory mode, once at the end otherwise);
//   * curvature_t without the reference's cancellation: sin(t+e)-sin(t) = s(cos e - 1) + c sin e.
// Every deviation from the faithful kernel is rounding-level (<= 1e-13 over a trajectory, see
// tests/test_gpu_trace.py) except where the reference itself is noise-dominated.
#pragma once
#include <type_traits>

#include "fastmath.cuh"
#include "golden_table.cuh"
#include "step.cuh"

namespace rtb {

typedef RayState FastState; // ang is stale (lazy) for theta_cost_t / finite_diff methods when LAZY

#ifndef RTB_TIE_EPS
#define RTB_TIE_EPS 3.6e-15 // near-tie band of the golden comparisons (see golden_trig)
#endif
#define RTB_RESYNC_MASK 15 // exact sincos every 16 steps bounds the drift of the rotated (sin, cos)

// cos(k) - 1 for |k| <= RTB_SMALL_ANGLE_MAX without cancellation
__device__ __forceinline__ double cosm1_small(double k2)
{
    double p = kCos[0];
    p = fma(p, k2, kCos[1]);
    p = fma(p, k2, kCos[2]);
    p = fma(p, k2, kCos[3]);
    p = fma(p, k2, kCos[4]);
    p = fma(p, k2, kCos[5]);
    return fma(k2 * k2, p, -0.5 * k2);
}

// Cost functions of the golden-section angle search (RT:595,697 and RT:728,761) taking (sin, cos) of the probe.
struct CostIsoFast {
    double n1, ax, ay; // ax = n0*ux + impulse_x
    __device__ __forceinline__ double operator()(double s, double c) const
    {
        const double a = fma(n1, c, -ax), b = fma(n1, s, -ay);
        return fma(a, a, b * b);
    }
};

struct CostAnisoFast {
    double n1, n1g2, g2m1, mx, my, gx1h, gy1h; // n1g2 = n1 gamma^2; mx = mi_x + h*coef_i*gx0/2 ; gx1h = h*gx1/2
    __device__ __forceinline__ double operator()(double s, double c) const
    {
        // anisotropy^2 = gamma^2 s^2 + c^2 = 1 + (gamma^2 - 1) s^2 =: v, coef = sqrt(v) = v y with y = 1/sqrt(v).
        // The momentum of RT:217-230 simplifies exactly: p_x = n coef c (1 - s^2 (gamma^2-1)/v) = n c y (the bracket is
        // 1/v), p_y = n coef s (1 + c^2 (gamma^2-1)/v) = n gamma^2 s y (v + (gamma^2-1) c^2 = gamma^2): the probe's
        // momentum lies on the ellipse p_x^2 + p_y^2 / gamma^2 = n^2.
        const double v = fma(g2m1, s * s, 1.0);
        double y; // MUFU seed + one cubically convergent step
        asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(v));
        const double e = fma(-v * y, y, 1.0);
        y = fma(y, fma(e, 0.375, 0.5) * e, y);
        const double coef = v * y;
        const double a = fma(n1, c * y, -fma(coef, gx1h, mx));
        const double b = fma(n1g2, s * y, -fma(coef, gy1h, my));
        return fma(a, a, b * b);
    }
};

// ------------------------------------------------------------------------------------------------------------------
// golden (RT:175-199) without its bracket arithmetic.
//
// Every call starts from [theta - pi/2, theta + pi/2] (DELTA_G, RT:64) and shrinks the bracket by GOLD_RATIO per pass
// whatever the cost does, so (i) the loop `while abs(c-d) > GOLD_TOL` runs exactly RTB_GOLDEN_ITERS = 37 times,
// (ii) relative to theta every bracket end, probe and midpoint is a sum of +- constants picked by the decisions
// `f(c) < f(d)` taken so far.  The searches below therefore carry only what a decision needs -- the kept and the new
// probe's (sin, cos, cost), or for the symmetric cost just the signed distance of the bracket midpoint from the
// minimum -- plus the decision bits; a, b, c, d themselves are never formed.  The returned angle theta + (sum of the
// midpoint moves) differs from the reference's (a+b)/2 by the rounding of its bracket arithmetic (< 1e-15).
//
// Near-ties are where the rounding of *any* evaluation decides (among them the exact tie of a ray running along
// grad n): |f(c) - f(d)| within ~16 ulp-level perturbations of the residuals.  They are decided the reference's way,
// out of line: the search so far is replayed next to the reference's own un-fused bracket arithmetic (GoldenBracket),
// and both costs of the pass are evaluated afresh at the reference's probes c, d in the reference's expression order
// (CostIso / CostAniso of step.cuh) with libdevice's sincos -- no cached value, no angle-addition chain.  About one
// step in a thousand (symmetric cost) to a few in a hundred (anisotropic cost) per warp on the benchmark fans (ncu
// source counters, profiles/).
// ------------------------------------------------------------------------------------------------------------------
#ifndef RTB_TIE_FLOOR
#define RTB_TIE_FLOOR 2.0e-14 // covers the rounding drift of the reference's own a, b (37 passes x ulp(theta +- pi/2))
#endif

// What a near-tie needs to rebuild the reference's cost lambda: the step's inputs as they sit in registers anyway
// (nothing is computed for it unless a tie occurs).
struct TieIso { // RT:595,697
    double n0, ux, uy, gx0, gy0, n1, gx1, gy1, h;
    __device__ __forceinline__ CostIso make() const
    {
        CostIso cf;
        cf.n1 = n1;
        cf.n0ux = n0 * ux; cf.n0uy = n0 * uy;
        cf.ix = h * (gx0 + gx1) / 2.0; cf.iy = h * (gy0 + gy1) / 2.0;
        return cf;
    }
};
struct TieAniso { // RT:725-729, 758-762
    double n0, coef0, ux, uy, gx0, gy0, n1, gx1, gy1, h, gamma, g2m1;
    __device__ __forceinline__ CostAniso make() const
    {
        CostAniso cf;
        cf.n1 = n1; cf.gamma = gamma; cf.g2m1 = g2m1; cf.step = h;
        cf.mi_x = moment_c(n0, coef0, g2m1, ux, -(uy * uy));
        cf.mi_y = moment_c(n0, coef0, g2m1, uy, ux * ux);
        cf.cgx0 = coef0 * gx0; cf.cgy0 = coef0 * gy0;
        cf.gx1 = gx1; cf.gy1 = gy1;
        return cf;
    }
};

// The reference's bracket (RT:184-194), un-fused.
struct GoldenBracket {
    double a, b, c, d;
    __device__ __forceinline__ void probes()
    {
        c = __dsub_rn(b, __dmul_rn(__dsub_rn(b, a), RTB_GOLD_RATIO));
        d = __dadd_rn(a, __dmul_rn(__dsub_rn(b, a), RTB_GOLD_RATIO));
    }
    __device__ __forceinline__ void init(double theta)
    {
        a = __dsub_rn(theta, RTB_DELTA_G); b = __dadd_rn(theta, RTB_DELTA_G);
        probes();
    }
    __device__ __forceinline__ void shrink(bool left) // RT:189-194
    {
        if (left) b = d; else a = c;
        probes();
    }
};

// Evaluating search (anisotropic cost; isotropic cost when the turn is too large for golden_iso_sym).  State of a pass:
// the kept probe W and the newest probe N (sin, cos, cost), on which side N lies, and the bracket midpoint - theta.
// Sides and decisions are sign masks (0 or 0x80000000), so a pass takes them from and applies them to the high words
// of its doubles with integer logic: no FP64 compare, no 64-bit select for a sign.
struct TrigState {
    double Ws, Wc, Wf, Ns, Nc, Nf, m;
    int nl; // sign mask: N is the left probe (c)
};

#define RTB_SIGN ((int)0x80000000)

template <class F>
__device__ __forceinline__ void trig_init(const F &cost, double st, double ct, TrigState &g)
{
    const double S0 = kGolden0[0], C0 = kGolden0[1];
    // the first probes: c = theta - P0 (left, kept), d = theta + P0 (right, newest)
    g.Ws = fma(-ct, S0, st * C0); g.Wc = fma(st, S0, ct * C0);
    g.Ns = fma(ct, S0, st * C0); g.Nc = fma(-st, S0, ct * C0);
    g.Wf = cost(g.Ws, g.Wc); g.Nf = cost(g.Ns, g.Nc);
    g.nl = 0;
    g.m = 0.0;
}

// f(c) - f(d) inside the rounding noise of the residuals?  df^2 <= tie2 * (f(c) + f(d)), compared on the high words
// (a superset of the exact comparison; f(c) + f(d) >= max f).
__device__ __forceinline__ bool trig_in_band(const TrigState &g, double tie2, double &df)
{
    df = g.Nf - g.Wf;
    return __double2hiint(df * df) <= __double2hiint(tie2 * (g.Nf + g.Wf));
}

// RT:189 `if func(c) < func(d)` as a sign mask, from the sign of f(N) - f(W): c is N when nl, so
// left <=> (df < 0) == nl <=> the sign bits of df and nl agree.
__device__ __forceinline__ int trig_decide(const TrigState &g, double df)
{
    return ((__double2hiint(df) ^ g.nl) & RTB_SIGN) ^ RTB_SIGN;
}

// After the decision of pass k (`left`: sign mask): the kept probe (c when b moves, d when a moves), the midpoint move,
// and the new probe at the kept one -+ pi rho^(k+4): one rotation by table constants, one cost evaluation.
template <class F>
__device__ __forceinline__ void trig_advance(const F &cost, TrigState &g, int left, int k)
{
    if (left == g.nl) { g.Ws = g.Ns; g.Wc = g.Nc; g.Wf = g.Nf; }
    const GoldenPass t = kGoldenPass[k];
    g.m += __hiloint2double(__double2hiint(t.mid) ^ left, __double2loint(t.mid)); // left: the midpoint moves by -mid
    const double sr = __hiloint2double(__double2hiint(t.sr) ^ left, __double2loint(t.sr)); // the new probe lies left of the kept one when b moved
    g.Ns = fma(g.Wc, sr, g.Ws * t.cr);
    g.Nc = fma(-g.Ws, sr, g.Wc * t.cr);
    g.Nf = cost(g.Ns, g.Nc);
    g.nl = left;
}

// Decision of pass k the reference's way, for a pass whose costs are inside the band.  Nothing about the earlier
// passes is kept by the fast loop, so they are replayed here -- same fast arithmetic, same decisions -- next to the
// reference's own bracket; a pass inside the band (this one, and any earlier one) is decided by the costs of the
// reference's probes c, d in the reference's expression order with libdevice's sincos.
template <class F, class TIE>
static __device__ __noinline__ bool golden_trig_tie(const F cost, const TIE t, double tie2, double theta, double st, double ct, int k)
{
    const auto cf = t.make();
    GoldenBracket br;
    br.init(theta);
    TrigState g;
    trig_init(cost, st, ct, g);
    for (int j = 0;; ++j) {
        double df;
        int left;
        if (trig_in_band(g, tie2, df)) left = cf(br.c) < cf(br.d) ? RTB_SIGN : 0; // RT:189
        else left = trig_decide(g, df);
        if (j == k) return left != 0;
        br.shrink(left != 0);
        trig_advance(cost, g, left, j);
    }
}

// Passes [k, kend) of the evaluating search: a call-free inner loop (its state stays in registers); a pass inside the
// band leaves it, is decided out of line, and the loop resumes behind it.
template <class F, class TIE>
__device__ __forceinline__ void golden_trig_passes(const F &cost, const TIE &ref, double tie2, double theta, double st,
                                                   double ct, TrigState &g, int &k, int kend)
{
    for (;;) {
        bool band = false;
        double df;
#pragma unroll 1
        for (; k < kend; ++k) {
            if (trig_in_band(g, tie2, df)) { band = true; break; }
            trig_advance(cost, g, trig_decide(g, df), k);
        }
        if (!band) break;
        const bool left = golden_trig_tie(cost, ref, tie2, theta, st, ct, k);
        trig_advance(cost, g, left ? RTB_SIGN : 0, k);
        ++k;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Anisotropic cost, passes k >= RTB_GOLDEN_TAIL_K1 without cost evaluations.
//
// After K1 passes the bracket is pi rho^K1 wide (9.8e-3 for K1 = 12) and the cost f is, to the accuracy a decision
// needs, its cubic Taylor polynomial about the minimiser x*: with e = (midpoint of the probes) - x*, q = half their
// distance, kappa = f3/f2 and lambda = f4/f2 (fn = n-th derivative of f at x*),
//     f(d) - f(c) = 2 q f2 [ e + kappa (e^2/2 + q^2/6) + O(lambda (|e| + q)^3) ],
// so RT:189 `f(c) < f(d)` <=> t = e + kappa (e^2/2 + q^2/6) > 0.  x* comes from three or four Newton steps on the first
// derivative started at the kept probe (first and second derivative in closed form from the residuals,
// aniso_cost_d12), kappa from the secant of the second derivative over the first step.  A pass is decided by the sign
// of t unless |t| is within a band that covers
//   - the band of the evaluating search in the same units (|f(c) - f(d)| <= TIE sqrt(f(c) + f(d)): there the
//     reference's own rounding decides), widened by a quarter,
//   - the cubic's truncation error lambda_max (|e| + q)^3 / 3 with lambda_max = 16 max(gamma, 1/gamma)^2 + 8 (measured
//     maximum 118 at gamma = 3: tools/proto/golden_tail.py), the error of the secant kappa, the last Newton step squared;
// a pass inside the band is decided exactly as before, by golden_trig_tie (replay of the evaluating search next to the
// reference's bracket).  Outside the band the decision equals the evaluating search's (prototype: 0 differences in
// 2e5 random steps x 25 passes for gamma = 0.5, 1.5, 3), hence the reference's.  Steps where the expansion is not
// trustworthy (second derivative <= 0, Newton not converged, x* outside the bracket, |kappa| > 64) return false and
// take the evaluating passes.  5 FP64 instructions per pass instead of 19, straight-line; a step with a pass inside
// the band repeats its tail in golden_aniso_tail_ties (rolled, with the out-of-line decisions).
// ------------------------------------------------------------------------------------------------------------------
#ifndef RTB_GOLDEN_TAIL_K1
#define RTB_GOLDEN_TAIL_K1 12
#endif

// F1 = half the first, F2 = half the second derivative, and the value f of the anisotropic cost at the angle with
// (sin, cos) = (s, c).  With y = 1/sqrt(v), v = 1 + (gamma^2 - 1) s^2: (c y)' = -gamma^2 s y^3, (s y)' = c y^3,
// (v y)' = (gamma^2 - 1) s c y.
__device__ __forceinline__ void aniso_cost_d12(const CostAnisoFast &cf, double s, double c, double &F1, double &F2, double &f)
{
    const double v = fma(cf.g2m1, s * s, 1.0);
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(v));
    const double er = fma(-v * y, y, 1.0);
    y = fma(y, fma(er, 0.375, 0.5) * er, y);
    const double w = v * y;
    const double Rx = fma(cf.n1, c * y, -fma(w, cf.gx1h, cf.mx));
    const double Ry = fma(cf.n1g2, s * y, -fma(w, cf.gy1h, cf.my));
    const double y2 = y * y, By3 = cf.n1g2 * (y2 * y), z = cf.g2m1 * (s * c), zy = z * y;
    const double Rx1 = -fma(By3, s, zy * cf.gx1h);
    const double Ry1 = fma(By3, c, -(zy * cf.gy1h));
    F1 = fma(Rx, Rx1, Ry * Ry1);
    const double z1 = cf.g2m1 * fma(c, c, -(s * s));
    const double t1 = z * y2;
    const double q2 = fma(z1, y, -(zy * t1));
    const double u = 3.0 * By3 * t1;
    const double Rx2 = fma(u, s, -fma(By3, c, q2 * cf.gx1h));
    const double Ry2 = -fma(u, c, fma(By3, s, q2 * cf.gy1h));
    F2 = __dadd_rn(fma(Rx1, Rx1, Ry1 * Ry1), fma(Rx, Rx2, Ry * Ry2));
    f = fma(Rx, Rx, Ry * Ry);
}

// What the tail passes need besides the cost: the expansion and its bands.
struct AnisoTail {
    double kaph;   // kappa / 2
    double apx;    // error bound of the expansion, x rho^(2k) in pass k
    double na, nb; // band of the evaluating search in units of t: na / q + nb
    double e0;     // midpoint of the probes of pass K1 minus x*
};

__device__ __forceinline__ bool aniso_tail_in_band(const AnisoTail &a, const GoldenTail &t, double e, double &tt)
{
    tt = fma(a.kaph, fma(e, e, t.q23), e);
    const double bd = fma(a.apx, t.r2, fma(a.na, t.invq, a.nb));
    return (__double2hiint(tt) & 0x7fffffff) <= __double2hiint(bd);
}

// RT:189: f(c) < f(d), i.e. t > 0: b = d, the midpoint moves left
__device__ __forceinline__ double aniso_tail_move(const GoldenTail &t, double e, double tt)
{
    return e - __hiloint2double(__double2hiint(t.mid) | (__double2hiint(tt) & 0x80000000), __double2loint(t.mid));
}

// The tail of a step that has a pass inside the band: same passes, rolled, a pass inside the band decided out of line.
template <class TIE>
static __device__ __noinline__ double golden_aniso_tail_ties(const CostAnisoFast cost, const TIE ref, const AnisoTail a,
                                                             double tie2, double theta, double st, double ct)
{
    double e = a.e0;
    int k = RTB_GOLDEN_TAIL_K1;
    for (;;) {
        bool band = false;
#pragma unroll 1
        for (; k < RTB_GOLDEN_ITERS; ++k) {
            double tt;
            if (aniso_tail_in_band(a, kGoldenTail[k], e, tt)) { band = true; break; }
            e = aniso_tail_move(kGoldenTail[k], e, tt);
        }
        if (!band) break;
        const bool left = golden_trig_tie(cost, ref, tie2, theta, st, ct, k);
        e = __dadd_rn(e, left ? -kGoldenTail[k].mid : kGoldenTail[k].mid);
        ++k;
    }
    return e;
}

// Passes K .. RTB_GOLDEN_ITERS - 1 in a straight line; `tie` collects whether any of them was inside the band.  |t| is
// compared as a float: its high word re-biased (truncated towards zero: the band only grows); a |t| below the float
// range counts as inside.
template <int K>
__device__ __forceinline__ void aniso_tail_unrolled(double kaph, float apx, float na, float nb, double &e, bool &tie)
{
    if constexpr (K < RTB_GOLDEN_ITERS) {
        typedef GoldenTailC<K> C;
        const double tt = fma(kaph, fma(e, e, C::q23), e);
        const float bd = fmaf(apx, C::r2, fmaf(na, C::invq, nb));
        const int hi = __double2hiint(tt);
        tie |= (((hi & 0x7fffffff) - 0x38000000) << 3) <= __float_as_int(bd);
        const double mid = C::mid; // RT:189: f(c) < f(d), i.e. t > 0: b = d, the midpoint moves left
        e -= __hiloint2double(__double2hiint(mid) | (hi & 0x80000000), __double2loint(mid));
        aniso_tail_unrolled<K + 1>(kaph, apx, na, nb, e, tie);
    }
}

template <class TIE>
__device__ __forceinline__ bool golden_aniso_tail(const CostAnisoFast &cost, const TIE &ref, double scale, double tie2,
                                                  double theta, double st, double ct, const TrigState &g, double &th_out,
                                                  double &s_out, double &c_out)
{
    constexpr int K1 = RTB_GOLDEN_TAIL_K1;
    const double qK = kGoldenQ[K1], WK = qK * (1.0 / (RTB_GOLDEN_RHO - 0.5));
    // Newton on the first derivative from the kept probe W: the error goes e -> (kappa / 2) e^2 per step, from at most
    // 0.62 WK; three steps reach 1e-8 unless kappa is large AND x* sits at the far end of the bracket: then a fourth
    double F1, F2a, F2b, F2c, f, s1, c1, s2, c2;
    aniso_cost_d12(cost, g.Ws, g.Wc, F1, F2a, f);
    // (products that feed a sum are formed with __dmul_rn: a contraction into the sum that differs between the
    // instantiations of the kernel would break the bit-equality of the output modes)
    const double d0 = __dmul_rn(-F1, rcp_fast3(F2a));
    rotate(g.Ws, g.Wc, d0, s1, c1);
    aniso_cost_d12(cost, s1, c1, F1, F2b, f);
    const double d1 = __dmul_rn(-F1, rcp_fast3(F2b));
    rotate(s1, c1, d1, s2, c2);
    aniso_cost_d12(cost, s2, c2, F1, F2c, f);
    double iA = rcp_fast3(F2c);
    double dl = __dmul_rn(-F1, iA);                 // the last step
    double delta = __dadd_rn(d0, __dadd_rn(d1, dl)); // x* - W
    if (!(fabs(dl) <= 1e-8) && fabs(dl) <= 1e-3) {
        rotate(g.Ws, g.Wc, delta, s1, c1);
        aniso_cost_d12(cost, s1, c1, F1, F2c, f);
        iA = rcp_fast3(F2c);
        dl = __dmul_rn(-F1, iA);
        delta = __dadd_rn(delta, dl);
    }
    const double kap = 2.0 * __dsub_rn(F2b, F2a) * rcp_fast3(d0 * __dadd_rn(F2a, F2b));   // third / second derivative
    AnisoTail a;
    a.e0 = __dsub_rn(g.nl ? -qK : qK, delta);                                      // N left: W is d, the midpoint is W - q
    // every test fails on a NaN
    const bool ok = F2a > 0.0 && F2b > 0.0 && F2c > 0.0 && fabs(d0) <= WK && fabs(d0) >= 1e-10 && fabs(dl) <= 1e-7 &&
                    fabs(kap) <= 64.0 && fabs(a.e0) <= 0.5000005 * WK;
    if (!ok) return false;
    const double g2 = cost.g2m1 + 1.0;
    const double lam = fma(16.0, fmax(g2, rcp_fast3(g2)), 8.0);
    const double hw = 0.5 * WK, hq = __dadd_rn(hw, qK);
    a.kaph = 0.5 * kap;
    a.apx = lam * fma(fabs(d0), fma(0.5 * hw, hw, 0.5 * kGoldenTail[K1].q23), (hq * hq * hq) * (1.0 / 3.0)) *
            rcp_fast3(kGoldenTail[K1].r2);
    const double ties = RTB_TIE_EPS * scale;
    a.na = (0.3125 * ties) * sqrt_fast(2.0 * f) * iA;                      // 1.25 TIE sqrt(2 f*) / (4 F2)
    a.nb = fma((1.25 * 1.556) * ties, rsqrt_fast3(F2c), fma(fabs(kap) + 2.0, dl * dl, RTB_TIE_FLOOR));
    double e = a.e0;
    bool tie = false;
    // the band in single precision, rounded up (a bound): its FFMAs take their constants as immediates
    const float up = 1.000001f;
    aniso_tail_unrolled<K1>(a.kaph, __double2float_ru(a.apx) * up, __double2float_ru(a.na) * up, __double2float_ru(a.nb) * up, e, tie);
    if (tie) e = golden_aniso_tail_ties(cost, ref, a, tie2, theta, st, ct);
    th_out = __dadd_rn(theta, __dadd_rn(g.m, __dsub_rn(e, a.e0)));
    rotate(g.Ws, g.Wc, __dadd_rn(delta, e), s_out, c_out);                           // the final midpoint is x* + e = W + delta + e
    return true;
}

template <class F, class TIE>
__device__ __forceinline__ double golden_trig(const F &cost, const TIE &ref, double scale, double theta, double st,
                                              double ct, double &s_out, double &c_out)
{
    const double tie2 = (RTB_TIE_EPS * scale) * (RTB_TIE_EPS * scale);
    TrigState g;
    trig_init(cost, st, ct, g);
    int k = 0;
#ifndef RTB_NO_GOLDEN_TAIL
    if constexpr (std::is_same<F, CostAnisoFast>::value) {
        golden_trig_passes(cost, ref, tie2, theta, st, ct, g, k, RTB_GOLDEN_TAIL_K1);
        double th;
        if (golden_aniso_tail(cost, ref, scale, tie2, theta, st, ct, g, th, s_out, c_out)) return th;
    }
#endif
    golden_trig_passes(cost, ref, tie2, theta, st, ct, g, k, RTB_GOLDEN_ITERS);
    // (a+b)/2 lies half a probe distance from the kept probe: to its left when the kept probe is d (N is c)
    const double e = g.nl ? -RTB_GOLDEN_FINAL_HALF : RTB_GOLDEN_FINAL_HALF, e2h = 0.5 * e * e;
    s_out = fma(-g.Ws, e2h, fma(g.Wc, e, g.Ws));
    c_out = fma(-g.Wc, e2h, fma(-g.Ws, e, g.Wc));
    return theta + g.m;
}

// Isotropic cost without evaluating it: f(t) = |n1 (cos t, sin t) - A|^2 = n1^2 + |A|^2 - 2 n1 |A| cos(t - ts) is
// symmetric about ts = arg A, so f(c) < f(d) <=> the midpoint of the probes -- which is the bracket midpoint --
// lies right of ts.  One signed distance e = midpoint - ts is carried; a pass subtracts copysign(kGoldenMid[k], e):
// one integer sign transfer and one DADD, no comparison.  ts - theta comes from the cross and dot products of A with
// the ray direction.  Where the reference's computed f(c) - f(d) would be within its own rounding noise --
// |e| <= TIE/2 (|D| / (|A| q) + sqrt(n1 / |A|)) + floor, q = half the probe distance of the pass, D = n1 - |A|; tested
// on the high words, a superset -- the search of that step is redone out of line by golden_sym_slow with the reference's
// own bracket arithmetic, both probes of a pass inside the band evaluated the reference's way.
// Returns false when ts is more than 2^-6 from theta (a turn no step of the tested scenarios makes): the caller then
// runs golden_trig.
struct SymBand { double sa, sb; }; // tie band of pass k: |e| <= sa * kGoldenInvQ[k] + sb

__device__ __forceinline__ bool sym_in_band(double e, const SymBand &bd, int k)
{
    const double thr = fma(bd.sa, kGoldenInvQ[k], bd.sb);
    return (__double2hiint(e) & 0x7fffffff) <= __double2hiint(thr);
}

// The whole search, the reference's way, for the rare step that has a pass inside the band: the bracket with the
// reference's un-fused expressions (RT:184-194), the decisions from the same midpoint rule as the fast loop -- or,
// inside the band, from the costs of both probes in the reference's expression order with libdevice's sincos.
// Returns the final e = (a+b)/2 - ts in the fast loop's convention (sum of the midpoint moves).
static __device__ __noinline__ double golden_sym_slow(const TieIso t, double theta, double delta, const SymBand bd)
{
    const CostIso cf = t.make();
    double a = __dsub_rn(theta, RTB_DELTA_G), b = __dadd_rn(theta, RTB_DELTA_G);
    double c = __dsub_rn(b, __dmul_rn(__dsub_rn(b, a), RTB_GOLD_RATIO));
    double d = __dadd_rn(a, __dmul_rn(__dsub_rn(b, a), RTB_GOLD_RATIO));
    double e = -delta;
    for (int j = 0; j < RTB_GOLDEN_ITERS; ++j) {
        bool left = e > 0.0;
        if (sym_in_band(e, bd, j)) left = cf(c) < cf(d); // RT:189
        if (left) b = d; else a = c;
        c = __dsub_rn(b, __dmul_rn(__dsub_rn(b, a), RTB_GOLD_RATIO));
        d = __dadd_rn(a, __dmul_rn(__dsub_rn(b, a), RTB_GOLD_RATIO));
        const double q = kGoldenMid[j];
        e += left ? -q : q;
    }
    return e;
}

__device__ __forceinline__ bool golden_iso_sym(double n1, double ax, double ay, const TieIso &ref, double theta,
                                               double st, double ct, double &th_out, double &s_out, double &c_out)
{
    const double A2 = fma(ax, ax, ay * ay);
    const double rA = rsqrt_fast3(A2);
    const double cs = ax * rA, ss = ay * rA;     // (cos, sin)(ts)
    const double sd = fma(ct, ss, -(st * cs));   // sin(ts - theta)
    const double cd = fma(ct, cs, st * ss);      // cos(ts - theta)
    if (!(cd > 0.99988)) return false;           // |ts - theta| < 2^-6 (cos 2^-6 = 0.999877933)
    const double t = sd * rcp_fast3(cd), t2 = t * t;
    double p = fma(t2, 1.0 / 9.0, -1.0 / 7.0);   // atan t = t - t^3/3 + t^5/5 - t^7/7 + t^9/9, |t| < 2^-6
    p = fma(p, t2, 0.2);
    p = fma(p, t2, -1.0 / 3.0);
    const double delta = fma(t * t2, p, t);
    const double D = fma(-A2, rA, n1);
    SymBand bd;
    bd.sa = fabs((0.5 * RTB_TIE_EPS) * D * rA);
    // sqrt(n1 / |A|) <= (1 + n1 / |A|) / 2; the floor covers the rounding drift of the reference's own a, b
    bd.sb = fma(0.25 * RTB_TIE_EPS, fma(n1, rA, 1.0), RTB_TIE_FLOOR);
    double e = -delta;                           // midpoint of [theta - pi/2, theta + pi/2] minus ts
    bool tie = false;
#pragma unroll
    for (int k = 0; k < RTB_GOLDEN_ITERS; ++k) { // straight-line: 37 x (band test on the high words, sign transfer, DADD)
        tie |= sym_in_band(e, bd, k);
        const double q = kGoldenMid[k];          // RT:189: f(c) < f(d), i.e. e > 0: b = d, the midpoint moves left
        e -= __hiloint2double(__double2hiint(q) | (__double2hiint(e) & 0x80000000), __double2loint(q));
    }
    if (tie) e = golden_sym_slow(ref, theta, delta, bd); // about one step in a thousand: redone the reference's way
    th_out = theta + (delta + e);
    const double e2h = 0.5 * e * e;              // |e| <= half the final bracket ~3e-8
    s_out = fma(-ss, e2h, fma(cs, e, ss));
    c_out = fma(-cs, e2h, fma(-ss, e, cs));
    return true;
}

// One step.  Updates the state in place; returns dist = |r1 - r0|.  LAZY: theta of atan2-type
// updaters is not evaluated (s.ang stale); `i` is the step index (resync schedule).
template <int METHOD, bool ISO, bool LAZY>
__device__ __forceinline__ double step_fast(const DevField &f, const DevField *fd, const CellBox &in, FastState &s, double h,
                                            double gamma, double g2m1, double (&hist)[3][2], int i, bool bootstrap,
                                            bool &inner, unsigned mask = 0u, double2 *wbuf = nullptr)
{
    constexpr int ADV = advancer_of(METHOD);
    constexpr int UPD = updater_of(METHOD);
    double fx, fy;
    bool ignore = true;
    // grad n minus its tangential part is (u x grad n) n^, n^ = (-uy, ux) (|u| = 1 to rounding): one signed scalar
    // `cross` carries it, k1 = cross * h/n is the curvature times the step, and the second-order displacement is
    // n^ * (h/2) k1.  The step length follows from the same identity: |u h + n^ w|^2 = h^2 + w^2.
    const double cross = fma(s.ux, s.gy, -(s.uy * s.gx));
    const double k1 = cross * s.hn;
    double dist = h; // ADV == 1: the chord is the step itself
    if (ADV == 1) {
        fx = fma(s.ux, h, s.x);
        fy = fma(s.uy, h, s.y);
    } else if (ADV == 2 || bootstrap) {
        const double w = (0.5 * h) * k1;
        fx = fma(-s.uy, w, fma(s.ux, h, s.x));
        fy = fma(s.ux, w, fma(s.uy, h, s.y));
        if (UPD != 2 || bootstrap) { // UPD == 2 takes it from the k1^2 of its rotation below
            const double z = k1 * k1;  // |dx| = h sqrt(1 + (k1/2)^2) = h (1 + z/8 - z^2/128 + O(z^3 2^-10)), z < 2^-16
            dist = abs_lt_hi(k1, HI_2M8) ? fma(h, z * fma(z, -0.0078125, 0.125), h) : sqrt_pos(fma(w, w, h * h));
        }
    } else {
        // curvature_t RT:342-365: the arc of radius R = n / cross through the turn e = k1.  Written without the
        // radius: R (sin(t+e) - sin t) = h (ux sin(e)/e + uy (cos e - 1)/e), so the reference's cancelling difference,
        // its division by the curvature and the square root of the chord (2 R sin(e/2) = h (1 - e^2/24 + e^4/1920))
        // all become short even/odd series in e for |e| < 2^-8; larger turns take the literal form below.
        const double curv = fabs(cross) * s.inv_n;
        const bool straight = curv < RTB_GOLD_TOL; // RT:349: first-order step, angle left alone
        if (straight) ignore = false;
        if (abs_lt_hi(k1, HI_2M8)) {
            const double z = k1 * k1;
            const double S = fma(z, fma(z, K_1_120, -K_SIXTH), 1.0); // sin(e)/e
            const double eC = k1 * fma(z, K_1_24, -0.5);               // (cos e - 1)/e
            double tx = fma(s.uy, eC, s.ux * S), ty = fma(-s.ux, eC, s.uy * S);
            if (straight) { tx = s.ux; ty = s.uy; }
            fx = fma(h, tx, s.x);
            fy = fma(h, ty, s.y);
            dist = straight ? h : h * fma(z, fma(z, 1.0 / 1920.0, -1.0 / 24.0), 1.0);
        } else {
            const double d = curv * h;
            const double sg = (cross < 0) ? -1.0 : 1.0; // RT:360: -1 where gx*uy - gy*ux > 0
            const double e = sg * d;
            double ds, dc; // sin(t+e)-sin(t), cos(t+e)-cos(t)
            if (d <= RTB_SMALL_ANGLE_MAX) {
                const double e2 = e * e;
                const double se = sin_kernel(e, e2), cm = cosm1_small(e2);
                ds = fma(s.uy, cm, s.ux * se);
                dc = fma(s.ux, cm, -(s.uy * se));
            } else { // huge step * curvature: plain evaluation like the reference
                const double ang = (s.ang != s.ang) ? atan2(s.uy, s.ux) : s.ang; // NaN = lazy theta
                double s1, c1;
                sincos_cw(ang + e, s1, c1);
                ds = s1 - s.uy;
                dc = c1 - s.ux;
            }
            const double icurv = s.n * rcp_fast(cross); // signed radius: sg / curvature
            fx = fma(ds, icurv, s.x);
            fy = fma(-dc, icurv, s.y);
            const double dx = s.x - fx, dy = s.y - fy;
            dist = sqrt_pos(fma(dx, dx, dy * dy));
        }
    }
    double fn, fgx, fgy;
    inner = n_gradient_cells(f, fd, in, fx, fy, fn, fgx, fgy, mask, wbuf); // true: new position strictly inside the box
    const double inv_fn = rcp_fast3(fn);
    const double hn1 = h * inv_fn;


    if (UPD == 3 || bootstrap) { // finite differences RT:370-372, bootstrap RT:843-857
        double vx, vy;
        if (bootstrap) {
            if (i == 1) {
                hist[1][0] = fx; hist[1][1] = fy;
                vx = fx - hist[0][0]; vy = fy - hist[0][1];
            } else {
                hist[2][0] = fx; hist[2][1] = fy;
                vx = 3 * fx - 4 * hist[1][0] + hist[0][0];
                vy = 3 * fy - 4 * hist[1][1] + hist[0][1];
            }
        } else {
            vx = 11 * fx - 18 * hist[2][0] + 9 * hist[1][0] - 2 * hist[0][0];
            vy = 11 * fy - 18 * hist[2][1] + 9 * hist[1][1] - 2 * hist[0][1];
            hist[0][0] = hist[1][0]; hist[0][1] = hist[1][1];
            hist[1][0] = hist[2][0]; hist[1][1] = hist[2][1];
            hist[2][0] = fx; hist[2][1] = fy;
        }
        const double r = rsqrt_fast3(fma(vx, vx, vy * vy));
        s.ux = vx * r; s.uy = vy * r;
        s.ang = LAZY ? __longlong_as_double(0x7ff8000000000000ll) : atan2(s.uy, s.ux); // same value in every output mode
    } else if (ADV == 3 && !ignore) {
        // curvature below GOLD_TOL: angle and unit vector unchanged (op3/4/5/10)
    } else if (UPD == 1) { // theta_cost_t RT:393-407
        const double a = fma(0.5 * h, s.gx + fgx, s.n * s.ux);
        const double b = fma(0.5 * h, s.gy + fgy, s.n * s.uy);
        const double r = rsqrt_fast3(fma(a, a, b * b));
        s.ux = a * r; s.uy = b * r;
        s.ang = LAZY ? __longlong_as_double(0x7ff8000000000000ll) : atan2(s.uy, s.ux);
    } else if (UPD == 2) { // tfinal_2o RT:374-391
        // Straight-line form: both rotations use the fixed mid-degree polynomial (valid for |k| < 2^-8, every
        // per-step increment in practice) so the whole update is one basic block the scheduler can interleave
        // with the reciprocal / square-root chains; the rare cases (large curvature*step) and the periodic exact
        // resync of (sin, cos) with the accumulated angle are repaired in ONE conditional block afterwards.
        // k2 = grad n(new) x u(theta + k1) * h/n: with (sin, cos)(theta + k1) = u + (u cm + n^ sk), sk = sin k1,
        // cm = cos k1 - 1, the cross product is crossF (1 + cm) - dotF sk -- no intermediate unit vector -- and the
        // final direction is ONE rotation of u by (k1 + k2)/2.
        const double z = k1 * k1;
        const double sk = fma(k1 * z, fma(z, K_1_120, -K_SIXTH), k1);
        const double cm = z * fma(z, K_1_24, -0.5);
        const double crossF = fma(s.ux, fgy, -(s.uy * fgx)), dotF = fma(s.ux, fgx, s.uy * fgy);
        double k2 = fma(-dotF, sk, fma(crossF, cm, crossF)) * hn1;
        double half = 0.5 * (k1 + k2);
        double ang = s.ang + half;
        double nuy, nux;
        rotate_small(s.uy, s.ux, half, nuy, nux);
        const bool small = abs_lt_hi(k1, HI_2M8) && abs_lt_hi(k2, HI_2M8);
        if (ADV == 2) // |dx| = h sqrt(1 + (k1/2)^2) = h (1 + z/8 - z^2/128 + O(z^3 2^-10)), z = k1^2 < 2^-16
            dist = fma(h, z * fma(z, -0.0078125, 0.125), h);
        if (!small || (i & RTB_RESYNC_MASK) == 0) {
            if (!small) {
                if (ADV == 2) {
                    const double w = (0.5 * h) * k1;
                    dist = sqrt_pos(fma(w, w, h * h));
                }
                double s1, c1;
                sincos_cw(s.ang + k1, s1, c1);
                k2 = (c1 * fgy - s1 * fgx) * hn1;
                ang = s.ang + 0.5 * (k1 + k2);
            }
            sincos_cw(ang, nuy, nux);
        }
        s.ang = ang; s.uy = nuy; s.ux = nux;
    } else if (UPD == 4) { // golden, isotropic cost RT:595,697
        CostIsoFast cf;
        cf.n1 = fn;
        cf.ax = fma(0.5 * h, s.gx + fgx, s.n * s.ux);
        cf.ay = fma(0.5 * h, s.gy + fgy, s.n * s.uy);
        const TieIso ref = {s.n, s.ux, s.uy, s.gx, s.gy, fn, fgx, fgy, h}; // near-ties: the reference's own expression
        double ns, nc, th;
#ifndef RTB_NO_GOLDEN_SYM
        if (!golden_iso_sym(fn, cf.ax, cf.ay, ref, s.ang, s.uy, s.ux, th, ns, nc))
#endif
            th = golden_trig(cf, ref, fn, s.ang, s.uy, s.ux, ns, nc);
        s.ang = th;
        if ((i & RTB_RESYNC_MASK) == 0) sincos_cw(s.ang, ns, nc); // periodic exact resync of the carried (sin, cos)
        s.uy = ns; s.ux = nc;
    } else { // golden, anisotropic cost RT:728,761
        CostAnisoFast cf;
        cf.n1 = fn; cf.n1g2 = fn * (g2m1 + 1.0); cf.g2m1 = g2m1;
        const double ic2 = rcp_fast(s.coef * s.coef);
        const double nc = s.n * s.coef;
        cf.mx = fma(0.5 * h * s.coef, s.gx, nc * s.ux * fma(-(s.uy * s.uy), g2m1 * ic2, 1.0));
        cf.my = fma(0.5 * h * s.coef, s.gy, nc * s.uy * fma(s.ux * s.ux, g2m1 * ic2, 1.0));
        cf.gx1h = 0.5 * h * fgx; cf.gy1h = 0.5 * h * fgy;
        const TieAniso ref = {s.n, s.coef, s.ux, s.uy, s.gx, s.gy, fn, fgx, fgy, h, gamma, g2m1}; // near-ties
        double nsin, ncos;
        s.ang = golden_trig(cf, ref, fn * fmax(gamma, 1.0), s.ang, s.uy, s.ux, nsin, ncos);
        if ((i & RTB_RESYNC_MASK) == 0) sincos_cw(s.ang, nsin, ncos);
        s.uy = nsin; s.ux = ncos;
    }
    s.x = fx; s.y = fy; s.n = fn; s.inv_n = inv_fn; s.hn = hn1; s.gx = fgx; s.gy = fgy;
    if (!ISO) {
        const double gs = gamma * s.uy;
        s.coef = sqrt_fast(fma(gs, gs, s.ux * s.ux));
    }
    return dist;
}

// moments RT:232-245 from the state
template <bool ISO>
__device__ __forceinline__ void moments_fast(const FastState &s, double g2m1, double &px, double &py)
{
    if (ISO) {
        px = s.n * s.ux;
        py = s.n * s.uy;
    } else {
        const double k = g2m1 * rcp_fast(s.coef * s.coef);
        const double nc = s.n * s.coef;
        px = nc * s.ux * fma(-(s.uy * s.uy), k, 1.0);
        py = nc * s.uy * fma(s.ux * s.ux, k, 1.0);
    }
}

} // namespace rtb
