// step.cuh -- the eleven step methods op1..op11 (RT_bench.py:469-764) as one device
// function template.  METHOD is a compile-time constant so every kernel instantiation
// contains only its own advancer/updater; sin/cos of the current angle are carried in the
// ray state (the reference recomputes the same values up to four times per step).
#pragma once
#include "field.cuh"

namespace rtb {

#define RTB_PI 3.141592653589793
// Golden search parameters RT_bench.py:64-66
#define RTB_DELTA_G (RTB_PI / 2)
#define RTB_GOLD_RATIO 0.6180339887498949  // (sqrt(5)-1)/2
#define RTB_GOLD_TOL 1.4901161193847656e-08 // sqrt(DBL_EPSILON)

struct RayState {
    double x, y, ang;   // position, group angle
    double ux, uy;      // cos(ang), sin(ang)
    double n, gx, gy;   // field at (x,y)
    double coef;        // anisotropy(ang, gamma)
    double inv_n;       // 1/n (default kernels only)
    double hn;          // h/n (default kernels only): h*(u x grad n)/n and h^2/(2n) share it
};

// Position advancer of each method: 1 = first_order_taylor, 2 = second_order_taylor, 3 = curvature_t
__host__ __device__ constexpr int advancer_of(int m)
{
    return (m == 1 || m == 2) ? 1 : (m == 3 || m == 4 || m == 5 || m == 10) ? 3 : 2;
}
// Angle updater: 1 = theta_cost_t, 2 = tfinal_2o, 3 = finite_diff, 4 = golden iso, 5 = golden aniso
__host__ __device__ constexpr int updater_of(int m)
{
    return (m == 1 || m == 4 || m == 8) ? 1 : (m == 2 || m == 3 || m == 6) ? 2 : (m == 7) ? 3
         : (m == 5 || m == 9) ? 4 : 5;
}

// anisotropy RT_bench.py:118-119 from sin/cos of the angle
__device__ __forceinline__ double anisotropy_sc(double s, double c, double gamma)
{
    const double a = gamma * s;
    return sqrt(a * a + c * c);
}

// moment RT_bench.py:217-230 with coef precomputed; g2m1 = gamma^2-1
__device__ __forceinline__ double moment_c(double n, double coef, double g2m1, double o0, double o1)
{
    return n * coef * o0 * (1.0 + o1 * g2m1 / (coef * coef));
}

// moments RT_bench.py:232-245.  ISO (gamma == 1): g2m1 == 0 so the bracket is exactly 1.
template <bool ISO>
__device__ __forceinline__ void moments_sc(double n, double coef, double g2m1, double c, double s, double &px,
                                           double &py)
{
    if (ISO) {
        px = n * coef * c;
        py = n * coef * s;
    } else {
        px = moment_c(n, coef, g2m1, c, -(s * s));
        py = moment_c(n, coef, g2m1, s, c * c);
    }
}

struct CostIso { // lambda RT_bench.py:595,697
    double n1, n0ux, n0uy, ix, iy; // n0*u and the impulse integrals, subtracted in the reference's order
    __device__ __forceinline__ double operator()(double th) const
    {
        double s, c;
        sincos(th, &s, &c);
        return eval_sc(s, c);
    }
    __device__ __forceinline__ double eval_sc(double s, double c) const // reference evaluation order
    {
        const double a = n1 * c - n0ux - ix;
        const double b = n1 * s - n0uy - iy;
        return a * a + b * b;
    }
};

struct CostAniso { // lambda RT_bench.py:728,761
    double n1, gamma, g2m1, mi_x, mi_y, cgx0, cgy0, gx1, gy1, step;
    __device__ __forceinline__ double operator()(double th) const
    {
        double s, c;
        sincos(th, &s, &c);
        return eval_sc(s, c);
    }
    __device__ __forceinline__ double eval_sc(double s, double c) const // reference evaluation order
    {
        const double coef = anisotropy_sc(s, c, gamma);
        const double a = moment_c(n1, coef, g2m1, c, -(s * s)) - mi_x - step * (cgx0 + coef * gx1) / 2.0;
        const double b = moment_c(n1, coef, g2m1, s, c * c) - mi_y - step * (cgy0 + coef * gy1) / 2.0;
        return a * a + b * b;
    }
};

// golden RT_bench.py:175-199 (no caching of f(c)/f(d), exactly like the reference)
template <class F>
__device__ __forceinline__ double golden(const F &func, double a, double b)
{
    double c = b - (b - a) * RTB_GOLD_RATIO;
    double d = a + (b - a) * RTB_GOLD_RATIO;
    while (fabs(c - d) > RTB_GOLD_TOL) {
        if (func(c) < func(d)) b = d; else a = c;
        c = b - (b - a) * RTB_GOLD_RATIO;
        d = a + (b - a) * RTB_GOLD_RATIO;
    }
    return (b + a) / 2.0;
}

// One step of opMETHOD.  hist = the three previous positions (VECTOR_LIST, op7 only).
// Outputs the new position, angle and field; the caller does store_update_results.
template <int METHOD, bool EXACT>
__device__ __forceinline__ void step_method(const DevField &f, const RayState &s, double h, double gamma,
                                            double g2m1, double (&hist)[3][2], double &fx, double &fy,
                                            double &fang, double &fn, double &fgx, double &fgy)
{
    constexpr int ADV = advancer_of(METHOD);
    constexpr int UPD = updater_of(METHOD);
    bool ignore = true;
    if (ADV == 1) { // first_order_taylor RT:300-312
        fx = s.x + s.ux * h;
        fy = s.y + s.uy * h;
    } else if (ADV == 2) { // second_order_taylor RT:314-333
        const double dot = s.gx * s.ux + s.gy * s.uy;
        const double s2 = h * h;
        const double den = 2.0 * s.n;
        fx = (s.x + s.ux * h) + (s.gx - dot * s.ux) * s2 / den;
        fy = (s.y + s.uy * h) + (s.gy - dot * s.uy) * s2 / den;
    } else { // curvature_t RT:335-365
        const double dot = s.gx * s.ux + s.gy * s.uy;
        const double qx = s.gx - dot * s.ux, qy = s.gy - dot * s.uy;
        const double curv = sqrt(qx * qx + qy * qy) / s.n;
        if (curv < RTB_GOLD_TOL) {
            fx = s.x + s.ux * h;
            fy = s.y + s.uy * h;
            ignore = false;
        } else {
            const double d = curv * h;
            double sd, cd;
            if (s.gx * s.uy - s.gy * s.ux > 0) {
                sincos(s.ang - d, &sd, &cd);
                fx = s.x + (s.uy - sd) / curv;
                fy = s.y + (cd - s.ux) / curv;
            } else {
                sincos(s.ang + d, &sd, &cd);
                fx = s.x + (sd - s.uy) / curv;
                fy = s.y + (-cd + s.ux) / curv;
            }
        }
    }
    n_gradient<EXACT>(f, fx, fy, fn, fgx, fgy);

    if (UPD == 3) { // finite_diff RT:370-372 with VECTOR_LIST append/pop RT:646-648
        const double vx = 11 * fx - 18 * hist[2][0] + 9 * hist[1][0] - 2 * hist[0][0];
        const double vy = 11 * fy - 18 * hist[2][1] + 9 * hist[1][1] - 2 * hist[0][1];
        fang = atan2(vy, vx);
        hist[0][0] = hist[1][0]; hist[0][1] = hist[1][1];
        hist[1][0] = hist[2][0]; hist[1][1] = hist[2][1];
        hist[2][0] = fx; hist[2][1] = fy;
        return;
    }
    if (ADV == 3 && !ignore) { // curvature is low: keep the angle (op3/4/5/10)
        fang = s.ang;
        return;
    }
    if (UPD == 1) { // theta_cost_t RT:393-407
        fang = atan2(s.n * s.uy + h * (s.gy + fgy) / 2.0, s.n * s.ux + h * (s.gx + fgx) / 2.0);
    } else if (UPD == 2) { // tfinal_2o RT:374-391
        const double k1 = h * (s.ux * s.gy - s.uy * s.gx) / s.n;
        double s1, c1;
        sincos(s.ang + k1, &s1, &c1);
        const double k2 = h * (c1 * fgy - s1 * fgx) / fn;
        fang = s.ang + (k1 + k2) / 2.0;
    } else if (UPD == 4) {
        CostIso cf;
        cf.n1 = fn;
        cf.n0ux = s.n * s.ux; cf.n0uy = s.n * s.uy;
        cf.ix = h * (s.gx + fgx) / 2.0; cf.iy = h * (s.gy + fgy) / 2.0;
        fang = golden(cf, s.ang - RTB_DELTA_G, s.ang + RTB_DELTA_G);
    } else {
        CostAniso cf;
        cf.n1 = fn; cf.gamma = gamma; cf.g2m1 = g2m1; cf.step = h;
        cf.mi_x = moment_c(s.n, s.coef, g2m1, s.ux, -(s.uy * s.uy)); // RT:725-726 (coef of i_angle == s.coef)
        cf.mi_y = moment_c(s.n, s.coef, g2m1, s.uy, s.ux * s.ux);
        cf.cgx0 = s.coef * s.gx; cf.cgy0 = s.coef * s.gy;
        cf.gx1 = fgx; cf.gy1 = fgy;
        fang = golden(cf, s.ang - RTB_DELTA_G, s.ang + RTB_DELTA_G);
    }
}

} // namespace rtb
