/*
 * TEST INFRASTRUCTURE ONLY -- CPU oracle for the RT_bench.py ray-stepping hot path.
 *
 * A plain-C restatement of the reference algorithm, statement by statement, in the
 * reference's own evaluation order (compile with -ffp-contract=off: numpy never fuses
 * a*b+c).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library, and only as the checker / timed CPU baseline.
 * The product path (raytracing_b200/) never links, imports or calls it.
 *
 * Parity pin: validated against the unmodified reference run in the build container
 * (oracle/ref_loader.py) -- see tests/golden/gen_golden.py and tests/test_oracle_golden.py.
 *
 * Citations "RT:a-b" are line ranges of /root/reference/RT_bench.py.  The spline
 * evaluator restates FITPACK (Dierckx) bispev/fpbisp/fpbspl, the third-party code
 * scipy.interpolate.RectBivariateSpline.__call__ runs (scipy 1.18.1 here; not vendored
 * in the reference; call sites RT:153-155).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define RTO_PI 3.141592653589793

/*
 * Sensitivity probe (tests only): when enabled, every sin/cos result is moved by +-1 ulp with
 * probability 1/2, i.e. replaced by the answer another correct (<= 1 ulp) libm could give.  The
 * spread of the outputs under this noise is the reference ALGORITHM's own conditioning and
 * bounds the parity any independent implementation (different libm, FMA contraction) can reach.
 */
static int g_noise = 0;
static _Thread_local uint64_t g_rng = 0x9E3779B97F4A7C15ull;
void rto_set_trig_noise(int seed) { g_noise = seed; }
static inline double noisy(double r)
{
    if (!g_noise) return r;
    g_rng ^= g_rng << 13; g_rng ^= g_rng >> 7; g_rng ^= g_rng << 17;
    unsigned b = (unsigned)(g_rng >> 33) & 3u;
    if (b == 0) return nextafter(r, INFINITY);
    if (b == 1) return nextafter(r, -INFINITY);
    return r;
}
/* Seeds >= 1000 additionally move the ARGUMENT of every sin/cos by +-1 ulp: the angle another correct implementation
 * carries differs from the reference's by the rounding of its own additions, and where the reference evaluates
 * sin(theta + d) - sin(theta) for a tiny d (curvature_t, RT:361-363) the rounding of theta + d alone moves the step
 * by ulp(theta)/curvature -- noise the result-only probe cannot see because oracle and reference round alike. */
static inline double rsin(double x) { return noisy(sin(g_noise >= 1000 ? noisy(x) : x)); }
static inline double rcos(double x) { return noisy(cos(g_noise >= 1000 ? noisy(x) : x)); }

typedef struct {
    const double *tu; /* knots of the first spline axis (physical y, RT:455-457) */
    const double *tv; /* knots of the second spline axis (physical x)            */
    const double *c;  /* (nu-ku-1)*(nv-kv-1) coefficients, first axis major      */
    int nu, nv, ku, kv;
} rto_spline;

typedef struct {
    int method;        /* 1..11 = op1..op11 (RT:469-764)                                   */
    int is_interface;  /* op_interface: compute the Snell error metric (RT:896-919)        */
    int64_t max_size;  /* rows of s_ray (RT:796-799)                                        */
    double gamma;      /* anisotropy parameter (RT:258,266,274,285)                         */
    double step;       /* DELTA_S                                                           */
    double limx_i, limx_s, limy_i, limy_s; /* box, strict tests (RT:878)                    */
} rto_params;

/* ---- FITPACK fpbspl: the k+1 non-zero B-splines at t[l] <= x < t[l+1] (l is 1-based) ---- */
static void fpbspl(const double *t, int k, double x, int l, double *h)
{
    double hh[5];
    h[0] = 1.0;
    for (int j = 1; j <= k; ++j) {
        for (int i = 0; i < j; ++i) hh[i] = h[i];
        h[0] = 0.0;
        for (int i = 1; i <= j; ++i) {
            int li = l + i, lj = li - j;
            double tli = t[li - 1], tlj = t[lj - 1];
            if (tli == tlj) { h[i] = 0.0; continue; }
            double f = hh[i - 1] / (tli - tlj);
            h[i - 1] = h[i - 1] + f * (tli - x);
            h[i] = f * (x - tlj);
        }
    }
}

/* FITPACK fpbisp for a single point: clamp, interval search, basis, contraction. */
static int interval(const double *t, int n, int k, double *arg)
{
    int k1 = k + 1, nk1 = n - k1;
    double tb = t[k1 - 1], te = t[nk1];
    if (*arg < tb) *arg = tb;
    if (*arg > te) *arg = te;
    int l = k1;
    while (!(*arg < t[l]) && l != nk1) ++l;   /* t[l] is 1-based t(l+1) */
    return l;
}

double rto_bispev(const rto_spline *s, double u, double v)
{
    double wu[4], wv[4];
    int ku1 = s->ku + 1, kv1 = s->kv + 1;
    int nkv1 = s->nv - kv1;
    int lu = interval(s->tu, s->nu, s->ku, &u);
    fpbspl(s->tu, s->ku, u, lu, wu);
    int lv = interval(s->tv, s->nv, s->kv, &v);
    fpbspl(s->tv, s->kv, v, lv, wv);
    int l = (lu - ku1) * nkv1;
    int l1 = l + (lv - kv1);
    double sp = 0.0;
    for (int i1 = 0; i1 < ku1; ++i1) {
        int l2 = l1;
        for (int j1 = 0; j1 < kv1; ++j1) {
            sp = sp + s->c[l2] * wu[i1] * wv[j1];
            ++l2;
        }
        l1 += nkv1;
    }
    return sp;
}

void rto_bispev_many(const rto_spline *s, const double *u, const double *v, int64_t n, double *out)
{
    for (int64_t i = 0; i < n; ++i) out[i] = rto_bispev(s, u[i], v[i]);
}

typedef struct { const rto_spline *z, *gu, *gv; } field_t; /* gu = grd[0] (d/dy), gv = grd[1] (d/dx) */

/* n_gradient RT:141-156 */
static void n_gradient(const field_t *f, double x, double y, double *n, double *gx, double *gy)
{
    *n = rto_bispev(f->z, y, x);
    *gx = rto_bispev(f->gv, y, x);
    *gy = rto_bispev(f->gu, y, x);
}

/* anisotropy RT:118-119 */
static double anisotropy(double theta, double gamma)
{
    double a = gamma * rsin(theta), b = rcos(theta);
    return sqrt(a * a + b * b);
}

/* moment RT:217-230; g2m1 = gamma**2-1 */
static double moment(double n, double theta, double gamma, double o0, double o1)
{
    double coef = anisotropy(theta, gamma);
    double g2m1 = gamma * gamma - 1.0;
    return n * coef * o0 * (1.0 + o1 * g2m1 / (coef * coef));
}

/* moments RT:232-245 */
static void moments(double theta, double n, double ux, double uy, double gamma, double *px, double *py)
{
    *px = moment(n, theta, gamma, ux, -(uy * uy));
    *py = moment(n, theta, gamma, uy, ux * ux);
}

/* impulse_t RT:202-214 */
static double impulse_t(double a, double b, double step) { return step * (a + b) / 2.0; }

/* second_order_taylor RT:314-333 */
static void second_order_taylor(double x, double y, double ux, double uy, double step, double n0,
                                double gx, double gy, double *fx, double *fy)
{
    double dot = gx * ux + gy * uy;
    double s2 = step * step;
    *fx = (x + ux * step) + (gx - dot * ux) * s2 / (2.0 * n0);
    *fy = (y + uy * step) + (gy - dot * uy) * s2 / (2.0 * n0);
}

/* Second sensitivity probe (test infrastructure): curvature_t switches between a straight step that keeps the angle and
 * an arc at curv < GOLD_TOL (RT:349).  A ray whose curvature sits on that threshold takes one branch or the other
 * depending on the last bits of curv; scaling the threshold by 1 +- 1e-14 shows which rays those are. */
static double g_curv_tol_scale = 1.0;
void rto_set_curv_tol_scale(double s) { g_curv_tol_scale = s; }


/* curvature_t RT:335-365; returns `ignore` */
static int curvature_t(double ang, double gx, double gy, double ux, double uy, double n0, double x,
                       double y, double step, double gold_tol, double *fx, double *fy)
{
    double dot = gx * ux + gy * uy;
    double qx = gx - dot * ux, qy = gy - dot * uy;
    double curv = sqrt(qx * qx + qy * qy) / n0;
    if (curv < gold_tol * g_curv_tol_scale) { /* scale == 1 unless the threshold probe is on (rto_set_curv_tol_scale) */
        *fx = x + ux * step;
        *fy = y + uy * step;
        return 0;
    }
    double d = curv * step;
    if (gx * uy - gy * ux > 0) {
        *fx = x + (rsin(ang) - rsin(ang - d)) / curv;
        *fy = y + (rcos(ang - d) - rcos(ang)) / curv;
    } else {
        *fx = x + (rsin(ang + d) - rsin(ang)) / curv;
        *fy = y + (-rcos(ang + d) + rcos(ang)) / curv;
    }
    return 1;
}

/* tfinal_2o RT:374-391 */
static double tfinal_2o(double ang, double step, double n0, double n1, double gx0, double gy0,
                        double gx1, double gy1)
{
    double k1 = step * (rcos(ang) * gy0 - rsin(ang) * gx0) / n0;
    double k2 = step * (rcos(ang + k1) * gy1 - rsin(ang + k1) * gx1) / n1;
    return ang + (k1 + k2) / 2.0;
}

/* theta_cost_t RT:393-407 */
static double theta_cost_t(double n0, double ang, double gx0, double gy0, double gx1, double gy1, double step)
{
    return atan2(n0 * rsin(ang) + impulse_t(gy0, gy1, step), n0 * rcos(ang) + impulse_t(gx0, gx1, step));
}

typedef struct {
    int aniso;
    double n0, n1, ux, uy, gx0, gy0, gx1, gy1, step, gamma, coef_i, mi_x, mi_y;
} cost_t;

/* iso lambda RT:595,697; aniso lambda RT:728,761 */
static double cost(const cost_t *c, double th)
{
    if (!c->aniso) {
        double a = c->n1 * rcos(th) - c->n0 * c->ux - impulse_t(c->gx0, c->gx1, c->step);
        double b = c->n1 * rsin(th) - c->n0 * c->uy - impulse_t(c->gy0, c->gy1, c->step);
        return a * a + b * b;
    }
    double st = rsin(th), ct = rcos(th);
    double a = moment(c->n1, th, c->gamma, ct, -(st * st)) - c->mi_x -
               impulse_t(c->coef_i * c->gx0, anisotropy(th, c->gamma) * c->gx1, c->step);
    double b = moment(c->n1, th, c->gamma, st, ct * ct) - c->mi_y -
               impulse_t(c->coef_i * c->gy0, anisotropy(th, c->gamma) * c->gy1, c->step);
    return a * a + b * b;
}

/* golden RT:175-199 */
static double golden(const cost_t *cf, double a, double b, double ratio, double tol)
{
    double c = b - (b - a) * ratio;
    double d = a + (b - a) * ratio;
    while (fabs(c - d) > tol) {
        if (cost(cf, c) < cost(cf, d)) b = d; else a = c;
        c = b - (b - a) * ratio;
        d = a + (b - a) * ratio;
    }
    return (b + a) / 2.0;
}

typedef struct { double x, y, ang, n, gx, gy, ux, uy, coef; } state_t;

/* op1..op11 RT:469-764.  hist = VECTOR_LIST for op7 (3 entries on entry). */
static void step_method(int m, const field_t *f, const rto_params *p, const state_t *s, double (*hist)[2],
                        double gold_ratio, double gold_tol, double *fx, double *fy, double *fang,
                        double *fn, double *fgx, double *fgy)
{
    double h = p->step;
    int ignore = 1;
    int adv = (m == 1 || m == 2) ? 1 : (m == 3 || m == 4 || m == 5 || m == 10) ? 3 : 2;
    if (adv == 1) { *fx = s->x + s->ux * h; *fy = s->y + s->uy * h; }
    else if (adv == 2) second_order_taylor(s->x, s->y, s->ux, s->uy, h, s->n, s->gx, s->gy, fx, fy);
    else ignore = curvature_t(s->ang, s->gx, s->gy, s->ux, s->uy, s->n, s->x, s->y, h, gold_tol, fx, fy);
    n_gradient(f, *fx, *fy, fn, fgx, fgy);
    if (!ignore) { *fang = s->ang; return; }
    switch (m) {
    case 1: case 4: case 8:
        *fang = theta_cost_t(s->n, s->ang, s->gx, s->gy, *fgx, *fgy, h); break;
    case 2: case 3: case 6:
        *fang = tfinal_2o(s->ang, h, s->n, *fn, s->gx, s->gy, *fgx, *fgy); break;
    case 7: { /* finite_diff RT:370-372 with VECTOR_LIST append/pop RT:646-648 */
        double vx = 11 * (*fx) - 18 * hist[2][0] + 9 * hist[1][0] - 2 * hist[0][0];
        double vy = 11 * (*fy) - 18 * hist[2][1] + 9 * hist[1][1] - 2 * hist[0][1];
        *fang = atan2(vy, vx);
        hist[0][0] = hist[1][0]; hist[0][1] = hist[1][1];
        hist[1][0] = hist[2][0]; hist[1][1] = hist[2][1];
        hist[2][0] = *fx; hist[2][1] = *fy;
        break; }
    default: { /* 5, 9 iso golden; 10, 11 aniso golden */
        cost_t c;
        c.aniso = (m >= 10);
        c.n0 = s->n; c.n1 = *fn; c.ux = s->ux; c.uy = s->uy;
        c.gx0 = s->gx; c.gy0 = s->gy; c.gx1 = *fgx; c.gy1 = *fgy;
        c.step = h; c.gamma = p->gamma; c.coef_i = s->coef;
        if (c.aniso) {
            c.mi_x = moment(s->n, s->ang, p->gamma, s->ux, -(s->uy * s->uy));
            c.mi_y = moment(s->n, s->ang, p->gamma, s->uy, s->ux * s->ux);
        } else c.mi_x = c.mi_y = 0;
        double dg = RTO_PI / 2; /* DELTA_G RT:64 */
        *fang = golden(&c, s->ang - dg, s->ang + dg, gold_ratio, gold_tol);
        break; }
    }
}

/* Snell error metric RT:896-919; px,py = positions of rows 0..i */
static double snell_error(double theta0, const double *px, const double *py, int64_t i)
{
    double angreal;
    if (theta0 < RTO_PI / 4) angreal = 90 - 180 * theta0 / RTO_PI;
    else if (theta0 == RTO_PI / 4) angreal = 0;
    else angreal = 180 * asin(sqrt(2.0) * rsin(RTO_PI / 2 - theta0)) / RTO_PI;
    int64_t ib = (int64_t)(9.5 * (double)i / 10);
    int64_t ia = (int64_t)((double)(9 * i) / 10);
    double distx = px[ib] - px[ia], disty = py[ib] - py[ia];
    double angsim = 180 * atan(fabs(distx / disty)) / RTO_PI;
    return fabs(angsim - angreal);
}

/*
 * trazar RT:766-948 for rays k=0..R-1 with explicit fans (theta0,x0,y0 replace
 * theta_v/pos_x of constants(), RT:809-813).
 *   s_ray : (max_size,6,R) C-order or NULL      d_ray : (3,R)      errors : (R) or NULL
 *   final : (6,R) or NULL = the last written row of s_ray per ray (FINAL mode check)
 * Returns the total number of ray-steps (sum of d_ray[2]).
 */
int64_t rto_trace(const rto_params *p, const rto_spline *z, const rto_spline *gu, const rto_spline *gv,
                  const double *theta0, const double *x0, const double *y0, int64_t R, double *s_ray,
                  double *d_ray, double *errors, double *final, int nthreads)
{
    const double gold_ratio = (sqrt(5.0) - 1) / 2; /* RT:65 */
    const double gold_tol = sqrt(2.220446049250313e-16); /* RT:66 */
    field_t f = { z, gu, gv };
    int64_t M = p->max_size, total = 0;
    int m = p->method;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : total)
    for (int64_t k = 0; k < R; ++k) {
        double *px = NULL, *py = NULL;
        if (p->is_interface && errors) {
            px = (double *)malloc(sizeof(double) * (size_t)M);
            py = (double *)malloc(sizeof(double) * (size_t)M);
        }
        g_rng = 0x9E3779B97F4A7C15ull ^ ((uint64_t)(g_noise + 1) * 0xD1B54A32D192ED03ull) ^ ((uint64_t)k << 20);
        state_t s;
        double hist[3][2] = {{0, 0}, {0, 0}, {0, 0}};
        double row[6];
        s.x = x0[k]; s.y = y0[k]; s.ang = theta0[k];
        s.ux = rcos(s.ang); s.uy = rsin(s.ang);
        n_gradient(&f, s.x, s.y, &s.n, &s.gx, &s.gy);
        s.coef = anisotropy(s.ang, p->gamma);
        double mx, my;
        moments(s.ang, s.n, s.ux, s.uy, p->gamma, &mx, &my);
        row[0] = s.x; row[1] = s.y; row[2] = mx; row[3] = my; row[4] = 0.0; row[5] = s.ang;
        if (s_ray) for (int c = 0; c < 6; ++c) s_ray[(0 * 6 + c) * R + k] = row[c];
        if (px) { px[0] = s.x; py[0] = s.y; }
        double dist_sim = 0, dist_real = 0, nray_prev = s.coef * s.n, T = 0.0;
        int64_t i = 0, loop_iter = 1;
        if (m == 7) loop_iter = 3; /* RT:837 */
        for (i = 1; i < M; ++i) {
            double fx, fy, fang, fn, fgx, fgy;
            if (m == 7 && i < 3) { /* bootstrap RT:838-864 */
                if (i == 1) { hist[0][0] = s.x; hist[0][1] = s.y; }
                second_order_taylor(s.x, s.y, s.ux, s.uy, p->step, s.n, s.gx, s.gy, &fx, &fy);
                n_gradient(&f, fx, fy, &fn, &fgx, &fgy);
                hist[i][0] = fx; hist[i][1] = fy;
                double vx, vy;
                if (i == 1) { vx = hist[1][0] - hist[0][0]; vy = hist[1][1] - hist[0][1]; }
                else { vx = 3 * hist[2][0] - 4 * hist[1][0] + hist[0][0]; vy = 3 * hist[2][1] - 4 * hist[1][1] + hist[0][1]; }
                fang = atan2(vy, vx);
            } else {
                step_method(m, &f, p, &s, hist, gold_ratio, gold_tol, &fx, &fy, &fang, &fn, &fgx, &fgy);
            }
            /* store_update_results RT:783-790 */
            double dx = s.x - fx, dy = s.y - fy;
            double dist = sqrt(dx * dx + dy * dy);
            dist_sim += dist;
            dist_real += p->step;
            s.ux = rcos(fang); s.uy = rsin(fang);
            s.coef = anisotropy(fang, p->gamma);
            moments(fang, fn, s.ux, s.uy, p->gamma, &mx, &my);
            s.x = fx; s.y = fy; s.ang = fang; s.n = fn; s.gx = fgx; s.gy = fgy;
            /* RT:871-875 */
            double nray = s.coef * s.n;
            T = T + dist * (nray_prev + nray) / 2.0;
            nray_prev = nray;
            row[0] = s.x; row[1] = s.y; row[2] = mx; row[3] = my; row[4] = T; row[5] = s.ang;
            if (s_ray) for (int c = 0; c < 6; ++c) s_ray[((size_t)i * 6 + c) * R + k] = row[c];
            if (px) { px[i] = s.x; py[i] = s.y; }
            if (m == 7 && i < loop_iter) continue; /* the bootstrap rows are written before the loop that tests the box, RT:838-866 */
            if (s.x > p->limx_s || s.x < p->limx_i || s.y > p->limy_s || s.y < p->limy_i) break; /* RT:878 */
        }
        if (i >= M) i = M - 1; /* loop ran to completion: python leaves i = max_size-1 */
        d_ray[0 * R + k] = dist_real; d_ray[1 * R + k] = dist_sim; d_ray[2 * R + k] = (double)i; /* RT:888-890 */
        if (final) for (int c = 0; c < 6; ++c) final[c * R + k] = row[c];
        if (px) { errors[k] = snell_error(theta0[k], px, py, i); free(px); free(py); }
        else if (errors) errors[k] = 0.0;
        total += i;
    }
    return total;
}

int rto_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
