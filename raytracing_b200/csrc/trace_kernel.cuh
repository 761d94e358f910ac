// trace_kernel.cuh -- the integrator: trazar's two nested loops (RT_bench.py:807 x 866) as
// ONE kernel launch.  One thread owns one ray; its whole state lives in registers and the
// thread runs the step loop itself (persistent step loop, no per-step launch).  Output is
// the reference's ray-minor SoA s_ray[(row*6+col)*R + k] (RT_bench.py:802): for a fixed
// row/column a warp writes 32 consecutive doubles = one 256-byte segment per store
// instruction, so trajectory write-back is coalesced by construction; streaming (.cs)
// stores keep the 48 B/ray-step of output from evicting the spline tables out of L2.
#pragma once
#include "step_fast.cuh"

namespace rtb {

#ifndef RTB_MIN_BLOCKS
#define RTB_MIN_BLOCKS 5 // default kernels: <= 96 registers -> 5 blocks of 128 threads per SM
#endif
#ifndef RTB_BLOCK
#define RTB_BLOCK 128
#endif

#define RTB_MAX_PEERS 8

struct TraceArgs {
    DevField f;
    const DevField *f_dev;     // the same struct in global memory (rare out-of-line paths)
    CellBox in;                // cells of the fast region strictly inside the box (make_cell_box)
    const double *theta0, *x0, *y0;
    const double *steps;       // per-ray step or nullptr
    const int64_t *max_sizes;  // per-ray max_size or nullptr
    int64_t R;
    int max_size;
    int out_stride;
    int tail_zero;
    int lane_shift;            // small fans: one ray per 2^lane_shift lanes (the others idle), see pick_lane_shift
    int stage;                 // 1: warp-staged cell records (field.cuh) -- set for large fans, where throughput counts
    double step, gamma;
    double limx_i, limx_s, limy_i, limy_s;
    double *s_ray;             // TRAJ only
    double *d_ray;             // (3,R)
    double *last_row;          // (6,R) or nullptr
    double *px_stats;          // (3,R) or nullptr
    const int *cap_rows;       // (2,R) rows to capture (pass 2 of the Snell metric) or nullptr
    double *cap_out;           // (4,R): x_a, y_a, x_b, y_b
    double *ckpt;              // (ckpt_rows, RTB_CKPT_FIELDS, R) or nullptr: ray state after every 2^ckpt_shift steps --
    int ckpt_shift, ckpt_rows; // written by pass 1 of the Snell metric, read by pass 2 (cap_rows set) to start near row a
    double *exit_out;          // (3,R) or nullptr: point where the last segment crosses the box, travel time there
    unsigned long long *total_steps;
    // Multi-GPU result gather fused into the kernel (rtb_set_gather): every finished ray also stores its d_ray column
    // into the (3, gather_R) array of each peer GPU, at ray offset gather_off, over NVLink peer mappings.
    double *gather[RTB_MAX_PEERS];
    int gather_world;
    int64_t gather_R, gather_off;
};

constexpr int RTB_CKPT_FIELDS = 17; // x y ang ux uy n gx gy coef inv_n hn + op7's three past positions

__device__ __forceinline__ void store_row(double *p, int64_t R, double x, double y, double px, double py,
                                          double T, double ang)
{
    __stcs(p, x);
    __stcs(p + R, y);
    __stcs(p + 2 * R, px);
    __stcs(p + 3 * R, py);
    __stcs(p + 4 * R, T);
    __stcs(p + 5 * R, ang);
}

// EXACT = faithful restatement (reference evaluation order, libdevice sincos, IEEE div/sqrt, FITPACK
// recurrence): the parity anchor.  !EXACT = default B200 form (step_fast.cuh).
// EXTRA = 1: p_x statistics, Snell row capture or exit points requested (kept out of the plain kernels: their
// accumulators cost 15 registers); EXTRA = 2: the same plus saving / resuming from ray-state checkpoints (the two
// passes of the Snell metric; its own flavour because the checkpoint code costs the others 130 B of spills).
template <int METHOD, bool ISO, bool TRAJ, bool EXACT, int EXTRA>
__global__ void __launch_bounds__(RTB_BLOCK, EXACT ? (4 * 128) / RTB_BLOCK : RTB_MIN_BLOCKS) trace_kernel(const TraceArgs a)
{
    constexpr int UPD = updater_of(METHOD);
    constexpr bool LAZY = !EXACT && !TRAJ && (UPD == 1 || UPD == 3); // theta evaluated once at the end
    // Small fans are spread one ray per 2^lane_shift lanes (up to one ray per warp): the step loop of a lone ray is a
    // dependent chain, and 32 rays that sit in 32 different cells make every table load of their warp 32 cache
    // lines wide; spread out, each ray gets a scheduler slot of its own and the fan covers all SMs.
    const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t r = gt >> a.lane_shift;
    const bool valid = r < a.R && (gt & ((1 << a.lane_shift) - 1)) == 0;
    const int64_t rr = r < a.R ? r : a.R - 1; // idle and padded lanes shadow a ray, never step or store
    const double h = a.steps ? a.steps[rr] : a.step;
    int M = a.max_sizes ? (int)a.max_sizes[rr] : a.max_size;
    const double gamma = a.gamma;
    const double g2m1 = gamma * gamma - 1.0;

    int cap_a = -1, cap_b = -1;
    if (EXTRA && !TRAJ && a.cap_rows) {
        cap_a = a.cap_rows[rr];
        cap_b = a.cap_rows[a.R + rr];
        M = min(M, cap_b + 1); // pass 2 stops at the later captured row
    }

    // INITIAL CONDITIONS RT:809-826
    RayState s;
    s.x = a.x0[rr]; s.y = a.y0[rr]; s.ang = a.theta0[rr];
    if (EXACT) sincos(s.ang, &s.uy, &s.ux);
    else sincos_cw(s.ang, s.uy, s.ux);
    if (EXACT) n_gradient<true>(a.f, s.x, s.y, s.n, s.gx, s.gy);
    else n_gradient_fast(a.f, a.f_dev, s.x, s.y, s.n, s.gx, s.gy);
    s.inv_n = EXACT ? 0.0 : rcp_fast(s.n);
    s.hn = h * s.inv_n;
    s.coef = ISO ? 1.0 : anisotropy_sc(s.uy, s.ux, gamma);
    double px, py;
    moments_sc<ISO>(s.n, s.coef, g2m1, s.ux, s.uy, px, py);
    double T = 0.0, dist_sim = 0.0, dist_real = 0.0, nray_prev = s.coef * s.n;
    double hist[3][2] = {{s.x, s.y}, {0, 0}, {0, 0}};

    // p_x statistics over the non-zero entries (np.ma.masked_equal(...,0), RT:1358): shifted sums
    double st_shift = 0.0, st_s1 = 0.0, st_s2 = 0.0;
    int st_cnt = 0;
    if (EXTRA && px != 0.0) { st_shift = px; st_cnt = 1; }

    double capv[4] = {0, 0, 0, 0};
    if (EXTRA && cap_a == 0) { capv[0] = s.x; capv[1] = s.y; }
    if (EXTRA && cap_b == 0) { capv[2] = s.x; capv[3] = s.y; }

    // Pass 2 of the Snell metric resumes from the last checkpoint at or before its first captured row: the state
    // is the one pass 1 had there, the steps that follow are the same instructions on the same values.
    int i_start = 1;
    if (EXTRA == 2 && !TRAJ && a.ckpt && a.cap_rows) {
        const int c = min(cap_a >> a.ckpt_shift, a.ckpt_rows);
        if (c > 0) {
            const double *q = a.ckpt + ((size_t)(c - 1) * RTB_CKPT_FIELDS) * a.R + rr;
            s.x = q[0]; s.y = q[a.R]; s.ang = q[2 * a.R]; s.ux = q[3 * a.R]; s.uy = q[4 * a.R]; s.n = q[5 * a.R];
            s.gx = q[6 * a.R]; s.gy = q[7 * a.R]; s.coef = q[8 * a.R]; s.inv_n = q[9 * a.R]; s.hn = q[10 * a.R];
#pragma unroll
            for (int k = 0; k < 3; ++k) { hist[k][0] = q[(11 + 2 * k) * a.R]; hist[k][1] = q[(12 + 2 * k) * a.R]; }
            const int at = c << a.ckpt_shift;
            if (at == cap_a) { capv[0] = s.x; capv[1] = s.y; }
            if (at == cap_b) { capv[2] = s.x; capv[3] = s.y; }
            i_start = at + 1;
        }
    }

    double *sp = nullptr;
    int next_keep = 0; // next row index that is kept in s_ray
    if (TRAJ) {
        sp = a.s_ray + rr;
        if (valid) store_row(sp, a.R, s.x, s.y, px, py, T, s.ang);
        sp += 6 * a.R;
        next_keep = a.out_stride;
    }

    // exit point (two-point shooting search): start of the last segment and the crossing found at exit
    double pvx = s.x, pvy = s.y, pvT = 0.0, ex_t = 1.0;

    bool alive = valid && i_start < M;
    int last_i = 0;
    int i = i_start;
    const int Mloop = TRAJ ? a.max_size : M;
    // The loop condition must be warp-uniform (lanes may have different M with per-ray max_sizes):
    // every lane iterates until no lane of the warp is alive; `alive` drops at i+1 >= M.
    __shared__ double2 stage[EXACT ? 1 : RTB_BLOCK / 32][RTB_CELL_REC]; // one cell record per warp (field.cuh)
    double2 *const wbuf = (EXACT || !a.stage) ? nullptr : stage[threadIdx.x >> 5];
    for (;; ++i) {
        const unsigned mask = __ballot_sync(0xffffffffu, alive);
        if (!mask) break;
        if (alive) {
            double dist;
            bool inner = false; // new position known to be strictly inside the box (no FP64 compares needed)
            if (EXTRA) { pvx = s.x; pvy = s.y; pvT = T; }
            if (EXACT) {
                double fx, fy, fang, fn, fgx, fgy;
                if (METHOD == 7 && i < 3) { // bootstrap RT:838-864
                    const double dot = s.gx * s.ux + s.gy * s.uy;
                    const double s2 = h * h, den = 2.0 * s.n;
                    fx = (s.x + s.ux * h) + (s.gx - dot * s.ux) * s2 / den;
                    fy = (s.y + s.uy * h) + (s.gy - dot * s.uy) * s2 / den;
                    n_gradient<true>(a.f, fx, fy, fn, fgx, fgy);
                    double vx, vy; // static indices keep VECTOR_LIST in registers
                    if (i == 1) {
                        hist[1][0] = fx; hist[1][1] = fy;
                        vx = hist[1][0] - hist[0][0]; vy = hist[1][1] - hist[0][1];
                    } else {
                        hist[2][0] = fx; hist[2][1] = fy;
                        vx = 3 * hist[2][0] - 4 * hist[1][0] + hist[0][0];
                        vy = 3 * hist[2][1] - 4 * hist[1][1] + hist[0][1];
                    }
                    fang = atan2(vy, vx);
                } else {
                    step_method<METHOD, true>(a.f, s, h, gamma, g2m1, hist, fx, fy, fang, fn, fgx, fgy);
                }
                // store_update_results RT:783-790
                const double dx = s.x - fx, dy = s.y - fy;
                dist = sqrt(dx * dx + dy * dy);
                sincos(fang, &s.uy, &s.ux);
                s.coef = ISO ? 1.0 : anisotropy_sc(s.uy, s.ux, gamma);
                moments_sc<ISO>(fn, s.coef, g2m1, s.ux, s.uy, px, py);
                s.x = fx; s.y = fy; s.ang = fang; s.n = fn; s.gx = fgx; s.gy = fgy;
            } else {
                dist = step_fast<METHOD, ISO, LAZY>(a.f, a.f_dev, a.in, s, h, gamma, g2m1, hist, i, METHOD == 7 && i < 3, inner, mask, wbuf);
                if (TRAJ || EXTRA) moments_fast<ISO>(s, g2m1, px, py); // else only the last row's, after the loop
            }
            dist_sim += dist;
            dist_real += h;
            // RT:871-875
            const double nray = s.coef * s.n;
            T = T + dist * (nray_prev + nray) / 2.0;
            nray_prev = nray;
            last_i = i;
            if (EXTRA && px != 0.0) {
                if (st_cnt == 0) st_shift = px;
                const double d = px - st_shift;
                st_s1 += d;
                st_s2 = fma(d, d, st_s2);
                ++st_cnt;
            }
            if (EXTRA && !TRAJ) {
                if (i == cap_a) { capv[0] = s.x; capv[1] = s.y; }
                if (i == cap_b) { capv[2] = s.x; capv[3] = s.y; }
                if (EXTRA == 2 && a.ckpt && !a.cap_rows && (i & ((1 << a.ckpt_shift) - 1)) == 0 && (i >> a.ckpt_shift) <= a.ckpt_rows) {
                    double *q = a.ckpt + ((size_t)((i >> a.ckpt_shift) - 1) * RTB_CKPT_FIELDS) * a.R + r;
                    q[0] = s.x; q[a.R] = s.y; q[2 * a.R] = s.ang; q[3 * a.R] = s.ux; q[4 * a.R] = s.uy; q[5 * a.R] = s.n;
                    q[6 * a.R] = s.gx; q[7 * a.R] = s.gy; q[8 * a.R] = s.coef; q[9 * a.R] = s.inv_n; q[10 * a.R] = s.hn;
#pragma unroll
                    for (int k = 0; k < 3; ++k) { q[(11 + 2 * k) * a.R] = hist[k][0]; q[(12 + 2 * k) * a.R] = hist[k][1]; }
                }
            }
            // boundary test RT:878 (strict)
            const bool outside = !inner && (s.x > a.limx_s || s.x < a.limx_i || s.y > a.limy_s || s.y < a.limy_i);
            if (outside) {
                if (EXTRA) { // fraction of the last segment inside the box (first limit crossed)
                    double t = 1.0;
                    if (s.x > a.limx_s) t = fmin(t, (a.limx_s - pvx) / (s.x - pvx));
                    if (s.x < a.limx_i) t = fmin(t, (a.limx_i - pvx) / (s.x - pvx));
                    if (s.y > a.limy_s) t = fmin(t, (a.limy_s - pvy) / (s.y - pvy));
                    if (s.y < a.limy_i) t = fmin(t, (a.limy_i - pvy) / (s.y - pvy));
                    ex_t = fmax(t, 0.0);
                }
            }
            alive = !outside && i + 1 < M;
            if (TRAJ && i == next_keep) {
                store_row(sp, a.R, s.x, s.y, px, py, T, s.ang);
                sp += 6 * a.R;
                next_keep += a.out_stride;
            }
        } else if (TRAJ) {
            if (i == next_keep) {
                if (valid && a.tail_zero) store_row(sp, a.R, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0);
                sp += 6 * a.R;
                next_keep += a.out_stride;
            }
        }
    }
    if (TRAJ && a.tail_zero && valid) { // rows after the last step of every ray of this warp stay 0 (np.zeros RT:802)
        for (; i < Mloop; ++i) {
            if (i == next_keep) {
                store_row(sp, a.R, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0);
                sp += 6 * a.R;
                next_keep += a.out_stride;
            }
        }
    }
    if (LAZY && s.ang != s.ang) s.ang = atan2(s.uy, s.ux); // theta of the last row (NaN = not evaluated yet)
    if (!EXACT && !TRAJ && !EXTRA && last_i > 0) moments_fast<ISO>(s, g2m1, px, py);

    // per-ray results RT:888-890
    if (valid) {
        if (a.d_ray) {
            a.d_ray[r] = dist_real;
            a.d_ray[a.R + r] = dist_sim;
            a.d_ray[2 * a.R + r] = (double)last_i;
            for (int pe = 0; pe < a.gather_world; ++pe) { // peer stores: 24 B per ray per GPU, hidden behind the stepping
                double *g = a.gather[pe] + a.gather_off + r;
                g[0] = dist_real;
                g[a.gather_R] = dist_sim;
                g[2 * a.gather_R] = (double)last_i;
            }
        }
        if (a.last_row) {
            a.last_row[r] = s.x; a.last_row[a.R + r] = s.y;
            a.last_row[2 * a.R + r] = px; a.last_row[3 * a.R + r] = py;
            a.last_row[4 * a.R + r] = T; a.last_row[5 * a.R + r] = s.ang;
        }
        if (EXTRA && a.px_stats) {
            double mean = 0.0, sd = 0.0;
            if (st_cnt > 0) {
                const double m1 = st_s1 / st_cnt;
                mean = st_shift + m1;
                const double var = st_s2 / st_cnt - m1 * m1;
                sd = var > 0.0 ? sqrt(var) : 0.0;
            }
            a.px_stats[r] = (double)st_cnt;
            a.px_stats[a.R + r] = mean;
            a.px_stats[2 * a.R + r] = sd;
        }
        if (EXTRA && a.exit_out) { // a ray that never left the box reports its last position (ex_t == 1)
            a.exit_out[r] = pvx + ex_t * (s.x - pvx);
            a.exit_out[a.R + r] = pvy + ex_t * (s.y - pvy);
            a.exit_out[2 * a.R + r] = pvT + ex_t * (T - pvT);
        }
        if (EXTRA && !TRAJ && a.cap_out) {
            a.cap_out[r] = capv[0]; a.cap_out[a.R + r] = capv[1];
            a.cap_out[2 * a.R + r] = capv[2]; a.cap_out[3 * a.R + r] = capv[3];
        }
    }
    // ray-step count = sum of d_ray[2]; one atomic per warp
    if (a.total_steps) {
        unsigned v = valid ? (unsigned)last_i : 0u;
        v = __reduce_add_sync(0xffffffffu, v);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(a.total_steps, (unsigned long long)v);
    }
}

// Host-side launcher, one per METHOD (instantiated in trace_inst.cu).
typedef cudaError_t (*launch_fn)(const TraceArgs &, bool iso, bool traj, bool exact, int block, cudaStream_t);

template <int METHOD>
cudaError_t launch_trace(const TraceArgs &a, bool iso, bool traj, bool exact, int block, cudaStream_t st)
{
    const unsigned grid = (unsigned)(((a.R << a.lane_shift) + block - 1) / block);
    const int extra = (a.ckpt != nullptr && !traj) ? 2
                    : (a.px_stats != nullptr || a.cap_rows != nullptr || a.exit_out != nullptr) ? 1 : 0;
#define RTB_L(I, T, E)                                                                  \
    do {                                                                                \
        if (extra == 2) trace_kernel<METHOD, I, false, E, 2><<<grid, block, 0, st>>>(a); \
        else if (extra) trace_kernel<METHOD, I, T, E, 1><<<grid, block, 0, st>>>(a);    \
        else trace_kernel<METHOD, I, T, E, 0><<<grid, block, 0, st>>>(a);               \
    } while (0)
    if (iso) {
        if (traj) { if (exact) RTB_L(true, true, true); else RTB_L(true, true, false); }
        else      { if (exact) RTB_L(true, false, true); else RTB_L(true, false, false); }
    } else {
        if (traj) { if (exact) RTB_L(false, true, true); else RTB_L(false, true, false); }
        else      { if (exact) RTB_L(false, false, true); else RTB_L(false, false, false); }
    }
#undef RTB_L
    return cudaGetLastError();
}

} // namespace rtb
