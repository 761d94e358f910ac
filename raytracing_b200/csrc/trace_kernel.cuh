// trace_kernel.cuh -- the integrator: trazar's two nested loops (RT_bench.py:807 x 866) as
// ONE kernel launch.  One thread owns one ray; its whole state lives in registers and the
// thread runs the step loop itself (persistent step loop, no per-step launch).  Output is
// the reference's ray-minor SoA s_ray[(row*6+col)*R + k] (RT_bench.py:802): for a fixed
// row/column a warp owns 32 consecutive doubles = one 256-byte segment.
//
// Two kernels share the per-ray code (ray_init / ray_step / ray_finish):
//   trace_kernel        static mapping thread <-> ray.  Every output mode; the only one that writes trajectories:
//                       per-lane streaming stores (a warp's 32 doubles of one (row, column) are one 256-byte segment),
//                       or rows staged per warp in shared memory that leave as 2-D TMA stores, one per tile of
//                       RTB_TRAJ_ROWS rows x 6 columns x 32 rays (cp.async.bulk.tensor shared -> global; traj_stage.cuh);
//   trace_queue_kernel  persistent warps that pull rays from a global queue (atomic counter): when lanes of a warp
//                       have finished their rays (RT:878 break, or per-ray max_size) the warp compacts -- the dead
//                       lanes take the next rays of the queue while the others keep stepping -- so a divergent fan
//                       (unsorted launch angles, mixed DELTA_S candidates) keeps all 32 lanes busy.  FINAL-type
//                       outputs only (a trajectory needs the static mapping for its coalesced rows).
// Both produce bit-identical results: rays are independent and a ray's arithmetic does not depend on its lane.
#pragma once
#include "step_fast.cuh"
#include "traj_stage.cuh"

namespace rtb {

#ifndef RTB_MIN_BLOCKS
#define RTB_MIN_BLOCKS 5 // default kernels: <= 96 registers -> 5 blocks of 128 threads per SM
#endif
// The golden-section steppers (op8-op11) carry the search state on top of the ray's: with 128 registers (4 blocks per
// SM) their pass loop has no spill traffic, worth more than the fifth block (measured +2..5 %, op9 / op10 / op11).
constexpr int min_blocks_of(int method) { return updater_of(method) >= 4 ? 4 : RTB_MIN_BLOCKS; }
#ifndef RTB_QUEUE_MIN_BLOCKS
#define RTB_QUEUE_MIN_BLOCKS 5
#endif
#ifndef RTB_TRAJ_MIN_BLOCKS
#define RTB_TRAJ_MIN_BLOCKS 5 // trajectory kernels
#endif
#ifndef RTB_BLOCK
#define RTB_BLOCK 128
#endif
#ifndef RTB_QUEUE_WAIT
#define RTB_QUEUE_WAIT 16 // step-loop iterations a warp keeps going with dead lanes before it compacts (see the kernel)
#endif

#define RTB_MAX_PEERS 8

// EXTRA: which per-step extras a kernel instantiation carries (their accumulators cost registers, so the plain
// kernels have none): p_x statistics | exit point of the last segment | Snell row capture + ray-state checkpoints.
constexpr int X_STATS = 1, X_EXIT = 2, X_SNELL = 4, X_ALL = 7;

struct TraceArgs {
    DevField f;
    const DevField *f_dev;     // the same struct in global memory (rare out-of-line paths)
    CellBox in;                // cells of the fast region strictly inside the box (make_cell_box)
    const double *theta0, *x0, *y0;
    const double *steps;       // per-ray step or nullptr
    const int64_t *max_sizes;  // per-ray max_size or nullptr
    int64_t R;
    int64_t cand_R;            // > 0: ray r is fan ray r % cand_R of candidate r / cand_R -- theta0/x0/y0 have cand_R
                               // entries, steps/max_sizes one entry per candidate (rtb_sweep: no per-ray copies)
    int max_size;
    int out_stride;
    int tail_zero;
    int lane_shift;            // small fans: one ray per 2^lane_shift lanes (the others idle), see pick_lane_shift
    int stage;                 // 1: warp-staged cell records (field.cuh) -- set for large fans, where throughput counts
    double step, gamma;
    double limx_i, limx_s, limy_i, limy_s;
    double *s_ray;             // TRAJ only
    int traj_bulk;             // TRAJ: rows leave through shared memory as 2-D TMA stores (traj_stage.cuh), else per-lane stores
    TrajMaps tmaps;            // tensor maps of s_ray for the staged path (encoded by rtb_launch)
    double *d_ray;             // (3,R)
    double *last_row;          // (6,R) or nullptr
    double *px_stats;          // (3,R) or nullptr
    const int *cap_rows;       // (2,R) rows to capture (pass 2 of the Snell metric) or nullptr
    double *cap_out;           // (4,R): x_a, y_a, x_b, y_b
    double *ckpt;              // (ckpt_rows, RTB_CKPT_FIELDS, R) or nullptr: ray state after every 2^ckpt_shift steps --
    int ckpt_shift, ckpt_rows; // written by pass 1 of the Snell metric, read by pass 2 (cap_rows set) to start near row a
    double *exit_out;          // (3,R) or nullptr: point where the last segment crosses the box, travel time there
    unsigned long long *total_steps;
    unsigned long long *nonfinite;  // rays that ended with a non-finite position or travel time (rtb_stats.nonfinite_rays)
    unsigned long long *queue_head; // trace_queue_kernel: next unclaimed ray (zeroed before the launch)
    // Multi-GPU result gather fused into the kernel (rtb_set_gather*): every finished ray also stores its d_ray column
    // into the (3, gather_R) array of each peer GPU over NVLink peer mappings, at the ray's GLOBAL index:
    // gather_off + r for contiguous shards, ((r >> b) * nranks + rank) << b | (r & (2^b - 1)) for block-cyclic ones.
    double *gather[RTB_MAX_PEERS];
    int gather_world;
    int gather_bshift;         // log2 of the block-cyclic block, -1 = contiguous
    int gather_nranks, gather_rank;
    int64_t gather_R, gather_off;
};

constexpr int RTB_CKPT_FIELDS = 17; // x y ang ux uy n gx gy coef inv_n hn + op7's three past positions

// x != 0 (np.ma.masked_equal(..., 0), RT:1358) without an FP64 compare: +-0 are the only doubles whose bits below the
// sign are all zero
__device__ __forceinline__ bool nonzero_bits(double x)
{
    return (((unsigned)__double2hiint(x) << 1) | (unsigned)__double2loint(x)) != 0u;
}

// Everything one ray carries besides its RayState; all indices are compile-time, so it lives in registers.
struct RayRegs {
    RayState s;
    double h;
    int M;                      // this ray's max_size (rows), capped at the later Snell sample row in pass 2
    int i_start;                // first step index (1, or after the checkpoint a pass-2 ray resumes from)
    double T, dist_sim, dist_real, nray_prev, px, py;
    double hist[3][2];          // op7: VECTOR_LIST
    double st_shift, st_s1, st_s2; // p_x statistics over the non-zero entries: shifted sums
    int st_cnt;
    int cap_a, cap_b;
    double capv[4];
    double pvx, pvy, pvT, ex_t; // start of the last segment and the fraction of it inside the box
    int last_i;
};

// INITIAL CONDITIONS RT:809-826 of ray rr.
template <int METHOD, bool ISO, bool TRAJ, bool EXACT, int EXTRA>
__device__ __forceinline__ void ray_init(const TraceArgs &a, int64_t rr, double gamma, double g2m1, RayRegs &q)
{
    RayState &s = q.s;
    int64_t fi = rr, ci = rr; // fan entry / candidate entry of this ray
    if (a.cand_R > 0) { ci = rr / a.cand_R; fi = rr - ci * a.cand_R; }
    q.h = a.steps ? a.steps[ci] : a.step;
    q.M = a.max_sizes ? (int)a.max_sizes[ci] : a.max_size;
    q.cap_a = -1; q.cap_b = -1;
    if ((EXTRA & X_SNELL) && !TRAJ && a.cap_rows) {
        q.cap_a = a.cap_rows[rr];
        q.cap_b = a.cap_rows[a.R + rr];
        q.M = min(q.M, q.cap_b + 1); // pass 2 stops at the later captured row
    }
    s.x = a.x0[fi]; s.y = a.y0[fi]; s.ang = a.theta0[fi];
    if (EXACT) sincos(s.ang, &s.uy, &s.ux);
    else sincos_cw(s.ang, s.uy, s.ux);
    if (EXACT) n_gradient<true>(a.f, s.x, s.y, s.n, s.gx, s.gy);
    else n_gradient_fast(a.f, a.f_dev, s.x, s.y, s.n, s.gx, s.gy);
    s.inv_n = EXACT ? 0.0 : rcp_fast(s.n);
    s.hn = q.h * s.inv_n;
    s.coef = (ISO && !EXACT) ? 1.0 : anisotropy_sc(s.uy, s.ux, gamma); // EXACT: sqrt(sin^2 + cos^2) like RT:118 (1 +- 1 ulp)
    moments_sc<ISO>(s.n, s.coef, g2m1, s.ux, s.uy, q.px, q.py);
    q.T = 0.0; q.dist_sim = 0.0; q.dist_real = 0.0; q.nray_prev = s.coef * s.n;
    q.hist[0][0] = s.x; q.hist[0][1] = s.y;
    q.hist[1][0] = 0; q.hist[1][1] = 0; q.hist[2][0] = 0; q.hist[2][1] = 0;
    q.st_shift = 0.0; q.st_s1 = 0.0; q.st_s2 = 0.0; q.st_cnt = 0;
    if ((EXTRA & X_STATS) && q.px != 0.0) { q.st_shift = q.px; q.st_cnt = 1; }
    q.capv[0] = q.capv[1] = q.capv[2] = q.capv[3] = 0.0;
    if ((EXTRA & X_SNELL) && q.cap_a == 0) { q.capv[0] = s.x; q.capv[1] = s.y; }
    if ((EXTRA & X_SNELL) && q.cap_b == 0) { q.capv[2] = s.x; q.capv[3] = s.y; }
    q.pvx = s.x; q.pvy = s.y; q.pvT = 0.0; q.ex_t = 1.0;
    q.last_i = 0;
    // Pass 2 of the Snell metric resumes from the last checkpoint at or before its first captured row: the state
    // is the one pass 1 had there, the steps that follow are the same instructions on the same values.
    q.i_start = 1;
    if ((EXTRA & X_SNELL) && !TRAJ && a.ckpt && a.cap_rows) {
        const int c = min(q.cap_a >> a.ckpt_shift, a.ckpt_rows);
        if (c > 0) {
            const double *p = a.ckpt + ((size_t)(c - 1) * RTB_CKPT_FIELDS) * a.R + rr;
            s.x = p[0]; s.y = p[a.R]; s.ang = p[2 * a.R]; s.ux = p[3 * a.R]; s.uy = p[4 * a.R]; s.n = p[5 * a.R];
            s.gx = p[6 * a.R]; s.gy = p[7 * a.R]; s.coef = p[8 * a.R]; s.inv_n = p[9 * a.R]; s.hn = p[10 * a.R];
#pragma unroll
            for (int k = 0; k < 3; ++k) { q.hist[k][0] = p[(11 + 2 * k) * a.R]; q.hist[k][1] = p[(12 + 2 * k) * a.R]; }
            const int at = c << a.ckpt_shift;
            if (at == q.cap_a) { q.capv[0] = s.x; q.capv[1] = s.y; }
            if (at == q.cap_b) { q.capv[2] = s.x; q.capv[3] = s.y; }
            q.i_start = at + 1;
        }
    }
}

// One execution of the loop body RT:867-879 for step index i of ray r (r indexes the checkpoint rows).
// Returns whether the ray goes on (inside the box and i + 1 < max_size).
template <int METHOD, bool ISO, bool TRAJ, bool EXACT, int EXTRA, bool LAZY>
__device__ __forceinline__ bool ray_step(const TraceArgs &a, int64_t r, double gamma, double g2m1, RayRegs &q, int i,
                                         unsigned mask, double2 *wbuf)
{
    RayState &s = q.s;
    const double h = q.h;
    double dist;
    bool inner = false; // new position known to be strictly inside the box (no FP64 compares needed)
    if (EXTRA & X_EXIT) { q.pvx = s.x; q.pvy = s.y; q.pvT = q.T; }
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
                q.hist[1][0] = fx; q.hist[1][1] = fy;
                vx = q.hist[1][0] - q.hist[0][0]; vy = q.hist[1][1] - q.hist[0][1];
            } else {
                q.hist[2][0] = fx; q.hist[2][1] = fy;
                vx = 3 * q.hist[2][0] - 4 * q.hist[1][0] + q.hist[0][0];
                vy = 3 * q.hist[2][1] - 4 * q.hist[1][1] + q.hist[0][1];
            }
            fang = atan2(vy, vx);
        } else {
            step_method<METHOD, true>(a.f, s, h, gamma, g2m1, q.hist, fx, fy, fang, fn, fgx, fgy);
        }
        // store_update_results RT:783-790
        const double dx = s.x - fx, dy = s.y - fy;
        dist = sqrt(dx * dx + dy * dy);
        sincos(fang, &s.uy, &s.ux);
        s.coef = anisotropy_sc(s.uy, s.ux, gamma);
        moments_sc<ISO>(fn, s.coef, g2m1, s.ux, s.uy, q.px, q.py);
        s.x = fx; s.y = fy; s.ang = fang; s.n = fn; s.gx = fgx; s.gy = fgy;
    } else {
        dist = step_fast<METHOD, ISO, LAZY>(a.f, a.f_dev, a.in, s, h, gamma, g2m1, q.hist, i, METHOD == 7 && i < 3, inner,
                                            mask, wbuf);
        if (TRAJ || (EXTRA & X_STATS)) moments_fast<ISO>(s, g2m1, q.px, q.py); // else only the last row's, after the loop
    }
    q.dist_sim += dist;
    q.dist_real += h;
    // RT:871-875
    const double nray = __dmul_rn(s.coef, s.n); // used twice: no contraction into the sum below (bit-equal output modes)
    q.T = q.T + dist * (q.nray_prev + nray) / 2.0;
    q.nray_prev = nray;
    q.last_i = i;
    if ((EXTRA & X_STATS) && nonzero_bits(q.px)) {
        if (q.st_cnt == 0) q.st_shift = q.px;
        const double d = q.px - q.st_shift;
        q.st_s1 += d;
        q.st_s2 = fma(d, d, q.st_s2);
        ++q.st_cnt;
    }
    if ((EXTRA & X_SNELL) && !TRAJ) {
        if (i == q.cap_a) { q.capv[0] = s.x; q.capv[1] = s.y; }
        if (i == q.cap_b) { q.capv[2] = s.x; q.capv[3] = s.y; }
        if (a.ckpt && !a.cap_rows && (i & ((1 << a.ckpt_shift) - 1)) == 0 && (i >> a.ckpt_shift) <= a.ckpt_rows) {
            double *p = a.ckpt + ((size_t)((i >> a.ckpt_shift) - 1) * RTB_CKPT_FIELDS) * a.R + r;
            p[0] = s.x; p[a.R] = s.y; p[2 * a.R] = s.ang; p[3 * a.R] = s.ux; p[4 * a.R] = s.uy; p[5 * a.R] = s.n;
            p[6 * a.R] = s.gx; p[7 * a.R] = s.gy; p[8 * a.R] = s.coef; p[9 * a.R] = s.inv_n; p[10 * a.R] = s.hn;
#pragma unroll
            for (int k = 0; k < 3; ++k) { p[(11 + 2 * k) * a.R] = q.hist[k][0]; p[(12 + 2 * k) * a.R] = q.hist[k][1]; }
        }
    }
    // boundary test RT:878 (strict); op7's two bootstrap steps are taken before the loop that holds the test
    // (RT:838-864), so a ray that leaves the box on step 1 or 2 is still stepped to row 3
    const bool outside = !inner && !(METHOD == 7 && i < 3) &&
                         (s.x > a.limx_s || s.x < a.limx_i || s.y > a.limy_s || s.y < a.limy_i);
    if ((EXTRA & X_EXIT) && outside) { // fraction of the last segment inside the box (first limit crossed)
        double t = 1.0;
        if (s.x > a.limx_s) t = fmin(t, (a.limx_s - q.pvx) / (s.x - q.pvx));
        if (s.x < a.limx_i) t = fmin(t, (a.limx_i - q.pvx) / (s.x - q.pvx));
        if (s.y > a.limy_s) t = fmin(t, (a.limy_s - q.pvy) / (s.y - q.pvy));
        if (s.y < a.limy_i) t = fmin(t, (a.limy_i - q.pvy) / (s.y - q.pvy));
        q.ex_t = fmax(t, 0.0);
    }
    return !outside && i + 1 < q.M;
}

// Global index of local ray r in the fused gather arrays (contiguous or block-cyclic shards).
__device__ __forceinline__ int64_t gather_index(const TraceArgs &a, int64_t r)
{
    if (a.gather_bshift < 0) return a.gather_off + r;
    const int b = a.gather_bshift;
    return ((((r >> b) * a.gather_nranks + a.gather_rank) << b) | (r & (((int64_t)1 << b) - 1)));
}

// per-ray results RT:888-890 (+ the optional outputs of include/rtb.h)
template <bool ISO, bool TRAJ, bool EXACT, int EXTRA, bool LAZY>
__device__ __forceinline__ void ray_finish(const TraceArgs &a, int64_t r, double g2m1, RayRegs &q)
{
    RayState &s = q.s;
    if (LAZY && s.ang != s.ang) s.ang = atan2(s.uy, s.ux); // theta of the last row (NaN = not evaluated yet)
    if (!EXACT && !TRAJ && !(EXTRA & X_STATS) && q.last_i > 0) moments_fast<ISO>(s, g2m1, q.px, q.py);
    if (a.d_ray) {
        a.d_ray[r] = q.dist_real;
        a.d_ray[a.R + r] = q.dist_sim;
        a.d_ray[2 * a.R + r] = (double)q.last_i;
        if (a.gather_world > 0) { // peer stores: 24 B per ray per GPU, hidden behind the stepping
            const int64_t gi = gather_index(a, r);
            if (gi < a.gather_R) {
                for (int pe = 0; pe < a.gather_world; ++pe) {
                    double *g = a.gather[pe] + gi;
                    g[0] = q.dist_real;
                    g[a.gather_R] = q.dist_sim;
                    g[2 * a.gather_R] = (double)q.last_i;
                }
            }
        }
    }
    if (a.last_row) {
        a.last_row[r] = s.x; a.last_row[a.R + r] = s.y;
        a.last_row[2 * a.R + r] = q.px; a.last_row[3 * a.R + r] = q.py;
        a.last_row[4 * a.R + r] = q.T; a.last_row[5 * a.R + r] = s.ang;
    }
    if ((EXTRA & X_STATS) && a.px_stats) {
        double mean = 0.0, sd = 0.0;
        if (q.st_cnt > 0) {
            const double m1 = q.st_s1 / q.st_cnt;
            mean = q.st_shift + m1;
            const double var = q.st_s2 / q.st_cnt - m1 * m1;
            sd = var > 0.0 ? sqrt(var) : 0.0;
        }
        a.px_stats[r] = (double)q.st_cnt;
        a.px_stats[a.R + r] = mean;
        a.px_stats[2 * a.R + r] = sd;
    }
    if ((EXTRA & X_EXIT) && a.exit_out) { // a ray that never left the box reports its last position (ex_t == 1)
        a.exit_out[r] = q.pvx + q.ex_t * (s.x - q.pvx);
        a.exit_out[a.R + r] = q.pvy + q.ex_t * (s.y - q.pvy);
        a.exit_out[2 * a.R + r] = q.pvT + q.ex_t * (q.T - q.pvT);
    }
    if ((EXTRA & X_SNELL) && !TRAJ && a.cap_out) {
        a.cap_out[r] = q.capv[0]; a.cap_out[a.R + r] = q.capv[1];
        a.cap_out[2 * a.R + r] = q.capv[2]; a.cap_out[3 * a.R + r] = q.capv[3];
    }
    // non-finite guard: the reference steps a NaN ray on to max_size (its comparisons are all false, RT:878) and so
    // does this kernel; the caller learns about it from rtb_stats.nonfinite_rays
    if (a.nonfinite) {
        const unsigned ex = 0x7ff00000u;
        const bool bad = ((unsigned)__double2hiint(s.x) & ex) == ex || ((unsigned)__double2hiint(s.y) & ex) == ex ||
                         ((unsigned)__double2hiint(q.T) & ex) == ex;
        if (bad) atomicAdd(a.nonfinite, 1ull);
    }
}

// EXACT = faithful restatement (reference evaluation order, libdevice sincos, IEEE div/sqrt, FITPACK
// recurrence): the parity anchor.  !EXACT = default B200 form (step_fast.cuh).
// TMODE: 0 = no trajectory, 1 = trajectory by per-lane stores, 2 = trajectory staged in shared memory + TMA stores.
template <int METHOD, bool ISO, int TMODE, bool EXACT, int EXTRA>
__global__ void __launch_bounds__(RTB_BLOCK, EXACT ? (4 * 128) / RTB_BLOCK : TMODE ? RTB_TRAJ_MIN_BLOCKS : min_blocks_of(METHOD))
trace_kernel(const __grid_constant__ TraceArgs a)
{
    constexpr bool TRAJ = TMODE != 0;
    constexpr int UPD = updater_of(METHOD);
    constexpr bool LAZY = !EXACT && !TRAJ && (UPD == 1 || UPD == 3); // theta evaluated once at the end
    // Small fans are spread one ray per 2^lane_shift lanes (up to one ray per warp): the step loop of a lone ray is a
    // dependent chain, and 32 rays that sit in 32 different cells make every table load of their warp 32 cache
    // lines wide; spread out, each ray gets a scheduler slot of its own and the fan covers all SMs.
    const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t r = gt >> a.lane_shift;
    const bool valid = r < a.R && (gt & ((1 << a.lane_shift) - 1)) == 0;
    const int64_t rr = r < a.R ? r : a.R - 1; // idle and padded lanes shadow a ray, never step or store
    const double gamma = a.gamma;
    const double g2m1 = gamma * gamma - 1.0;

    RayRegs q;
    ray_init<METHOD, ISO, TRAJ, EXACT, EXTRA>(a, rr, gamma, g2m1, q);

    // Trajectory write-back (traj_stage.cuh): row 0 now, then one row per kept step; rows after a ray's last step
    // stay 0 (np.zeros RT:802) -- written as zeros unless the caller cleared the array (RTB_F_NO_TAIL_ZERO).
    extern __shared__ __align__(128) unsigned char dyn_smem[];
    TrajWriter<TMODE> tw;
    if (TRAJ) {
        tw.init(a.s_ray, &a.tmaps, a.R, r, valid, dyn_smem);
        tw.row(true, q.s.x, q.s.y, q.px, q.py, q.T, q.s.ang);
    }
    int next_keep = a.out_stride; // next step index whose row is kept in s_ray

    bool alive = valid && q.i_start < q.M;
    if (TMODE == 2 && !alive) tw.ended(); // a ray without steps, or a lane past the end of the fan
    int i = q.i_start;
    // The loop condition must be warp-uniform (lanes may have different M with per-ray max_sizes):
    // every lane iterates until no lane of the warp is alive; `alive` drops at i+1 >= M.
    __shared__ double2 stage[EXACT ? 1 : RTB_BLOCK / 32][RTB_CELL_REC]; // one cell record per warp (field.cuh)
    double2 *const wbuf = (EXACT || !a.stage) ? nullptr : stage[threadIdx.x >> 5];
    for (;; ++i) {
        const unsigned mask = __ballot_sync(0xffffffffu, alive);
        if (!mask) break;
        const bool stepped = alive;
        if (alive) {
            alive = ray_step<METHOD, ISO, TRAJ, EXACT, EXTRA, LAZY>(a, r, gamma, g2m1, q, i, mask, wbuf);
            if (TMODE == 1 && i == next_keep) { // per-lane stores need no convergence: each lane writes its own row
                tw.row_live(q.s.x, q.s.y, q.px, q.py, q.T, q.s.ang);
                next_keep += a.out_stride;
            }
        } else if (TMODE == 1) {
            if (i == next_keep) {
                tw.row(a.tail_zero != 0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0);
                next_keep += a.out_stride;
            }
        }
        if (TMODE == 2) {
            if (i == next_keep) { // warp-uniform: the whole warp stages its row (lanes whose ray has ended store nothing)
                tw.row(stepped, q.s.x, q.s.y, q.px, q.py, q.T, q.s.ang);
                next_keep += a.out_stride;
            }
            if (stepped && !alive) tw.ended(); // rows this ray will not reach read as zeros (np.zeros RT:802)
        }
    }
    if (TRAJ) { // rows after the last step of every ray of this warp
        const int n_tail = next_keep < a.max_size ? (a.max_size - 1 - next_keep) / a.out_stride + 1 : 0;
        tw.zero_tail(n_tail, a.tail_zero != 0);
        tw.finish();
    }
    if (valid) ray_finish<ISO, TRAJ, EXACT, EXTRA, LAZY>(a, r, g2m1, q);
    // ray-step count = sum of d_ray[2]; one atomic per warp
    if (a.total_steps) {
        unsigned v = valid ? (unsigned)q.last_i : 0u;
        v = __reduce_add_sync(0xffffffffu, v);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(a.total_steps, (unsigned long long)v);
    }
}

// Ray queue + warp compaction (FINAL-type outputs, default arithmetic).  A persistent grid of 148 x blocks-per-SM
// blocks; every warp claims rays from `queue_head`.  Lanes whose ray has ended wait at most RTB_QUEUE_WAIT iterations
// for the rest of the warp: neighbouring rays of a sorted fan end within a step or two of each other, and refilling
// the whole warp at once hands it 32 consecutive rays again, which keeps them in one field cell (the warp-staged
// record path of field.cuh); in an unsorted fan the wait costs < 1 % of a ray's steps and every refill is a
// compaction -- dead lanes take the next rays of the queue, live lanes keep their state in registers (a refill is
// an initialisation, not a shuffle).  Same per-ray instructions as trace_kernel, hence the same bits.
template <int METHOD, bool ISO, int EXTRA>
__global__ void __launch_bounds__(RTB_BLOCK, RTB_QUEUE_MIN_BLOCKS) trace_queue_kernel(const __grid_constant__ TraceArgs a)
{
    constexpr int UPD = updater_of(METHOD);
    constexpr bool LAZY = (UPD == 1 || UPD == 3);
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const double gamma = a.gamma;
    const double g2m1 = gamma * gamma - 1.0;
    __shared__ double2 stage[RTB_BLOCK / 32][RTB_CELL_REC];
    double2 *const wbuf = a.stage ? stage[threadIdx.x >> 5] : nullptr;

    RayRegs q;
    int64_t r = -1;            // ray held by this lane (-1: none)
    int i = 0;                 // its next step index
    bool alive = false;
    bool drained = false;      // the queue is empty (warp-uniform)
    unsigned long long steps_done = 0;
    for (;;) {
        // compaction: finished rays leave their lanes, dead lanes take the next rays of the queue
        unsigned mask = __ballot_sync(FULL, alive);
        if (!alive && r >= 0) {
            ray_finish<ISO, false, false, EXTRA, LAZY>(a, r, g2m1, q);
            steps_done += (unsigned)q.last_i;
            r = -1;
        }
        if (!drained) {
            const unsigned dead = ~mask;
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(a.queue_head, (unsigned long long)__popc(dead));
            base = __shfl_sync(FULL, base, 0);
            drained = base >= (unsigned long long)a.R;
            if (!drained && !alive) {
                const int64_t nr = (int64_t)base + __popc(dead & ((1u << lane) - 1u));
                if (nr < a.R) {
                    r = nr;
                    ray_init<METHOD, ISO, false, false, EXTRA>(a, r, gamma, g2m1, q);
                    i = q.i_start;
                    alive = i < q.M; // a ray without steps (max_size 1) is finished at the next compaction
                }
            }
            mask = __ballot_sync(FULL, alive);
        }
        if (mask == 0u) {
            if (drained && !__any_sync(FULL, r >= 0)) break;
            continue;
        }
        // the step loop, same shape as trace_kernel's; left when the warp has been running with dead lanes for
        // RTB_QUEUE_WAIT iterations (or has none alive)
        int wait = 0;
        for (;;) {
            mask = __ballot_sync(FULL, alive);
            if (mask != FULL && (mask == 0u || ++wait > RTB_QUEUE_WAIT)) break;
            if (alive) {
                alive = ray_step<METHOD, ISO, false, false, EXTRA, LAZY>(a, r, gamma, g2m1, q, i, mask, wbuf);
                ++i;
            }
        }
    }
    if (a.total_steps) { // one atomic per warp
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) steps_done += __shfl_xor_sync(FULL, steps_done, o);
        if (lane == 0 && steps_done) atomicAdd(a.total_steps, steps_done);
    }
}

// Host-side launcher, one per METHOD (instantiated in trace_inst.cu).
struct LaunchCfg {
    bool iso, traj, exact, queue;
    int block;        // threads per block (32 for spread small fans)
    int queue_blocks; // queue kernel: persistent blocks (SMs x blocks per SM)
};
typedef cudaError_t (*launch_fn)(const TraceArgs &, const LaunchCfg &, cudaStream_t);

inline int extra_mask(const TraceArgs &a, bool traj)
{
    int e = 0;
    if (a.px_stats) e |= X_STATS;
    if (a.exit_out) e |= X_EXIT;
    if (!traj && (a.cap_rows || a.ckpt)) e |= X_SNELL;
    if (e != 0 && e != X_STATS && e != X_EXIT && e != X_SNELL) e = X_ALL; // unusual combinations share one flavour
    if (traj && e != 0 && e != X_STATS) e = X_ALL;
    return e;
}

template <int METHOD>
cudaError_t launch_trace(const TraceArgs &a, const LaunchCfg &c, cudaStream_t st)
{
    const unsigned grid = (unsigned)(((a.R << a.lane_shift) + c.block - 1) / c.block);
    const int e = extra_mask(a, c.traj);
    const size_t dyn = (c.traj && a.traj_bulk && !c.exact) ? traj_stage_bytes(c.block) : 0;
#define RTB_LX(I, T, E, X) trace_kernel<METHOD, I, T, E, X><<<grid, c.block, dyn, st>>>(a)
#define RTB_LT(I, E) /* trajectory flavours: staged + TMA stores (default kernels) or per-lane stores */ \
    do {                                                                      \
        if (a.traj_bulk && !E) {                                              \
            if (e == 0) RTB_LX(I, 2, false, 0);                               \
            else if (e == X_STATS) RTB_LX(I, 2, false, X_STATS);              \
            else RTB_LX(I, 2, false, X_ALL);                                  \
        } else {                                                              \
            if (e == 0) RTB_LX(I, 1, E, 0);                                   \
            else if (e == X_STATS) RTB_LX(I, 1, E, X_STATS);                  \
            else RTB_LX(I, 1, E, X_ALL);                                      \
        }                                                                     \
    } while (0)
#define RTB_LF(I, E) /* FINAL flavours */                                     \
    do {                                                                      \
        if (e == 0) RTB_LX(I, 0, E, 0);                                       \
        else if (e == X_STATS) RTB_LX(I, 0, E, X_STATS);                      \
        else if (e == X_EXIT) RTB_LX(I, 0, E, X_EXIT);                        \
        else if (e == X_SNELL) RTB_LX(I, 0, E, X_SNELL);                      \
        else RTB_LX(I, 0, E, X_ALL);                                          \
    } while (0)
#define RTB_LQ(I, X) trace_queue_kernel<METHOD, I, X><<<c.queue_blocks, RTB_BLOCK, 0, st>>>(a)
#define RTB_Q(I)                                                              \
    do {                                                                      \
        if (e == 0) RTB_LQ(I, 0);                                             \
        else if (e == X_STATS) RTB_LQ(I, X_STATS);                            \
        else if (e == X_EXIT) RTB_LQ(I, X_EXIT);                              \
        else if (e == X_SNELL) RTB_LQ(I, X_SNELL);                            \
        else RTB_LQ(I, X_ALL);                                                \
    } while (0)
    if (c.queue && !c.traj && !c.exact && a.lane_shift == 0) {
        if (c.iso) RTB_Q(true); else RTB_Q(false);
    } else if (c.iso) {
        if (c.traj) { if (c.exact) RTB_LT(true, true); else RTB_LT(true, false); }
        else        { if (c.exact) RTB_LF(true, true); else RTB_LF(true, false); }
    } else {
        if (c.traj) { if (c.exact) RTB_LT(false, true); else RTB_LT(false, false); }
        else        { if (c.exact) RTB_LF(false, true); else RTB_LF(false, false); }
    }
#undef RTB_Q
#undef RTB_LQ
#undef RTB_LF
#undef RTB_LT
#undef RTB_LX
    return cudaGetLastError();
}

} // namespace rtb
