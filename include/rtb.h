/*
 * rtb.h -- C ABI of librtb.so, the B200 (sm_100a) ray-stepping engine that replaces the
 * hot path of neyuru/RayTracing's RT_bench.py.
 *
 * The reference has no FFI layer: its boundary for this path is the Python call
 *     trazar(selected_func, z, grd, show, step, divisor, user_choice)
 *         -> (s_ray, d_ray, compute_times, errors)                  RT_bench.py:766-948
 *     search_delta(option, z, grd, step, divisor, user_choice)      RT_bench.py:950-958
 *     n_gradient(vector, grd, z) -> (n, [dn/dx, dn/dy])             RT_bench.py:141-156
 * plus the module globals those functions read (gamma RT:725, N RT:797, GOLD_* RT:64-66).
 * Each entry point below states which of these it replaces.  The Python shim that keeps
 * the reference signatures lives in raytracing_b200/engine.py and binds this header with
 * ctypes (raytracing_b200/_cabi.py); INTEGRATION.md shows the stub a maintainer of the
 * reference would add.
 *
 * Conventions: every function returns 0 on success and a negative rtb_status otherwise
 * (never exits the process, cf. sys.exit() RT:1406); rtb_last_error() gives the text.
 * One context per GPU, one CUDA stream per context; calls are synchronous on return.
 * Thread-safe across contexts, not within one.  The caller owns every buffer.
 * All arrays are float64 / int64 in C order.  There is no CPU fallback: without a CUDA
 * device rtb_create fails with RTB_ERR_CUDA.
 */
#ifndef RTB_H
#define RTB_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTB_ABI_VERSION 2

typedef struct rtb_ctx rtb_ctx;

enum rtb_status {
    RTB_OK = 0,
    RTB_ERR_ARG = -1,     /* bad argument (method out of range, NULL pointer, ...)   */
    RTB_ERR_CUDA = -2,    /* CUDA runtime error, text in rtb_last_error              */
    RTB_ERR_NOFIELD = -3, /* rtb_trace before rtb_set_field                           */
    RTB_ERR_NOMEM = -4
};

/*
 * A scipy RectBivariateSpline as FITPACK stores it (RT_bench.py:455-457):
 * `tu` are the knots of the FIRST spline axis -- physical y, because the reference builds
 * RectBivariateSpline(y, x, Z) -- `tv` the knots of the second axis (physical x), `c` the
 * (nu-ku-1)*(nv-kv-1) coefficients, first-axis major.  Host pointers; copied.
 */
typedef struct {
    const double *tu;
    const double *tv;
    const double *c;
    int32_t nu, nv, ku, kv;
} rtb_spline;

/* Where caller buffers live. */
#define RTB_PTR_HOST 0   /* host memory: the library stages H2D / D2H inside the call */
#define RTB_PTR_DEVICE 1 /* device memory on the context's GPU (e.g. torch data_ptr()) */

/* rtb_params.flags */
#define RTB_F_INTERFACE 1u   /* op_interface: compute the Snell error metric (RT:896-919) into `errors` */
#define RTB_F_EXACT_BASIS 2u /* faithful kernels: FITPACK's Cox-de Boor recurrence with the real knots, libm-grade
                                sin/cos/atan2, IEEE division and sqrt, the reference's evaluation order.
                                Default: per-cell expanded polynomials on the uniform part of the grid and the
                                re-associated step arithmetic of step_fast.cuh; results differ by < 1e-12 */
#define RTB_F_NO_TAIL_ZERO 4u /* s_ray was zero-filled by the caller: do not write the rows after a ray's
                                 last step (np.zeros, RT:802) */

#define RTB_F_STATIC_GRID 8u  /* one thread per ray for the whole launch (the only mapping for trajectories) */
#define RTB_F_RAY_QUEUE 16u   /* FINAL-type outputs of fans >= 2^16 rays: persistent warps pull rays from a queue and
                                 compact when lanes finish (warp ballot + refill of the dead lanes).  Opt-in: meant for
                                 fans whose neighbouring rays differ widely in length, but on every fan measured on B200
                                 the static grid is as fast or faster; results are bit-identical either way */
#define RTB_F_SORT_FAN 32u    /* rtb_trace, FINAL-type outputs (s_ray NULL): the library orders the rays by (x0, y0, theta0)
                                 before the launch and hands every result back at the caller's ray index -- the remedy for
                                 unsorted fans (neighbouring rays share a warp: shuffled 2^20-ray fan 73.5 -> 15.9 ms).  The
                                 reference has no counterpart (it traces ray by ray, RT:807); results are bit-identical */

typedef struct {
    int32_t method;     /* 1..11 = op1..op11 (RT:469-764); the shim maps function objects to ids  */
    uint32_t flags;     /* RTB_F_*                                                                  */
    int32_t ptr_kind;   /* RTB_PTR_HOST / RTB_PTR_DEVICE for ALL array arguments of the call       */
    int32_t out_stride; /* rows kept in s_ray: row i is stored at i/out_stride when i%out_stride==0;
                           1 = the reference's full (max_size,6,R) array; ignored if s_ray==NULL  */
    int64_t max_size;   /* rows of s_ray: N*divisor (fisheye) or ceil(s/step)+1 (RT:796-799)        */
    double gamma;       /* anisotropy parameter, module global `gamma` (RT:725,758)                 */
    double step;        /* DELTA_S                                                                  */
    double limx_i, limx_s, limy_i, limy_s; /* box of constants() (RT:263..), strict tests RT:878    */
    /* Optional per-ray step and max_size (length R; NULL = uniform `step`/`max_size`).  This is how
       the DELTA_S finder batches candidates x rays into one launch (RT:1303-1318).  When given,
       `max_size` must be the maximum of max_sizes[] and s_ray must be NULL.                       */
    const double *steps;
    const int64_t *max_sizes;
} rtb_params;

typedef struct {
    double kernel_ms;   /* CUDA-event time of the integration kernel(s) on the context's stream  */
    double total_ms;    /* CUDA-event time of the whole call incl. staging copies                 */
    int64_t ray_steps;  /* sum over rays of d_ray[2] = executions of the loop body RT:867-879      */
    int64_t launches;   /* kernels launched by the call                                            */
    int64_t h2d_bytes, d2h_bytes; /* staged bytes (RTB_PTR_HOST)                                   */
    int64_t nonfinite_rays; /* rays whose last position or travel time is NaN/Inf.  Like the reference (whose box
                               tests are all false for NaN, RT:878) such a ray is stepped on to max_size; this is
                               the guard flag that tells the caller it happened                               */
} rtb_stats;

/* Create / destroy the per-GPU context (replaces the reference's process-global state). */
int rtb_create(int device, rtb_ctx **out);
void rtb_destroy(rtb_ctx *ctx);
/* Launch on an existing stream (e.g. torch.cuda.current_stream().cuda_stream); NULL = own stream.
   The legacy default stream must be named explicitly as cudaStreamLegacy ((void*)0x1). */
int rtb_set_stream(rtb_ctx *ctx, void *cuda_stream);
const char *rtb_last_error(rtb_ctx *ctx);
int rtb_abi_version(void);
/* Hex digest of the CUDA sources the library was compiled from (written by csrc/Makefile): lets a caller check that a
   prebuilt librtb.so matches the source tree it sits in. */
const char *rtb_build_id(void);
/* The same for the sources of the trace kernels alone: identifies the kernels an ncu capture (profiles/ncu_latest.json)
   was taken of, whatever changed on the host side of the library. */
const char *rtb_kernel_id(void);
/* Tuning / test switches of a context.  RTB_OPT_QUEUE: 0 = per call (RTB_F_RAY_QUEUE / RTB_F_STATIC_GRID; default static),
   1 = always where possible, 2 = never (static grid).  RTB_OPT_TRAJ_STORE: 0 = per-lane st.global.cs (default: faster
   on B200), 1 = trajectory rows staged in shared memory, 2-D TMA stores (needs R even). */
#define RTB_OPT_QUEUE 1
#define RTB_OPT_TRAJ_STORE 2
#define RTB_OPT_SHOOT_GROUPS 3 /* rtb_shoot_refine: groups of pairs iterating on their own streams, 1..16; 0 = auto */
#define RTB_OPT_SHOOT_SPECULATE 4 /* rtb_shoot_refine: two golden passes per launch (both outcomes of a pass are traced
                                     before it is decided): 0 / 1 = off (default: it is slower on every problem measured), 2 = on; same results */
int rtb_set_option(rtb_ctx *ctx, int32_t option, int32_t value);

/*
 * Upload the field: z = bilinear spline of n, gu = grd[0] (d n/d y), gv = grd[1] (d n/d x),
 * the objects interpolacion() returns (RT:455-458).  Replaces passing `z, grd` to trazar.
 */
int rtb_set_field(rtb_ctx *ctx, const rtb_spline *z, const rtb_spline *gu, const rtb_spline *gv);

/*
 * interpolacion (RT:435-464) on the device: from the samples Z (my, mx) of the index field on the grid y[my] x x[mx]
 * (genZ RT:412-433; Z[i][j] = f(x[j], y[i])) to the same resident field rtb_set_field produces, without scipy:
 *   GradX, GradY = np.gradient(Z, delta, edge_order=2)                      (RT:449; bit-exact)
 *   z    = RectBivariateSpline(y, x, Z, kx=1, ky=1)                         (RT:455; coefficients = samples)
 *   grd  = RectBivariateSpline(y, x, GradX | GradY, kx=3, ky=3)             (RT:456-457; FITPACK s=0 knots,
 *          coefficients from the same collocation system by banded LU: equal to scipy's to a few ulp of max|Grad|)
 * `delta` is what the reference feeds np.gradient -- the global DELTA, not the true spacing of linspace (RT:449).
 * y, x, Z are host or device pointers per ptr_kind.  stats: kernel_ms = gradient + substitution sweeps,
 * total_ms = whole call including the cell tables.
 */
int rtb_fit_field(rtb_ctx *ctx, int32_t ptr_kind, const double *y, int64_t my, const double *x, int64_t mx,
                  const double *Z, double delta, rtb_stats *stats);

/* np.gradient(Z, delta, edge_order=2) alone (RT:449): g0 = d/d(axis 0), g1 = d/d(axis 1), both (my, mx). */
int rtb_gradient(rtb_ctx *ctx, int32_t ptr_kind, const double *Z, int64_t my, int64_t mx, double delta,
                 double *g0, double *g1);

/*
 * Read the resident field back in scipy's tck form (host pointers): which = 0 (z), 1 (grd[0]), 2 (grd[1]).
 * Any of nu, nv, tu, tv, c may be NULL; call once with the arrays NULL to learn nu, nv
 * (c has (nu-k-1)*(nv-k-1) entries, k = 1 for z and 3 for grd).
 */
int rtb_get_field(rtb_ctx *ctx, int32_t which, int32_t *nu, int32_t *nv, double *tu, double *tv, double *c);

/*
 * n_gradient (RT:141-156) on `n` points: out_n[i], out_gx[i] = grd[1](y,x), out_gy[i] = grd[0](y,x).
 */
int rtb_field_eval(rtb_ctx *ctx, int32_t ptr_kind, uint32_t flags, const double *x, const double *y,
                   int64_t n, double *out_n, double *out_gx, double *out_gy);

/*
 * trazar (RT:766-948) for an explicit fan of R rays: ray k starts at (x0[k], y0[k]) with group
 * angle theta0[k] (the reference takes these from constants(): theta_v, pos_x, y=-2, RT:809-813).
 *   s_ray   (ceil(max_size/out_stride), 6, R) or NULL : x, y, p_x, p_y, T, theta per kept row (RT:802,871-875)
 *   d_ray   (3, R)                                    : sum h, sum dist, last i (RT:888-890)
 *   errors  (R) or NULL                               : Snell error in degrees (RTB_F_INTERFACE), else 0
 *   last_row(6, R) or NULL                            : the last row written for each ray (row d_ray[2])
 *   px_stats(3, R) or NULL                            : over rows 0..d_ray[2] with p_x != 0 (RT:1358):
 *                                                       count, mean(p_x), population std(p_x)
 * compute_times (RT:800,882) is not produced per ray -- per-ray wall time has no meaning on a GPU;
 * the shim fills it with kernel_ms/R.
 */
int rtb_trace(rtb_ctx *ctx, const rtb_params *p, const double *theta0, const double *x0,
              const double *y0, int64_t R, double *s_ray, double *d_ray, double *errors,
              double *last_row, double *px_stats, rtb_stats *stats);

/*
 * The DELTA_S sweep (RT:1303-1318: executor.map(search_delta, ...) over the candidate steps) as ONE batched trace of
 * C candidates x R rays with the per-candidate reductions done on the device, so a C caller gets what the selection
 * rules (find_index RT:1323-1373) consume without copying per-ray data:
 *   RTB_SWEEP_SNELL   (interface): m[c][k] = Snell error of ray k (RT:896-919);   mean = sum/count, max (RT:954)
 *   RTB_SWEEP_CLOSURE (fisheye)  : m[c]    = 100 |(1,0) - rays[-1, 0:2, 0]| / 2pi of the fan's FIRST ray (RT:956;
 *                                  rays[-1] is row max_size-1, zero if the ray ended earlier); count 1
 *   RTB_SWEEP_PX_CV   (vert/aniso): m[c][k] = 100 std/mean of the non-zero p_x entries of ray k (RT:1358-1359) over
 *                                  rays skip_lo <= k < R - skip_hi (the reference leaves out its first and last ray)
 * theta0/x0/y0 (R) is the fan every candidate traces; steps/max_sizes (C) the candidates.  Outputs (C each; host or
 * device per p->ptr_kind): sum[c], max[c], count[c].  The reductions use a fixed tree (bitwise reproducible); a fan
 * sharded over several GPUs all-reduces sum/count (SUM) and max (MAX).  p->step, p->max_size, p->steps, p->max_sizes
 * are ignored; p->method/flags/gamma/box as in rtb_trace.
 */
#define RTB_SWEEP_SNELL 1
#define RTB_SWEEP_CLOSURE 2
#define RTB_SWEEP_PX_CV 3
typedef struct {
    const double *steps;      /* (C) DELTA_S candidates                        */
    const int64_t *max_sizes; /* (C) rows per candidate (RT:796-799)            */
    int64_t C;
    int32_t metric;           /* RTB_SWEEP_*                                    */
    int64_t skip_lo, skip_hi; /* RTB_SWEEP_PX_CV: local rays left out at both ends */
} rtb_sweep_spec;
int rtb_sweep(rtb_ctx *ctx, const rtb_params *p, const double *theta0, const double *x0, const double *y0, int64_t R,
              const rtb_sweep_spec *spec, double *sum, double *max, int64_t *count, rtb_stats *stats);

/*
 * rtb_sweep for callers that cannot allocate.  No counterpart in the reference: added
 * (code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc).
 *
 * Same kernels, arguments and results as rtb_sweep, but the per-ray scratch (d_ray, last rows / p_x statistics of all
 * C * R rays) lives in one file-scope array in device memory -- RTB_NOALLOC_MAX_RAYS rays x 9 doubles = 9 MiB held by
 * every GPU that loads the library until the process exits -- and the call makes no cudaMalloc / cudaFree or any other
 * allocating call.  Limits, each rejected with RTB_ERR_ARG and a message in rtb_last_error:
 *   - p->ptr_kind must be RTB_PTR_DEVICE (theta0, x0, y0, spec->steps, spec->max_sizes, sum, max, count on the device);
 *   - spec->metric RTB_SWEEP_CLOSURE or RTB_SWEEP_PX_CV (the Snell metric needs checkpoint buffers sized by the call);
 *   - spec->C <= RTB_NOALLOC_MAX_CANDIDATES and spec->C * R <= RTB_NOALLOC_MAX_RAYS.
 * All calls on one GPU share the array: calls that overlap on the GPU (two contexts / streams) race.  The array keeps
 * the previous call's contents; the part a call uses is cleared at its start.
 */
#define RTB_NOALLOC_MAX_RAYS 131072
#define RTB_NOALLOC_MAX_CANDIDATES 256
int rtb_sweep_noalloc(rtb_ctx *ctx, const rtb_params *p, const double *theta0, const double *x0, const double *y0,
                      int64_t R, const rtb_sweep_spec *spec, double *sum, double *max, int64_t *count, rtb_stats *stats);

/*
 * Batched two-point shooting search (BASELINE.json configs[4]).  The reference has NO two-point
 * search -- README.md:6 "shooting method" is the forward fan of trazar, and scipy.optimize.minimize
 * is imported (RT:48) but never called -- so these two calls are defined from reference parts
 * (trazar RT:766-948 for every ray, golden RT:175-199 for the launch-angle refinement) and are
 * parity-unpinned; DESIGN.md section 7 gives the definition, oracle/shoot_oracle.py the CPU restatement.
 *
 * All rays start at (src_x, src_y).  C DELTA_S candidates (steps[c], max_sizes[c]) are searched at
 * once (the sweep of RT:1303-1318 as a batch axis).  p->method/flags/ptr_kind/gamma/box are used as in
 * rtb_trace; p->step, p->max_size, p->steps, p->max_sizes are ignored.
 * "Exit point" of a ray: where its last segment crosses the box (linear interpolation between the
 * last row inside and the first row outside, RT:878), travel time (s_ray column 4) interpolated alike;
 * a ray that never leaves reports its last row.
 */
typedef struct {
    double src_x, src_y;
    const double *steps;       /* (C) DELTA_S candidates                                    */
    const int64_t *max_sizes;  /* (C) rows per candidate, ceil(s/step)+1 (RT:799)           */
    int64_t C;
    const double *target_x;    /* (Tn) target points                                        */
    const double *target_y;
    int64_t Tn;
} rtb_shoot_problem;

/*
 * Coarse phase: integrate the fan angles[0..A) for every candidate (C*A rays, one launch) and, for
 * each (candidate, target), find the first angle whose exit point is closest to the target.
 *   exit_xyT (3, C, A) or NULL : exit x, y and travel time of every coarse ray
 *   miss2    (C, Tn)           : smallest squared miss distance
 *   index    (C, Tn)           : index0 + argmin_j  (index0 = global index of angles[0] when the fan is
 *                                sharded by angle over several GPUs; reduce (miss2, index) with MIN)
 */
int rtb_shoot_coarse(rtb_ctx *ctx, const rtb_params *p, const rtb_shoot_problem *q, const double *angles,
                     int64_t A, int64_t index0, double *exit_xyT, double *miss2, int64_t *index,
                     rtb_stats *stats);

/*
 * Refinement: golden(f, lo, hi) of RT:175-199 for every (candidate, target) pair, with
 * f(theta) = squared miss distance of the ray launched at theta; every iteration re-traces both
 * probes of all pairs in one launch (2*C*Tn rays); pairs whose |c-d| <= GOLD_TOL stop tracing.
 *   lo, hi   (C, Tn) : brackets (e.g. the coarse neighbours angles[j-1], angles[j+1])
 *   theta    (C, Tn) : (a+b)/2 after convergence
 *   result   (5, C, Tn): miss distance, travel time, exit x, exit y, steps taken, of the ray at theta
 *   iterations        : golden iterations executed (max over pairs)
 */
int rtb_shoot_refine(rtb_ctx *ctx, const rtb_params *p, const rtb_shoot_problem *q, const double *lo,
                     const double *hi, double *theta, double *result, int64_t *iterations,
                     rtb_stats *stats);

/*
 * Iso-travel-time ("wavefront") points of a traced fan: the arithmetic of the reference's wavefront plot
 * (graficar RT:987-1007).  For every ray k and time times[j]: PchipInterpolator(t_ray, x)(tau), ...(y), ...(angle)
 * over rows 0..d_ray[2,k] of the full s_ray (out_stride 1), for rays with max(t_ray) >= tau.
 *   out (4, nt, R): x, y, group angle, valid flag (1.0 / 0.0; x = y = angle = 0 where invalid)
 * The cross-ray part of the plot (sorting the points, PCHIP of x over y and its normals, RT:1009-1024) has at most
 * ray_count points per time and stays on the host (raytracing_b200.wavefront).
 */
int rtb_wavefront(rtb_ctx *ctx, int32_t ptr_kind, const double *s_ray, const double *d_ray, int64_t rows,
                  int64_t R, const double *times, int32_t nt, double *out);

/*
 * Multi-GPU result gather fused into the integration kernel.  The reference is one process that owns every ray's
 * d_ray (RT:888-890); with the rays sharded over the GPUs of a node (one process per GPU) every rank wants the whole
 * (3, R_total) array.  Instead of an all-gather after the kernel, each rank's kernel stores a finished ray's column
 * straight into every peer's array over NVLink (24 B per ray per GPU, hidden behind the stepping):
 *   rtb_gather_alloc  allocates this rank's (3, R_total) array and returns its 64-byte CUDA IPC handle;
 *   rtb_gather_open   maps a peer's array from the handle it published (same node);
 *   rtb_set_gather    world (<= 8) base pointers in rank order -- own array included --, the total ray count and
 *                     the offset of this rank's fan; world = 0 switches the feature off;
 *   rtb_gather_close  unmaps (own = 0) or frees (own = 1).
 * While set, every rtb_trace call with DEVICE pointers and a d_ray output also writes the peers (calls with host
 * pointers do not: their results go back to the host).  The arrays are complete once every rank's call has returned;
 * ranks synchronise among themselves (e.g. a 1-element all-reduce).  A consumer that reads its array while peers may
 * already run their NEXT trace must alternate between two sets of arrays (raytracing_b200.dist.FusedGather does).
 */
int rtb_gather_alloc(rtb_ctx *ctx, int64_t bytes, void **ptr, void *handle64);
int rtb_gather_open(rtb_ctx *ctx, const void *handle64, void **ptr);
int rtb_gather_close(rtb_ctx *ctx, void *ptr, int32_t own);
int rtb_set_gather(rtb_ctx *ctx, int32_t world, double *const *bases, int64_t R_total, int64_t offset);
/* The same for block-cyclic shards (raytracing_b200.dist.cyclic_indices): local ray r of rank `rank` is global ray
   ((r / block) * nranks + rank) * block + r % block; block must be a power of two. */
int rtb_set_gather_cyclic(rtb_ctx *ctx, int32_t world, double *const *bases, int64_t R_total, int64_t block,
                          int32_t nranks, int32_t rank);

/*
 * Measured FP64 pipe peak (dependent-free DFMA streams on every SM), the roofline denominator
 * of SURVEY.md 8(d); returns FLOP/s (2 per DFMA).
 */
int rtb_fp64_peak(rtb_ctx *ctx, double *flops_per_s);

#ifdef __cplusplus
}
#endif
#endif /* RTB_H */
