// ctx.cuh -- the per-GPU context behind the C ABI (include/rtb.h), shared by rtb.cu and shoot.cu:
// stream, field tables, staging-buffer pool, error text, per-method kernel launchers.
#pragma once
#include "../../include/rtb.h"
#include "trace_kernel.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>

namespace rtb {
#define RTB_DECL(K) cudaError_t launch_trace_m##K(const TraceArgs &, const LaunchCfg &, cudaStream_t);
RTB_DECL(1) RTB_DECL(2) RTB_DECL(3) RTB_DECL(4) RTB_DECL(5) RTB_DECL(6)
RTB_DECL(7) RTB_DECL(8) RTB_DECL(9) RTB_DECL(10) RTB_DECL(11)
#undef RTB_DECL
static const launch_fn kLaunch[12] = {nullptr, launch_trace_m1, launch_trace_m2, launch_trace_m3,
                                      launch_trace_m4, launch_trace_m5, launch_trace_m6, launch_trace_m7,
                                      launch_trace_m8, launch_trace_m9, launch_trace_m10, launch_trace_m11};
} // namespace rtb

using namespace rtb;

#define RTB_COUNTER_SLOTS 256

struct rtb_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    bool has_field = false;
    DevField f;
    DevField *f_dev = nullptr;
    std::vector<void *> field_allocs;
    unsigned long long *d_total = nullptr; // [0] ray-steps of the call, [1] rays that ended non-finite, [2..] queue heads
    int block = RTB_BLOCK;
    int sm_count = 148;
    int queue_slot = 2; // d_total slot of the last ray-queue launch (round robin over RTB_COUNTER_SLOTS - 2)
    int queue_mode = 0; // rtb_set_option(RTB_OPT_QUEUE): 0 = auto, 1 = always the ray-queue kernel, 2 = never
    int traj_mode = 0;  // rtb_set_option(RTB_OPT_TRAJ_STORE): 0 = per-lane streaming stores, 1 = rows staged in shared memory + TMA stores
    // host-buffer calls on large fans are cut into chunks whose H2D copy, kernel and D2H copy overlap (rtb_trace)
    cudaStream_t pipe[3] = {nullptr, nullptr, nullptr};
    cudaEvent_t pipe_ev[4] = {nullptr, nullptr, nullptr, nullptr}; // fork + one join per pipe stream
    // the shooting refinement iterates groups of pairs on their own streams (shoot.cu)
    static constexpr int kShootStreams = 15;
    cudaStream_t shoot_st[kShootStreams] = {};
    cudaEvent_t shoot_ev[kShootStreams + 1] = {};
    int shoot_groups = 0; // rtb_set_option(RTB_OPT_SHOOT_GROUPS): 0 = auto
    int shoot_spec = 0;   // rtb_set_option(RTB_OPT_SHOOT_SPECULATE): 0 = auto, 1 = off, 2 = on
    // pinned bounce buffers of the large pageable device-to-host copies (Staging::copy_back)
    static constexpr int kBounce = 3;
    static constexpr size_t kBounceBytes = (size_t)64 << 20;
    void *bounce[kBounce] = {};
    cudaEvent_t bounce_ev[kBounce] = {};
    // fused result gather (rtb_set_gather): peer base pointers, total rays, this rank's ray offset
    double *gather[RTB_MAX_PEERS] = {};
    int gather_world = 0;
    int gather_bshift = -1, gather_nranks = 1, gather_rank = 0;
    int64_t gather_R = 0, gather_off = 0;
    std::string err;
    // grow-only pool of staging buffers (RTB_PTR_HOST calls): cudaMalloc/cudaFree per call cost ms
    struct PoolBuf { void *p; size_t bytes; bool used; };
    std::vector<PoolBuf> pool;
};

static inline int fail(rtb_ctx *c, int code, const char *what, cudaError_t e = cudaSuccess)
{
    if (c) {
        c->err = what;
        if (e != cudaSuccess) { c->err += ": "; c->err += cudaGetErrorString(e); }
    }
    return code;
}
#define CK(call)                                                              \
    do {                                                                      \
        cudaError_t e_ = (call);                                              \
        if (e_ != cudaSuccess) return fail(ctx, RTB_ERR_CUDA, #call, e_);     \
    } while (0)

// Lanes per ray for a launch of R rays: 2^shift, the largest power of two (<= 32) that still fits the whole fan
// into one wave of 16 warps per SM (148 SMs); 0 from ~76 k rays on.  Spread launches use 32-thread blocks.
static inline int pick_lane_shift(int64_t R)
{
    int shift = 0;
    while (shift < 5 && (R << (shift + 1)) <= (int64_t)148 * 16 * 32) ++shift;
    return shift;
}

struct Staging;
// Pieces of rtb_trace (rtb.cu) shared with shoot.cu and sweep.cu: common TraceArgs, one kernel launch (static grid or
// ray queue), and integration + Snell metric on a prepared device fan.
TraceArgs rtb_base_args(rtb_ctx *ctx, const rtb_params *p, int stride);
int rtb_launch(rtb_ctx *ctx, cudaStream_t st, const rtb_params *p, TraceArgs a, bool traj);
int rtb_trace_core(rtb_ctx *ctx, cudaStream_t st, const rtb_params *p, const TraceArgs &a, double *d_err,
                   bool has_errors, bool want_err, bool traj, int stride, Staging &sg, int64_t &launches);

// rtb_trace with RTB_F_SORT_FAN (sortfan.cu): order the fan, trace, scatter the results back.
int rtb_trace_sorted(rtb_ctx *ctx, const rtb_params *p, const double *theta0, const double *x0, const double *y0, int64_t R,
                     double *d_ray, double *errors, double *last_row, double *px_stats, rtb_stats *stats);

// Field construction pieces defined in rtb.cu and shared with fit.cu.
void rtb_free_field(rtb_ctx *ctx);
int rtb_make_axis(rtb_ctx *ctx, const double *t, int n, int k, DevAxis *ax);
int rtb_finish_field(rtb_ctx *ctx, DevField &f, const double *tzu, int nzu, const double *tzv, int nzv,
                     const double *tgu, const double *tgv);

// Cells of the default evaluator's fast region that lie strictly inside the box (two cells of margin cover the
// half-ulp slop of the device-side index): a position indexed into them needs no FP64 box compares.
static inline CellBox make_cell_box(const DevField &f, double limx_i, double limx_s, double limy_i, double limy_s)
{
    CellBox b = {1, 0, 1, 0};
    if (!f.fast_ok) return b;
    auto coord = [](const DevAxis &a, double x) { return std::floor(x * a.inv_h + a.c0 + 0.5); };
    auto clampi = [](double v) { return (int)std::max(-1.0e9, std::min(1.0e9, v)); };
    b.lo_u = std::max(f.flo_u, clampi(coord(f.gu_, limy_i) + 2));
    b.hi_u = std::min(f.fhi_u, clampi(coord(f.gu_, limy_s) - 2));
    b.lo_v = std::max(f.flo_v, clampi(coord(f.gv_, limx_i) + 2));
    b.hi_v = std::min(f.fhi_v, clampi(coord(f.gv_, limx_s) - 2));
    return b;
}

// ---------------------------------------------------------------------------------------
// Staging: with RTB_PTR_HOST the library owns temporary device buffers for the call.
// ---------------------------------------------------------------------------------------
struct Staging {
    rtb_ctx *ctx;
    bool host;
    cudaStream_t stream;
    std::vector<void *> tmp;
    struct Back { void *host; void *dev; size_t bytes; size_t rows, host_pitch, dev_pitch; };
    std::vector<Back> back;
    int64_t h2d = 0, d2h = 0;
    cudaError_t err = cudaSuccess;
    Staging(rtb_ctx *c, bool h) : ctx(c), host(h), stream(c->stream) {}
    Staging(rtb_ctx *c, bool h, cudaStream_t s) : ctx(c), host(h), stream(s) {}
    Staging(const Staging &) = delete; // the destructor hands the buffers back to the pool
    Staging &operator=(const Staging &) = delete;
    ~Staging()
    {
        for (void *p : tmp)
            for (auto &b : ctx->pool) if (b.p == p) b.used = false;
    }
    // Scratch of one round of a loop: buffers allocated after mark() go back to the pool at release() and may be
    // handed out again by the next round -- safe because every round runs on this Staging's one stream.
    size_t mark() const { return tmp.size(); }
    void release(size_t m)
    {
        for (size_t k = m; k < tmp.size(); ++k)
            for (auto &b : ctx->pool) if (b.p == tmp[k]) b.used = false;
        tmp.resize(m);
    }
    void *alloc(size_t bytes)
    {
        if (!bytes) bytes = 8;
        rtb_ctx::PoolBuf *best = nullptr;
        for (auto &b : ctx->pool)
            if (!b.used && b.bytes >= bytes && (!best || b.bytes < best->bytes)) best = &b;
        if (best && best->bytes <= 2 * bytes + (1 << 20)) {
            best->used = true;
            tmp.push_back(best->p);
            return best->p;
        }
        void *d = nullptr;
        cudaError_t e = cudaMalloc(&d, bytes);
        if (e != cudaSuccess) { // release idle pool memory and retry once
            for (auto it = ctx->pool.begin(); it != ctx->pool.end();)
                if (!it->used) { cudaFree(it->p); it = ctx->pool.erase(it); } else ++it;
            e = cudaMalloc(&d, bytes);
        }
        if (e != cudaSuccess) { err = e; return nullptr; }
        ctx->pool.push_back({d, bytes, true});
        tmp.push_back(d);
        return d;
    }
    template <class T> const T *in(const T *p, size_t count)
    {
        if (!p || !host) return p;
        void *d = alloc(count * sizeof(T));
        if (!d) return nullptr;
        cudaError_t e = cudaMemcpyAsync(d, p, count * sizeof(T), cudaMemcpyHostToDevice, stream);
        if (e != cudaSuccess) err = e;
        h2d += (int64_t)(count * sizeof(T));
        return (const T *)d;
    }
    template <class T> T *out(T *p, size_t count)
    {
        if (!p || !host) return p;
        void *d = alloc(count * sizeof(T));
        if (!d) return nullptr;
        back.push_back({p, d, count * sizeof(T), 1, 0, 0});
        return (T *)d;
    }
    // Columns [col0, col0+cols) of a host array of `rows` rows of `pitch` elements: a dense (rows, cols) device array.
    template <class T> T *out_cols(T *p, size_t rows, size_t pitch, size_t col0, size_t cols)
    {
        if (!p) return nullptr;
        void *d = alloc(rows * cols * sizeof(T));
        if (!d) return nullptr;
        back.push_back({p + col0, d, cols * sizeof(T), rows, pitch * sizeof(T), cols * sizeof(T)});
        return (T *)d;
    }
    // A large result going to PAGEABLE host memory (the drop-in trazar() returns a fresh numpy array: 45.8 GB for the
    // 2^20-ray fisheye fan).  cudaMemcpy stages such a copy through the driver's own pinned buffer on one host thread and
    // takes the first-touch page faults of the fresh array on that same thread (measured 4 GB/s).  Here the copy is cut
    // into 64 MB chunks that DMA into three pinned bounce buffers while a team of host threads moves -- and first-touches --
    // the previous chunk into the caller's array.
    bool copy_back_pageable(void *host, const void *dev, size_t bytes)
    {
        for (int k = 0; k < rtb_ctx::kBounce; ++k) {
            if (!ctx->bounce[k] && cudaMallocHost(&ctx->bounce[k], rtb_ctx::kBounceBytes) != cudaSuccess) { (void)cudaGetLastError(); return false; }
            if (!ctx->bounce_ev[k] && cudaEventCreateWithFlags(&ctx->bounce_ev[k], cudaEventDisableTiming) != cudaSuccess) return false;
        }
        const size_t CH = rtb_ctx::kBounceBytes;
        const size_t nch = (bytes + CH - 1) / CH;
        const unsigned hw = std::thread::hardware_concurrency();
        const int T = (int)std::max(1u, std::min(16u, hw ? hw : 4u));
        auto issue = [&](size_t i) {
            const size_t off = i * CH, n = std::min(CH, bytes - off);
            const int k = (int)(i % rtb_ctx::kBounce);
            cudaError_t e = cudaMemcpyAsync(ctx->bounce[k], (const char *)dev + off, n, cudaMemcpyDeviceToHost, stream);
            if (e == cudaSuccess) e = cudaEventRecord(ctx->bounce_ev[k], stream);
            if (e != cudaSuccess) err = e;
        };
        for (size_t i = 0; i < std::min<size_t>(nch, rtb_ctx::kBounce); ++i) issue(i);
        for (size_t i = 0; i < nch && err == cudaSuccess; ++i) {
            const size_t off = i * CH, n = std::min(CH, bytes - off);
            const int k = (int)(i % rtb_ctx::kBounce);
            cudaError_t e = cudaEventSynchronize(ctx->bounce_ev[k]);
            if (e != cudaSuccess) { err = e; break; }
            std::vector<std::thread> team;
            const size_t per = ((n + T - 1) / T + 4095) & ~(size_t)4095;
            for (int t = 0; t < T; ++t) {
                const size_t a = std::min(n, (size_t)t * per), b = std::min(n, a + per);
                if (b > a) team.emplace_back([=] { std::memcpy((char *)host + off + a, (const char *)ctx->bounce[k] + a, b - a); });
            }
            for (auto &th : team) th.join();
            if (i + rtb_ctx::kBounce < nch) issue(i + rtb_ctx::kBounce);
        }
        d2h += (int64_t)bytes;
        return err == cudaSuccess;
    }
    void copy_back()
    {
        for (auto &b : back) {
            if (b.rows == 1 && b.bytes >= ((size_t)16 << 20)) { // large and contiguous: pageable destination?  (61 MB of
                // configs[0]'s s_ray: 15 ms by cudaMemcpy into fresh pages, 3 ms through the bounce buffers)
                cudaPointerAttributes at;
                const bool pageable = cudaPointerGetAttributes(&at, b.host) != cudaSuccess || at.type == cudaMemoryTypeUnregistered;
                (void)cudaGetLastError();
                if (pageable && copy_back_pageable(b.host, b.dev, b.bytes)) continue;
                if (err != cudaSuccess) return;
            }
            cudaError_t e = b.rows == 1
                ? cudaMemcpyAsync(b.host, b.dev, b.bytes, cudaMemcpyDeviceToHost, stream)
                : cudaMemcpy2DAsync(b.host, b.host_pitch, b.dev, b.dev_pitch, b.bytes, b.rows, cudaMemcpyDeviceToHost, stream);
            if (e != cudaSuccess) err = e;
            d2h += (int64_t)(b.bytes * b.rows);
        }
    }
};

