// traj_stage.cuh -- trajectory write-back of trace_kernel (TRAJ mode): s_ray[(row*6 + col)*R + k], RT_bench.py:802.
//
// For a fixed (row, column) the 32 rays of a warp own one 256-byte segment of s_ray.  Two ways to get a row there:
//   per-lane : six st.global.cs.f64 per thread per row (coalesced: one 256-byte segment per store instruction), with
//              the 64-bit address arithmetic of six column pointers in the step loop;
//   staged   : the warp parks RTB_TRAJ_ROWS rows x 6 columns x 32 lanes in its slice of shared memory -- six STS with
//              immediate offsets per row -- and ONE lane issues ONE 2-D TMA store (cp.async.bulk.tensor.2d shared ->
//              global, UTMASTG in SASS) for the whole tile: s_ray seen as the matrix (rows*6, R), the tile is the box
//              (6*RTB_TRAJ_ROWS) x 32 at (row*6, first ray of the warp).  Rays past the end of the fan are clipped by
//              the tensor map.  (Per-segment 1-D bulk copies would not do: UBLKCP takes warp-uniform addresses, so 24
//              lanes issuing one each are serialised into 24 issue rounds -- tools/hbm_write_microbench.cu.)
// The staged path needs 16-byte aligned rows (R even, s_ray 16-byte aligned) and one ray per lane; rtb_launch encodes the
// two tensor maps when it is asked for (RTB_OPT_TRAJ_STORE 1).  Measured on the 45.8 GB fisheye trajectory it is 5-7 %
// SLOWER than the per-lane stores: the write-back was never limited by the store instructions --
// a bare store kernel with this access pattern reaches the HBM copy peak either way (tools/hbm_write_microbench.cu) --
// but by the instruction count of the step loop, and the proxy fence + warp synchronisation of every flush cost more
// than the per-column address additions they replace.  So the per-lane path is the default and the staged one is kept
// selectable, bit-identical, for parts where the load/store pipe is the limiter.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace rtb {

#ifndef RTB_TRAJ_ROWS
#define RTB_TRAJ_ROWS 4 // rows per tile: 6 KB of shared memory per warp, one TMA store per 4 steps
#endif

__host__ __device__ constexpr size_t traj_stage_bytes(int block_threads)
{
    return (size_t)(block_threads / 32) * RTB_TRAJ_ROWS * 6 * 32 * sizeof(double);
}

struct TrajMaps {
    alignas(64) CUtensorMap box; // (6*RTB_TRAJ_ROWS) x 32 tiles
    alignas(64) CUtensorMap row; // 6 x 32: single rows (the last rows of a trajectory)
};

// MODE 0: no trajectory, 1: per-lane stores, 2: staged rows + TMA stores
template <int MODE> struct TrajWriter;

template <> struct TrajWriter<0> {
    __device__ __forceinline__ void init(double *, const TrajMaps *, int64_t, int64_t, bool, unsigned char *) {}
    __device__ __forceinline__ void row(bool, double, double, double, double, double, double) {}
    __device__ __forceinline__ void row_live(double, double, double, double, double, double) {}
    __device__ __forceinline__ void zero_tail(int, bool) {}
    __device__ __forceinline__ void ended() {}
    __device__ __forceinline__ void finish() {}
};

template <> struct TrajWriter<1> { // per-lane stores
    char *gp;    // this lane's entry of the next row
    int64_t Rb;  // bytes from one column to the next: 8 R.  The six column addresses of a row and the next row's first
                 // are formed by repeated 64-bit addition (12 instructions per row; indexing gp[c * R] costs 27)
    bool valid;
    __device__ __forceinline__ void init(double *s_ray, const TrajMaps *, int64_t R, int64_t r, bool valid_, unsigned char *)
    {
        Rb = R * (int64_t)sizeof(double); valid = valid_; gp = reinterpret_cast<char *>(s_ray + r);
    }
    __device__ __forceinline__ void put(double x, double yv, double px, double py, double T, double ang)
    {
        char *p = gp;
        __stcs(reinterpret_cast<double *>(p), x); p += Rb;
        __stcs(reinterpret_cast<double *>(p), yv); p += Rb;
        __stcs(reinterpret_cast<double *>(p), px); p += Rb;
        __stcs(reinterpret_cast<double *>(p), py); p += Rb;
        __stcs(reinterpret_cast<double *>(p), T); p += Rb;
        __stcs(reinterpret_cast<double *>(p), ang);
        gp = p + Rb;
    }
    // a row of a stepping ray (`live`: the lane is valid by construction) or, for a lane whose ray has ended, zeros
    // when the caller wants the tail written (store)
    __device__ __forceinline__ void row(bool store, double x, double yv, double px, double py, double T, double ang)
    {
        if (valid && store) put(x, yv, px, py, T, ang);
        else gp += 6 * Rb;
    }
    __device__ __forceinline__ void row_live(double x, double yv, double px, double py, double T, double ang)
    {
        put(x, yv, px, py, T, ang);
    }
    __device__ __forceinline__ void zero_tail(int n, bool tail_zero)
    {
        if (valid && tail_zero)
            for (int j = 0; j < n; ++j) put(0.0, 0.0, 0.0, 0.0, 0.0, 0.0);
    }
    __device__ __forceinline__ void ended() {}
    __device__ __forceinline__ void finish() {}
};

template <> struct TrajWriter<2> { // staged rows + TMA stores
    const TrajMaps *maps;
    double *buf;           // the warp's staging slice [RTB_TRAJ_ROWS][6][32]
    double *mine;          // this lane's column of it
    int x0, y;             // first ray of the warp, (next unflushed row) * 6
    int k;                 // rows staged and not yet flushed
    int lane;
    bool any;              // the warp holds at least one ray of the fan
    bool zero_pending;     // this lane's ray has ended: its column is cleared at the start of the next tile

    __device__ __forceinline__ void init(double *, const TrajMaps *maps_, int64_t R, int64_t r, bool, unsigned char *smem)
    {
        k = 0; maps = maps_; zero_pending = false;
        lane = threadIdx.x & 31;
        const int64_t r0 = r - lane;
        x0 = (int)r0; y = 0;
        any = r0 < R;
        buf = reinterpret_cast<double *>(smem) + (size_t)(threadIdx.x >> 5) * (RTB_TRAJ_ROWS * 6 * 32);
        mine = buf + lane;
    }

    // the TMA stores of the previous flush have finished READING the staging slice
    __device__ __forceinline__ void reads_done()
    {
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
    }

    __device__ __forceinline__ void tma_store(const CUtensorMap *m, const double *src, int yy)
    {
        const unsigned sa = (unsigned)__cvta_generic_to_shared(src);
        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];"
                     :: "l"(m), "r"(x0), "r"(yy), "r"(sa) : "memory");
    }

    // rows [0, n) of the staging slice -> s_ray; called by the whole warp
    __device__ __forceinline__ void flush(int n)
    {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // generic-proxy writes visible to the TMA unit
        __syncwarp();
        if (lane == 0 && any) {
            if (n == RTB_TRAJ_ROWS) tma_store(&maps->box, buf, y);
            else for (int j = 0; j < n; ++j) tma_store(&maps->row, buf + j * (6 * 32), y + 6 * j);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        y += 6 * n;
    }

    __device__ __forceinline__ void clear_rows(int from)
    {
#pragma unroll
        for (int j = 0; j < RTB_TRAJ_ROWS; ++j)
            if (j >= from) {
                double *b = mine + j * (6 * 32);
                b[0] = 0.0; b[32] = 0.0; b[64] = 0.0; b[96] = 0.0; b[128] = 0.0; b[160] = 0.0;
            }
    }

    // One kept row; called by all 32 lanes of the warp together.  A lane whose ray has ended (live == false) stores
    // nothing: its column of the slice was cleared when the ray ended (ended()), and zeros are what np.zeros left in
    // the rows after a ray's last step (RT:802).
    __device__ __forceinline__ void row(bool live, double x, double yv, double px, double py, double T, double ang)
    {
        if (k == 0) {
            reads_done();
            if (zero_pending) { clear_rows(0); zero_pending = false; }
        }
        if (live) {
            double *b = mine + k * (6 * 32);
            b[0] = x; b[32] = yv; b[64] = px; b[96] = py; b[128] = T; b[160] = ang;
        }
        if (++k == RTB_TRAJ_ROWS) { flush(RTB_TRAJ_ROWS); k = 0; }
    }

    __device__ __forceinline__ void row_live(double x, double yv, double px, double py, double T, double ang)
    {
        row(true, x, yv, px, py, T, ang);
    }

    // This lane's ray has just stored its last row.  The rows of the current tile it will not reach are cleared now:
    // they were last read by the previous flush, and reads_done ran when this tile was begun.  With k == 0 the row
    // just stored completed a tile whose TMA store is in flight and still reading the slice -- nothing may be
    // touched; the whole column is cleared at the start of the next tile, after reads_done (zero_pending).
    __device__ __forceinline__ void ended()
    {
        if (k > 0) clear_rows(k);
        zero_pending = true;
    }

    // n more rows of zeros (np.zeros RT:802) after the last step of every ray of the warp.
    __device__ __forceinline__ void zero_tail(int n, bool tail_zero)
    {
        if (k > 0) { flush(k); k = 0; }
        if (!tail_zero || n <= 0) return;
        reads_done();
        clear_rows(0);
        // the same zero tile serves every remaining flush: no shared-memory writes in between
        for (; n > 0; n -= RTB_TRAJ_ROWS) flush(min(n, RTB_TRAJ_ROWS));
    }

    __device__ __forceinline__ void finish()
    {
        if (k > 0) { flush(k); k = 0; }
        if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); // shared memory must outlive the stores
        __syncwarp();
    }
};

} // namespace rtb
