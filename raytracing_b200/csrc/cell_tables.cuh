// cell_tables.cuh -- expands the three splines of the field into per-cell polynomials (DevField::pg).
//
// On the uniform part of the grid the bicubic gradient splines (FITPACK form: one coefficient per knot, 16
// of them weighted by two cubic B-spline bases per evaluation, RT_bench.py:153-155) and the bilinear n are,
// inside one cell, plain polynomials of the offsets (tu, tv) from the cell centre.  Expanding them once per
// field (one thread per cell, on the device) trades 288 B of table per cell for half of the FP64 work of
// every evaluation in the step loop: the B200 has the HBM/L2 to spare (interface: 1.07 M cells = 311 MB) and
// the path is FP64-pipe bound.  The polynomials are the same functions; only rounding differs (<= a few ulp
// of max|coefficient|, like any re-association).
#pragma once
#include "field.cuh"

namespace rtb {

// Uniform cubic B-spline weights as polynomials of t = s - 1/2 (s = fraction of the cell):
//   w_a(t) = sum_p kW[a][p] t^p;  rows sum to (1, 0, 0, 0) (partition of unity).
__constant__ double kW[4][4] = {{1.0 / 48.0, -1.0 / 8.0, 1.0 / 4.0, -1.0 / 6.0},
                                {23.0 / 48.0, -5.0 / 8.0, -1.0 / 4.0, 1.0 / 2.0},
                                {23.0 / 48.0, 5.0 / 8.0, -1.0 / 4.0, -1.0 / 2.0},
                                {1.0 / 48.0, 1.0 / 8.0, 1.0 / 4.0, 1.0 / 6.0}};

// One thread per cell (lu, lv) of the fast region.
__global__ void cell_table_kernel(DevField f, double2 *pg)
{
    const int nv = f.fhi_v - f.flo_v + 1, nu = f.fhi_u - f.flo_u + 1;
    const size_t ci = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (ci >= (size_t)nu * nv) return;
    const int lu = f.flo_u + (int)(ci / nv), lv = f.flo_v + (int)(ci % nv);
    // bilinear n: coefficients (lu-2..lu-1, lv-2..lv-1) (the bilinear interval is the cubic one minus 1)
    {
        const double *c = f.cz + (size_t)(lu - 2) * f.ncz + (lv - 2);
        const double c00 = c[0], c01 = c[1], c10 = c[f.ncz], c11 = c[f.ncz + 1];
        pg[RTB_CELL_REC * ci + 16] = make_double2(0.25 * ((c00 + c01) + (c10 + c11)),
                                                   0.5 * ((c10 - c00) + (c11 - c01))); // n0, d/dtu (u = physical y)
        pg[RTB_CELL_REC * ci + 17] = make_double2(0.5 * ((c01 - c00) + (c11 - c10)),     // d/dtv
                                                   (c00 - c01) - (c10 - c11));
    }
    // bicubic pair: m[p][q] = sum_a sum_b kW[a][p] c[a][b] kW[b][q]
    const double2 *c = f.cg + (size_t)(lu - 3) * f.ncg + (lv - 3);
    double2 cq[4][4]; // cq[a][q] = sum_b c[a][b] kW[b][q]
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        const double2 c0 = c[(size_t)a * f.ncg], c1 = c[(size_t)a * f.ncg + 1], c2 = c[(size_t)a * f.ncg + 2],
                      c3 = c[(size_t)a * f.ncg + 3];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            cq[a][q].x = fma(c3.x, kW[3][q], fma(c2.x, kW[2][q], fma(c1.x, kW[1][q], c0.x * kW[0][q])));
            cq[a][q].y = fma(c3.y, kW[3][q], fma(c2.y, kW[2][q], fma(c1.y, kW[1][q], c0.y * kW[0][q])));
        }
    }
#pragma unroll
    for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            double2 m;
            m.x = fma(cq[3][q].x, kW[3][p], fma(cq[2][q].x, kW[2][p], fma(cq[1][q].x, kW[1][p], cq[0][q].x * kW[0][p])));
            m.y = fma(cq[3][q].y, kW[3][p], fma(cq[2][q].y, kW[2][p], fma(cq[1][q].y, kW[1][p], cq[0][q].y * kW[0][p])));
            pg[RTB_CELL_REC * ci + 4 * p + q] = m;
        }
}

} // namespace rtb
