"""TEST INFRASTRUCTURE ONLY -- CPU restatement of interpolacion() (RT_bench.py:435-464) without scipy on the path:
np.gradient(Z, DELTA, edge_order=2) and the s=0 RectBivariateSpline fits (FITPACK regrid: fpregr knot placement,
fpgrre coefficients).  FITPACK is a third-party dependency of the reference (scipy, any version with
RectBivariateSpline; 1.x in this image); its published algorithm for s=0 is interpolation with the knots
    t = [x0]*(k+1) + x[(k+1)//2 : m-(k+1)//2] + [x[m-1]]*(k+1)
and coefficients solving A_u C A_v^T = Z with A the B-spline collocation matrix at the samples (FITPACK solves it by
Givens QR, this file by banded LU -- the same system).  PINNED: tests/test_fit_oracle.py checks knots exactly and
coefficients to a few ulp against scipy itself (the reference's own call, RT:455-457) on the four scenario grids and
on ragged random grids.
"""
import numpy as np


def gradient(Z, delta):
    """np.gradient(Z, delta, edge_order=2) written out (numpy/lib/function_base.py, uniform spacing branch)."""
    Z = np.asarray(Z, dtype=np.float64)
    out = []
    for ax in range(2):
        f = np.moveaxis(Z, ax, 0)
        g = np.empty_like(f)
        g[1:-1] = (f[2:] - f[:-2]) / (2. * delta)
        a, b, c = -1.5 / delta, 2. / delta, -0.5 / delta
        g[0] = a * f[0] + b * f[1] + c * f[2]
        a, b, c = 0.5 / delta, -2. / delta, 1.5 / delta
        g[-1] = a * f[-3] + b * f[-2] + c * f[-1]
        out.append(np.moveaxis(g, 0, ax))
    return out


def interp_knots(x, k):
    """fpregr with s=0: every sample is a knot except (k=3) the second and the second-last."""
    x = np.asarray(x, dtype=np.float64)
    s = (k + 1) // 2
    return np.concatenate([[x[0]] * (k + 1), x[s:len(x) - s], [x[-1]] * (k + 1)])


def bspl3(t, l, x):
    """FITPACK fpbspl, degree 3: B_{l-3..l}(x) for t[l] <= x <= t[l+1]."""
    h = [1.0, 0.0, 0.0, 0.0]
    for j in range(1, 4):
        hh = h[:j]
        h[0] = 0.0
        for i in range(1, j + 1):
            tli, tlj = t[l + i], t[l + i - j]
            if tli == tlj:
                h[i] = 0.0
                continue
            f = hh[i - 1] / (tli - tlj)
            h[i - 1] = h[i - 1] + f * (tli - x)
            h[i] = f * (x - tlj)
    return h


def colloc_band(t, x):
    """A[i, d+2] = A_full[i, i+d], d = -2..2: cubic B-splines on knots t at the samples x."""
    m = len(x)
    A = np.zeros((m, 5))
    for i in range(m):
        l = min(max(i + 2, 3), m - 1)
        h = bspl3(t, l, x[i])
        for q in range(4):
            d = (l - 3 + q) - i
            if -2 <= d <= 2:
                A[i, d + 2] = h[q]
            else:
                assert h[q] == 0.0
    return A


def band_lu(A):
    """In-place banded LU without pivoting (totally positive matrix); returns (l2, l1, 1/u0, u1, u2) columns."""
    A = A.copy()
    m = A.shape[0]
    for k in range(m):
        for i in range(k + 1, min(k + 2, m - 1) + 1):
            l = A[i, k - i + 2] / A[k, 2]
            A[i, k - i + 2] = l
            for j in range(k + 1, min(k + 2, m - 1) + 1):
                A[i, j - i + 2] -= l * A[k, j - k + 2]
    A[:, 2] = 1.0 / A[:, 2]
    return A


def band_solve(fac, B):
    """Solve along axis 0 of B for every column (the device's band_solve_kernel)."""
    m = B.shape[0]
    Y = np.array(B, dtype=np.float64)
    for i in range(m):
        if i >= 1:
            Y[i] -= fac[i, 1] * Y[i - 1]
        if i >= 2:
            Y[i] -= fac[i, 0] * Y[i - 2]
    for i in range(m - 1, -1, -1):
        if i + 1 < m:
            Y[i] -= fac[i, 3] * Y[i + 1]
        if i + 2 < m:
            Y[i] -= fac[i, 4] * Y[i + 2]
        Y[i] *= fac[i, 2]
    return Y


def fit_cubic(y, x, G):
    """(tu, tv, c) of RectBivariateSpline(y, x, G, kx=3, ky=3) (s=0)."""
    tu, tv = interp_knots(y, 3), interp_knots(x, 3)
    fu, fv = band_lu(colloc_band(tu, y)), band_lu(colloc_band(tv, x))
    C = band_solve(fu, np.asarray(G, dtype=np.float64))
    C = band_solve(fv, C.T).T
    return tu, tv, np.ascontiguousarray(C).ravel()


def interpolacion(x, y, Z, delta):
    """RT_bench.py:435-464 -> ((tu, tv, c, 1, 1), ((tu, tv, c, 3, 3), (tu, tv, c, 3, 3)))."""
    g0, g1 = gradient(Z, delta)
    z = (interp_knots(y, 1), interp_knots(x, 1), np.asarray(Z, dtype=np.float64).ravel().copy(), 1, 1)
    a = fit_cubic(y, x, g0) + (3, 3)
    b = fit_cubic(y, x, g1) + (3, 3)
    return z, (a, b)
