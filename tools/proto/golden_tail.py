"""Dev prototype (numpy): passes k >= K1 of the golden-section search on the anisotropic cost (RT:725-729, 758-762) decided
from the cubic Taylor expansion of the cost about its minimiser instead of cost evaluations.  Checks, on random steps,
that outside a band every decision equals the evaluating search's, and reports how often the band is hit.
usage: golden_tail.py [K1] [gamma] [residual scale]"""
import sys
import numpy as np
rho = (np.sqrt(5) - 1) / 2
ITERS = 37
TIE_EPS = 3.6e-15


def run(K1=12, gamma=3.0, RS=1.0, N=200000, verbose=True):
    """Returns (decisions outside the band that differ, evaluating-search band passes the tail's band misses,
    share of steps with a tail pass in the band: new, old)."""
    say = print if verbose else (lambda *a, **k: None)
    rng = np.random.default_rng(1)
    g = gamma * gamma - 1

    th0 = rng.uniform(-np.pi, np.pi, N)
    n0 = rng.uniform(1.0, 1.6, N); h = rng.uniform(0.002, 0.03, N) * RS
    gx0 = rng.normal(0, 0.3, N); gy0 = rng.normal(0, 0.3, N)
    n1 = n0 + h * rng.normal(0, 0.3, N); gx1 = gx0 + rng.normal(0, 0.02, N); gy1 = gy0 + rng.normal(0, 0.02, N)
    ux, uy = np.cos(th0), np.sin(th0)
    coef0 = np.sqrt(1 + g * uy * uy)
    ncf = n0 * coef0
    mx = 0.5 * h * coef0 * gx0 + ncf * ux * (1 - uy * uy * g / coef0**2)
    my = 0.5 * h * coef0 * gy0 + ncf * uy * (1 + ux * ux * g / coef0**2)
    Gx, Gy = 0.5 * h * gx1, 0.5 * h * gy1
    A_, B_ = n1, n1 * gamma * gamma
    scale = n1 * max(gamma, 1.0)

    def parts(s, c):
        v = 1 + g * s * s; y = 1 / np.sqrt(v); w = v * y
        Rx = A_ * c * y - (w * Gx + mx); Ry = B_ * s * y - (w * Gy + my)
        return y, Rx, Ry
    def cost(th):
        y, Rx, Ry = parts(np.sin(th), np.cos(th))
        return Rx * Rx + Ry * Ry
    def d12(s, c):
        """F1 = f'/2, F2 = f''/2 at the angle with (sin, cos) = (s, c)."""
        y, Rx, Ry = parts(s, c)
        y2 = y * y; y3 = y2 * y; z = g * s * c; zy = z * y; By3 = B_ * y3
        Rx1 = -(By3 * s + zy * Gx); Ry1 = By3 * c - zy * Gy
        F1 = Rx * Rx1 + Ry * Ry1
        z1 = g * (c * c - s * s); t1 = z * y2; q2 = z1 * y - zy * t1; u = 3 * By3 * t1
        Rx2 = -By3 * c + u * s - q2 * Gx; Ry2 = -By3 * s - u * c - q2 * Gy
        F2 = Rx1 * Rx1 + Ry1 * Ry1 + Rx * Rx2 + Ry * Ry2
        return F1, F2, Rx * Rx + Ry * Ry

    # the evaluating search (what the shipped kernel decides outside its own band)
    a = th0 - np.pi / 2; b = th0 + np.pi / 2
    c = b - (b - a) * rho; d = a + (b - a) * rho
    dec = np.zeros((ITERS, N), bool); rel = np.zeros((ITERS, N))
    for k in range(ITERS):
        if k == K1: cK, dK = c.copy(), d.copy()
        fc, fd = cost(c), cost(d)
        left = fc < fd
        dec[k] = left; rel[k] = np.abs(fc - fd) / np.sqrt(fc + fd)      # old band: rel <= TIE_EPS * scale
        b = np.where(left, d, b); a = np.where(left, a, c)
        c = b - (b - a) * rho; d = a + (b - a) * rho
    th_plain = (a + b) / 2

    W = np.pi * rho ** np.arange(ITERS + 1)
    q = (rho - 0.5) * W                        # half the probe distance of pass k
    midmove = 0.5 * (1 - rho) * W              # |midpoint move| of pass k

    # tail: Newton from the kept probe of pass K1 (one of c, d: take d, as if N were the left probe)
    x0 = dK
    s, cc = np.sin(x0), np.cos(x0)
    def rot(s, c, dl):                          # small rotation by series
        d2 = dl * dl
        sn = dl * (1 + d2 * (-1 / 6 + d2 * (1 / 120)))
        cm = d2 * (-0.5 + d2 * (1 / 24 - d2 / 720))
        return s + (s * cm + c * sn), c + (c * cm - s * sn)
    F1, F2a, _ = d12(s, cc); dl0 = -F1 / F2a
    s1, c1 = rot(s, cc, dl0)
    F1, F2b, _ = d12(s1, c1); dl1 = -F1 / F2b
    s2, c2 = rot(s1, c1, dl1)
    F1, F2c, fst = d12(s2, c2); dl2 = -F1 / F2c
    delta = dl0 + dl1 + dl2                    # x* - x0
    kap = 2 * (F2b - F2a) / dl0 / (F2a + F2b)  # f'''/f'' by the secant over the first Newton step
    Gm = max(gamma, 1 / gamma) ** 2
    lam_b = 16 * Gm + 8                        # bound on |f''''/f''|
    ok = (F2c > 0) & (np.abs(dl2) < 1e-7) & (np.abs(dl0) > 1e-10) & (np.abs(delta) < 0.01)
    say("usable", ok.mean(), "newton |dl| max", np.abs(dl0)[ok].max(), np.abs(dl1)[ok].max(), np.abs(dl2)[ok].max())
    # exact kappa / lambda for the report
    xs = x0 + delta; eps = 1e-4
    f2 = lambda x: d12(np.sin(x), np.cos(x))[1]
    kap_ex = (f2(xs + eps) - f2(xs - eps)) / (2 * eps) / F2c; lam_ex = (f2(xs + eps) - 2 * f2(xs) + f2(xs - eps)) / eps**2 / F2c
    say("kappa err", np.abs(kap - kap_ex)[ok].max(), "bound used", (lam_b * np.abs(dl0))[ok].max(), " lambda max", np.abs(lam_ex)[ok].max(), "bound", lam_b)

    e_ = (x0 - q[K1]) - xs                      # midpoint of the probes (x0 is d) minus the minimiser
    kap_err = lam_b * np.abs(dl0)               # secant midpoint is dl0/2 from x0, x* is dl0 away: <= lam |dl0|
    noise_a = 1.25 * TIE_EPS * scale * np.sqrt(2 * fst) / (4 * F2c)          # / q
    noise_b = 1.25 * TIE_EPS * scale * 4.4 * np.sqrt(2 * F2c) / (4 * F2c) + 2e-14 + (np.abs(kap) + 2) * dl2 * dl2
    tie_apx = np.zeros(N, bool); tie_noise = np.zeros(N, bool); first_apx = np.zeros(ITERS)
    mism = np.zeros(N, bool); tie_new = np.zeros(N, bool); tie_old = np.zeros(N, bool); old_in_new = 0
    for k in range(K1, ITERS):
        qq = q[k]
        m2 = e_ * e_ / 2 + qq * qq / 6
        t = e_ + kap * m2
        band = noise_a / qq + noise_b + kap_err * m2 + lam_b * (np.abs(e_) + qq) ** 3 / 3
        left = t > 0
        tie = np.abs(t) <= band
        old = rel[k] <= TIE_EPS * scale
        ta = np.abs(t) <= kap_err * m2 + lam_b * (np.abs(e_) + qq) ** 3 / 3; tie_apx |= ta & ok; first_apx[k] = (ta & ok).mean()
        tie_noise |= (np.abs(t) <= noise_a / qq + noise_b) & ok
        mism |= (left != dec[k]) & ~tie & ok
        old_in_new += (old & ~tie & ok).sum()
        tie_new |= tie & ok; tie_old |= old & ok
        e_ = e_ - np.where(dec[k], midmove[k], -midmove[k])
    say("K1 %d gamma %.2f: decisions outside the band that differ: %d ; old-band passes the new band misses: %d" % (K1, gamma, mism.sum(), old_in_new))
    say("steps with a tail pass in the band: new %.5f  old %.5f ; extra %.5f" % (tie_new.mean(), tie_old.mean(), (tie_new & ~tie_old).mean()))
    say("approx-error band alone %.5f (per pass %s) ; noise band alone %.5f" % (tie_apx.mean(), np.array2string(first_apx[K1:K1+6], precision=5), tie_noise.mean()))
    say("final |th - plain| max", np.max(np.abs((xs + e_) - th_plain)[ok]))
    return int(mism.sum()), int(old_in_new), float(tie_new.mean()), float(tie_old.mean())


if __name__ == "__main__":
    run(int(sys.argv[1]) if len(sys.argv) > 1 else 12, float(sys.argv[2]) if len(sys.argv) > 2 else 3.0,
        float(sys.argv[3]) if len(sys.argv) > 3 else 1.0)
