"""goofit.landau — the helper functions upstream exposes from python/extras/landau.cpp (landauPDF, gaussPDF, landauGaussPDF,
landau_quantile, landau_quantile_c; its examples/simpleFit.py draws Landau toys with them), written in numpy.

landauPDF is the CERNLIB G110 DENLAN density (Koelbig & Schorr, Comput. Phys. Commun. 31 (1984) 97): the rational approximations the
device-side LandauPdf evaluates (csrc/gf_ops.cuh gf_landau), as a function of (x - x0) / xi, divided by xi.
landau_quantile inverts the cumulative integral of that density numerically (tabulated once on 4x10^5 points and interpolated)
instead of interpolating CERNLIB's RANLAN table as upstream does: the two agree to the accuracy of that table (a few 1e-4 in lambda),
not bit for bit. Host-side numpy only: nothing here touches the GPU."""
import numpy as np

_NUM = np.array([[0.4259894875, -0.1249762550, 0.03984243700, -0.006298287635, 0.001511162253],
                 [0.1788541609, 0.1173957403, 0.01488850518, -0.001394989411, 0.0001283617211],
                 [0.1788544503, 0.09359161662, 0.006325387654, 0.00006611667319, -0.000002031049101],
                 [0.9874054407, 118.6723273, 849.2794360, -743.7792444, 427.0262186],
                 [1.003675074, 167.5702434, 4789.711289, 21217.86767, -22324.94910],
                 [1.000827619, 664.9143136, 62972.92665, 475554.6998, -5743609.109]])
_DEN = np.array([[1.0, -0.3388260629, 0.09594393323, -0.01608042283, 0.003778942063],
                 [1.0, 0.7428795082, 0.3153932961, 0.06694219548, 0.008790609714],
                 [1.0, 0.6097809921, 0.2560616665, 0.04746722384, 0.006957301675],
                 [1.0, 106.8615961, 337.6496214, 2016.712389, 1597.063511],
                 [1.0, 156.9424537, 3745.310488, 9834.698876, 66924.28357],
                 [1.0, 651.4101098, 56974.73333, 165917.4725, -2815759.939]])


def _quartic(k, t):
    return k[0] + (k[1] + (k[2] + (k[3] + k[4] * t) * t) * t) * t


def _denlan(v):
    v = np.asarray(v, dtype=np.float64)
    out = np.zeros_like(v)
    with np.errstate(over="ignore", under="ignore", divide="ignore", invalid="ignore"):
        m = v < -5.5
        if m.any():
            u = np.exp(v[m] + 1.0)
            d = 0.3989422803 * (np.exp(-1 / u) / np.sqrt(u)) * (1 + (0.04166666667 + (-0.01996527778 + 0.02709538966 * u) * u) * u)
            out[m] = np.where(u < 1e-10, 0.0, d)
        m = (v >= -5.5) & (v < -1)
        if m.any():
            u = np.exp(-v[m] - 1)
            out[m] = np.exp(-u) * np.sqrt(u) * _quartic(_NUM[0], v[m]) / _quartic(_DEN[0], v[m])
        for r, lo, hi in ((1, -1, 1), (2, 1, 5)):
            m = (v >= lo) & (v < hi)
            if m.any():
                out[m] = _quartic(_NUM[r], v[m]) / _quartic(_DEN[r], v[m])
        for r, lo, hi in ((3, 5, 12), (4, 12, 50), (5, 50, 300)):
            m = (v >= lo) & (v < hi)
            if m.any():
                u = 1 / v[m]
                out[m] = u * u * _quartic(_NUM[r], u) / _quartic(_DEN[r], u)
        m = v >= 300
        if m.any():
            u = 1 / (v[m] - v[m] * np.log(v[m]) / (v[m] + 1))
            out[m] = u * u * (1 + (-1.845568670 + -4.284640743 * u) * u)
    return out


def landauPDF(x, x0=0.0, xi=1.0):
    """Landau density in x with location x0 and scale xi: DENLAN((x - x0) / xi) / xi; 0 for xi <= 0"""
    x, x0, xi = np.broadcast_arrays(np.asarray(x, dtype=np.float64), x0, xi)
    with np.errstate(divide="ignore", invalid="ignore"):
        d = _denlan((x - x0) / np.where(xi > 0, xi, 1.0)) / np.where(xi > 0, xi, 1.0)
    d = np.where(xi > 0, d, 0.0)
    return d if d.ndim else float(d)


def gaussPDF(x, mu=0.0, sigma=1.0):
    x = np.asarray(x, dtype=np.float64)
    d = 0.3989422804014 / sigma * np.exp(-((x - mu) ** 2) / (2 * np.asarray(sigma, dtype=np.float64) ** 2))
    return d if np.ndim(d) else float(d)


def landauGaussPDF(x, mu=0.0, eta=1.0, sigma=1.0):
    """Landau (location mu, width eta) folded with a Gaussian of width sigma: midpoint rule over +-8 sigma in 100 steps (more, up to 10^5,
    once sigma > 3 eta), the discretisation upstream uses (including its second division of the Landau density by eta)"""
    def one(xv, m, e, s):
        n = 100
        if s > 3 * e:
            n *= int(s / e / 3.0)
        n = min(n, 100000)
        lo, hi = xv - 8 * s, xv + 8 * s
        step = (hi - lo) / n
        i = np.arange(1, n // 2 + 1)
        xx = np.concatenate([lo + (i - 0.5) * step, hi - (i - 0.5) * step])
        return float(step * np.sum(landauPDF(xx, m, e) / e * gaussPDF(xv, xx, s)))
    out = np.vectorize(one, otypes=[np.float64])(x, mu, eta, sigma)
    return out if out.ndim else float(out)


_TABLE = None


def _cdf_table():
    """lambda grid and cumulative integral of the density: uniform below 5, logarithmic above (the tail falls like 1 / lambda^2)"""
    global _TABLE
    if _TABLE is None:
        v = np.concatenate([np.linspace(-8.0, 5.0, 200001)[:-1], np.geomspace(5.0, 1e7, 200001)])
        d = _denlan(v)
        c = np.concatenate([[0.0], np.cumsum(0.5 * (d[1:] + d[:-1]) * np.diff(v))])
        total = c[-1] + 1.0 / v[-1]  # beyond the grid the density is 1 / lambda^2 to leading order
        _TABLE = (v, c / total, total)
    return _TABLE


def landau_quantile(z, xi=1.0):
    """lambda with P(Landau < lambda) = z, times the scale xi (0 for xi <= 0, -inf / +inf for z <= 0 / z >= 1, as upstream)"""
    v, c, total = _cdf_table()
    z, xi = np.broadcast_arrays(np.asarray(z, dtype=np.float64), xi)
    zz = np.clip(z, 1e-300, 1 - 1e-16)
    q = np.interp(zz, c, v)
    # beyond the table the survival function is 1 / (total lambda) to leading order
    q = np.where(zz > c[-1], 1.0 / (np.maximum(1 - zz, 1e-300) * total), q)
    out = np.where(z <= 0, -np.inf, np.where(z >= 1, np.inf, q)) * xi
    out = np.where(xi > 0, out, 0.0)
    return out if out.ndim else float(out)


def landau_quantile_c(z, xi=1.0):
    return landau_quantile(1.0 - np.asarray(z, dtype=np.float64), xi)
