"""tools/make_zachfit_toy.py DIR [N] — synthetic stand-ins for the two text files examples/zachFit/zachFit.cpp reads in its default
mode (dataFiles/dstwidth_kpi_resMC.dat, dataFiles/dstwidth_kpi_data.dat: one delta-m value per line; upstream downloads the BaBar
samples at configure time, they are not part of the checkout). The resolution sample is the three-Gaussian mixture of the example's
own start values plus a threshold background; the data sample is a Breit-Wigner at 0.1454 GeV of width 1e-4 GeV smeared with that
resolution, on a square-root threshold background."""
import os
import sys

import numpy as np


def resolution(rng, n):
    k = rng.choice(3, size=n, p=[0.65, 0.02, 0.33])
    mean = np.array([0.145402, 0.145465, 0.145404])[k]
    sigma = np.array([0.00010, 0.00075, 0.00020])[k]
    return rng.normal(mean, sigma)


def threshold_background(rng, n, lo=0.13957, hi=0.1665):
    u = rng.random(n)  # density ~ sqrt(x - lo): x = lo + (hi - lo) u^(2/3)
    return lo + (hi - lo) * u ** (2.0 / 3.0)


def main():
    out = sys.argv[1]
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
    rng = np.random.default_rng(20231105)
    os.makedirs(os.path.join(out, "dataFiles"), exist_ok=True)
    mc = np.concatenate([resolution(rng, n - n // 200), threshold_background(rng, n // 200)])
    sig = 0.1454 + 0.5e-4 * np.tan(np.pi * (rng.random(n - n // 5) - 0.5)) + (resolution(rng, n - n // 5) - 0.145402)
    data = np.concatenate([sig, threshold_background(rng, n // 5)])
    for name, v in (("dstwidth_kpi_resMC.dat", mc), ("dstwidth_kpi_data.dat", data)):
        v = v[(v > 0.1396) & (v < 0.1664)]
        rng.shuffle(v)
        np.savetxt(os.path.join(out, "dataFiles", name), v, fmt="%.9f")


if __name__ == "__main__":
    main()
