"""Host-side float64 tables (numpy), computed exactly as the reference computes them and
uploaded to the kernels: FFT wavenumbers, the separable CIC-window factor, the
P(k) bin edges, and the k²-space decision thresholds that make device binning bit-exact.
"""
from __future__ import annotations

import numpy as np


def k_vals(N, L):
    """`2*np.pi*np.fft.fftfreq(N, d=L/N)` — FraP_kTilDelta.py:74-75, Power_compare.py:179."""
    return 2 * np.pi * np.fft.fftfreq(N, d=L / N)


def sinc_table(N, L):
    """`np.sinc(k_vals/(2*k_quist))`, `k_quist = np.pi*(d)**(-1)` — FraP_kTilDelta.py:77-80.
    The 3-D window is W[i,j,k] = ((s[i]*s[j])*s[k])**2."""
    d = L / N
    k_quist = np.pi * (d) ** (-1)
    return np.sinc(k_vals(N, L) / (2 * k_quist))


def sinc_table_kq(N, L):
    """Same factor with `k_quist = np.pi * N / L` (Power_compare.py:436-437, Plotter.py:157)."""
    k_quist = np.pi * N / L
    return np.sinc(k_vals(N, L) / (2 * k_quist))


def pk_edges(N, L, n_edges=300):
    """`np.logspace(np.log10(2π/L), np.log10(k_mag.max()), n_edges)` — Power_compare.py:
    187-190; k_mag.max() formed as sqrt((kx²+ky²)+kz²) like numpy does."""
    kv = k_vals(N, L)
    k2max = (kv ** 2).max()
    k_max = np.sqrt((k2max + k2max) + k2max)
    k_min = 2 * np.pi / L
    return np.logspace(np.log10(k_min), np.log10(k_max), n_edges)


def sqrt_thresholds(edges):
    """For each edge e: the smallest float64 x with np.sqrt(x) >= e.  Also returns the
    smallest x with np.sqrt(x) > edges[-1].  sqrt is monotone and correctly rounded, so
    `x >= thresh[e]`  <=>  `np.sqrt(x) >= edges[e]` for every float64 x >= 0."""
    e = np.asarray(edges, dtype=np.float64)

    def smallest(pred_ge):
        x = e * e
        for _ in range(64):                      # walk down while the predicate still holds
            lower = np.nextafter(x, 0.0)
            move = pred_ge(lower)
            if not move.any():
                break
            x = np.where(move, lower, x)
        for _ in range(64):                      # walk up until it holds
            bad = ~pred_ge(x)
            if not bad.any():
                break
            x = np.where(bad, np.nextafter(x, np.inf), x)
        return x

    thr = smallest(lambda x: np.sqrt(x) >= e)
    thr_gt = smallest(lambda x: np.sqrt(x) > e)
    return thr, float(thr_gt[-1])


def split_poly(u_max, degree=10):
    """Monomial coefficients (low -> high, in t = 2u²/u_max² - 1) of Q with
    erfc(u) + 2u/sqrt(pi)·exp(-u²) = 1 - u³·Q(u²): the short-range fraction of the Newtonian
    force for the P3M split.  Chebyshev interpolation on [0, u_max²]; |error| ~ 1e-8."""
    import math

    from numpy.polynomial import chebyshev as Ch

    def Q(w):
        w = np.asarray(w, dtype=np.float64)
        u = np.sqrt(np.maximum(w, 1e-300))
        with np.errstate(all="ignore"):
            from math import pi
            erf = np.vectorize(math.erf)(u)
            q = (erf - 2 * u / math.sqrt(pi) * np.exp(-w)) / u ** 3
        series = 4 / (3 * math.sqrt(math.pi)) * (1 - 3 * w / 5 + 3 * w * w / 14 - w ** 3 / 18)
        return np.where(w < 1e-3, series, q)

    wmax = u_max * u_max
    c = Ch.chebinterpolate(lambda t: Q((t + 1) * 0.5 * wmax), degree)
    return Ch.cheb2poly(c)
