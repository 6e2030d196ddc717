"""TEST INFRASTRUCTURE ONLY — float64 numpy restatement of the CONCEPT-style power-spectrum
estimator and the derivation of its bins from the reference's own files.

Pinned by: the five `powerspec_a=*` files the reference ships (PPPMP3M/{PP,PM,P3M}/powerspec/
powerspec_a=1.00, SnapVsPK/powerspec/powerspec_a={0.17,1.00}; byte copies under
tests/golden/concept_powerspec/).  Their `modes` and `k` columns depend on the grid only:
`bins_from_modes_column` recovers the integer-k² range of every bin from the modes column by
brute-force enumeration of the half Fourier space, and `geometry_bruteforce` recomputes modes and the
geometric-mean k — an independent check of the product's lattice-count tables
(nbody_b200/powerspec.py).  The estimator itself (`powerspec`) restates the options of
SnapVsPK/Uncertain.param:29-62 — CIC deposit (the reference's cic_density, FraP_kTilDelta.py:93-123),
interlacing with the half-cell-shifted deposit, deconvolution by the reference's window
(cic_window_ft, :73-84), P = |δ_k|² V / N⁶ averaged over the modes of each bin.  CONCEPT's source is
absent, so the P column of a given particle set is PARITY UNPINNED at the single-evaluation level; it
is pinned statistically by the full-run comparison (tools/pin_concept.py).
"""
from __future__ import annotations

import numpy as np

from . import restate


def half_space_k2(N):
    """Integer k² of every mode of the half Fourier space CONCEPT loops over (kz > 0, or kz = 0 and
    ky > 0, or kz = ky = 0 and kx > 0), Nyquist planes excluded; returns (k2, index arrays)."""
    m = np.fft.fftfreq(N, 1.0 / N).astype(np.int64)
    mx, my, mz = np.meshgrid(m, m, m, indexing="ij")
    half = (mz > 0) | ((mz == 0) & (my > 0)) | ((mz == 0) & (my == 0) & (mx > 0))
    nyq = N // 2
    inner = (np.abs(mx) < nyq) & (np.abs(my) < nyq) & (np.abs(mz) < nyq)
    sel = half & inner
    return (mx * mx + my * my + mz * mz)[sel], sel


def bins_from_modes_column(N, modes):
    """[(first k², last k²)] per bin such that the half-space mode counts equal `modes`, walking up
    the occupied integer k² below Nyquist.  Raises if the column cannot be produced by contiguous
    k² ranges."""
    k2, _ = half_space_k2(N)
    k2 = k2[k2 < (N // 2) ** 2]
    cnt = np.bincount(k2)
    s, out = 1, []
    for want in modes:
        lo, acc = s, 0
        while acc < want:
            if s >= cnt.size:
                raise ValueError("modes column runs past the Nyquist frequency")
            acc += int(cnt[s])
            s += 1
        if acc != want:
            raise ValueError(f"modes count {want} is not a contiguous k² range starting at {lo}")
        last = s - 1
        while cnt[last] == 0:
            last -= 1
        while s < cnt.size and cnt[s] == 0:
            s += 1
        out.append((lo, last))
    return out


def geometry_bruteforce(N, ranges):
    """(modes, geometric-mean |k| in units of k_min) per bin by explicit enumeration."""
    k2, _ = half_space_k2(N)
    modes, kgeo = [], []
    for lo, hi in ranges:
        sel = k2[(k2 >= lo) & (k2 <= hi)]
        modes.append(sel.size)
        kgeo.append(np.exp(np.mean(0.5 * np.log(sel.astype(np.float64)))))
    return np.array(modes), np.array(kgeo)


def powerspec(pos, L, N, ranges, interlace=True, deconvolve=True):
    """Binned P(k) of a particle set: returns (modes, P) for the integer-k² `ranges`."""
    cell = L / N
    npart = pos.shape[0]
    dk = np.fft.fftn(restate.cic_density(pos, L, N))
    if interlace:
        m = np.fft.fftfreq(N, d=1.0 / N)
        ph = np.exp(1j * np.pi * m / N)                       # exp(+i k_a cell/2)
        if N % 2 == 0:
            ph[N // 2] = 0.0
        phase = ph[:, None, None] * ph[None, :, None] * ph[None, None, :]
        dk = 0.5 * (dk + np.fft.fftn(restate.cic_density(pos + 0.5 * cell, L, N)) * phase)
    dk = dk / (npart / N ** 3)
    if deconvolve:
        s = restate.sinc_table(N, L)
        W = ((s[:, None, None] * s[None, :, None]) * s[None, None, :]) ** 2
        dk = dk / W
    P3 = np.abs(dk) ** 2 * L ** 3 / float(N) ** 6
    k2, sel = half_space_k2(N)
    p = P3[sel]
    modes, out = [], []
    for lo, hi in ranges:
        w = (k2 >= lo) & (k2 <= hi)
        modes.append(int(w.sum()))
        out.append(p[w].mean())
    return np.array(modes), np.array(out)


def noise_ratio(R, L, N, ranges, p_lin):
    """Σ P_lin / Σ |R_k|² P_lin over the half-space modes of each bin (realisation correction)."""
    k2, sel = half_space_k2(N)
    a2 = (np.abs(R) ** 2)[sel]
    pl = p_lin(2 * np.pi / L * np.sqrt(k2.astype(np.float64)))
    return np.array([pl[(k2 >= lo) & (k2 <= hi)].sum() / (a2 * pl)[(k2 >= lo) & (k2 <= hi)].sum() for lo, hi in ranges])
