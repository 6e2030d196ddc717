"""CONCEPT-style power-spectrum estimator.

The reference's simulations write `powerspec_a=*` text files through CONCEPT's own estimator,
configured by `powerspec_options` (SnapVsPK/Uncertain.param:29-62, TheFeedingPipeline/NN.param:28-50):
CIC interpolation on an upstream grid of cbrt(N) cells, interlaced, deconvolved, k_max = Nyquist,
"bins per decade" running from 4 (at 4 k_min) to 40 (at 100 k_min), a tophat σ₈ and three power
columns (data, realisation-corrected, linear).  The five files the reference ships
(PPPMP3M/{PP,PM,P3M}/powerspec/powerspec_a=1.00, SnapVsPK/powerspec/powerspec_a={0.17,1.00}) are the
known answers of this module: their `modes` and `k` columns are functions of the grid alone and are
reproduced exactly (tests/test_concept_powerspec.py).

Device work (nb_cic_deposit ×2, cuFFT, the interlace combine and the K6 shell-binning kernel) goes
through `Engine.power_spectrum_custom`; everything here is host geometry on integer k² — all modes
of equal integer k² = mx²+my²+mz² share a bin, so a bin is a range of integer k².

Bin edges.  CONCEPT's source is not part of the reference, and its rule for placing the edges is only
partly recoverable from five files of one grid size: the first three edges sit at 10^((j−½)/4)·k_min
(4 bins per decade below 4 k_min) and from ≈8.6 k_min upward consecutive edges follow the forward
recursion e_{j+1} = e_j·10^(1/b(e_j)), b(k) linear in log k between (4 k_min, 4) and (100 k_min, 40),
to within the ±0.3 % the integer lattice resolves; the two edges in between (6.40–6.48 and
8.54–8.60 k_min) fit neither.  Because bins are ranges of integer k² in units of k_min², the files
pin the bins of EVERY grid size below 32 k_min: `PINNED_EDGES` holds one representative per
admissible interval, and above the last pinned edge the recursion continues (grids > 64).
Non-default `bins_per_decade` use the recursion from 10^(−1/(2 b(k_min)))·k_min throughout.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

#: default `powerspec_options['bins per decade']` (Uncertain.param:50-55): {k / k_min: bins per decade}
DEFAULT_BINS_PER_DECADE = {4.0: 4.0, 100.0: 40.0}

#: Bin edges in units of k_min = 2π/L for the default options, one value inside each interval the
#: reference's five powerspec files allow (edge j lies between sqrt(last k² of bin j−1) and
#: sqrt(first k² of bin j); k² ranges 1, 2–5, 6–17, 18–41, 42–73, 74–106, 107–146, 147–193,
#: 194–250, 251–317, 318–395, 396–486, 488–594, 595–718, 720–861, 862–1022 reproduce the files'
#: modes column 3, 25, 124, 418, 730, 1012, …, 15506).
PINNED_EDGES = (
    10 ** (-1 / 8), 10 ** (1 / 8), 10 ** (3 / 8), 10 ** (5 / 8), 6.44, 8.57, 10.32, 12.10, 13.91, 15.827,
    17.82, 19.887, 22.07, 24.382, 26.815, 29.351)


def bins_per_decade_at(k_over_kmin, table=None):
    """b(k): linear in log10 k between the table's knots, constant outside (Uncertain.param:50-55)."""
    table = DEFAULT_BINS_PER_DECADE if table is None else table
    ks = np.array(sorted(table), dtype=np.float64)
    bs = np.array([table[k] for k in sorted(table)], dtype=np.float64)
    return float(np.interp(math.log10(k_over_kmin), np.log10(ks), bs))


def bin_edges(k_max_over_kmin, bins_per_decade=None):
    """Edges in units of k_min, the last one ≥ k_max (see the module docstring for what is pinned)."""
    default = bins_per_decade is None or dict(bins_per_decade) == DEFAULT_BINS_PER_DECADE
    if default:
        e = [x for x in PINNED_EDGES]
    else:
        e = [10 ** (-0.5 / bins_per_decade_at(1.0, bins_per_decade))]
    while e[-1] < k_max_over_kmin:
        e.append(e[-1] * 10 ** (1.0 / bins_per_decade_at(e[-1], bins_per_decade)))
    e = np.array(e)
    n = int(np.searchsorted(e, k_max_over_kmin, side="left"))     # first edge >= k_max closes the last bin
    return e[:n + 1]


def lattice_r3(k2_end):
    """r3[s] = number of integer triples (mx,my,mz) with mx²+my²+mz² = s, for 0 <= s < k2_end."""
    k2_end = int(k2_end)
    m = int(math.isqrt(max(k2_end - 1, 0)))
    sq = np.arange(-m, m + 1, dtype=np.int64) ** 2
    two = np.add.outer(sq, sq).ravel()
    r2 = np.bincount(two[two < k2_end], minlength=k2_end).astype(np.int64)
    r3 = np.zeros(k2_end, dtype=np.int64)
    for a in range(m + 1):
        s = a * a
        r3[s:] += (1 if a == 0 else 2) * r2[:k2_end - s]
    return r3


@dataclass
class BinGeometry:
    """Bins of the estimator for one grid: everything that does not depend on the particle data."""
    gridsize: int
    first_k2: np.ndarray          # first integer k² of each bin (bins are contiguous ranges)
    k2_end: int                   # exclusive upper bound (k < k_max)
    modes: np.ndarray             # half-space mode count per bin — the files' `modes` column
    k_grid: np.ndarray            # geometric mean of |k| over the bin's modes, units of k_min
    r3: np.ndarray = field(repr=False, default=None)
    bin_of_k2: np.ndarray = field(repr=False, default=None)   # integer k² -> bin index (−1: none)


def bin_geometry(gridsize, k_max="nyquist", bins_per_decade=None) -> BinGeometry:
    """Integer-k² ranges, `modes` and `k` columns for a grid (CONCEPT: modes counted over half of
    Fourier space, k = geometric mean of |k| over the modes of the bin — reproduces every k value of
    the reference's files to their 8 significant figures)."""
    N = int(gridsize)
    nyq = N // 2
    kmax = float(nyq) if k_max == "nyquist" else float(k_max)
    if not 1.0 < kmax <= nyq:
        raise ValueError("k_max must be 'nyquist' or a multiple of k_min in (1, gridsize/2]")
    k2_end = int(math.ceil(kmax * kmax - 1e-9))              # modes with k < k_max
    edges = bin_edges(kmax, bins_per_decade)
    r3 = lattice_r3(k2_end)
    s = np.arange(k2_end)
    occupied = s[(r3 > 0) & (s > 0)]
    which = np.searchsorted(edges, np.sqrt(occupied.astype(np.float64)), side="right") - 1
    ok = (which >= 0) & (which < edges.size - 1)
    occupied, which = occupied[ok], which[ok]
    used = np.unique(which)
    first = np.array([occupied[which == b][0] for b in used], dtype=np.int64)
    bin_of = np.full(k2_end, -1, dtype=np.int64)
    lut = {b: i for i, b in enumerate(used)}
    bin_of[occupied] = [lut[b] for b in which]
    nb = first.size
    w = r3[occupied].astype(np.float64)
    idx = bin_of[occupied]
    full = np.bincount(idx, weights=w, minlength=nb)
    lnk = np.bincount(idx, weights=w * 0.5 * np.log(occupied.astype(np.float64)), minlength=nb)
    modes = np.rint(full / 2).astype(np.int64)
    return BinGeometry(N, first, k2_end, modes, np.exp(lnk / full), r3, bin_of)


def k2_thresholds(geom: BinGeometry, boxsize):
    """Decision thresholds in physical k² for the K6 kernel: halfway below the first integer k² of
    every bin and halfway above the last one (the kernel forms k² from float64 (2π·fftfreq)² sums,
    which are within 1e-15 of k_min²·integer)."""
    kf2 = (2.0 * math.pi / float(boxsize)) ** 2
    last = int(np.nonzero(geom.bin_of_k2 >= 0)[0].max())
    return kf2 * (np.concatenate([geom.first_k2, [last + 1]]).astype(np.float64) - 0.5)


def _binned_mean(geom: BinGeometry, per_k2):
    """Mode-weighted mean over each bin of a quantity given per integer k²."""
    s = np.nonzero(geom.bin_of_k2 >= 0)[0]
    w = geom.r3[s].astype(np.float64)
    nb = geom.first_k2.size
    return np.bincount(geom.bin_of_k2[s], weights=w * per_k2[s], minlength=nb) / np.bincount(
        geom.bin_of_k2[s], weights=w, minlength=nb)


def linear_column(geom: BinGeometry, boxsize, linear):
    """`P_linear` column: mean of the linear spectrum `linear(k)` over the modes of each bin."""
    kf = 2.0 * math.pi / float(boxsize)
    s = np.arange(geom.k2_end)
    p = np.zeros(geom.k2_end)
    p[1:] = np.asarray(linear(kf * np.sqrt(s[1:].astype(np.float64))), dtype=np.float64)
    return _binned_mean(geom, p), p


def realization_ratio(geom: BinGeometry, noise_abs2, p_lin_k2):
    """Realisation correction factor per bin, Σ P_lin / Σ |R_k|² P_lin over the modes of the bin:
    `noise_abs2` is |R_k|² of the primordial noise on the full (n,n,n) FFT grid, n = gridsize
    (`powerspec_options['realization correction']`, Uncertain.param:46-48)."""
    n = geom.gridsize
    a2 = np.asarray(noise_abs2, dtype=np.float64)
    if a2.shape != (n, n, n):
        raise ValueError("noise_abs2 must live on the estimator grid")
    m = np.fft.fftfreq(n, 1.0 / n).astype(np.int64) ** 2
    k2 = (m[:, None, None] + m[None, :, None] + m[None, None, :]).ravel()
    keep = k2 < geom.k2_end
    tot = np.bincount(k2[keep], weights=a2.ravel()[keep], minlength=geom.k2_end)
    s = np.nonzero(geom.bin_of_k2 >= 0)[0]
    nb = geom.first_k2.size
    num = np.bincount(geom.bin_of_k2[s], weights=geom.r3[s] * p_lin_k2[s], minlength=nb)
    den = np.bincount(geom.bin_of_k2[s], weights=tot[s] * p_lin_k2[s], minlength=nb)
    return num / den


def sigma_tophat(k, P, R):
    """rms fluctuation in tophat spheres of radius R from a binned spectrum: cubic spline of ln P in
    ln k between the first and last bin, σ² = ∫ k³ P W²(kR) dln k / 2π².  Reproduces the σ₈ values in
    the headers of the reference's files from their own columns to ±0.002 (their quadrature and
    range extension are CONCEPT-internal)."""
    from scipy.interpolate import CubicSpline
    k, P = np.asarray(k, dtype=np.float64), np.asarray(P, dtype=np.float64)
    cs = CubicSpline(np.log(k), np.log(P))
    lk = np.linspace(math.log(k[0]), math.log(k[-1]), 4097)
    kk = np.exp(lk)
    x = kk * R
    W = 3.0 * (np.sin(x) - x * np.cos(x)) / x ** 3
    f = kk ** 3 * np.exp(cs(lk)) * W * W / (2.0 * math.pi ** 2)
    return math.sqrt(float(np.sum(0.5 * (f[1:] + f[:-1]) * np.diff(lk))))


@dataclass
class PowerSpec:
    k: np.ndarray
    modes: np.ndarray
    power: np.ndarray
    corrected: np.ndarray | None = None
    linear: np.ndarray | None = None
    sigma: tuple = ()
    gridsize: int = 0
    boxsize: float = 0.0
    a: float = 1.0
    t: float | None = None
    tophat: float | None = None

    def columns(self):
        cols = [self.power]
        if self.corrected is not None:
            cols.append(self.corrected)
        if self.linear is not None:
            cols.append(self.linear)
        return cols


def concept_powerspec(positions, boxsize, gridsize=None, *, interlace=True, deconvolve=True, k_max="nyquist",
                      bins_per_decade=None, linear=None, noise_abs2=None, tophat=None, a=1.0, t=None,
                      precision="f32", engine=None) -> PowerSpec:
    """Power spectrum of a particle set the way the reference's CONCEPT runs compute it
    (Uncertain.param:29-62).  positions: (Np,3) float64 numpy array or device tensor.
    gridsize: upstream grid, default cbrt(Np) (`'particles': 'cbrt(Ñ)'`).  linear: callable k -> P_lin
    for the `linear` column; with `noise_abs2` (|R_k|² of the realisation on the estimator grid) also
    the realisation-corrected column.  tophat: radius of the σ in Mpc (`'tophat': 8*Mpc/h`)."""
    import torch
    from .engine import default_engine
    eng = engine or default_engine()
    pos = positions if isinstance(positions, torch.Tensor) else eng.to_device(
        np.ascontiguousarray(positions, dtype=np.float64))
    npart = int(pos.shape[0])
    if gridsize is None:
        gridsize = round(npart ** (1.0 / 3.0))
    N, L = int(gridsize), float(boxsize)
    geom = bin_geometry(N, k_max, bins_per_decade)
    dtype = {"f32": torch.float32, "f64": torch.float64}[precision]
    cell = L / N
    g1 = eng.cic_deposit(pos, cell, N, dtype=dtype)
    g2 = eng.cic_deposit(pos, cell, N, dtype=dtype, shift=0.5 * cell) if interlace else None
    thr = k2_thresholds(geom, L)
    # delta_k = FFT(counts) / (Np/N³) for k != 0, P = |delta_k|² V / N⁶  =>  norm = V / Np²
    counts, sums = eng.power_spectrum_custom(g1, g2, L, N, thr, window_power=2 if deconvolve else 0,
                                             norm=L ** 3 / float(npart) ** 2)
    counts = counts.cpu().numpy()
    if not np.array_equal(counts, 2 * geom.modes):
        raise RuntimeError("device bin counts differ from the lattice count of the bins")
    power = sums.cpu().numpy() / counts
    kf = 2.0 * math.pi / L
    out = PowerSpec(kf * geom.k_grid, geom.modes, power, gridsize=N, boxsize=L, a=float(a), t=t, tophat=tophat)
    if linear is not None:
        out.linear, p_k2 = linear_column(geom, L, linear)
        if noise_abs2 is not None:
            out.corrected = power * realization_ratio(geom, noise_abs2, p_k2)
    if tophat is not None:
        out.sigma = tuple(sigma_tophat(out.k, c, float(tophat)) for c in out.columns())
    return out


# ------------------------------------------------------------------------------- text files
def format_rows(ps: PowerSpec):
    """Data rows in CONCEPT's layout (`'significant figures': 8`, Uncertain.param:59-61)."""
    cols = ps.columns()
    return ["%.7e%7d" % (ps.k[i], ps.modes[i]) + "".join("  %.7e" % c[i] for c in cols) for i in range(ps.k.size)]


def write_powerspec(path, ps: PowerSpec, job="nbody_b200"):
    """Text file with the header and column layout of the reference's `powerspec_a=*` files
    (PPPMP3M/P3M/powerspec/powerspec_a=1.00:1-11)."""
    cols = ps.columns()
    names = ["component 0", "(corrected)", "(linear)"]
    used = [names[0]] + ([names[1]] if ps.corrected is not None else []) + ([names[2]] if ps.linear is not None else [])
    a_txt = ("%.8g" % ps.a)
    t_txt = "" if ps.t is None else "t = %.6f Gyr, " % ps.t
    width = 20 + 15 * len(cols)
    lines = [f"# Power spectrum from {job} at {t_txt}a = {a_txt}", "#", "# Components:",
             f"#   0: matter  (upstream grid size {ps.gridsize})", "#",
             "#" + f"grid size {ps.gridsize}".center(width), "# ╭" + "─" * width + "╮",
             "#" + " " * 21 + "  ".join(n.center(13) for n in used)]
    if ps.sigma:
        lines.append("#" + " " * 21 + "  ".join(("σ₈ = %.2e" % s).center(13) for s in ps.sigma))
    lines.append("#  k [Mpc⁻¹]   modes  " + "  ".join("P [Mpc³]".center(13) for _ in cols))
    with open(path, "w", encoding="utf-8") as f:
        f.write("\n".join(lines + format_rows(ps)) + "\n")


def read_powerspec(path) -> PowerSpec:
    """Reads a CONCEPT `powerspec_a=*` file (or one written by write_powerspec)."""
    import re
    sig, a, t, grid = (), 1.0, None, 0
    with open(path, encoding="utf-8") as f:
        head = [ln for ln in f if ln.startswith("#")]
    for ln in head:
        m = re.search(r"a = ([0-9.eE+-]+)", ln)
        if m and "Power spectrum" in ln:
            a = float(m.group(1))
            mt = re.search(r"t = ([0-9.eE+-]+) Gyr", ln)
            t = float(mt.group(1)) if mt else None
        if "σ₈" in ln:
            sig = tuple(float(x) for x in re.findall(r"σ₈ = ([0-9.eE+-]+)", ln))
        m = re.search(r"upstream grid size (\d+)", ln)
        if m:
            grid = int(m.group(1))
    d = np.loadtxt(path, ndmin=2)
    ps = PowerSpec(d[:, 0], d[:, 1].astype(np.int64), d[:, 2], gridsize=grid, a=a, t=t, sigma=sig)
    if d.shape[1] >= 5:
        ps.corrected, ps.linear = d[:, 3], d[:, 4]
    elif d.shape[1] == 4:
        ps.linear = d[:, 3]
    return ps
