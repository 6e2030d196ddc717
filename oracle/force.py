"""TEST INFRASTRUCTURE ONLY — float64 CPU restatement of the PP / PM / P3M force step.

PARITY UNPINNED: the reference has no source for this path.  The force solver is the
external code CONCEPT (github.com/jmd-dk/concept, version unpinned — SURVEY §8c), called
as a subprocess at TheFeedingPipeline/NeuNetværk_test.py:155-171 and SnapVsPK/Ucertain.py:
89-97.  What follows is OUR specification (SURVEY §8a'), assembled from
  * the reference's own primitives: cic_density (FraP_kTilDelta.py:93-123), cic_window_ft
    (:73-84), the k grid 2π·fftfreq (CiC/3D startbetingelser1.py:69-72) and the
    i·k/k² operator (:105-107);
  * the reference's .param choices: Fourier differentiation, deconvolve (True, True),
    interlace (True, True) (Different_Sizes/Giant.param:45-58, NN.param:55-61), P3M with
    r_s = 1.25 cells and r_cut = 4.5 r_s (corroborated by the 5x5x5 tiling in the logs);
  * the standard erfc / Gaussian force split and Plummer softening.
The CUDA kernels are checked against this file; this file is checked against analytic
properties in tests/test_force_oracle.py (Newton's third law, 1/r² at long range,
PM + short-range = softened Newton, interlacing removes odd aliases).  The periodic PP force
(`ewald_force`, `ewald_table`, `pp_accel_ewald`: Ewald summation after Hernquist, Bouchet & Suto
1991, the method CONCEPT documents for its PP gravity) is checked in tests/test_ewald_cpu.py
(independence of the splitting parameter, periodicity, symmetry zeros, small-separation limit).
"""
from __future__ import annotations

import math

import numpy as np

from . import restate


# ------------------------------------------------------------------------------- PM
def pm_kernel_tables(N, L, r_s=0.0, deconv_order=2):
    """Per-axis tables: signed k (Nyquist zeroed for the gradient), k², and the separable
    factor t = exp(-k² r_s²) / s^(2·deconv_order) with s the reference's sinc factor."""
    kv = restate.k_vals(N, L)
    kgrad = kv.copy()
    if N % 2 == 0:
        kgrad[N // 2] = 0.0            # a real field has no gradient at the Nyquist plane
    s = restate.sinc_table(N, L)
    t = np.exp(-(kv ** 2) * r_s ** 2) / s ** (2 * deconv_order)
    return kv, kgrad, t


def pm_accel_mesh(pos, L, N, G=1.0, mass=1.0, r_s=0.0, deconv_order=2, interlace=False,
                  interlace_down=False):
    """Long-range acceleration on the mesh, 3 x (N,N,N) float64.
      rho      = mass * cic_density / cell³                        (a1)
      g_k      = 4πG · (i k / k²) · rho_k · exp(-k² r_s²) / W^deconv_order,  g_0 = 0
    (φ_k = -4πG ρ_k / k², g = -∇φ; i·k/k² is the reference's ψ operator.)
    interlace: second deposit of positions shifted by +cell/2 in x,y,z, phase-corrected by
    exp(+i k·s), averaged (upstream interlacing).
    interlace_down: additionally returns the force sampled on the shifted mesh,
    g2_k = g_k·exp(-i k·s) (downstream interlacing: gather from both, average)."""
    cell = L / N
    kv, kgrad, t = pm_kernel_tables(N, L, r_s, deconv_order)
    nzh = N // 2 + 1
    rho_k = np.fft.rfftn(restate.cic_density(pos, L, N))
    m = np.fft.fftfreq(N, d=1.0 / N)                          # signed integer frequencies
    ph = np.exp(1j * np.pi * m / N)                            # exp(+i k_a cell/2)
    if N % 2 == 0:
        ph[N // 2] = 0.0          # Hermitian part of exp(±iπ/2): keeps the combined mesh real
    phase = ph[:, None, None] * ph[None, :, None] * ph[None, None, :nzh]
    if interlace:
        rho2 = np.fft.rfftn(restate.cic_density(pos + 0.5 * cell, L, N))
        rho_k = 0.5 * (rho_k + rho2 * phase)
    k2 = (kv ** 2)[:, None, None] + (kv ** 2)[None, :, None] + (kv ** 2)[None, None, :nzh]
    k2[0, 0, 0] = 1.0
    filt = (t[:, None, None] * t[None, :, None] * t[None, None, :nzh]) / k2
    filt[0, 0, 0] = 0.0
    pref = 4.0 * math.pi * G * mass / cell ** 3
    base = pref * filt * rho_k
    gx = np.fft.irfftn(1j * kgrad[:, None, None] * base, s=(N, N, N), axes=(0, 1, 2))
    gy = np.fft.irfftn(1j * kgrad[None, :, None] * base, s=(N, N, N), axes=(0, 1, 2))
    gz = np.fft.irfftn(1j * kgrad[None, None, :nzh] * base, s=(N, N, N), axes=(0, 1, 2))
    if not interlace_down:
        return gx, gy, gz
    b2 = base * np.conj(phase)
    hx = np.fft.irfftn(1j * kgrad[:, None, None] * b2, s=(N, N, N), axes=(0, 1, 2))
    hy = np.fft.irfftn(1j * kgrad[None, :, None] * b2, s=(N, N, N), axes=(0, 1, 2))
    hz = np.fft.irfftn(1j * kgrad[None, None, :nzh] * b2, s=(N, N, N), axes=(0, 1, 2))
    return (gx, gy, gz), (hx, hy, hz)


def pm_accel(pos, L, N, interlace_down=False, **kw):
    """Long-range acceleration at the particles: CIC gather (adjoint of the deposit); with
    interlace_down the average of the gathers from the plain and the half-cell-shifted mesh."""
    if not interlace_down:
        g = pm_accel_mesh(pos, L, N, **kw)
        return np.column_stack([restate.trilinear_gather(c, pos, L, N) for c in g])
    g, h = pm_accel_mesh(pos, L, N, interlace_down=True, **kw)
    a1 = np.column_stack([restate.trilinear_gather(c, pos, L, N) for c in g])
    a2 = np.column_stack([restate.trilinear_gather(c, pos + 0.5 * L / N, L, N) for c in h])
    return 0.5 * (a1 + a2)


def accel(cfg, pos, periodic_pp=False, i_subset=None):
    """Acceleration under a ForceConfig-like object (fields method, boxsize, gridsize, G,
    mass, eps, r_s, r_cut, interlace, deconv_order) — mirrors nbody_b200.force."""
    if cfg.method == "pp":
        return pp_accel(pos, cfg.boxsize, cfg.G, cfg.mass, cfg.eps, periodic_pp, i_subset)
    a = pm_accel(pos, cfg.boxsize, cfg.gridsize, G=cfg.G, mass=cfg.mass, r_s=cfg.r_s,
                 deconv_order=cfg.deconv_order, interlace=bool(cfg.interlace[0]),
                 interlace_down=bool(cfg.interlace[1]))
    if i_subset is not None:
        a = a[np.asarray(i_subset)]
    if cfg.method == "p3m":
        a = a + p3m_short_accel(pos, cfg.boxsize, cfg.r_s, cfg.r_cut, cfg.G, cfg.mass, cfg.eps, i_subset)
    return a


# ------------------------------------------------------------------------------- pair forces
def split_factor(r, r_s):
    """Short-range fraction of the Newtonian force: erfc(u) + 2u/sqrt(pi) exp(-u²), u = r/(2 r_s)."""
    from scipy.special import erfc
    u = r / (2.0 * r_s)
    return erfc(u) + 2.0 * u / math.sqrt(math.pi) * np.exp(-u * u)


def pp_accel(pos, L=None, G=1.0, mass=1.0, eps=0.0, periodic=False, i_subset=None, chunk=256):
    """Softened direct sum: a_i = -G m Σ_j r_ij / (r² + ε²)^{3/2}, r_ij = x_i - x_j
    (minimum image if periodic).  i_subset: indices of the i-particles to evaluate."""
    idx = np.arange(pos.shape[0]) if i_subset is None else np.asarray(i_subset)
    out = np.zeros((idx.size, 3))
    for s in range(0, idx.size, chunk):
        ii = idx[s:s + chunk]
        d = pos[ii, None, :] - pos[None, :, :]
        if periodic:
            d -= L * np.round(d / L)
        r2 = (d * d).sum(-1) + eps * eps
        r2[r2 == 0.0] = np.inf
        w = r2 ** -1.5
        out[s:s + chunk] = -G * mass * (d * w[..., None]).sum(1)
    return out


def ewald_force(d, L, alpha_box=2.0, nreal=4, nfour=5, subtract_nearest=False):
    """Periodic force per unit G·m on a particle from a source at separation d (M,3) (pointing to
    the source), Ewald sum of Hernquist, Bouchet & Suto 1991:
      Σ_n (d-nL)/|d-nL|³ [erfc(α|d-nL|) + 2α|d-nL|/√π exp(-α²|d-nL|²)]
      + 4π/L³ Σ_{h≠0} h/h² exp(-h²/4α²) sin(h·d).
    subtract_nearest: minus d/|d|³ of the n = 0 term (the correction the tables hold)."""
    from scipy.special import erfc
    d = np.atleast_2d(np.asarray(d, dtype=np.float64))
    alpha = alpha_box / L
    out = np.zeros_like(d)
    rng = np.arange(-nreal, nreal + 1)
    for nx in rng:
        for ny in rng:
            for nz in rng:
                r = d - L * np.array([nx, ny, nz], dtype=np.float64)
                r2 = (r * r).sum(-1)
                ok = r2 > 0
                rr = np.sqrt(np.where(ok, r2, 1.0))
                w = (erfc(alpha * rr) + 2 * alpha * rr / math.sqrt(math.pi) * np.exp(-alpha * alpha * rr * rr)) / rr ** 3
                if subtract_nearest and nx == 0 and ny == 0 and nz == 0:
                    w = w - 1.0 / rr ** 3
                out += np.where(ok, w, 0.0)[:, None] * r
    mrng = np.arange(-nfour, nfour + 1)
    for mx in mrng:
        for my in mrng:
            for mz in mrng:
                if mx == 0 and my == 0 and mz == 0:
                    continue
                h = 2 * math.pi / L * np.array([mx, my, mz], dtype=np.float64)
                h2 = (h * h).sum()
                out += (4 * math.pi / L ** 3 * math.exp(-h2 / (4 * alpha * alpha)) / h2 * np.sin(d @ h))[:, None] * h
    return out


def ewald_table(L, g, alpha_box=2.0, nreal=4, nfour=5):
    """(g+1,g+1,g+1,3) correction table on the octant [0, L/2]³.  The component odd in a coordinate
    is exactly zero on that coordinate's plane through the origin (NOT on the far face d = L/2: the
    periodic force vanishes there, so the correction equals minus the nearest-image term)."""
    x = np.arange(g + 1) * (0.5 * L / g)
    X, Y, Z = np.meshgrid(x, x, x, indexing="ij")
    d = np.stack([X.ravel(), Y.ravel(), Z.ravel()], axis=1)
    t = ewald_force(d, L, alpha_box, nreal, nfour, subtract_nearest=True).reshape(g + 1, g + 1, g + 1, 3)
    t[0, :, :, 0] = 0.0
    t[:, 0, :, 1] = 0.0
    t[:, :, 0, 2] = 0.0
    return t


def ewald_correction(d, L, table):
    """Trilinear interpolation of the octant table at nearest-image separations d (...,3): the value
    the pair kernel adds to the softened nearest-image force."""
    g = table.shape[0] - 1
    u = np.minimum(np.abs(d) * (2.0 * g / L), g)
    c = np.minimum(u.astype(np.int64), g - 1)
    t = u - c
    cx, cy, cz = c[..., 0], c[..., 1], c[..., 2]
    tx, ty, tz = t[..., 0:1], t[..., 1:2], t[..., 2:3]

    def at(ox, oy, oz):
        return table[cx + ox, cy + oy, cz + oz]
    c00 = at(0, 0, 0) + tz * (at(0, 0, 1) - at(0, 0, 0))
    c01 = at(0, 1, 0) + tz * (at(0, 1, 1) - at(0, 1, 0))
    c10 = at(1, 0, 0) + tz * (at(1, 0, 1) - at(1, 0, 0))
    c11 = at(1, 1, 0) + tz * (at(1, 1, 1) - at(1, 1, 0))
    c0 = c00 + ty * (c01 - c00)
    c1 = c10 + ty * (c11 - c10)
    return (c0 + tx * (c1 - c0)) * np.where(d < 0, -1.0, 1.0)


def pp_accel_ewald(pos, L, table, G=1.0, mass=1.0, eps=0.0, i_subset=None, chunk=64):
    """Periodic direct sum: softened nearest-image force + interpolated Ewald correction."""
    idx = np.arange(pos.shape[0]) if i_subset is None else np.asarray(i_subset)
    out = np.zeros((idx.size, 3))
    for s in range(0, idx.size, chunk):
        ii = idx[s:s + chunk]
        d = pos[None, :, :] - pos[ii, None, :]                 # towards the source
        d -= L * np.round(d / L)
        r2 = (d * d).sum(-1) + eps * eps
        r2[r2 == 0.0] = np.inf
        f = d * (r2 ** -1.5)[..., None] + ewald_correction(d, L, table)
        out[s:s + chunk] = G * mass * f.sum(1)
    return out


def p3m_short_accel(pos, L, r_s, r_cut, G=1.0, mass=1.0, eps=0.0, i_subset=None, chunk=128):
    """Short-range pair acceleration, minimum image, r < r_cut:
       a_i = -G m Σ_j r_ij / (r²+ε²)^{3/2} · split_factor(r)."""
    idx = np.arange(pos.shape[0]) if i_subset is None else np.asarray(i_subset)
    out = np.zeros((idx.size, 3))
    for s in range(0, idx.size, chunk):
        ii = idx[s:s + chunk]
        d = pos[ii, None, :] - pos[None, :, :]
        d -= L * np.round(d / L)
        r2 = (d * d).sum(-1)
        inr = (r2 < r_cut * r_cut) & (r2 > 0)
        r = np.sqrt(np.where(inr, r2, 1.0))
        w = np.where(inr, split_factor(r, r_s) * (np.where(inr, r2, 1.0) + eps * eps) ** -1.5, 0.0)
        out[s:s + chunk] = -G * mass * (d * w[..., None]).sum(1)
    return out


# ------------------------------------------------------------------------------- kick / drift
def kick_drift(pos, mom, acc, L, mass=1.0, a=1.0, dt_kick=0.0, dt_drift=0.0):
    """mom += m·g·Δt_kick ; pos += mom/(m a²)·Δt_drift ; pos %= L (Power_compare.py:391)."""
    mom = mom + mass * acc * dt_kick
    pos = pos + mom * (dt_drift / (mass * a * a))
    pos %= L
    return pos, mom
