"""TEST INFRASTRUCTURE ONLY — the CPU baseline of the PP / PM / P3M step: the same float64
arithmetic as oracle/force.py, with the particle loops in C (oracle/oracle_c.c, OpenMP where
the loop is parallel; the deposit stays serial like the reference's numba loop) and the
FFTs in scipy.fft with `workers` threads.  Timed by bench.py (`cpu_baseline`,
`--impl reference`); checked against oracle/force.py in tests/test_oracle_c.py.
"""
from __future__ import annotations

import math

import numpy as np
import scipy.fft as sfft

from . import cport, force, restate


def pm_accel(cfg, pos, workers=-1):
    L, N = cfg.boxsize, cfg.gridsize
    cell = L / N
    nzh = N // 2 + 1
    kv, kgrad, t = force.pm_kernel_tables(N, L, cfg.r_s, cfg.deconv_order)
    rho_k = sfft.rfftn(cport.cic_density(pos, L, N), workers=workers)
    m = np.fft.fftfreq(N, d=1.0 / N)
    ph = np.exp(1j * np.pi * m / N)
    if N % 2 == 0:
        ph[N // 2] = 0.0
    phase = ph[:, None, None] * ph[None, :, None] * ph[None, None, :nzh]
    if cfg.interlace[0]:
        rho2 = sfft.rfftn(cport.cic_density(pos, L, N, shift=0.5 * cell), workers=workers)
        rho_k = 0.5 * (rho_k + rho2 * phase)
    k2 = (kv ** 2)[:, None, None] + (kv ** 2)[None, :, None] + (kv ** 2)[None, None, :nzh]
    k2[0, 0, 0] = 1.0
    filt = (t[:, None, None] * t[None, :, None] * t[None, None, :nzh]) / k2
    filt[0, 0, 0] = 0.0
    base = (4.0 * math.pi * cfg.G * cfg.mass / cell ** 3) * filt * rho_k
    ks = (kgrad[:, None, None], kgrad[None, :, None], kgrad[None, None, :nzh])
    acc = np.empty_like(pos)
    for c in range(3):
        g = sfft.irfftn(1j * ks[c] * base, s=(N, N, N), axes=(0, 1, 2), workers=workers)
        a = cport.cic_gather(g, pos, L, N)
        if cfg.interlace[1]:
            h = sfft.irfftn(1j * ks[c] * base * np.conj(phase), s=(N, N, N), axes=(0, 1, 2), workers=workers)
            a = 0.5 * (a + cport.cic_gather(h, pos, L, N, shift=0.5 * cell))
        acc[:, c] = a
    return acc


def accel(cfg, pos, workers=-1):
    if cfg.method == "pp":
        return cport.pp_accel(pos, cfg.G, cfg.mass, cfg.eps, 0.0)
    a = pm_accel(cfg, pos, workers)
    if cfg.method == "p3m":
        a = a + cport.p3m_short_accel(pos, cfg.boxsize, cfg.r_s, cfg.r_cut, cfg.G, cfg.mass, cfg.eps)
    return a


def step(cfg, pos, mom, dt_kick, dt_drift, workers=-1):
    acc = accel(cfg, pos, workers)
    return force.kick_drift(pos, mom, acc, cfg.boxsize, cfg.mass, cfg.a, dt_kick, dt_drift)


def snapshot_pk(pos, L, Nmesh, workers=1):
    """compute_power_spectrum_from_snapshot body (Power_compare.py:150-200) with the C
    deposit and scipy FFT — the BASELINE config-2 CPU baseline."""
    delta = restate.overdensity_cellvol(cport.cic_density(pos, L, Nmesh), pos.shape[0], L, Nmesh)
    dk = sfft.fftn(delta, workers=workers)
    Pk = np.abs(dk) ** 2 * L ** 3 / Nmesh ** 6
    e = restate.pk_edges(Nmesh, L)
    c, s = cport.shell_bin(Pk, Nmesh, restate.k_vals(Nmesh, L), e, True)
    v = c > 0
    return 0.5 * (e[:-1] + e[1:])[v], s[v] / c[v]
