"""Host side of the time loop (nbody_b200/cosmo.py, timeloop.py): closed
forms of the background integrals, linear-theory sanity, and the leapfrog bookkeeping driven by
a numpy stand-in for the force step built on the float64 oracle (oracle/force.py)."""
import math

import numpy as np

from nbody_b200 import timeloop
from nbody_b200.cosmo import Cosmology, scale_factor_edges
from nbody_b200.params import ForceConfig, G_NEWTON
from oracle import force as oforce
from oracle import restate


class OracleStepper:
    """ForceStepper.step with the oracle's accelerations (numpy arrays, returns new arrays)."""

    def __init__(self, cfg):
        self.cfg = cfg

    def step(self, pos, mom, dt_kick, dt_drift):
        acc = oforce.accel(self.cfg, pos)
        return oforce.kick_drift(pos, mom, acc, self.cfg.boxsize, mass=self.cfg.mass, a=1.0,
                                 dt_kick=dt_kick, dt_drift=dt_drift)


def numpy_zeldovich(cosmo, L, n, a_begin, mass, seed, A_s=2.1e-9, n_s=0.9649, lattice_offset=0.0):
    """Numpy restatement of timeloop.zeldovich_ics (same draws, same conventions)."""
    pk = timeloop.linear_pk_grid(cosmo, L, n, A_s, n_s)
    rs = np.random.RandomState(seed)
    c1 = rs.normal(0, 1 / np.sqrt(2), size=(n, n, n))
    c2 = rs.normal(0, 1 / np.sqrt(2), size=(n, n, n))
    R = restate.hermitian_R(c1, c2, guard=True)
    _, psi = restate.grf_delta_psi(pk, R, L, n)
    psi = np.stack([p.ravel() for p in psi], axis=1)
    d_rel = cosmo.growth(a_begin) / cosmo.growth(1.0)
    q = restate.lattice(L, n) + lattice_offset * L / n
    pos = (q + d_rel * psi) % L
    mom = mass * a_begin ** 2 * float(cosmo.H(a_begin)) * cosmo.growth_rate(a_begin) * d_rel * psi
    return pos, mom, psi, q, (c1, c2)


def test_einstein_de_sitter_closed_forms():
    c = Cosmology(H0=0.07, Omega_b=0.2, Omega_cdm=0.8)
    for a in (0.02, 0.3, 1.0):
        assert abs(c.growth(a) / a - 1) < 1e-10 and abs(c.growth_rate(a) - 1) < 1e-10
    a0, a1 = 0.05, 0.4
    assert abs(c.kick_factor(a0, a1) / (2 * (math.sqrt(a1) - math.sqrt(a0)) / c.H0) - 1) < 1e-12
    assert abs(c.drift_factor(a0, a1) / (2 * (a0 ** -0.5 - a1 ** -0.5) / c.H0) - 1) < 1e-12
    assert abs(c.cosmic_time(1.0) / (2 / (3 * c.H0)) - 1) < 1e-10


def test_lcdm_growth_and_spectrum_sanity():
    c = Cosmology()                      # the reference snapshots' H0, Ωb, Ωcdm attributes
    assert abs(c.h - 0.6736) < 1e-6 and abs(c.Omega_m - 0.31377) < 1e-4
    assert abs(c.growth(0.02) / 0.02 - 1) < 1e-4                 # matter dominated at a_begin
    assert 0.77 < c.growth(1.0) < 0.80
    e = 1e-4
    f_num = (math.log(c.growth(1 + e)) - math.log(c.growth(1 - e))) / (math.log(1 + e) - math.log(1 - e))
    assert abs(c.growth_rate(1.0) - f_num) < 1e-6
    assert 13.5 < c.cosmic_time(1.0) < 14.1                      # Gyr
    assert 0.78 < c.sigma_R(A_s=2.1e-9, n_s=0.9649) < 0.88       # sigma_8 of the Giant.param cosmology
    k = np.array([0.0, 1e-3, 0.02, 1.0])
    p = c.linear_power(k)
    assert p[0] == 0 and p[2] > p[1] > 0 and p[2] > p[3]
    assert abs(c.linear_power(k[1:], a=0.5) / p[1:] - (c.growth(0.5) / c.growth(1.0)) ** 2).max() < 1e-12


def test_step_edges_and_factors_tile_the_interval():
    c = Cosmology()
    edges, idx = timeloop.edges_through(0.02, [1.0, 0.17, 0.01], dlna_max=0.05)
    assert edges[0] == 0.02 and abs(edges[idx[0]] - 0.17) < 1e-15 and abs(edges[idx[1]] - 1.0) < 1e-15
    assert np.all(np.diff(np.log(edges)) <= 0.05 + 1e-12) and np.all(np.diff(edges) > 0)
    K, D, Kc = timeloop.leapfrog_factors(c, edges)
    # kicks (with the last closing half-kick) and drifts each cover [a_begin, a_end] exactly once
    assert abs((K.sum() + Kc[-1]) / c.kick_factor(0.02, 1.0) - 1) < 1e-10
    assert abs(D.sum() / c.drift_factor(0.02, 1.0) - 1) < 1e-10
    assert len(scale_factor_edges(0.5, 1.0, 1.0)) == 2


def test_leapfrog_reproduces_linear_growth_with_the_oracle_force():
    """16^3 particles, PM on the cbrt(N) mesh of the reference's .param files, a = 0.02 -> 0.2: the
    low-k modes of the displacement field grow by D(0.2)/D(0.02) and the momenta follow
    a^2 H f times it.  (A mesh of 2 cbrt(N) cells puts the first harmonic of the particle lattice
    on the mesh Nyquist frequency, where the deconvolution amplifies it: the force on a perturbed
    lattice is then 3-7x the fluid value — measured with a plane-wave displacement — so lattice
    ICs want mesh = cbrt(N), which is what Giant.param:28-44 configures.  The lattice sits a quarter
    of a cell off the mesh nodes: ON the nodes every particle is at the kink of the CIC weight, the
    response to a displacement is |dx|-like instead of linear and the low modes come out 1-3 % off
    with mode-dependent phases — measured)"""
    c = Cosmology()
    L, n, N = 1000.0, 16, 16
    rho_m = 3 * c.H0 ** 2 / (8 * math.pi * G_NEWTON) * c.Omega_m
    mass = rho_m * L ** 3 / n ** 3
    cfg = ForceConfig(method="pm", npart=n ** 3, boxsize=L, gridsize=N, interlace=(True, True),
                      deconvolve=(True, True), G=G_NEWTON, mass=mass)
    a0, a1 = 0.02, 0.2
    # A_s 100x below Giant.param's keeps second-order mode coupling out of the comparison
    # quarter-cell offset: the cleanest linear response (see the docstring); at the product's default (particles
    # ON the mesh nodes, CONCEPT's placement by the full-run pin) the same low modes are 2.8 % off at this
    # tiny amplitude
    pos, mom, psi, q, _ = numpy_zeldovich(c, L, n, a0, mass, seed=5, A_s=2.1e-11, lattice_offset=0.25)
    outs = []
    p1, m1 = timeloop.evolve(OracleStepper(cfg), c, pos, mom, scale_factor_edges(a0, a1, 0.05),
                             on_output=lambda a, p, m: outs.append(a))
    assert outs == [a1]
    disp = (p1 - q + 0.5 * L) % L - 0.5 * L

    def low_modes(field):                       # |k| <= 2 k_f Fourier amplitudes of a lattice field
        fk = np.fft.fftn(field.reshape(n, n, n, 3), axes=(0, 1, 2))
        sel = [(1, 0, 0), (0, 1, 0), (0, 0, 1), (1, 1, 0), (0, 1, 1), (1, 0, 1), (1, 1, 1), (2, 0, 0), (0, 0, 2)]
        return np.array([fk[s] for s in sel])
    g = c.growth(a1) / c.growth(1.0)
    want = low_modes(psi) * g
    got = low_modes(disp)
    err = np.abs(got - want).max() / np.abs(want).max()
    assert err < 0.01, err
    want_m = low_modes(psi) * (mass * a1 ** 2 * float(c.H(a1)) * c.growth_rate(a1) * g)
    err_m = np.abs(low_modes(m1) - want_m).max() / np.abs(want_m).max()
    assert err_m < 0.01, err_m


def numpy_second_order(delta, L):
    """Numpy restatement of timeloop.second_order_displacement."""
    n = delta.shape[0]
    kv = restate.k_vals(n, L)
    kx, ky, kz = kv[:, None, None], kv[None, :, None], kv[None, None, :]
    k2 = kx * kx + ky * ky + kz * kz
    k2[0, 0, 0] = 1.0
    phi = -np.fft.fftn(delta) / k2
    phi[0, 0, 0] = 0.0
    d = lambda ka, kb: np.fft.ifftn(-(ka * kb) * phi).real      # noqa: E731
    pxx, pyy, pzz = d(kx, kx), d(ky, ky), d(kz, kz)
    src = pxx * pyy + pxx * pzz + pyy * pzz - d(kx, ky) ** 2 - d(kx, kz) ** 2 - d(ky, kz) ** 2
    phi2 = -np.fft.fftn(src) / k2
    phi2[0, 0, 0] = 0.0
    return np.stack([np.fft.ifftn(1j * ka * phi2).real.ravel() for ka in (kx, ky, kz)], axis=1)


def test_second_order_growth_factors():
    """D2 = -3/7 D1^2, f2 = 2 in Einstein-de Sitter (the displacement field itself runs on the device:
    tests/test_gpu_timeloop.py)."""
    eds = Cosmology(H0=0.07, Omega_b=0.2, Omega_cdm=0.8)
    for a in (0.02, 0.5):
        d2, f2 = timeloop.second_order_growth(eds, a)
        assert abs(d2 / (-3.0 / 7.0 * a * a) - 1) < 1e-9 and abs(f2 - 2.0) < 1e-9
    c = Cosmology()
    d2, f2 = timeloop.second_order_growth(c, 1.0)
    assert -3.0 / 7.0 * 1.02 < d2 < -3.0 / 7.0 and 1.0 < f2 < 1.1          # Ωm(1)^(-1/143) ≈ 1.008, 2·Ωm^(6/11) ≈ 1.06


def test_loop_argument_errors():
    import pytest
    c = Cosmology()
    with pytest.raises(ValueError):
        scale_factor_edges(0.5, 0.5)
    with pytest.raises(ValueError):
        timeloop.edges_through(0.5, [0.1, 0.5])                 # nothing after a_begin
    cfg = ForceConfig(method="pm", npart=8, boxsize=10.0, gridsize=4, a=0.5)
    with pytest.raises(ValueError):                             # the a-dependence lives in the integrals
        timeloop.evolve(OracleStepper(cfg), c, np.zeros((8, 3)), np.zeros((8, 3)), scale_factor_edges(0.1, 0.2))
    # outputs in the middle of the loop are synchronised copies: the loop state is not disturbed
    cfg = ForceConfig(method="pm", npart=64, boxsize=100.0, gridsize=4, G=G_NEWTON, mass=1e3)
    rs = np.random.RandomState(0)
    pos, mom = rs.random_sample((64, 3)) * 100.0, rs.normal(size=(64, 3))
    edges = scale_factor_edges(0.1, 0.2, 0.1)
    seen = []
    a = timeloop.evolve(OracleStepper(cfg), c, pos, mom, edges, output_at=[1, 3],
                        on_output=lambda s, p, m: seen.append(s))
    b = timeloop.evolve(OracleStepper(cfg), c, pos, mom, edges)
    assert len(seen) == 3 and np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


class SplitOracleStepper(OracleStepper):
    """numpy stand-in with ForceStepper's split interface (in-place updates like the device tensors)."""

    def accel_long(self, pos):
        c = self.cfg
        if c.method == "pp":
            return None
        return oforce.pm_accel(pos, c.boxsize, c.gridsize, G=c.G, mass=c.mass, r_s=c.r_s, deconv_order=c.deconv_order,
                               interlace=bool(c.interlace[0]), interlace_down=bool(c.interlace[1]))

    def accel_short(self, pos):
        c = self.cfg
        if c.method == "pm":
            return None
        if c.method == "pp":
            return oforce.pp_accel(pos, c.boxsize, c.G, c.mass, c.eps, True)
        return oforce.p3m_short_accel(pos, c.boxsize, c.r_s, c.r_cut, c.G, c.mass, c.eps)

    def kick(self, pos, mom, acc, k):
        mom += acc * k

    def kick_drift(self, pos, mom, acc, k, d):
        if acc is not None:
            mom += acc * k
        pos += mom * d
        pos %= self.cfg.boxsize

    def max_norm2(self, acc):
        return float((acc * acc).sum(1).max())


def test_substep_count_criterion():
    """n = smallest power of two with Δt_base/n <= sqrt(2 η ε a³ / |g|max): closed-form cases."""
    c = Cosmology()
    a0, a1, eps, eta = 0.5, 0.5 * math.exp(0.025), 0.2, 0.025
    dt_base = 0.025 / float(c.H(a0))
    g1 = 2 * eta * eps * a0 ** 3 / dt_base ** 2            # acceleration that allows exactly the base step
    assert timeloop.substep_count(c, a0, a1, 0.0, eps) == 1
    assert timeloop.substep_count(c, a0, a1, (0.99 * g1) ** 2, eps, eta) == 1
    assert timeloop.substep_count(c, a0, a1, (1.01 * g1) ** 2, eps, eta) == 2
    assert timeloop.substep_count(c, a0, a1, (15.0 * g1) ** 2, eps, eta) == 4       # sqrt(15) < 4
    assert timeloop.substep_count(c, a0, a1, (17.0 * g1) ** 2, eps, eta) == 8
    assert timeloop.substep_count(c, a0, a1, (1e9 * g1) ** 2, eps, eta, max_sub=16) == 16
    import pytest
    with pytest.raises(FloatingPointError):
        timeloop.substep_count(c, a0, a1, float("inf"), eps)


def _clustered_set(n, L, seed):
    rs = np.random.RandomState(seed)
    centres = rs.random_sample((6, 3)) * L
    pos = centres[rs.randint(0, 6, n)] + rs.normal(scale=0.05 * L, size=(n, 3))
    return pos % L


def test_subcycled_loop_equals_plain_loop_without_substeps_and_converges_with_them():
    """(1) max_sub = 1 reproduces `evolve` exactly for P3M (same kicks, same drifts); a PM run never
    sub-cycles.  (2) On a clustered set the sub-cycled loop is closer to a much finer uniform run than the
    plain loop with the same base steps (reference: 64x finer uniform steps), and it takes more than one sub-step where the short-range
    acceleration demands it."""
    c = Cosmology()
    L, n, N = 60.0, 216, 8
    rho_m = 3 * c.H0 ** 2 / (8 * math.pi * G_NEWTON) * c.Omega_m
    mass = rho_m * L ** 3 / n * 3.0                       # overdense: short dynamical times
    pos = _clustered_set(n, L, 2)
    mom = np.zeros_like(pos)
    for method in ("p3m", "pm"):
        cfg = ForceConfig(method=method, npart=n, boxsize=L, gridsize=N, interlace=(True, True), deconvolve=(True, True),
                          G=G_NEWTON, mass=mass, softening_frac=0.05)
        edges = scale_factor_edges(0.5, 0.6, 0.03)
        p0, m0 = timeloop.evolve(OracleStepper(cfg), c, pos.copy(), mom.copy(), edges)
        stats = []
        p1, m1 = timeloop.evolve_subcycled(SplitOracleStepper(cfg), c, pos.copy(), mom.copy(), edges, max_sub=1,
                                           stats=stats)
        assert stats == [1] * (len(edges) - 1)
        assert np.abs((p1 - p0 + L / 2) % L - L / 2).max() < 1e-9 * L and np.abs(m1 - m0).max() < 1e-9 * np.abs(m0).max()
    cfg = ForceConfig(method="p3m", npart=n, boxsize=L, gridsize=N, interlace=(True, True), deconvolve=(True, True),
                      G=G_NEWTON, mass=mass, softening_frac=0.05)
    edges = scale_factor_edges(0.5, 0.6, 0.03)
    fine = scale_factor_edges(0.5, 0.6, 0.03 / 64)
    pf, mf = timeloop.evolve(OracleStepper(cfg), c, pos.copy(), mom.copy(), fine)
    p0, m0 = timeloop.evolve(OracleStepper(cfg), c, pos.copy(), mom.copy(), edges)
    stats = []
    ps, ms = timeloop.evolve_subcycled(SplitOracleStepper(cfg), c, pos.copy(), mom.copy(), edges, eta=0.025, max_sub=32,
                                       stats=stats)
    assert max(stats) > 1 and all(s & (s - 1) == 0 for s in stats)
    err_plain = np.abs(m0 - mf).max()
    err_sub = np.abs(ms - mf).max()
    assert err_sub < 0.5 * err_plain, (err_sub, err_plain, stats)
