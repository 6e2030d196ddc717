"""Time loop on the device (nbody_b200/timeloop.py): initial conditions and
leapfrog steps against the numpy / oracle loop of test_timeloop_cpu.py, and a `.param`-driven run
from a_begin = 0.02 to a = 1 checked through linear growth of the low-k power and through the
snapshot files it writes."""
import math
import os

import numpy as np
import pytest

from test_timeloop_cpu import OracleStepper, numpy_second_order, numpy_zeldovich

pytestmark = pytest.mark.gpu


def _physical_cfg(method, n, N, L, cosmo):
    from nbody_b200.params import ForceConfig, G_NEWTON
    rho_m = 3 * cosmo.H0 ** 2 / (8 * math.pi * G_NEWTON) * cosmo.Omega_m
    return ForceConfig(method=method, npart=n ** 3, boxsize=L, gridsize=N, interlace=(True, True),
                       deconvolve=(True, True), G=G_NEWTON, mass=rho_m * L ** 3 / n ** 3)


def _pdiff(a, b, L):
    d = np.abs(a - b)
    return np.minimum(d, L - d).max()


def test_zeldovich_ics_match_the_numpy_restatement():
    from nbody_b200 import timeloop
    from nbody_b200.cosmo import Cosmology
    c = Cosmology()
    L, n, a0 = 1000.0, 16, 0.02
    cfg = _physical_cfg("pm", n, n, L, c)
    pos_ref, mom_ref, psi, _q, (c1, c2) = numpy_zeldovich(c, L, n, a0, cfg.mass, seed=8)
    pos, mom = timeloop.zeldovich_ics(c, L, n, a0, cfg.mass, c1=c1, c2=c2)
    assert _pdiff(pos.cpu().numpy(), pos_ref, L) <= 1e-10 * L
    assert np.abs(mom.cpu().numpy() - mom_ref).max() <= 1e-10 * np.abs(mom_ref).max()
    # the default is the reference's own lattice (no offset); any other placement is a shift of it
    pos0, _ = timeloop.zeldovich_ics(c, L, n, a0, cfg.mass, c1=c1, c2=c2, lattice_offset=0.25)
    assert _pdiff(pos0.cpu().numpy(), (pos_ref + 0.25 * L / n) % L, L) <= 1e-10 * L


def test_second_order_ics_match_the_numpy_restatement():
    from nbody_b200 import timeloop
    from nbody_b200.cosmo import Cosmology
    from oracle import restate
    c = Cosmology()
    L, n, a0 = 1000.0, 16, 0.1
    cfg = _physical_cfg("pm", n, n, L, c)
    pos1, mom1, _psi, _q, (c1, c2) = numpy_zeldovich(c, L, n, a0, cfg.mass, seed=21)
    delta, _ = restate.grf_delta_psi(timeloop.linear_pk_grid(c, L, n, 2.1e-9, 0.9649),
                                     restate.hermitian_R(c1, c2, guard=True), L, n)
    psi2 = numpy_second_order(delta, L)
    d2, f2 = timeloop.second_order_growth(c, a0)
    pos_ref = (pos1 + d2 * psi2) % L
    mom_ref = mom1 + cfg.mass * a0 ** 2 * float(c.H(a0)) * f2 * d2 * psi2
    pos, mom = timeloop.zeldovich_ics(c, L, n, a0, cfg.mass, c1=c1, c2=c2, lpt=2)
    assert np.abs(d2 * psi2).max() > 1e-4                      # the term is there
    assert _pdiff(pos.cpu().numpy(), pos_ref, L) <= 1e-10 * L
    assert np.abs(mom.cpu().numpy() - mom_ref).max() <= 1e-10 * np.abs(mom_ref).max()
    with pytest.raises(ValueError):
        timeloop.zeldovich_ics(c, L, n, a0, cfg.mass, c1=c1, c2=c2, lpt=3)


@pytest.mark.parametrize("method,n,steps", [("pm", 16, 12), ("p3m", 20, 6)])
def test_leapfrog_steps_match_the_oracle_loop(method, n, steps):
    from nbody_b200 import force as F
    from nbody_b200 import timeloop
    from nbody_b200.cosmo import Cosmology, scale_factor_edges
    from nbody_b200.engine import default_engine
    c = Cosmology()
    L, a0 = 1000.0, 0.02
    cfg = _physical_cfg(method, n, n, L, c)
    pos0, mom0, _psi, _q, _ = numpy_zeldovich(c, L, n, a0, cfg.mass, seed=12)
    edges = scale_factor_edges(a0, a0 * math.exp(0.05 * steps), 0.05)
    assert edges.size == steps + 1
    outs = []
    p_ref, m_ref = timeloop.evolve(OracleStepper(cfg), c, pos0, mom0, edges, output_at=[steps // 2],
                                   on_output=lambda a, p, m: outs.append((a, p, m)))
    eng = default_engine()
    got = []
    p, m = timeloop.evolve(F.ForceStepper(cfg, eng), c, eng.to_device(pos0), eng.to_device(mom0), edges,
                           output_at=[steps // 2],
                           on_output=lambda a, p, m: got.append((a, p.cpu().numpy(), m.cpu().numpy())))
    assert [g[0] for g in got] == [o[0] for o in outs] and len(got) == 2
    for (_, pg, mg), (_, pr, mr) in zip(got, outs):
        assert _pdiff(pg, pr, L) <= 1e-5 * cfg.cell
        assert np.abs(mg - mr).max() <= 2e-5 * np.abs(mr).max()
    assert _pdiff(p.cpu().numpy(), p_ref, L) <= 1e-5 * cfg.cell
    assert np.abs(m.cpu().numpy() - m_ref).max() <= 2e-5 * np.abs(m_ref).max()


PARAM = '''
_size = 32
initial_conditions = {'species': 'matter', 'N': _size**3}
output_times = {'snapshot': [0.5, 1.0]}
boxsize = 1000*Mpc
potential_options = {
    'gridsize': {'global': {_size}, 'particles': {'gravity': {'pm': ['cbrt(N)', 'cbrt(N)']}}},
    'interlace': {'gravity': {'pm': (True, True)}},
    'deconvolve': {'gravity': {'pm': (True, True)}},
}
select_forces = {'matter': {'gravity': 'pm'}}
realization_options = {'LPT': 1}
H0 = 67.36*km/(s*Mpc)
Ωb = 0.02237/h**2
Ωcdm = 0.12/h**2
a_begin = 0.02
primordial_spectrum = {'A_s': 2.1e-9, 'n_s': 0.9649}
'''


def test_param_driven_run_grows_linearly_at_low_k_and_writes_snapshots(tmp_path):
    """Giant.param-style run at 32^3: outputs at a = 0.5 and 1, both written as snapshot files;
    the power below k = 0.03 / Mpc grows by (D(1)/D(0.5))^2 and sits on the linear spectrum the
    ICs were drawn from (same realisation at both times, so cosmic variance cancels in the ratio)."""
    from nbody_b200 import api, snapshot, timeloop
    from nbody_b200.cosmo import Cosmology
    out_dir = str(tmp_path / "snapshots")
    res = timeloop.simulate(PARAM, seed=3, output_dir=out_dir, dlna_max=0.025)
    assert [round(a, 12) for a, _p, _m in res] == [0.5, 1.0]
    L, n = 1000.0, 32
    c = Cosmology(H0=0.06888989143747800, Omega_b=0.02237 / 0.6736 ** 2, Omega_cdm=0.12 / 0.6736 ** 2)
    spectra = []
    for a, pos, mom in res:
        assert pos.shape == (n ** 3, 3) and np.isfinite(pos).all() and np.isfinite(mom).all()
        assert pos.min() >= 0 and pos.max() < L
        assert np.abs(mom.mean(axis=0)).max() <= 1e-6 * np.abs(mom).max()        # momentum conserved
        f = os.path.join(out_dir, f"snapshot_a={a:.2f}.hdf5")
        assert np.array_equal(snapshot.read_positions(f), pos) and np.array_equal(snapshot.read_momenta(f), mom)
        assert snapshot.read_boxsize(f) == L and snapshot.read_attrs(f)["a"] == a
        k, p = api.power_spectrum_from_positions(pos, L, n, deconvolve=True)
        spectra.append((k, p))
    (k, p_half), (_, p_one) = spectra
    low = k < 0.03
    assert low.sum() >= 8
    growth = (c.growth(1.0) / c.growth(0.5)) ** 2
    ratio = p_one[low] / p_half[low]
    print("P(a=1)/P(a=0.5) over linear growth:", np.round(ratio / growth, 3))
    # particles start ON the mesh nodes (CONCEPT's placement, the one the reference's powerspec files are
    # matched by): the lowest modes then grow 2-4 % less than linear theory in amplitude (median 0.94 in power,
    # measured); the quarter-cell placement is within 5 % but does not reproduce the reference's runs
    assert np.abs(np.median(ratio) / growth - 1) < 0.09 and np.abs(ratio / growth - 1).max() < 0.20
    res_q = timeloop.simulate(PARAM, seed=3, dlna_max=0.025, lattice_offset=0.25)
    pq = [api.power_spectrum_from_positions(pos, L, n, deconvolve=True)[1] for _a, pos, _m in res_q]
    assert np.abs(np.median(pq[1][low] / pq[0][low]) / growth - 1) < 0.05
    # against the linear spectrum itself: band-averaged over the low bins (few modes per bin)
    lin = c.linear_power(k[low])
    assert abs(np.mean(p_one[low] / lin) - 1) < 0.35


def test_periodic_pp_steps_use_the_ewald_force():
    """`pp` in the comoving loop is periodic: timeloop.PeriodicPP steps with nearest image + Ewald
    table; three steps against the same loop on the float64 restatement (table of 16 intervals)."""
    from nbody_b200 import force as F
    from nbody_b200 import timeloop
    from nbody_b200.cosmo import Cosmology, scale_factor_edges
    from nbody_b200.engine import default_engine
    from oracle import force as oforce
    c = Cosmology()
    L, n, a0 = 400.0, 6, 0.05
    cfg = _physical_cfg("pp", n, 0, L, c)
    pos0, mom0, _psi, _q, _ = numpy_zeldovich(c, L, n, a0, cfg.mass, seed=4)
    table = oforce.ewald_table(L, 16)

    class OracleEwald:
        def __init__(self):
            self.cfg = cfg

        def step(self, pos, mom, dt_kick, dt_drift):
            acc = oforce.pp_accel_ewald(pos, L, table, G=cfg.G, mass=cfg.mass, eps=cfg.eps)
            return oforce.kick_drift(pos, mom, acc, L, mass=cfg.mass, a=1.0, dt_kick=dt_kick, dt_drift=dt_drift)
    edges = scale_factor_edges(a0, a0 * math.exp(0.15), 0.05)
    p_ref, m_ref = timeloop.evolve(OracleEwald(), c, pos0, mom0, edges)
    eng = default_engine()
    st = timeloop.PeriodicPP(F.ForceStepper(cfg, eng), ewald_gridsize=16)
    p, m = timeloop.evolve(st, c, eng.to_device(pos0), eng.to_device(mom0), edges)
    assert _pdiff(p.cpu().numpy(), p_ref, L) <= 1e-7 * L
    assert np.abs(m.cpu().numpy() - m_ref).max() <= 1e-4 * np.abs(m_ref).max()


def test_second_order_displacement_analytic_and_restatement():
    """nb_lpt2_displacement: closed form for two crossed plane waves, zero for a single one (Zel'dovich is
    exact in 1-D), and the numpy restatement (complex FFTs, real part) on a random field."""
    import torch
    from nbody_b200 import timeloop
    L, n = 100.0, 16
    x = np.arange(n) * (L / n)
    k = 2 * np.pi / L * 2
    A, B = 0.3, -0.7
    X, Y, _Z = np.meshgrid(x, x, x, indexing="ij")
    delta = A * np.cos(k * X) + B * np.cos(k * Y)
    got = timeloop.second_order_displacement(torch.from_numpy(delta), L).cpu().numpy()
    want = np.stack([A * B / (2 * k) * np.sin(k * X) * np.cos(k * Y), A * B / (2 * k) * np.cos(k * X) * np.sin(k * Y),
                     np.zeros_like(X)], axis=-1).reshape(-1, 3)
    assert np.abs(got - want).max() <= 1e-12 * np.abs(want).max()
    rs = np.random.RandomState(4)
    delta = rs.normal(size=(n, n, n))
    delta -= delta.mean()
    got = timeloop.second_order_displacement(torch.from_numpy(delta), L).cpu().numpy()
    ref = numpy_second_order(delta, L)
    assert np.abs(got - ref).max() <= 1e-12 * np.abs(ref).max()
    delta = 0.4 * np.cos(2 * np.pi / L * 3 * X + 0.3)
    psi2 = timeloop.second_order_displacement(torch.from_numpy(delta), L).cpu().numpy()
    assert np.abs(psi2).max() < 1e-12
    for odd in (9, 12):                                   # odd mesh and a mesh that is not a power of two
        d = rs.normal(size=(odd, odd, odd))
        d -= d.mean()
        got = timeloop.second_order_displacement(torch.from_numpy(d), L).cpu().numpy()
        ref = numpy_second_order(d, L)
        assert np.abs(got - ref).max() <= 1e-11 * np.abs(ref).max()
