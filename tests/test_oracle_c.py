"""CPU: the C restatement (oracle/oracle_c.c) against the numpy restatement and the goldens."""
import numpy as np

from conftest import golden, zeldovich_particles
from oracle import cport, force, restate


def test_c_deposit_bit_identical_to_reference_loop_order(snap_test1):
    pos, L, gold = snap_test1
    g = cport.cic_density(pos, L, 32)
    assert np.array_equal(g, restate.cic_density(pos, L, 32))        # same accumulation order
    d = restate.deconvolve_cic(restate.compute_delta(g), L, 32)
    assert np.abs(d - gold).max() < 1e-12
    out = np.array([[-0.3, 5.2, -7.9], [500.0, 1e-9, 499.9999]]) * (L / 500.0)
    assert np.array_equal(cport.cic_density(out, L, 16), restate.cic_density(out, L, 16))


def test_c_gather_and_shell_bin():
    rs = np.random.RandomState(1)
    N, L = 16, 300.0
    f = rs.normal(size=(N, N, N))
    pos = rs.uniform(-50, 400, size=(2000, 3))
    assert np.abs(cport.cic_gather(f, pos, L, N) - restate.trilinear_gather(f, pos, L, N)).max() < 1e-13
    P = rs.uniform(size=(N, N, N))
    e = restate.pk_edges(N, L)
    for closed in (True, False):
        c, s = cport.shell_bin(P, N, restate.k_vals(N, L), e, closed)
        cr, sr = restate.shell_bin(restate.k_mag_grid(N, L).ravel(), P.ravel(), e, closed)
        assert np.array_equal(c, cr) and np.allclose(s, sr, rtol=1e-12)


def test_c_pair_forces():
    L = 100.0
    pos = zeldovich_particles(10, L, seed=4)
    for box in (0.0, L):
        a = cport.pp_accel(pos, 2.0, 3.0, 0.5, box)
        ref = force.pp_accel(pos, L, 2.0, 3.0, 0.5, periodic=box > 0)
        assert np.abs(a - ref).max() <= 1e-12 * np.abs(ref).max()
    r_s = 1.25 * L / 32
    a, npairs = cport.p3m_short_accel(pos, L, r_s, 4.5 * r_s, 1.0, 1.0, 0.1, return_pairs=True)
    ref = force.p3m_short_accel(pos, L, r_s, 4.5 * r_s, 1.0, 1.0, 0.1)
    assert np.abs(a - ref).max() <= 1e-12 * np.abs(ref).max()
    assert npairs > 0
