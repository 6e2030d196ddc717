"""CPU: analytic properties that pin OUR force-step restatement (oracle/force.py; the
reference has no source for this path — "parity unpinned")."""
import numpy as np

from oracle import force, restate


def test_pp_newton_third_law_and_two_body():
    rs = np.random.RandomState(0)
    pos = rs.uniform(0, 10, size=(64, 3))
    a = force.pp_accel(pos, G=2.0, mass=3.0, eps=0.1)
    assert np.abs(a.sum(0)).max() < 1e-12 * np.abs(a).max() * 64
    two = np.array([[0.0, 0, 0], [2.0, 0, 0]])
    a = force.pp_accel(two, G=1.0, mass=1.0, eps=0.0)
    assert np.allclose(a[0], [0.25, 0, 0]) and np.allclose(a[1], [-0.25, 0, 0])


def test_split_factor_limits():
    r_s = 1.0
    assert abs(force.split_factor(np.array([1e-6]), r_s)[0] - 1) < 1e-12
    assert force.split_factor(np.array([4.5]), r_s)[0] < 0.03


def test_pm_plus_short_range_is_newtonian_pair_force():
    """Two particles a few cells apart: PM long range + erfc short range ~ minimum-image
    Newton (+ periodic images, small at r << L)."""
    L, N = 64.0, 64
    r_s = 1.25 * L / N
    sep = np.array([2.3, 1.1, -2.2])
    pos = np.array([[20.1, 30.3, 25.7], [20.1, 30.3, 25.7]]) + np.vstack([np.zeros(3), sep])
    glr = force.pm_accel(pos, L, N, G=1.0, mass=1.0, r_s=r_s, deconv_order=2, interlace=True)
    gsr = force.p3m_short_accel(pos, L, r_s, 4.5 * r_s)
    r = np.linalg.norm(sep)
    newton = sep / r ** 3          # acceleration of particle 0 points to particle 1
    tot = glr[0] + gsr[0]
    # Ewald-image and mesh errors are a few per cent at r ~ 6.5 cells with a 64³ mesh
    assert np.linalg.norm(tot - newton) / np.linalg.norm(newton) < 0.05
    assert np.linalg.norm((glr[1] + gsr[1]) + tot) / np.linalg.norm(newton) < 0.05   # third law


def test_interlacing_reduces_grid_scatter_and_matches_analytic_long_range():
    """Pair force at a fixed separation, averaged over random placements w.r.t. the mesh:
    the mean equals the analytic long-range force Newton·(1 - split), and interlacing
    reduces the placement-dependent scatter (aliasing)."""
    L, N = 32.0, 32
    r_s = 1.25 * L / N
    rs = np.random.RandomState(3)
    sep = np.array([2.0, 1.5, -1.0])
    r = np.linalg.norm(sep)
    analytic = sep / r ** 3 * (1 - force.split_factor(r, r_s))
    scatter = {}
    for inter in (False, True):
        gs = []
        for _ in range(16):
            p0 = rs.uniform(8, 24, size=3)
            g = force.pm_accel(np.vstack([p0, p0 + sep]), L, N, r_s=r_s, deconv_order=2, interlace=inter)
            gs.append(0.5 * (g[0] - g[1]))
        gs = np.array(gs)
        scatter[inter] = np.linalg.norm(gs.std(0)) / np.linalg.norm(gs.mean(0))
        assert np.linalg.norm(gs.mean(0) - analytic) / np.linalg.norm(analytic) < 0.02
    assert scatter[True] < scatter[False]


def test_kick_drift_wraps_like_reference():
    pos = np.array([[499.5, 0.2, 250.0]])
    mom = np.array([[2.0, -1.0, 0.0]])
    p, m = force.kick_drift(pos, mom, np.zeros((1, 3)), 500.0, dt_drift=0.5)
    assert np.allclose(p, [[0.5, 499.7, 250.0]]) and np.array_equal(m, mom)
