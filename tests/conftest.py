import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


def surrogate_pk(k):
    """Same stand-in for CLASS as oracle/gen_golden.py."""
    k = np.asarray(k, dtype=np.float64)
    q = k / 0.02
    return 2.0e4 * q / (1.0 + q ** 2.2) ** 1.4


def zeldovich_particles(n, L, seed=20250611, amp=1.0):
    """Synthetic Zel'dovich set: the reference's recipe
    (CiC/3D startbetingelser1.py:62-173) with a surrogate P(k), lattice order."""
    from oracle import restate
    rs = np.random.RandomState(seed)
    c1 = rs.normal(0, 1 / np.sqrt(2), size=(n, n, n))
    c2 = rs.normal(0, 1 / np.sqrt(2), size=(n, n, n))
    R = restate.hermitian_R(c1, c2, guard=True)
    km = restate.k_mag_grid(n, L)
    km[0, 0, 0] = 1e-10
    _d, psi = restate.grf_delta_psi(surrogate_pk(km) * amp, R, L, n)
    lat = restate.lattice(L, n)
    pos = lat + np.column_stack([p.ravel() for p in psi])
    pos %= L
    return pos


@pytest.fixture(scope="session")
def snap_test1():
    g = golden("snap_test1.npz")
    return g["pos"], float(g["boxsize"]), g["delta_golden"]
