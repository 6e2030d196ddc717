"""The Ewald restatement (oracle/force.py) against properties of the periodic
force that do not depend on the implementation: independence of the splitting parameter,
periodicity, symmetry zeros, the small-separation limit of the correction, and agreement of the
table + nearest-image form with the direct Ewald sum."""
import math

import numpy as np

from oracle import force as oforce


def test_ewald_sum_is_independent_of_alpha_and_periodic():
    L = 100.0
    rs = np.random.RandomState(1)
    d = (rs.random_sample((40, 3)) - 0.5) * L
    f2 = oforce.ewald_force(d, L, alpha_box=2.0)
    f3 = oforce.ewald_force(d, L, alpha_box=2.6, nreal=4, nfour=7)
    scale = np.abs(f2).max()
    assert np.abs(f2 - f3).max() <= 1e-11 * scale
    shifted = oforce.ewald_force(d + np.array([L, -2 * L, 0.0]), L, nreal=6)
    assert np.abs(shifted - f2).max() <= 1e-10 * scale
    assert np.abs(oforce.ewald_force(-d, L) + f2).max() <= 1e-12 * scale


def test_symmetry_zeros_and_small_separation_limit():
    L = 50.0
    pts = np.array([[L / 2, 0, 0], [L / 2, L / 2, 0], [L / 2, L / 2, L / 2], [0, L / 2, 0]])
    assert np.abs(oforce.ewald_force(pts, L)).max() <= 1e-12 / L ** 2 * 1e3
    d = np.array([[1e-3, 2e-3, -1.5e-3]]) * L
    c = oforce.ewald_force(d, L, subtract_nearest=True)
    want = -4 * math.pi / 3 * d / L ** 3        # the neutralising background inside the sphere pushes away
    assert np.abs(c - want).max() <= 1e-4 * np.abs(want).max()


def test_table_plus_nearest_image_reproduces_the_direct_sum():
    L, g = 80.0, 32
    table = oforce.ewald_table(L, g)
    assert table.shape == (g + 1, g + 1, g + 1, 3)
    assert not table[0, :, :, 0].any() and not table[:, 0, :, 1].any() and not table[:, :, 0, 2].any()
    rs = np.random.RandomState(3)
    pos = rs.random_sample((24, 3)) * L
    a = oforce.pp_accel_ewald(pos, L, table, G=2.0, mass=1.5, eps=0.0)
    want = np.zeros_like(a)
    for i in range(pos.shape[0]):
        d = np.delete(pos, i, axis=0) - pos[i]
        want[i] = 2.0 * 1.5 * oforce.ewald_force(d, L).sum(0)
    # trilinear interpolation of a smooth correction on 32 intervals: ~1e-3 of the correction,
    # far less of the total force
    assert np.abs(a - want).max() <= 2e-4 * np.abs(want).max()
    # total momentum: pairwise antisymmetric
    assert np.abs(a.sum(0)).max() <= 1e-12 * np.abs(a).sum()
    # and it differs visibly from the nearest-image-only sum the table corrects
    mi = oforce.pp_accel(pos, L, G=2.0, mass=1.5, eps=0.0, periodic=True)
    assert np.abs(mi - want).max() > 20 * np.abs(a - want).max()
