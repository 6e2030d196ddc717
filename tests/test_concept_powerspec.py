"""CONCEPT-style power-spectrum estimator against the five `powerspec_a=*` files
the reference ships (byte copies under tests/golden/concept_powerspec/, made by oracle/gen_golden.py).

CPU part: the grid-only columns (`modes` exactly, `k` to the files' 8 significant figures), the σ₈
header values recomputed from the files' own columns, the linear column from the closed-form
spectrum, and the text layout.  GPU part: the device estimator against the float64 restatement."""
import math
import os

import numpy as np
import pytest

from nbody_b200 import powerspec as PS
from nbody_b200.cosmo import Cosmology
from oracle import concept_ps

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "concept_powerspec")
FILES = {"P3M_a=1.00": 500.0, "PM_a=1.00": 500.0, "PP_a=1.00": 500.0, "SnapVsPK_a=0.17": 1000.0,
         "SnapVsPK_a=1.00": 1000.0}


def _ref(tag):
    return PS.read_powerspec(os.path.join(GOLD, tag + ".txt"))


@pytest.mark.parametrize("tag", sorted(FILES))
def test_modes_and_k_columns_reproduce_the_reference_files(tag):
    ref, L = _ref(tag), FILES[tag]
    assert ref.gridsize == 64
    geom = PS.bin_geometry(64)
    assert np.array_equal(geom.modes, ref.modes)                       # 3, 25, 124, 418, ..., 15506
    k = 2 * math.pi / L * geom.k_grid
    assert ["%.7e" % x for x in k] == ["%.7e" % x for x in ref.k]      # 8 significant figures


def test_bins_are_the_ranges_the_modes_column_implies():
    """The integer-k² ranges recovered from the files' modes column by brute-force enumeration of half
    of Fourier space equal the product's bins, and the brute-force geometry equals the lattice-count
    tables (independent of how the bins were found)."""
    ref = _ref("P3M_a=1.00")
    ranges = concept_ps.bins_from_modes_column(64, ref.modes)
    geom = PS.bin_geometry(64)
    assert [r[0] for r in ranges] == geom.first_k2.tolist()
    assert ranges[-1][1] == 1022 and ranges[1] == (2, 5) and ranges[12] == (488, 594)
    modes, kgeo = concept_ps.geometry_bruteforce(64, ranges)
    assert np.array_equal(modes, geom.modes) and np.abs(kgeo / geom.k_grid - 1).max() < 1e-14
    # every pinned edge lies strictly inside the interval the files allow
    e = PS.bin_edges(32.0)
    for j, (lo, _hi) in enumerate(ranges):
        prev_last = 0 if j == 0 else ranges[j - 1][1]
        assert math.sqrt(prev_last) < e[j] <= math.sqrt(lo)
    assert e[-1] >= 32.0 > e[-2]


@pytest.mark.parametrize("N", [16, 32, 128])
def test_other_grids_share_the_low_bins_and_agree_with_brute_force(N):
    geom = PS.bin_geometry(N)
    g64 = PS.bin_geometry(64)
    nshared = min(geom.first_k2.size, g64.first_k2.size) - 1
    assert np.array_equal(geom.first_k2[:nshared], g64.first_k2[:nshared])
    assert np.array_equal(geom.modes[:nshared], g64.modes[:nshared])
    ranges = [(int(f), int(l)) for f, l in zip(geom.first_k2, list(geom.first_k2[1:] - 1) + [geom.k2_end - 1])]
    modes, kgeo = concept_ps.geometry_bruteforce(N, ranges)
    assert np.array_equal(modes, geom.modes) and np.abs(kgeo / geom.k_grid - 1).max() < 1e-13
    assert geom.k2_end == (N // 2) ** 2 and int(geom.modes.sum()) * 2 == int(geom.r3[1:].sum())


def test_non_default_bins_per_decade_and_kmax():
    g = PS.bin_geometry(64, k_max=20.0, bins_per_decade={1.0: 10.0, 10.0: 10.0})
    e = PS.bin_edges(20.0, {1.0: 10.0, 10.0: 10.0})
    assert np.allclose(np.diff(np.log10(e)), 0.1) and abs(e[0] - 10 ** -0.05) < 1e-12
    assert g.k2_end == 400 and g.modes.sum() * 2 == g.r3[1:400].sum()
    with pytest.raises(ValueError):
        PS.bin_geometry(64, k_max=40.0)


@pytest.mark.parametrize("tag", sorted(FILES))
def test_sigma8_of_the_file_columns_matches_the_header(tag):
    """σ₈ recomputed from each of the three P columns equals the header value to ±0.002 (the headers
    carry three digits; CONCEPT's own quadrature is not in the reference)."""
    ref = _ref(tag)
    R = 8.0 / Cosmology().h
    for col, want in zip(ref.columns(), ref.sigma):
        assert abs(PS.sigma_tophat(ref.k, col, R) - want) <= 0.002


@pytest.mark.parametrize("tag,a", [("P3M_a=1.00", 1.0), ("SnapVsPK_a=1.00", 1.0), ("SnapVsPK_a=0.17", 1.0 / 6.0)])
def test_closed_form_linear_spectrum_against_the_P_linear_column(tag, a):
    """cosmo.linear_power (Eisenstein-Hu with baryons, A_s / n_s of NN.param:92-95) stands in for
    CLASS: its mode-averaged bin means are within 9 % of the file's column in every bin and within
    2.5 % rms (the 5-9 % bins sit on the first acoustic trough, k ≈ 0.06-0.07 / Mpc)."""
    ref, L = _ref(tag), FILES[tag]
    c = Cosmology()
    lin, _ = PS.linear_column(PS.bin_geometry(64), L, lambda k: c.linear_power(k, A_s=2.105e-9, n_s=0.9649, a=a))
    r = lin / ref.linear - 1
    assert np.abs(r).max() < 0.09 and math.sqrt(np.mean(r * r)) < 0.03


def test_growth_between_the_two_snapvspk_files():
    """P_linear(a=1)/P_linear(a=1/6) of the reference's two files is D₊ ratio squared: 22.347 in every
    bin against 22.385 from cosmo.growth (0.17 %: CLASS carries radiation, cosmo.py does not)."""
    r = _ref("SnapVsPK_a=1.00").linear / _ref("SnapVsPK_a=0.17").linear
    c = Cosmology()
    assert np.ptp(r) < 2e-3 and abs((c.growth(1.0) / c.growth(1.0 / 6.0)) ** 2 / r.mean() - 1) < 3e-3


def test_text_layout_round_trip(tmp_path):
    ref = _ref("PM_a=1.00")
    rows = open(os.path.join(GOLD, "PM_a=1.00.txt"), encoding="utf-8").read().splitlines()
    data = [ln for ln in rows if not ln.startswith("#")]
    assert PS.format_rows(ref) == data                                  # byte-identical data rows
    ref.tophat = 8.0
    p = tmp_path / "powerspec_a=1.00"
    PS.write_powerspec(str(p), ref)
    back = PS.read_powerspec(str(p))
    assert np.array_equal(back.modes, ref.modes) and np.array_equal(back.power, ref.power)
    assert back.gridsize == 64 and back.a == 1.0 and abs(back.t - 13.814035) < 1e-9
    assert back.sigma == ref.sigma == (0.883, 0.884, 0.816)
    head = open(str(p), encoding="utf-8").read().splitlines()
    assert head[-17].split()[1:3] == ["k", "[Mpc⁻¹]"] and len(head) == len(rows)


def test_realization_ratio_matches_the_restatement():
    from oracle import restate
    n, L = 16, 300.0
    rs = np.random.RandomState(3)
    R = restate.hermitian_R(rs.normal(0, 1 / np.sqrt(2), (n, n, n)), rs.normal(0, 1 / np.sqrt(2), (n, n, n)), guard=True)
    geom = PS.bin_geometry(n)
    lin = lambda k: 1e4 * k / (1 + (k / 0.05) ** 2.5)      # noqa: E731
    ranges = [(int(f), int(l)) for f, l in zip(geom.first_k2, list(geom.first_k2[1:] - 1) + [geom.k2_end - 1])]
    _c, p_k2 = PS.linear_column(geom, L, lin)
    got = PS.realization_ratio(geom, np.abs(R) ** 2, p_k2)
    assert np.abs(got / concept_ps.noise_ratio(R, L, n, ranges, lin) - 1).max() < 1e-12


# ------------------------------------------------------------------------------------------- GPU
def _zeldovich_set(n, L, seed=4, amp=1.0):
    from oracle import restate
    rs = np.random.RandomState(seed)
    R = restate.hermitian_R(rs.normal(0, 1 / np.sqrt(2), (n, n, n)), rs.normal(0, 1 / np.sqrt(2), (n, n, n)), guard=True)
    km = restate.k_mag_grid(n, L)
    km[0, 0, 0] = 1e-10
    q = km / 0.02
    _d, psi = restate.grf_delta_psi(amp * 2.0e4 * q / (1.0 + q ** 2.2) ** 1.4, R, L, n)
    return (restate.lattice(L, n) + 0.25 * L / n + np.column_stack([c.ravel() for c in psi])) % L, R


@pytest.mark.gpu
@pytest.mark.parametrize("interlace,deconvolve", [(True, True), (False, True), (True, False)])
def test_device_estimator_matches_the_restatement(interlace, deconvolve):
    n, L = 32, 500.0
    pos, R = _zeldovich_set(n, L)
    geom = PS.bin_geometry(n)
    ranges = [(int(f), int(l)) for f, l in zip(geom.first_k2, list(geom.first_k2[1:] - 1) + [geom.k2_end - 1])]
    modes, P = concept_ps.powerspec(pos, L, n, ranges, interlace=interlace, deconvolve=deconvolve)
    lin = lambda k: 3e4 * (k / 0.02) / (1 + (k / 0.02) ** 2.2) ** 1.4      # noqa: E731
    for prec, tol in (("f32", 1e-5), ("f64", 1e-11)):
        ps = PS.concept_powerspec(pos, L, n, interlace=interlace, deconvolve=deconvolve, linear=lin,
                                  noise_abs2=np.abs(R) ** 2, tophat=11.9, precision=prec)
        assert np.array_equal(ps.modes, modes)                        # bin counts bit-exact
        assert np.abs(ps.power / P - 1).max() <= tol, prec           # per-bin P, FP32 tolerance 1e-5
    assert np.abs(ps.corrected / (P * concept_ps.noise_ratio(R, L, n, ranges, lin)) - 1).max() < 1e-10
    assert len(ps.sigma) == 3 and all(s > 0 for s in ps.sigma)


@pytest.mark.gpu
def test_device_estimator_on_the_64_grid_of_the_reference_files(tmp_path):
    """Same grid as the reference's files: the device run reproduces their modes and k columns and the
    written file has their layout."""
    n, L = 64, 500.0
    pos, _R = _zeldovich_set(n, L, seed=9)
    ps = PS.concept_powerspec(pos, L, tophat=8.0 / 0.6736, t=13.814035)
    ref = _ref("P3M_a=1.00")
    assert ps.gridsize == 64 and np.array_equal(ps.modes, ref.modes)
    assert ["%.7e" % x for x in ps.k] == ["%.7e" % x for x in ref.k]
    geom = PS.bin_geometry(n)
    ranges = [(int(f), int(l)) for f, l in zip(geom.first_k2, list(geom.first_k2[1:] - 1) + [geom.k2_end - 1])]
    _m, P = concept_ps.powerspec(pos, L, n, ranges)
    assert np.abs(ps.power / P - 1).max() <= 1e-5
    PS.write_powerspec(str(tmp_path / "powerspec_a=1.00"), ps)
    back = PS.read_powerspec(str(tmp_path / "powerspec_a=1.00"))
    assert np.array_equal(back.modes, ref.modes) and np.abs(back.power / ps.power - 1).max() < 1e-7
