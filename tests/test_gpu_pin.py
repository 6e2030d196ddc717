"""Statistical pin of the force path against the reference's own
full-run outputs: the 64³ CONCEPT runs whose `powerspec_a=*` files the reference ships are repeated
with this package (tools/pin_concept.py: ICs from the files' linear column, comoving leapfrog with
the sub-cycled short range, CONCEPT-style estimator) and compared bin by bin.

The reference cannot pin a single force evaluation (CONCEPT's source is absent), so the tolerances are
those of a two-realisation comparison with ONE realisation of the reference: measured with four seeds
(profiles/r2_pin_concept.json) the deviations are ≤ 2.3 % per bin at a = 0.17, ≤ 4 % (L = 1000) and
≤ 6.2 % (L = 500) at a = 1, ≤ 2.5 % on the P³M/PM ratio and ≤ 1.2 % on σ₈; the bounds asserted here
leave room for the second seed's cosmic variance."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def report():
    from tools import pin_concept
    return pin_concept.main(["--backend", "gpu", "--seeds", "2", "--methods", "pm,p3m,snapvspk"])


def test_grid_columns_are_exact(report):
    for tag, r in report["runs"].items():
        assert r["modes_exact"] and r["k_8_significant_figures"], tag
        # input spectrum = the file's P_linear column (a = 0.17: D+ of cosmo.py against CLASS, 0.17 % in D+²)
        assert r["linear_column_max_rel"] < 4e-3, tag


def test_pm_runs_match_the_reference_spectra(report):
    r17, r100 = report["runs"]["SnapVsPK_a=0.17"], report["runs"]["SnapVsPK_a=1.00"]
    assert np.abs(np.array(r17["P_corrected_over_ref_corrected"]) - 1).max() < 0.05       # measured 0.023
    assert np.abs(np.array(r100["P_corrected_over_ref_corrected"]) - 1).max() < 0.08      # measured 0.040
    pm = report["runs"]["PM_a=1.00"]
    assert np.abs(np.array(pm["P_corrected_over_ref_corrected"]) - 1).max() < 0.10        # measured 0.062
    for r in (r17, r100, pm):
        assert abs(r["sigma8_mean"][0] / r["sigma8_ref"][0] - 1) < 0.025                  # measured <= 0.012


def test_p3m_run_and_the_short_range_signature(report):
    p3m = report["runs"]["P3M_a=1.00"]
    assert p3m["short_range_evaluations"] > p3m["steps"]                                  # sub-cycling was active
    assert np.abs(np.array(p3m["P_corrected_over_ref_corrected"]) - 1).max() < 0.10        # measured 0.061
    assert abs(p3m["sigma8_mean"][0] / p3m["sigma8_ref"][0] - 1) < 0.025
    # P3M / PM per bin: same seeds on each side, cosmic variance cancels to first order.  The reference's ratio
    # falls from 1.01 at k = 0.1 / Mpc to 0.71 at the Nyquist frequency; ours follows it within 4 % (measured 2.5 %)
    ratio = np.array(report["p3m_over_pm"]["ratio"])
    assert np.abs(ratio - 1).max() < 0.04
    assert report["p3m_over_pm"]["reference"][-1] < 0.75 and report["p3m_over_pm"]["ours"][-1] < 0.75
