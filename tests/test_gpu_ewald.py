"""Periodic PP with Ewald summation on the device (nbody_b200/csrc/ewald.cu)
against the float64 restatement in oracle/force.py (parity unpinned like the rest of the force
path: the reference has no source for it; test_ewald_cpu.py checks the restatement against
implementation-independent properties of the periodic force)."""
import numpy as np
import pytest

from conftest import zeldovich_particles
from oracle import force as oforce

pytestmark = pytest.mark.gpu
TOL = 1e-5


def test_device_table_matches_the_restatement():
    from nbody_b200.engine import default_engine
    eng = default_engine()
    L, g = 250.0, 16
    t4, t8 = eng.ewald_table(L, g, want_f64=True)
    ref = oforce.ewald_table(L, g)
    got = t8.cpu().numpy()
    assert got.shape == ref.shape
    assert np.abs(got - ref).max() <= 1e-11 * np.abs(ref).max()
    f4 = t4.cpu().numpy().reshape(g + 1, g + 1, g + 1, 4)
    assert np.abs(f4[..., :3] - ref).max() <= 1e-6 * np.abs(ref).max() and not f4[..., 3].any()
    # cached: the same tensors come back
    assert eng.ewald_table(L, g) is t4


@pytest.mark.parametrize("n_part", [500, 1728])
def test_pp_ewald_accel_matches_the_restatement(n_part):
    """500 particles: one j-chunk per i-block (plain stores); 1728: split j range (atomic adds)."""
    from nbody_b200.engine import default_engine
    eng = default_engine()
    L, g = 500.0, 16
    pos = zeldovich_particles(12, L, seed=6)[:n_part]
    eps = 0.025 * L / 12
    t4 = eng.ewald_table(L, g)
    half = 0.5 * L
    p4 = eng.pp_pack(eng.to_device(pos), origin=(half, half, half), mass=1.0)
    a = eng.pp_accel_ewald(p4, p4, eps, L, t4, g, scale=3.0).cpu().numpy()
    ref = oforce.pp_accel_ewald(pos, L, oforce.ewald_table(L, g), G=2.0, mass=1.5, eps=eps)
    err = np.abs(a - ref).max() / np.sqrt((ref ** 2).mean())
    print(f"n = {n_part}: err {err:.2e}")
    assert err <= TOL
    # accumulate adds on top
    b = eng.pp_accel_ewald(p4, p4, eps, L, t4, g, scale=3.0, out=eng.to_device(np.ones_like(pos)), accumulate=True)
    assert np.abs(b.cpu().numpy() - 1.0 - a).max() <= 1e-12 * np.abs(a).max()


def test_stepper_ewald_option_and_direct_sum():
    """ForceStepper(...).accel(pos, periodic_pp="ewald"): equal to the restatement with the same
    table, and — with a negligible softening — within the interpolation error of the direct Ewald
    sum, where the nearest-image-only sum is several per cent off."""
    from nbody_b200 import force as F
    from nbody_b200.engine import default_engine
    from nbody_b200.params import ForceConfig
    eng = default_engine()
    L, n = 400.0, 8
    pos = zeldovich_particles(n, L, seed=2)
    cfg = ForceConfig(method="pp", npart=n ** 3, boxsize=L, gridsize=0, G=1.5, mass=2.0, softening_frac=1e-4)
    st = F.ForceStepper(cfg, eng)
    a = st.accel(eng.to_device(pos), periodic_pp="ewald").cpu().numpy().copy()      # 64^3 table
    sub = np.arange(0, n ** 3, 37)
    want = np.zeros((sub.size, 3))
    for k, i in enumerate(sub):
        d = np.delete(pos, i, axis=0) - pos[i]
        want[k] = cfg.G * cfg.mass * oforce.ewald_force(d, L).sum(0)
    rms = np.sqrt((want ** 2).mean())
    e_ewald = np.abs(a[sub] - want).max() / rms
    mi = st.accel(eng.to_device(pos), periodic_pp=True).cpu().numpy()
    e_mi = np.abs(mi[sub] - want).max() / rms
    print(f"Ewald table (g = 64) vs direct sum {e_ewald:.2e}; nearest image only {e_mi:.2e}")
    assert e_ewald <= 2e-4 and e_mi > 50 * e_ewald
    st.ewald_gridsize = 16
    a16 = st.accel(eng.to_device(pos), periodic_pp="ewald").cpu().numpy()
    ref16 = oforce.pp_accel_ewald(pos, L, oforce.ewald_table(L, 16), G=cfg.G, mass=cfg.mass, eps=cfg.eps)
    # a barely perturbed lattice: the net force is the small remainder of pair forces that cancel
    # across j-tiles, each tile summed in FP32 (partial sums ~50x the net force), so the error
    # relative to the rms NET force is ~1e-4 here; the 1e-5 parity bound is asserted on the set of
    # test_pp_ewald_accel_matches_the_restatement
    assert np.abs(a16 - ref16).max() / np.sqrt((ref16 ** 2).mean()) <= 3e-4


def test_particle_mesh_ewald_mode_against_the_direct_sum():
    """timeloop.PeriodicPP(mode="pme"): the periodic force through the PM mesh (reciprocal sum) + the cell-list
    erfc sum, with accuracy parameters (r_s = 3 mesh cells, r_cut = 6.5 r_s, mesh 4 cbrt(N)).  Against the direct
    Ewald sum (float64, oracle) on a subset: a few 1e-4 of the rms force, like the tabulated direct kernel; both
    periodic evaluations agree with each other at that level, the nearest-image-only sum does not."""
    from nbody_b200 import force as F
    from nbody_b200 import timeloop
    from nbody_b200.engine import default_engine
    from nbody_b200.params import ForceConfig
    eng = default_engine()
    L, n = 400.0, 16
    pos = zeldovich_particles(n, L, seed=3)
    cfg = ForceConfig(method="pp", npart=n ** 3, boxsize=L, gridsize=0, G=1.5, mass=2.0)
    st = F.ForceStepper(cfg, eng)
    dpos = eng.to_device(pos)
    direct = timeloop.PeriodicPP(st, mode="ewald").accel(dpos).cpu().numpy().copy()
    pme = timeloop.PeriodicPP(st, mode="pme").accel(dpos).cpu().numpy().copy()
    sub = np.arange(0, n ** 3, 97)
    want = np.zeros((sub.size, 3))
    for k, i in enumerate(sub):
        d = np.delete(pos, i, axis=0) - pos[i]
        dm = d - L * np.round(d / L)                       # nearest image
        r2 = (dm * dm).sum(1)
        # softened nearest-image term + Ewald correction of the unsoftened periodic sum (the product's definition)
        want[k] = cfg.G * cfg.mass * (oforce.ewald_force(d, L).sum(0)
                                      + (dm * ((r2 + cfg.eps ** 2) ** -1.5 - r2 ** -1.5)[:, None]).sum(0))
    rms = np.sqrt((want ** 2).mean())
    e_pme, e_dir = np.abs(pme[sub] - want).max() / rms, np.abs(direct[sub] - want).max() / rms
    both = np.abs(pme - direct).max() / np.sqrt((direct ** 2).mean())
    print(f"PME vs direct Ewald sum {e_pme:.2e}; table kernel {e_dir:.2e}; PME vs table kernel (all particles) {both:.2e}")
    assert e_pme <= 2e-3 and e_dir <= 5e-4 and both <= 2e-3
    with pytest.raises(ValueError):
        timeloop.PeriodicPP(st, mode="tree")
