"""Force path at sizes where a CTA of the short-range kernel handles several work units, tiles
overflow and cell ranges are partial — the small parity cases of test_gpu_force.py do not reach
these paths.  The float64 oracle (oracle/force.py, parity unpinned like the rest of the force
path) is evaluated on spatial SUBSETS (an interior cube and the periodic box corner), plus
size-independent properties on the whole set: no particle left unwritten, total momentum of the
pairwise / mesh force ~ 0."""
import numpy as np
import pytest

from oracle import force as oforce

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _glass(n, L, seed, jitter=0.5):
    rs = np.random.RandomState(seed)
    lat = (np.arange(n) * (L / n))
    pos = np.stack(np.meshgrid(lat, lat, lat, indexing="ij"), -1).reshape(-1, 3)
    return (pos + (rs.random_sample(pos.shape) - 0.5) * 2 * jitter * (L / n)) % L


def _subset_oracle(pos, L, centre, half, cfg):
    """Oracle short-range acceleration of the particles within `half` (max-norm, periodic) of
    `centre`, from all particles within half + r_cut."""
    rel = (pos - centre + 0.5 * L) % L - 0.5 * L
    dist = np.abs(rel).max(axis=1)
    i_glob = np.nonzero(dist < half)[0]
    j_glob = np.nonzero(dist < half + cfg.r_cut * 1.0001)[0]
    sub = rel[j_glob]
    i_loc = np.searchsorted(j_glob, i_glob)
    ref = oforce.p3m_short_accel(sub, 4.0 * L, cfg.r_s, cfg.r_cut, cfg.G, cfg.mass, cfg.eps, i_subset=i_loc, chunk=32)
    return i_glob, ref


@pytest.mark.parametrize("kind", ["glass", "clustered"])
def test_short_range_many_units_per_cta(kind):
    """128^3 particles, ncs = 45: 24 300 (column, z-run) units for at most 9 472 CTAs, a last run of
    one cell per column; the clustered variant overflows tiles (device-side list + second pass)."""
    import torch
    from nbody_b200 import force as F
    from nbody_b200.engine import default_engine
    from nbody_b200.params import ForceConfig
    eng = default_engine()
    n, N, L = 128, 256, 1000.0
    pos = _glass(n, L, seed=4)
    if kind == "clustered":   # pull 3 % of the particles into two tight clumps (one across the box corner)
        rs = np.random.RandomState(9)
        k = pos.shape[0] * 3 // 200
        pos[:k] = rs.normal(0.0, 6.0, size=(k, 3)) % L
        pos[k:2 * k] = (np.array([400.0, 610.0, 333.0]) + rs.normal(0.0, 9.0, size=(k, 3))) % L
    cfg = ForceConfig(method="p3m", npart=pos.shape[0], boxsize=L, gridsize=N, G=1.0, mass=1.0)
    st = F.ForceStepper(cfg, eng)
    dpos = eng.to_device(pos)
    acc = torch.full_like(dpos, float("nan"))
    st.short_range(dpos, acc, accumulate=False)
    a = acc.cpu().numpy()
    assert np.isfinite(a).all(), "a particle was left unwritten or picked up a NaN"
    # pairwise forces cancel: total momentum change ~ 0
    assert np.abs(a.sum(axis=0)).max() <= 1e-4 * np.abs(a).sum(axis=0).max()
    rms = np.sqrt((a ** 2).mean())
    # (centre, half width, tolerance): the two clumps hold ~2500 particles per sort cell with
    # separations down to the softening length; there the float32 tile coordinates (|z| < 4 cell
    # sides in the CTA frame) limit a pair force to ~3e-5 relative, so the clump regions carry 5e-5
    regions = [(np.array([520.0, 480.0, 505.0]), 14.0, TOL), (np.zeros(3), 14.0 if kind == "glass" else 3.0,
                                                              TOL if kind == "glass" else 5e-5)]
    if kind == "clustered":
        regions.append((np.array([400.0, 610.0, 333.0]), 4.0, 5e-5))
    for centre, half, tol in regions:
        i_glob, ref = _subset_oracle(pos, L, centre, half, cfg)
        assert i_glob.size > 20
        err = np.abs(a[i_glob] - ref).max() / max(rms, np.sqrt((ref ** 2).mean()))
        print(f"{kind}: centre {centre.tolist()} n_i {i_glob.size} err {err:.2e}")
        assert err <= tol


def test_pm_step_at_config4_size_properties():
    """BASELINE config 4 (256^3 particles on a 512^3 mesh, interlaced + deconvolved PM): finite
    accelerations, zero total momentum change (same CIC kernel for deposit and gather), positions
    wrapped into [0, L) after the drift."""
    import torch
    from nbody_b200 import force as F
    from nbody_b200.engine import default_engine
    from nbody_b200.params import ForceConfig
    eng = default_engine()
    n, N, L = 256, 512, 1000.0
    gen = torch.Generator(device="cuda").manual_seed(11)
    lat = torch.arange(n, device="cuda", dtype=torch.float64) * (L / n)
    pos = torch.stack(torch.meshgrid(lat, lat, lat, indexing="ij"), -1).reshape(-1, 3)
    pos = ((pos + (torch.rand(pos.shape, generator=gen, device="cuda", dtype=torch.float64) - 0.5) * (L / n)) % L)
    pos = pos.contiguous()
    cfg = ForceConfig(method="pm", npart=n ** 3, boxsize=L, gridsize=N, interlace=(True, True),
                      deconvolve=(True, True), G=1.0, mass=1.0)
    st = F.ForceStepper(cfg, eng)
    acc = st.accel(pos)
    assert bool(torch.isfinite(acc).all())
    tot = acc.sum(dim=0).abs().max().item()
    assert tot <= 1e-4 * acc.abs().sum(dim=0).max().item()
    mom = torch.zeros_like(pos)
    st.step(pos, mom, dt_kick=0.5, dt_drift=0.5)
    assert bool(((pos >= 0) & (pos < L)).all()) and bool(torch.isfinite(mom).all())
    assert mom.abs().max().item() > 0
