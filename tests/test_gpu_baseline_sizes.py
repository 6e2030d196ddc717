"""Parity at the sizes BASELINE.json names: the small parity cases of
test_gpu_force.py / test_gpu_density.py are repeated where the kernels run their production
configurations.

  C1  64³ particles / 64³ mesh, the whole `CiC/3D startbetingelser1.py` pipeline (:62-231, 274-318):
      GRF -> ψ -> displaced lattice -> CIC deposit -> both P(k)s, against the float64 restatement
  C3  PP at 128³: 4 096 random i-particles against all 2 097 152 j vs the C oracle
  C4  PM at 256³ / 512³: the whole acceleration field vs the C / scipy oracle (≈20 s of host time)
  C4' P³M at 128³ / 256³ in full vs the C oracle (the mesh + short-range composition at a size the
      oracle finishes in seconds)
  C5  P³M at 512³ / 1024³: short range of ≥ 10⁴ particles of a spatial subset vs the C oracle on a
      cut-out box; total force of 2 048 random particles vs the direct periodic (Ewald) sum over all
      134 M particles on the device — the one check of the 1024³ mesh force that does not go through
      a CPU FFT of that size.
Tolerances: 1e-5 (max |Δ| / rms) wherever the oracle computes the same quantity; the C5 Ewald check
compares P³M with the exact periodic force, i.e. it measures the P³M method error (stated there).
"""
import numpy as np
import pytest

from conftest import surrogate_pk, zeldovich_particles

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _err(a, ref):
    return np.abs(a - ref).max() / np.sqrt((ref ** 2).mean())


def test_c1_cic_script_end_to_end_64():
    """Config 1 end to end: same draws, same surrogate P(k) on the k grid of the script."""
    from nbody_b200 import api
    from oracle import restate
    n, L = 64, 1000.0
    rs = np.random.RandomState(20250611)
    c1 = rs.normal(0, 1 / np.sqrt(2), size=(n, n, n))
    c2 = rs.normal(0, 1 / np.sqrt(2), size=(n, n, n))
    km = restate.k_mag_grid(n, L)
    km[0, 0, 0] = 1e-10
    pk = surrogate_pk(km)
    # reference-side (float64 restatement of the script)
    R = restate.hermitian_R(c1, c2, guard=True)
    d_ref, psi_ref = restate.grf_delta_psi(pk, R, L, n)
    pos_ref = (restate.lattice(L, n) + np.column_stack([p.ravel() for p in psi_ref])) % L
    grid_ref = restate.cic_density(pos_ref, L, n) - 1.0
    # device side through the reference-named entry points
    d_x, psi = api.grf_delta_psi(pk, L, n, c1=c1, c2=c2, guard=True, precision="f64")
    assert np.abs(d_x - d_ref).max() <= 1e-10 * np.abs(d_ref).max()
    d_cic, pos = api.particle_assigment(psi, L, n, precision="f64", return_particles=True)
    dpos = np.abs(pos - pos_ref)
    assert np.minimum(dpos, L - dpos).max() <= 1e-9 * L
    assert np.abs(d_cic - grid_ref).max() <= 1e-9 * np.abs(grid_ref).max()
    idx = api.cic_density(pos_ref, L, n, return_index=True)[1]
    assert np.array_equal(idx, restate.cic_index_weight(pos_ref, L, n)[0])          # bit-exact cell indices
    for field, fromcic in ((d_ref, False), (grid_ref, True)):
        k, p, c = api.compute_power_spectrum(field, fromcic, L=L, N=n, withfac=True, variant="cic3d", return_counts=True)
        kr, pr, cr = restate.compute_power_spectrum(field, fromcic, L, n, withfac=True, variant="cic3d", return_counts=True)
        assert np.array_equal(c, cr) and np.array_equal(k, kr)                      # bin counts bit-exact
        assert np.abs(p / pr - 1).max() <= TOL


def test_c3_pp_128_random_i_subset_against_all_j():
    """The product PP force (all-pairs float32 kernel + float64 near-field fix) of 4 096 random particles
    against the float64 C oracle.  Without the fix the float32 coordinates (24 bits over the box) leave
    4.2e-4 on the worst particle (one pair at 0.06 Mpc, below the softening length) — asserted too, so the
    finding stays visible."""
    from nbody_b200 import force as F
    from nbody_b200.engine import default_engine
    from nbody_b200.params import ForceConfig
    from oracle import cport
    n, L = 128, 500.0
    pos = zeldovich_particles(n, L, seed=7)
    rs = np.random.RandomState(3)
    pos = np.ascontiguousarray(pos[rs.permutation(pos.shape[0])])      # the first 4096 rows are a random subset
    ni = 4096
    cfg = ForceConfig(method="pp", npart=n ** 3, boxsize=L, gridsize=0, G=1.0, mass=1.0)
    eng = default_engine()
    st = F.ForceStepper(cfg, eng)
    dpos = eng.to_device(pos)
    ref = cport.pp_accel(pos, 1.0, 1.0, cfg.eps, 0.0, 0, ni)
    acc = st.accel(dpos)[:ni].cpu().numpy()
    err = _err(acc, ref)
    st.pp_near_fix = False
    raw = _err(st.accel(dpos)[:ni].cpu().numpy(), ref)
    print(f"C3 PP 128^3, {ni} i x {pos.shape[0]} j: err {err:.2e} (float32 coordinates alone: {raw:.2e})")
    assert err <= TOL
    assert 1e-5 < raw < 2e-3


def test_c4_pm_256_512_full_field_against_the_c_oracle():
    from nbody_b200 import force as F
    from nbody_b200.engine import default_engine
    from nbody_b200.params import ForceConfig
    from oracle import cpu_step
    n, N, L = 256, 512, 1000.0
    pos = zeldovich_particles(n, L, seed=5)
    cfg = ForceConfig(method="pm", npart=n ** 3, boxsize=L, gridsize=N, interlace=(True, True), deconvolve=(True, True),
                      G=1.0, mass=1.0)
    eng = default_engine()
    acc = F.ForceStepper(cfg, eng).accel(eng.to_device(pos)).cpu().numpy()
    ref = cpu_step.pm_accel(cfg, pos)
    err = _err(acc, ref)
    print(f"C4 PM 256^3/512^3 full field: err {err:.2e}")
    assert err <= TOL


def _per_particle_stats(acc, ref, pos, L, eps, label, pos_all=None):
    """Per-particle error (max component, in units of the rms force) and its split by the distance to the
    nearest neighbour: the short-range kernel keeps float32 coordinates in a 4-cell CTA frame (quantum
    ≈ 2.4e-7 x 4 r_cut), which a pair closer than a few softening lengths feels as a relative error of its
    separation.  Returns (e, nn)."""
    from scipy.spatial import cKDTree
    rms = np.sqrt((ref ** 2).mean())
    e = np.abs(acc - ref).max(axis=1) / rms
    nn = cKDTree(pos if pos_all is None else pos_all, boxsize=L).query(pos, k=2)[0][:, 1]
    far = nn > 4.0 * eps
    print(f"{label}: median {np.median(e):.2e}  99% {np.percentile(e, 99):.2e}  99.9% {np.percentile(e, 99.9):.2e}  "
          f"max {e.max():.2e}; particles with a partner inside 4 eps: {(~far).sum()} of {e.size}, "
          f"max without one {e[far].max():.2e}")
    return e, far


def test_p3m_128_256_full_against_the_c_oracle():
    """Every particle of a 128^3 Zel'dovich set on a 256^3 mesh (mesh chain + cell sort + pair kernel) against
    the C oracle.  Stated tolerance of the float32 pair kernel: the tile coordinates are float32 in
    a CTA frame of 4 sort cells (quantum 2.4e-7 x 4 r_cut ≈ 2e-5 r_cut), which a pair at separation r feels as a
    relative error quantum/r of its own term.  Measured here: median 3e-7, 99 % 4e-6, 99.9 % 3e-5 of the rms
    force; the 0.4 % of the particles with a partner inside 4 softening lengths reach 1e-3.  The 1e-5 of
    north_star therefore holds for 99 % of the particles, not in the max norm — asserted as such."""
    from nbody_b200 import force as F
    from nbody_b200.engine import default_engine
    from nbody_b200.params import ForceConfig
    from oracle import cpu_step
    n, N, L = 128, 256, 1000.0
    pos = zeldovich_particles(n, L, seed=11)
    cfg = ForceConfig(method="p3m", npart=n ** 3, boxsize=L, gridsize=N, interlace=(True, True),
                      deconvolve=(True, True), G=1.0, mass=1.0)
    eng = default_engine()
    acc = F.ForceStepper(cfg, eng).accel(eng.to_device(pos)).cpu().numpy()
    ref = cpu_step.accel(cfg, pos)
    e, far = _per_particle_stats(acc, ref, pos, L, cfg.eps, "P3M 128^3/256^3 full")
    assert np.median(e) <= 1e-6 and np.percentile(e, 99) <= TOL
    assert np.percentile(e, 99.9) <= 1e-4 and e[far].max() <= 5e-4 and e.max() <= 3e-3


def test_p3m_128_exact_short_range_mode():
    """The same set with ForceStepper.p3m_near_fix (nb_p3m_short_exact): the pair kernel lists the in-range pairs
    closer than 20 softening lengths AND the pairs within 8e-6 (relative, in r^2) of the cut-off instead of evaluating
    them, and a float64 kernel evaluates the list from the original positions (applying the cut-off itself for the
    boundary shell, so a pair is inside r_cut for both partners or for neither).  north_star's 1e-5 then holds in the
    MAX norm over all 2.1 M particles: measured median 2.8e-7, 99.9 % 9.6e-7, max 2.5e-6 of the rms force (r_near =
    12 eps: max 5.9e-6; 8 eps: 1.3e-5)."""
    import torch
    from nbody_b200 import force as F
    from nbody_b200.engine import default_engine
    from nbody_b200.params import ForceConfig
    from oracle import cpu_step
    n, N, L = 128, 256, 1000.0
    pos = zeldovich_particles(n, L, seed=11)
    cfg = ForceConfig(method="p3m", npart=n ** 3, boxsize=L, gridsize=N, interlace=(True, True),
                      deconvolve=(True, True), G=1.0, mass=1.0)
    eng = default_engine()
    st = F.ForceStepper(cfg, eng)
    st.p3m_near_fix = True
    dpos = eng.to_device(pos)
    acc = st.accel(dpos).cpu().numpy()
    ref = cpu_step.accel(cfg, pos)
    e, far = _per_particle_stats(acc, ref, pos, L, cfg.eps, "P3M 128^3/256^3 full, near pairs in float64")
    assert np.median(e) <= 1e-6 and e.max() <= TOL
    # the list is in use: about half a partner per particle inside 20 eps
    srt = eng.cell_sort(dpos, L, st.ncs)
    cnt = torch.zeros(1, dtype=torch.int32, device=eng.device)
    eng.p3m_short(srt[0], srt[1], srt[2], L, st.ncs, cfg.r_s, cfg.r_cut, cfg.eps, scale=1.0,
                  out=torch.empty_like(dpos), pos=dpos, r_near=20.0 * cfg.eps, near_count=cnt)
    torch.cuda.synchronize()
    print("near pairs listed:", int(cnt.item()), "for", pos.shape[0], "particles")
    assert 0.1 * pos.shape[0] < int(cnt.item()) < 2 * pos.shape[0]


def test_c5_p3m_512_1024_subset_and_ewald_checks():
    import torch
    import bench
    from nbody_b200 import force as F
    from nbody_b200.engine import default_engine
    from nbody_b200.params import ForceConfig
    from oracle import cport
    eng = default_engine()
    n, N, L = 512, 1024, 1000.0
    pos = bench.zeldovich_on_device(eng, n, L, seed=20250611)            # the bench's config-5 particle set
    cfg = ForceConfig(method="p3m", npart=n ** 3, boxsize=L, gridsize=N, interlace=(True, True),
                      deconvolve=(True, True), G=1.0, mass=1.0)
    st = F.ForceStepper(cfg, eng)
    acc_s = torch.full_like(pos, float("nan"))
    st.short_range(pos, acc_s, accumulate=False)
    assert bool(torch.isfinite(acc_s).all()), "a particle was left unwritten or picked up a NaN"
    tot = acc_s.sum(dim=0).abs().max().item()
    assert tot <= 1e-4 * acc_s.abs().sum(dim=0).max().item()             # pair forces cancel
    rms_s = float(torch.sqrt((acc_s ** 2).mean()).item())
    # (1) short range of every particle of a (2 x 21 Mpc)^3 cube (and of one across the box corner) vs the C oracle
    #     on the cut-out that holds all their partners
    for centre, half in ((torch.tensor([500.0, 480.0, 505.0]), 21.0), (torch.zeros(3), 12.0)):
        rel = (pos - centre.to(pos) + 0.5 * L) % L - 0.5 * L
        dist = rel.abs().max(dim=1).values
        j_glob = torch.nonzero(dist < half + cfg.r_cut * 1.0001).squeeze(1)
        sub = rel[j_glob].cpu().numpy()
        inner = (np.abs(sub).max(axis=1) < half)
        ref = cport.p3m_short_accel(np.ascontiguousarray(sub + 2.0 * L), 4.0 * L, cfg.r_s, cfg.r_cut, cfg.G, cfg.mass,
                                    cfg.eps)[inner]
        got = acc_s[j_glob].cpu().numpy()[inner]
        e, far = _per_particle_stats(got, ref, sub[inner] + 2.0 * L, 4.0 * L, cfg.eps,
                                     f"C5 short range, centre {centre.tolist()}, {int(inner.sum())} particles",
                                     pos_all=sub + 2.0 * L)
        assert inner.sum() >= (8000 if half > 20 else 800)
        # measured: median 4e-7, 99.9 % 7e-6, max 3.4e-5 (no partner inside 4 eps in this region)
        assert np.median(e) <= 1e-6 and np.percentile(e, 99.9) <= TOL and e.max() <= 1e-4
    del acc_s, rel, dist
    # (2) total P3M force (1024^3 mesh chain + short range) of 2048 random particles vs the direct periodic sum
    #     over all 134 M particles (nearest image + Ewald table, itself checked against the float64 Ewald sum in
    #     test_gpu_ewald.py).  This is the P3M METHOD error (Gaussian split at r_s = 1.25 cells cut at 4.5 r_s,
    #     CIC mesh), not a rounding error: a few 1e-3 of the rms force.
    acc = st.accel(pos)
    g = torch.Generator(device="cuda").manual_seed(5)
    idx = torch.randint(0, pos.shape[0], (2048,), generator=g, device=pos.device)
    got = acc[idx].cpu().numpy()
    rms = float(torch.sqrt((acc ** 2).mean()).item())
    half = 0.5 * L
    p4 = eng.pp_pack(pos, origin=(half, half, half), mass=1.0)
    exact = eng.pp_accel_ewald(p4[idx].contiguous(), p4, cfg.eps, L, eng.ewald_table(L, 64), 64,
                               scale=cfg.G * cfg.mass).cpu().numpy()
    d = got - exact
    rel_rms, rel_max = np.sqrt((d ** 2).mean()) / rms, np.abs(d).max() / rms
    print(f"C5 P3M total vs direct Ewald sum over {pos.shape[0]} particles: rms {rel_rms:.2e} max {rel_max:.2e} of the rms force")
    assert rel_rms < 1e-2 and rel_max < 1e-1
