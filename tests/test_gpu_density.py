"""GPU parity: CIC deposit, overdensity, deconvolution and P(k) binning through the C-ABI
(nbody_b200.api -> ctypes -> libnbody_b200.so) against the float64 oracle and the
reference's golden vectors.

Tolerances (north_star / SURVEY §8a'):
  cell index (i,j,k)           exact
  shell-bin counts             exact (np.histogram and np.digitize semantics)
  CIC grid, delta              max|Δ| / max|ref| <= 1e-5   (float32 grid);  1e-12 (float64 grid)
  binned P(k)                  per non-empty bin |ΔP|/P <= 1e-5
"""
import numpy as np
import pytest

from conftest import golden, zeldovich_particles
from oracle import restate

pytestmark = pytest.mark.gpu

TOL = 1e-5


@pytest.fixture(scope="module")
def api():
    import torch
    assert torch.cuda.is_available()
    from nbody_b200 import api as _api
    return _api


def nrm(a, ref):
    return np.abs(a - ref).max() / np.abs(ref).max()


@pytest.mark.parametrize("tag", ["test1", "test57", "train1", "train200"])
@pytest.mark.parametrize("prec,tol", [("f32", TOL), ("f64", 1e-12)])
def test_golden_snapshot_to_delta(api, tag, prec, tol):
    g = golden(f"snap_{tag}.npz")
    d = api.snapshot_to_delta(g["pos"], float(g["boxsize"]), 32, subtract_mean=tag.startswith("train"),
                              precision=prec)
    assert d.dtype == np.float64 and d.shape == (32, 32, 32)
    assert nrm(d, g["delta_golden"]) <= tol


@pytest.mark.parametrize("N", [32, 64, 50])
def test_deposit_index_exact_and_grid(api, snap_test1, N):
    pos, L, _ = snap_test1
    grid, idx = api.cic_density(pos, L, N, return_index=True)
    ref_idx, _ = restate.cic_index_weight(pos, L, N)
    assert np.array_equal(idx, ref_idx)
    ref = restate.cic_density(pos, L, N)
    assert nrm(grid, ref) <= TOL
    assert abs(grid.sum() - pos.shape[0]) <= 1e-5 * pos.shape[0]
    g64 = api.cic_density(pos, L, N, precision="f64")
    assert nrm(g64, ref) <= 1e-13


def test_deposit_edge_cases(api):
    L, N = 500.0, 16
    cell = L / N
    # empty input
    g = api.cic_density(np.zeros((0, 3)), L, N)
    assert g.shape == (N, N, N) and not g.any()
    # exactly on nodes, on the upper box edge, out of the box on both sides
    pos = np.array([[0.0, 0.0, 0.0], [cell, 2 * cell, 3 * cell], [L, L, L], [L - 1e-9, 1e-9, L / 2],
                    [-0.3 * cell, 5.2 * cell, -7.9 * cell], [L + 0.25 * cell, 2 * L + 1.0, -L - 3.0],
                    [np.nextafter(L, 0), np.nextafter(cell, 0), np.nextafter(cell, L)]])
    grid, idx = api.cic_density(pos, L, N, precision="f64", return_index=True)
    ref_idx, _ = restate.cic_index_weight(pos, L, N)
    assert np.array_equal(idx, ref_idx)
    ref = restate.cic_density(pos, L, N)
    assert np.abs(grid - ref).max() <= 1e-14
    # ragged: a single particle
    one = api.cic_density(pos[3:4], L, N, precision="f64")
    assert np.abs(one - restate.cic_density(pos[3:4], L, N)).max() <= 1e-15


@pytest.mark.parametrize("L,N", [(500.0, 16), (1000.0, 48), (1000.0, 256), (500.0, 50), (123.456, 37)])
def test_deposit_division_free_index_adversarial(api, L, N):
    """The float32-grid deposit derives the cell index without IEEE divisions (nb_cell_split);
    it must still equal int(p/cell) % N bit for bit: nodes, +-1..3 ulp around nodes, box edges,
    out-of-box and negative coordinates, plus random points."""
    rng = np.random.default_rng(11)
    cell = L / N
    nodes = rng.integers(0, N + 1, size=4000).astype(np.float64) * cell
    near = nodes.copy()
    for _ in range(3):
        near = np.concatenate([np.nextafter(near, -np.inf), nodes, np.nextafter(near, np.inf)])[:40000]
    extra = np.array([0.0, -0.0, L, np.nextafter(L, 0), np.nextafter(L, 2 * L), -1e-300, 1e-300, -0.3 * cell,
                      -7.9 * cell, L + 0.25 * cell, 2 * L + 1.0, -L - 3.0, 5e-324, 1048575.5 * cell, 3e6 * cell])
    vals = np.concatenate([near, extra, rng.random(50000) * L, rng.random(2000) * 3 * L - L])
    vals = vals[: (vals.size // 3) * 3]
    pos = rng.permutation(vals).reshape(-1, 3)
    grid, idx = api.cic_density(pos, L, N, return_index=True)       # float32 grid: division-free path
    ref_idx, _ = restate.cic_index_weight(pos, L, N)
    assert np.array_equal(idx, ref_idx)
    ref = restate.cic_density(pos, L, N)
    assert nrm(grid, ref) <= TOL


def test_cic_assign_in_place_float32_and_float64(api, snap_test1):
    pos, L, _ = snap_test1
    N = 32
    for dt in (np.float32, np.float64):
        d = np.ones((N, N, N), dtype=dt)
        out = api.cic_assign(d, pos, L / N, N)
        assert out is d and d.dtype == dt
        ref = restate.cic_assign(np.ones((N, N, N), dtype=np.float64), pos, L / N, N)
        assert nrm(d.astype(np.float64), ref) <= (TOL if dt == np.float32 else 1e-13)


def test_compute_delta_and_deconvolve(api, snap_test1):
    pos, L, _ = snap_test1
    dens = restate.cic_density(pos, L, 32)
    d = api.compute_delta(dens, precision="f64")
    assert np.abs(d - restate.compute_delta(dens)).max() <= 1e-13
    assert abs(d.mean()) < 1e-12
    ref = restate.deconvolve_cic(restate.compute_delta(dens), L, 32)
    assert nrm(api.deconvolve_cic(d, L, 32, precision="f64"), ref) <= 1e-12
    assert nrm(api.deconvolve_cic(d, L, 32, precision="f32"), ref) <= TOL


@pytest.mark.parametrize("dec", [False, True])
@pytest.mark.parametrize("fac", [False, True])
def test_pk_snapshot_vs_reference_golden(api, snap_test1, dec, fac):
    pos, L, _ = snap_test1
    g = golden("ref_pk_snapshot.npz")
    k, p, c = api.power_spectrum_from_positions(pos, L, 32, deconvolve=dec, use_fac=fac, return_counts=True)
    kr, pr, cr = restate.power_spectrum_from_positions(pos, L, 32, deconvolve=dec, use_fac=fac,
                                                       return_counts=True)
    assert np.array_equal(c, cr)                       # bin counts bit-exact (histogram semantics)
    assert c.sum() == 32 ** 3 - 1
    assert np.array_equal(k, g[f"pc_dec{int(dec)}_fac{int(fac)}_k"])
    assert np.abs(p / g[f"pc_dec{int(dec)}_fac{int(fac)}_p"] - 1).max() <= TOL
    assert np.abs(p / pr - 1).max() <= TOL


def test_pk_variants_and_finer_mesh(api, snap_test1):
    pos, L, _ = snap_test1
    g = golden("ref_pk_snapshot.npz")
    k, p = api.power_spectrum_from_positions(pos, L, 64)
    assert np.array_equal(k, g["pc_mesh64_k"])
    assert np.abs(p / g["pc_mesh64_p"] - 1).max() <= TOL
    gp = golden("snap_giant32pm.npz")
    for variant in ("plotter", "lpt"):
        k, p, c = api.power_spectrum_from_positions(gp["pos"], 1000.0, 32, variant=variant, return_counts=True)
        kr, pr, cr = restate.power_spectrum_from_positions(gp["pos"], 1000.0, 32, variant=variant,
                                                           return_counts=True)
        assert np.array_equal(c, cr) and np.array_equal(k, kr)
        assert np.abs(p / pr - 1).max() <= TOL
        assert np.array_equal(k, g[f"{variant}_giant_k"])
        # the reference's own 'plotter' run is float32 end to end (see test_oracle_golden)
        assert np.abs(p / g[f"{variant}_giant_p"] - 1).max() <= (2e-3 if variant == "plotter" else TOL)


def test_pk_field_digitize_semantics(api):
    g = golden("ref_pk_field.npz")
    L, N = 500.0, 16
    for wf in (0, 1):
        k, p, c = api.compute_power_spectrum(g["delta_x"], False, L=L, N=N, withfac=bool(wf), return_counts=True)
        _, _, cr = restate.compute_power_spectrum(g["delta_x"], False, L, N, withfac=bool(wf), return_counts=True)
        assert np.array_equal(c, cr) and c.sum() == N ** 3 - 2       # digitize drops k == k_max (Q3)
        assert np.array_equal(k, g[f"pc_fft_fac{wf}_k"])
        assert np.abs(p / g[f"pc_fft_fac{wf}_p"] - 1).max() <= TOL
        k, p = api.compute_power_spectrum(g["delta_cic"], True, L=L, N=N, withfac=bool(wf))
        assert np.array_equal(k, g[f"pc_cic_fac{wf}_k"])
        assert np.abs(p / g[f"pc_cic_fac{wf}_p"] - 1).max() <= TOL
    k, p = api.compute_power_spectrum(g["delta_cic"], True, L=L, N=N, variant="plotter")
    assert np.array_equal(k, g["plotter_cic_k"])
    assert np.abs(p / g["plotter_cic_p"] - 1).max() <= TOL
    # float64 device path agrees to round-off
    k, p = api.compute_power_spectrum(g["delta_cic"], True, L=L, N=N, precision="f64")
    assert np.abs(p / g["pc_cic_fac0_p"] - 1).max() <= 1e-11


@pytest.mark.parametrize("n,N", [(32, 64), (48, 96), (33, 45)])
def test_zeldovich_set_deposit_and_pk(api, n, N):
    L = 1000.0
    pos = zeldovich_particles(n, L)
    grid, idx = api.cic_density(pos, L, N, return_index=True)
    ref_idx, _ = restate.cic_index_weight(pos, L, N)
    assert np.array_equal(idx, ref_idx)
    assert nrm(grid, restate.cic_density(pos, L, N)) <= TOL
    k, p, c = api.power_spectrum_from_positions(pos, L, N, return_counts=True)
    kr, pr, cr = restate.power_spectrum_from_positions(pos, L, N, return_counts=True)
    assert np.array_equal(c, cr)
    assert np.abs(p / pr - 1).max() <= TOL


def test_full_size_properties_c2(api):
    """BASELINE config 2 size (128^3 particles -> 256^3 mesh): size-independent properties."""
    import torch
    from nbody_b200.engine import default_engine
    eng = default_engine()
    n, N, L = 128, 256, 1000.0
    gen = torch.Generator(device="cuda").manual_seed(5)
    lat = torch.arange(n, device="cuda", dtype=torch.float64) * (L / n)
    pos = torch.stack(torch.meshgrid(lat, lat, lat, indexing="ij"), -1).reshape(-1, 3)
    pos = (pos + (torch.rand(pos.shape, generator=gen, device="cuda", dtype=torch.float64) - 0.5) * (L / n)) % L
    pos = pos.contiguous()
    grid = eng.cic_deposit(pos, L / N, N)
    total = float(eng.grid_sum(grid).item())
    assert abs(total - n ** 3) <= 1e-6 * n ** 3                    # mass conservation
    g2 = eng.cic_deposit(pos.flip(0).contiguous(), L / N, N)       # order independence
    assert float((grid - g2).abs().max()) <= 1e-5 * float(grid.abs().max())
    eng.grid_affine(grid, a=n ** 3 / N ** 3, c=1.0)
    edges, counts, sums = eng.power_spectrum_bins(grid, L, N, 300, norm=L ** 3 / N ** 6)
    assert int(counts.sum()) == N ** 3 - 1                         # every k>0 mode lands in a bin
    km = restate.k_mag_grid(N, L).ravel()
    cr, _ = restate.shell_bin(km, np.ones_like(km), restate.pk_edges(N, L), closed_last=True)
    assert np.array_equal(counts.cpu().numpy(), cr)                # counts are geometry only
    # Parseval: sum over k>0 of |delta_k|^2 = N^3 * sum(delta^2) - |delta_0|^2
    lhs = float(sums.sum()) / (L ** 3 / N ** 6)
    d = grid.double()
    rhs = float(N ** 3 * (d * d).sum() - d.sum() ** 2)
    assert abs(lhs / rhs - 1) <= 1e-5


# ------------------------------------------------------------- §8f row 4: batched snapshot -> delta
@pytest.mark.parametrize("prec,tol", [("f32", TOL), ("f64", 1e-12)])
def test_batched_snapshots_match_the_shipped_golden_deltas(api, prec, tol):
    """Feeder_pipe.run_frap_to_delta as ONE library call: the four committed CONCEPT snapshots
    (two of the test set, two of the training set with `delta -= delta.mean()`) plus a ragged
    random set; each against the reference's own shipped delta_*.npy / the float64 oracle."""
    tags = ["test1", "test57", "train1", "train200"]
    gs = [golden(f"snap_{t}.npz") for t in tags]
    rng = np.random.default_rng(3)
    extra = rng.random((777, 3)) * 250.0
    pos = [g["pos"] for g in gs] + [extra, np.zeros((0, 3)) + 1.0]
    box = [float(g["boxsize"]) for g in gs] + [250.0, 250.0]
    flags = [t.startswith("train") for t in tags] + [True, False]
    d = api.snapshots_to_delta(pos[:5], box[:5], 32, subtract_mean=flags[:5], precision=prec)
    assert d.shape == (5, 32, 32, 32) and d.dtype == np.float64
    for k, g in enumerate(gs):
        assert nrm(d[k], g["delta_golden"]) <= tol
    ref = restate.deconvolve_cic(restate.compute_delta(restate.cic_density(extra, 250.0, 32)), 250.0, 32)
    ref = ref - ref.mean()
    assert nrm(d[4], ref) <= tol
    # the batched call equals the one-snapshot call
    one = api.snapshot_to_delta(gs[2]["pos"], box[2], 32, subtract_mean=True, precision=prec)
    assert nrm(d[2], one) <= (1e-6 if prec == "f32" else 1e-13)


def test_run_frap_to_delta_directory_layout(api, tmp_path, monkeypatch):
    """Same input/output directory layout and file names as Feeder_pipe.py:163-195."""
    from nbody_b200 import snapshot
    gs = {t: golden(f"snap_{t}.npz") for t in ("test1", "test57", "train1")}
    files = {"Test_Data_concept": ["a_test1.hdf5", "b_test57.hdf5"], "TrainingVal_Concept": ["a_train1.hdf5"]}
    for sub, names in files.items():
        (tmp_path / sub).mkdir()
        for n in names:
            (tmp_path / sub / n).write_bytes(b"")
    tag_of = lambda f: f.split("_")[-1].split(".")[0]
    monkeypatch.setattr(snapshot, "read_positions", lambda f: gs[tag_of(f)]["pos"])
    monkeypatch.setattr(snapshot, "read_boxsize", lambda f: float(gs[tag_of(f)]["boxsize"]))
    written = api.run_frap_to_delta(str(tmp_path), 32, batch=2)
    assert [w.split("/")[-2:] for w in written] == [["Test data_CONCEPT", "delta_test_id-1.npy"],
                                                    ["Test data_CONCEPT", "delta_test_id-2.npy"],
                                                    ["Training & val data_CONCEPT", "delta_train_id-1.npy"]]
    for w, t in zip(written, ("test1", "test57", "train1")):
        d = np.load(w)
        assert d.dtype == np.float64 and nrm(d, gs[t]["delta_golden"]) <= TOL


@pytest.mark.parametrize("grid,N", [(30, 64), (11, 64), (7, 45), (1, 32)])
def test_pk_binning_split_tail_units(api, monkeypatch, grid, N):
    """The shell-binning kernel cuts the groups left after the last full round into kz pieces spread
    over all warps (1024^3 on a full B200: 208 groups in 14 pieces each).  NB_PK_GRID forces a small
    grid so that meshes of 32-64 cells reach that path — (30, 64): 8 left-over groups in 4 pieces,
    (11, 64): 40 in 1, (7, 45): odd mesh, (1, 32): one CTA: counts stay bit-exact, P within 1e-5."""
    monkeypatch.setenv("NB_PK_GRID", str(grid))
    L = 700.0
    rs = np.random.RandomState(grid * 100 + N)
    delta = rs.normal(size=(N, N, N))
    for fromcic in (False, True):
        k, p, c = api.compute_power_spectrum(delta, fromcic, L=L, N=N, return_counts=True)
        kr, pr, cr = restate.compute_power_spectrum(delta, fromcic, L, N, return_counts=True)
        assert np.array_equal(c, cr) and np.array_equal(k, kr)
        assert np.abs(p / pr - 1).max() <= TOL


def test_batched_snapshot_pk_matches_the_single_calls_and_the_reference(api, snap_test1, tmp_path):
    """nb_snapshot_pk_batch (Different_Sizes/Plotter.py:205-238 without the process pool): ragged particle
    sets, two box sizes, all three estimator variants; equal to the per-snapshot calls (same kernels) and to the
    reference's own output for the 3giant32pm snapshot.  The run-directory walker reproduces the reference's
    (method, N_in, ngrid, boxsize) bookkeeping on a directory of written snapshots."""
    from nbody_b200 import snapshot
    pos1, L1, _ = snap_test1
    gp = golden("snap_giant32pm.npz")
    g = golden("ref_pk_snapshot.npz")
    sets = [gp["pos"], pos1[:30011], pos1, gp["pos"][5:]]
    boxes = [1000.0, L1, L1, 1000.0]
    for variant in ("plotter", "lpt", "power_compare"):
        got = api.power_spectra_from_positions(sets, boxes, 32, variant=variant)
        assert len(got) == 4
        for (k, p), pos, L in zip(got, sets, boxes):
            ks, ps = api.power_spectrum_from_positions(pos, L, 32, variant=variant)
            assert np.array_equal(k, ks) and np.abs(p / ps - 1).max() <= TOL   # two deposit kernels: float32 sums in another order
        if variant != "power_compare":
            assert np.array_equal(got[0][0], g[f"{variant}_giant_k"])
            assert np.abs(got[0][1] / g[f"{variant}_giant_p"] - 1).max() <= (2e-3 if variant == "plotter" else TOL)
    assert api.power_spectra_from_positions([], [], 32) == []
    # directory walk: two runs with the reference's naming, one of each method
    for name, pos in (("3giant32pm", gp["pos"]), ("3giant32p3m", gp["pos"][::-1].copy())):
        d = tmp_path / name / "snapshots"
        d.mkdir(parents=True)
        snapshot.write_snapshot(str(d / "snapshot_a=1.00.hdf5"), pos, np.zeros_like(pos), boxsize=1000.0)
    (tmp_path / "notes").mkdir()
    data = api.power_spectra_of_runs(str(tmp_path))
    assert [m for m, _r in data] == [("p3m", 32, 32, 1000.0), ("pm", 32, 32, 1000.0)]
    for _m, (k, p) in data:
        assert np.array_equal(k, g["plotter_giant_k"]) and np.abs(p / g["plotter_giant_p"] - 1).max() <= 2e-3
