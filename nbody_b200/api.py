"""The reference's Python call surface, numpy in / numpy out, float64 at
the boundary, every function executed by the sm_100a kernels behind include/nbody_b200.h.

    cic_density(positions, boxsize, N)                      FraP_kTilDelta.py:94
    cic_assign(delta, coords, dx, Nmesh)                    Plotter.py:129
    compute_delta(density)                                  FraP_kTilDelta.py:126
    cic_window_ft(ngrid, boxsize)                           FraP_kTilDelta.py:73
    deconvolve_cic(delta_grid, boxsize, ngrid)              FraP_kTilDelta.py:87
    compute_power_spectrum_from_snapshot(path, L, Nmesh, deconvolve, use_fac)
                                                            Power_compare.py:135
    compute_power_spectrum(delta_x, fromcic, kx, ky, kz, L, N, withfac)
                                                            Power_compare.py:422
    read_positions / read_boxsize                           FraP_kTilDelta.py:55,63

`precision` selects the device storage of grids: "f32" (default, the performance path;
tolerances in DESIGN.md) or "f64" (parity mode; float64 grids, D2Z/Z2D cuFFT).
Errors are Python exceptions like the reference's (ValueError for bad input).
"""
from __future__ import annotations

import numpy as np
import torch

from . import tables
from .engine import default_engine
from .snapshot import read_boxsize, read_positions  # noqa: F401  (part of the surface)

_DT = {"f32": torch.float32, "f64": torch.float64}


def _dtype(precision):
    try:
        return _DT[precision]
    except KeyError:
        raise ValueError("precision must be 'f32' or 'f64'") from None


def _positions(positions):
    p = np.ascontiguousarray(positions, dtype=np.float64)
    if p.ndim != 2 or p.shape[1] != 3:
        raise ValueError("positions must have shape (Np, 3)")
    return p


def _out(t: torch.Tensor) -> np.ndarray:
    """device grid -> float64 numpy (the reference returns float64 arrays)."""
    return t.to(torch.float64).cpu().numpy()


# --------------------------------------------------------------------------------- a1
def cic_density(positions, boxsize, N, precision="f32", return_index=False):
    eng = default_engine()
    pos = eng.to_device(_positions(positions))
    dbg = torch.empty(pos.shape, dtype=torch.int32, device=eng.device) if return_index else None
    grid = eng.cic_deposit(pos, boxsize / N, int(N), dtype=_dtype(precision), dbg_index=dbg)
    if return_index:
        return _out(grid), dbg.cpu().numpy()
    return _out(grid)


def cic_assign(delta, coords, dx, Nmesh, precision=None):
    """In place like Plotter.py:128-145; the grid keeps its dtype (float32 or float64)."""
    if not isinstance(delta, np.ndarray) or delta.shape != (Nmesh,) * 3:
        raise ValueError("delta must be an (Nmesh,)*3 numpy array")
    if precision is None:
        precision = "f32" if delta.dtype == np.float32 else "f64"
    eng = default_engine()
    pos = eng.to_device(_positions(coords))
    g = eng.to_device(np.ascontiguousarray(delta), None).to(_dtype(precision))
    eng.cic_deposit(pos, float(dx), int(Nmesh), out=g, accumulate=True)
    delta[...] = g.cpu().numpy().astype(delta.dtype, copy=False)
    return delta


# --------------------------------------------------------------------------------- a2
def compute_delta(density, precision="f32"):
    eng = default_engine()
    g = eng.to_device(np.ascontiguousarray(density)).to(_dtype(precision))
    return _out(eng.compute_delta(g))


# --------------------------------------------------------------------------------- a3
def cic_window_ft(ngrid, boxsize, guard=True):
    """Host table (float64), separable form of FraP_kTilDelta.py:73-84."""
    s = tables.sinc_table(ngrid, boxsize)
    W = ((s[:, None, None] * s[None, :, None]) * s[None, None, :]) ** 2
    if guard:
        W[W == 0] = 1.0
    return W


def deconvolve_cic(delta_grid, boxsize, ngrid, precision="f32", guard=True):
    eng = default_engine()
    g = eng.to_device(np.ascontiguousarray(delta_grid)).to(_dtype(precision))
    if g.shape != (ngrid,) * 3:
        raise ValueError("delta_grid must have shape (ngrid,)*3")
    return _out(eng.deconvolve_cic(g, boxsize, int(ngrid), guard=guard))


def snapshot_to_delta(positions, boxsize, ngrid, subtract_mean=False, precision="f32"):
    """Feeder_pipe.py:174-195 in one resident pass: deposit -> delta -> deconvolve."""
    eng = default_engine()
    pos = eng.to_device(_positions(positions))
    g = eng.cic_deposit(pos, boxsize / ngrid, int(ngrid), dtype=_dtype(precision))
    eng.compute_delta(g)
    eng.deconvolve_cic(g, boxsize, int(ngrid), guard=True)
    if subtract_mean:
        # delta -= delta.mean()   (Feeder_pipe.py:193)
        eng.grid_affine(g, c=float(eng.grid_sum(g).item()) / g.numel())
    return _out(g)


def snapshots_to_delta(positions_list, boxsizes, ngrid, subtract_mean=False, precision="f32"):
    """Batched form of snapshot_to_delta for the snapshot corpus: every
    snapshot of `positions_list` goes through deposit -> delta -> deconvolve [-> minus mean] in ONE
    library call.  boxsizes: scalar or one per snapshot; subtract_mean: bool or one flag per
    snapshot.  Returns (B, ngrid, ngrid, ngrid) float64."""
    eng = default_engine()
    plist = [_positions(p) for p in positions_list]
    B = len(plist)
    if B == 0:
        return np.zeros((0, int(ngrid), int(ngrid), int(ngrid)))
    offsets = np.concatenate([[0], np.cumsum([p.shape[0] for p in plist])]).astype(np.int64)
    boxes = np.broadcast_to(np.asarray(boxsizes, dtype=np.float64), (B,)).copy()
    flags = np.broadcast_to(np.asarray(subtract_mean, dtype=bool), (B,)).astype(np.int32)
    pos = eng.to_device(np.concatenate(plist) if offsets[-1] else np.zeros((0, 3)))
    g = eng.snapshot_delta_batch(pos, offsets, boxes, int(ngrid), subtract_mean=flags if flags.any() else None,
                                 dtype=_dtype(precision))
    return _out(g)


def run_frap_to_delta(data_dir, ngrid, batch=64, precision="f32"):
    """Feeder_pipe.run_frap_to_delta (TheFeedingPipeline/Feeder_pipe.py:163-195) on the GPU: same
    directory layout in (`Test_Data_concept`, `TrainingVal_Concept`) and out (`Test data_CONCEPT/
    delta_test_id-<i>.npy`, `Training & val data_CONCEPT/delta_train_id-<i>.npy`, float64), the
    snapshots of each set processed `batch` at a time.  Returns the list of files written."""
    import os
    from . import snapshot
    written = []
    for sub_in, sub_out, stem, submean in (("Test_Data_concept", "Test data_CONCEPT", "delta_test_id-", False),
                                           ("TrainingVal_Concept", "Training & val data_CONCEPT", "delta_train_id-", True)):
        root, out = os.path.join(data_dir, sub_in), os.path.join(data_dir, sub_out)
        os.makedirs(out, exist_ok=True)
        files = sorted(snapshot.find_hdf5_files(root))
        for lo in range(0, len(files), int(batch)):
            chunk = files[lo:lo + int(batch)]
            pos = [snapshot.read_positions(f) for f in chunk]
            box = [snapshot.read_boxsize(f) for f in chunk]
            deltas = snapshots_to_delta(pos, box, ngrid, subtract_mean=submean, precision=precision)
            for k, d in enumerate(deltas):
                path = os.path.join(out, f"{stem}{lo + k + 1}.npy")
                np.save(path, d)
                written.append(path)
    return written


# --------------------------------------------------------------------------------- a4
def power_spectrum_from_positions(positions, L, Nmesh, deconvolve=False, use_fac=False,
                                  variant="power_compare", precision="f32", return_counts=False):
    """Body of compute_power_spectrum_from_snapshot after the snapshot read
    (Power_compare.py:150-200; variants Plotter.py:153-202, 1vs2tlptplotter.py:84-116)."""
    eng = default_engine()
    p = _positions(positions)
    npart = p.shape[0]
    pos = eng.to_device(p)
    Nmesh = int(Nmesh)
    g = eng.cic_deposit(pos, L / Nmesh, Nmesh, dtype=_dtype(precision))
    V = L ** 3
    if variant == "power_compare":
        if not use_fac:
            eng.grid_affine(g, a=(L / Nmesh) ** 3, b=npart / (L ** 3), c=1.0)
        else:
            eng.grid_affine(g, c=1.0)
        if deconvolve:
            eng.deconvolve_cic(g, L, Nmesh, guard=False)
        if use_fac:
            fac = Nmesh ** 3 * np.sqrt(2 * np.pi / V)
            norm = 1.0 / (fac ** 2)
        else:
            norm = V / (Nmesh ** 6)
        edges, counts, sums = eng.power_spectrum_bins(g, L, Nmesh, 300, norm=norm, closed_last=True)
    elif variant in ("plotter", "lpt"):
        eng.grid_affine(g, a=npart / Nmesh ** 3, c=1.0)
        edges, counts, sums = eng.power_spectrum_bins(
            g, L, Nmesh, 300 if variant == "plotter" else 100, window_power=2, clip_mode=2, clip_val=1e-3,
            norm=V / (Nmesh ** 6), closed_last=True, kq_form="piN_L")
    else:
        raise ValueError(f"unknown variant {variant!r}")
    counts = counts.cpu().numpy()
    sums = sums.cpu().numpy()
    centers = 0.5 * (edges[:-1] + edges[1:])
    valid = counts > 0
    out = (centers[valid], sums[valid] / counts[valid])
    return out + (counts,) if return_counts else out


def compute_power_spectrum_from_snapshot(snapshot_path, L, Nmesh, deconvolve=False, use_fac=False,
                                         variant="power_compare", precision="f32"):
    pos = read_positions(snapshot_path)
    return power_spectrum_from_positions(pos, L, Nmesh, deconvolve, use_fac, variant, precision)


def power_spectra_from_positions(positions_list, boxsizes, Nmesh, variant="plotter"):
    """Batched form of power_spectrum_from_positions for snapshots that share the mesh size (one library
    call, nb_snapshot_pk_batch): returns [(k_centers[valid], Pk[valid])] in input order.
    variant "plotter" = Different_Sizes/Plotter.py:148-202 (W clipped at 1e-3, δ_k / W), "lpt" =
    1LPTvs2LPT/1vs2tlptplotter.py:81-116 (100 edges), "power_compare" = Power_compare.py:135-200 with
    deconvolve=False, use_fac=False."""
    eng = default_engine()
    plist = [_positions(p) for p in positions_list]
    B = len(plist)
    if B == 0:
        return []
    Nmesh = int(Nmesh)
    offsets = np.concatenate([[0], np.cumsum([p.shape[0] for p in plist])]).astype(np.int64)
    boxes = np.broadcast_to(np.asarray(boxsizes, dtype=np.float64), (B,)).copy()
    pos = eng.to_device(np.concatenate(plist))
    if variant in ("plotter", "lpt"):
        kw = dict(n_edges=300 if variant == "plotter" else 100, window_power=2, clip_mode=2, clip_val=1e-3,
                  kq_form="piN_L")
    elif variant == "power_compare":
        kw = dict(n_edges=300)
    else:
        raise ValueError(f"unknown variant {variant!r}")
    edges, counts, sums = eng.snapshot_pk_batch(pos, offsets, boxes, Nmesh, closed_last=True, **kw)
    counts, sums = counts.cpu().numpy(), sums.cpu().numpy()
    out = []
    for b in range(B):
        centers = 0.5 * (edges[b][:-1] + edges[b][1:])
        v = counts[b] > 0
        out.append((centers[v], sums[b][v] / counts[b][v]))
    return out


def power_spectra_of_runs(root_dir, pattern="3giant*", variant="plotter"):
    """The main block of Different_Sizes/Plotter.py:205-242 without the process pool: every run directory
    `<root>/3giant<N>(pm|p3m)` contributes its first snapshot, mesh size = N of the directory name, boxsize from
    the snapshot's attributes; runs that share a mesh size go through ONE batched call.  Returns
    [((method, N_in, ngrid, boxsize), (k, Pk))] sorted by N_in like the reference's `plot_data`."""
    import glob
    import os
    import re
    from . import snapshot
    jobs = []
    for run in sorted(glob.glob(os.path.join(root_dir, pattern))):
        snaps = sorted(snapshot.find_hdf5_files(run))
        m = re.search(r"giant(\d+)(p3m|pm)", run)
        if not snaps or not m:
            continue
        n_in = int(m.group(1))
        jobs.append(((m.group(2), n_in, n_in, float(snapshot.read_boxsize(snaps[0]))), snaps[0]))
    results = [None] * len(jobs)
    for ngrid in sorted({j[0][2] for j in jobs}):
        idx = [i for i, j in enumerate(jobs) if j[0][2] == ngrid]
        pos = [snapshot.read_positions(jobs[i][1]) for i in idx]
        for i, r in zip(idx, power_spectra_from_positions(pos, [jobs[i][0][3] for i in idx], ngrid, variant)):
            results[i] = r
    data = sorted(zip([j[0] for j in jobs], results), key=lambda x: x[0][1])
    return data


# --------------------------------------------------------------------------------- a5
def compute_power_spectrum(delta_x, fromcic=False, kx=0, ky=0, kz=0, L=0, N=0, withfac=False,
                           variant="power_compare", precision="f32", return_counts=False):
    """kx, ky, kz are accepted for signature compatibility; the window is rebuilt from the
    separable table (it is a function of N and L only)."""
    eng = default_engine()
    N = int(N)
    g = eng.to_device(np.ascontiguousarray(delta_x)).to(_dtype(precision))
    if g.shape != (N, N, N):
        raise ValueError("delta_x must have shape (N,N,N)")
    norm = 1.0
    if withfac or variant == "cic3d":
        fac = np.sqrt(2 * np.pi / L ** 3) * N ** 3
        norm = 1.0 / fac ** 2
    wp, cm, cv = 0, 0, 0.0
    if fromcic:
        if variant == "plotter":
            wp, cm, cv = 1, 2, 0.1
        elif variant == "cic3d":
            wp, cm = 2, 1
        else:
            wp = 2
    edges, counts, sums = eng.power_spectrum_bins(g, L, N, 300, window_power=wp, clip_mode=cm, clip_val=cv,
                                                  norm=norm, closed_last=False, kq_form="piN_L")
    counts = counts.cpu().numpy()
    sums = sums.cpu().numpy()
    centers = 0.5 * (edges[1:] + edges[:-1])
    valid = counts > 0
    out = (centers[valid], sums[valid] / counts[valid])
    return out + (counts,) if return_counts else out


# --------------------------------------------------------------------------------- a6
def _draws(N, c1, c2, seed):
    """c1 then c2 from the legacy global RNG like the reference (c1 drawn fully before c2)."""
    if c1 is None or c2 is None:
        if seed is not None:
            np.random.seed(seed)
        c1 = np.random.normal(0, 1 / np.sqrt(2), size=[N] * 3)
        c2 = np.random.normal(0, 1 / np.sqrt(2), size=[N] * 3)
    c1 = np.ascontiguousarray(c1, dtype=np.float64)
    c2 = np.ascontiguousarray(c2, dtype=np.float64)
    if c1.shape != (N, N, N) or c2.shape != (N, N, N):
        raise ValueError("c1 and c2 must have shape (N,N,N)")
    return c1, c2


def createR(N, dim=3, mu=0, sig=1 / np.sqrt(2), c1=None, c2=None, seed=None, guard=False):
    """Cosmo data create (new).py:24-47: Hermitian complex noise R (N,N,N) complex128.  The
    draws come from numpy's legacy RNG on the host (bit-identical to the reference); the
    symmetrisation loop runs on the GPU.  guard=True gives the `if`-guarded variant of
    CiC/3D startbetingelser1.py:82-93."""
    if dim != 3:
        raise ValueError("only dim=3 is supported")
    if c1 is None or c2 is None:
        if seed is not None:
            np.random.seed(seed)
        c1 = np.random.normal(mu, sig, size=[N] * 3)
        c2 = np.random.normal(mu, sig, size=[N] * 3)
    c1, c2 = _draws(N, c1, c2, None)
    eng = default_engine()
    R = eng.hermitian_R(eng.to_device(c1), eng.to_device(c2), guard)
    return R.cpu().numpy()


# --------------------------------------------------------------------------------- a7
def find_delta(z=None, L=None, N=None, A_s=None, n_s=None, omega_b=None, omega_cdm=None, *, Pk_grid=None,
               c1=None, c2=None, seed=None, guard=False, use_fac=True, precision="f32"):
    """Cosmo data create (new).py:114-137: δ = ifftn(sqrt(P)·fac·R).real.
    CLASS is not part of this package: pass `Pk_grid`, P(k) evaluated on the (N,N,N) FFT grid
    `k_mag` (with k_mag[0,0,0] = 1e-10 like the reference, :304-307); the cosmological
    arguments are accepted for signature compatibility and ignored."""
    if Pk_grid is None:
        raise ValueError("find_delta needs Pk_grid (CLASS is not available in this package)")
    N = int(N)
    pk = np.ascontiguousarray(Pk_grid, dtype=np.float64).reshape(N, N, N)
    c1, c2 = _draws(N, c1, c2, seed)
    eng = default_engine()
    delta, _ = eng.grf_synthesize(eng.to_device(pk), eng.to_device(c1), eng.to_device(c2), L, N, guard,
                                  use_fac=use_fac, dtype=_dtype(precision), want_delta=True, want_psi=False)
    return _out(delta)


def grf_delta_psi(Pk_grid, L, N, c1=None, c2=None, seed=None, guard=True, precision="f32"):
    """CLASS_simualtor without CLASS (Power_compare.py:279-331, CiC/3D startbetingelser1.py:
    82-112): returns delta_x and (psi_x, psi_y, psi_z)."""
    N = int(N)
    pk = np.ascontiguousarray(Pk_grid, dtype=np.float64).reshape(N, N, N)
    c1, c2 = _draws(N, c1, c2, seed)
    eng = default_engine()
    delta, psi = eng.grf_synthesize(eng.to_device(pk), eng.to_device(c1), eng.to_device(c2), L, N, guard,
                                    dtype=_dtype(precision), want_delta=True, want_psi=True)
    psi = _out(psi)
    return _out(delta), (psi[0], psi[1], psi[2])


# --------------------------------------------------------------------------------- a8
def trilinear_gather(field, positions, L, N, precision="f64"):
    """Periodic CIC gather of `field` ((N,N,N) or (ncomp,N,N,N)) at `positions`."""
    eng = default_engine()
    f = eng.to_device(np.ascontiguousarray(field)).to(_dtype(precision)).contiguous()
    pos = eng.to_device(_positions(positions))
    out = eng.cic_gather(f, pos, L / N, int(N)).cpu().numpy()
    return out[:, 0] if np.ndim(field) == 3 else out


def particle_assigment(psi, L, N, precision="f32", return_particles=False):
    """Power_compare.py:336-418: lattice particles displaced by ψ (gathered at the lattice
    nodes), wrapped with `particles %= L`, CIC-deposited, minus 1."""
    N = int(N)
    eng = default_engine()
    x = np.linspace(0, L, N, endpoint=False)
    lattice = np.array(np.meshgrid(x, x, x, indexing="ij")).reshape(3, -1).T
    pos = eng.to_device(np.ascontiguousarray(lattice))
    f = eng.to_device(np.ascontiguousarray(np.stack([np.asarray(p, dtype=np.float64) for p in psi])))
    disp = eng.cic_gather(f.contiguous(), pos, L / N, N)            # ψ at the lattice nodes (exact)
    eng.displace_wrap(pos, disp, L)
    g = eng.cic_deposit(pos, L / N, N, dtype=_dtype(precision))
    eng.grid_affine(g, c=1.0)
    return (_out(g), pos.cpu().numpy()) if return_particles else _out(g)
