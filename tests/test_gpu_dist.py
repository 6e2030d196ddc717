"""GPU parity of the slab-decomposed (multi-GPU) path, with the ranks emulated as threads on
ONE device (nbody_b200.dist.ThreadComm): the sharded PM / P3M / PP accelerations, the kick-drift
step with particle migration, and the slab P(k) must equal the single-GPU path and the float64
oracle on the same inputs.  The real NCCL run (bench.py --gpus N) executes the same per-rank
code with TorchComm in place of ThreadComm."""
import numpy as np
import pytest
import torch

from conftest import zeldovich_particles
from oracle import force as oforce
from oracle import restate

pytestmark = pytest.mark.gpu
TOL = 1e-5


def cfg_for(method, n, N, L, **kw):
    from nbody_b200.params import ForceConfig
    return ForceConfig(method=method, npart=n ** 3, boxsize=L, gridsize=N, **kw)


def err(a, ref):
    return np.abs(a - ref).max() / np.sqrt((ref ** 2).mean())


def owner_np(pos, L, N, world):
    return ((pos[:, 0] / (L / N)).astype(np.int64) % N) // (N // world)


def run_ranks(world, cfg, pos, fn, method=None):
    """fn(stepper, pos_dev_shard, idx) on every emulated rank; returns list of results."""
    from nbody_b200.dist import DistForceStepper, run_threaded
    from nbody_b200.engine import Engine
    method = method or cfg.method
    if method == "pp":
        chunk = (pos.shape[0] + world - 1) // world
        idx = [np.arange(r * chunk, min(pos.shape[0], (r + 1) * chunk)) for r in range(world)]
    else:
        own = owner_np(pos, cfg.boxsize, cfg.gridsize, world)
        idx = [np.nonzero(own == r)[0] for r in range(world)]

    def body(comm):
        torch.cuda.set_device(0)
        eng = Engine(0)
        st = DistForceStepper(cfg, eng, comm=comm)
        p = eng.to_device(pos[idx[comm.rank]])
        out = fn(st, p, idx[comm.rank])
        torch.cuda.synchronize()
        return out

    return run_threaded(world, body), idx


@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("method,interlace", [("pm", (False, False)), ("pm", (True, True)), ("p3m", (True, True))])
def test_sharded_accel_matches_oracle(world, method, interlace):
    L, n, N = 1000.0, 16, 32
    pos = zeldovich_particles(n, L, seed=11)
    cfg = cfg_for(method, n, N, L, interlace=interlace, deconvolve=(True, True), G=1.0, mass=2.0)
    if method == "p3m" and world == 4:
        cfg = cfg_for(method, n, 64, L, interlace=interlace, deconvolve=(True, True), G=1.0, mass=2.0)  # slab >= 2 r_cut
    res, idx = run_ranks(world, cfg, pos, lambda st, p, i: st.accel(p).cpu().numpy())
    a = np.empty_like(pos)
    for r in range(world):
        a[idx[r]] = res[r]
    assert err(a, oforce.accel(cfg, pos)) <= TOL


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_pp(world):
    L, n = 500.0, 10
    pos = zeldovich_particles(n, L, seed=5)[:997]          # ragged: 997 is not divisible by the world size
    cfg = cfg_for("pp", n, 0, L, G=1.5, mass=2.0)
    cfg.npart = pos.shape[0]
    res, idx = run_ranks(world, cfg, pos, lambda st, p, i: st.accel(p, n_total=pos.shape[0]).cpu().numpy())
    a = np.vstack(res)
    assert err(a, oforce.accel(cfg, pos)) <= TOL


@pytest.mark.parametrize("device_exchange", [True, False])
@pytest.mark.parametrize("method,world,overlap", [("pm", 2, False), ("p3m", 2, False), ("p3m", 2, True), ("pm", 4, False)])
def test_sharded_step_with_migration(method, world, overlap, device_exchange):
    """A large drift moves particles across slab faces: after the step every rank holds exactly
    the particles of its slab and the union equals the oracle's step.  Both exchange implementations: the
    device kernels (nb_halo_select / nb_migrate_*, one host sync per step) and the eager torch bookkeeping.
    overlap: the pair kernel confined to an SM subset on a side stream beside the mesh chain (here two emulated ranks
    share one GPU, so the confined launch finds the SMs busy and the unconfined second launch has real work)."""
    L, n, N = 500.0, 12, 24
    pos = zeldovich_particles(n, L, seed=21)
    rs = np.random.RandomState(1)
    mom = rs.normal(scale=30.0, size=pos.shape)
    cfg = cfg_for(method, n, N, L, interlace=(True, True), G=1.0, mass=1.5, a=0.5)

    def fn(st, p, i):
        st.device_exchange = device_exchange
        st.halo_cap_floor = 64          # the capacity formula (identical on all ranks) decides, not the floor
        st.overlap, st.p3m_sms = overlap, 40
        m = st.eng.to_device(mom[i])
        p2, m2 = st.step(p, m, 0.3, 0.7)
        return p2.cpu().numpy(), m2.cpu().numpy()

    res, idx = run_ranks(world, cfg, pos, fn)
    acc = oforce.accel(cfg, pos)
    pr, mr = oforce.kick_drift(pos, mom, acc, L, mass=cfg.mass, a=cfg.a, dt_kick=0.3, dt_drift=0.7)
    got_p = np.vstack([r[0] for r in res])
    got_m = np.vstack([r[1] for r in res])
    assert got_p.shape == pr.shape
    moved = sum(int((owner_np(pos[idx[r]], L, N, world) != owner_np(pr[idx[r]], L, N, world)).sum())
                for r in range(world))
    assert moved > 0, "test must exercise migration"
    for r in range(world):     # ownership after migration
        assert (owner_np(res[r][0], L, N, world) == r).all()
    # match particles by momentum (unique floats), then compare
    o1 = np.lexsort((got_m[:, 2], got_m[:, 1], got_m[:, 0]))
    o2 = np.lexsort((mr[:, 2], mr[:, 1], mr[:, 0]))
    assert np.abs(got_m[o1] - mr[o2]).max() <= TOL * np.abs(mr).max()
    d = np.abs(got_p[o1] - pr[o2])
    d = np.minimum(d, L - d)
    assert d.max() <= 1e-5 * cfg.cell


@pytest.mark.parametrize("method", ["pm", "p3m", "pp"])
def test_sharded_step_host_matches_resident_step(method):
    """DistForceStepper.step_host (pinned host shard in, migrated shard out, momentum upload on a side stream) gives
    the particles of the resident sharded step."""
    L, n, N, world = 500.0, 12, 24, 2
    pos = zeldovich_particles(n, L, seed=22)
    mom = np.random.RandomState(2).normal(scale=30.0, size=pos.shape)
    cfg = cfg_for(method, n, N, L, interlace=(True, True), G=1.0, mass=1.5, a=0.5)

    def fn(st, p, i):
        m = st.eng.to_device(mom[i])
        hp = torch.from_numpy(pos[i]).pin_memory()
        hm = torch.from_numpy(mom[i]).pin_memory()
        cap = len(i) + 512
        op = torch.empty((cap, 3), dtype=torch.float64).pin_memory()
        om = torch.empty((cap, 3), dtype=torch.float64).pin_memory()
        k = st.step_host(hp, hm, op, om, 0.3, 0.7, n_total=pos.shape[0])
        torch.cuda.synchronize()
        host = (op[:k].numpy().copy(), om[:k].numpy().copy())
        st2 = type(st)(cfg, st.eng, comm=st.comm)
        p2, m2 = st2.step(p, m, 0.3, 0.7, n_total=pos.shape[0])
        return host, (p2.cpu().numpy(), m2.cpu().numpy())

    res, _ = run_ranks(world, cfg, pos, fn)
    for host, resident in res:
        assert host[0].shape == resident[0].shape
        o1 = np.lexsort((host[1][:, 2], host[1][:, 1], host[1][:, 0]))
        o2 = np.lexsort((resident[1][:, 2], resident[1][:, 1], resident[1][:, 0]))
        assert np.abs(host[1][o1] - resident[1][o2]).max() <= 2e-5 * np.abs(resident[1]).max()
        d = np.abs(host[0][o1] - resident[0][o2])
        assert np.minimum(d, L - d).max() <= 1e-5 * cfg.cell


@pytest.mark.parametrize("world", [1, 2, 4])
def test_sharded_power_spectrum(world):
    from nbody_b200.dist import DistPowerSpectrum, run_threaded
    from nbody_b200.engine import Engine
    L, n, N = 1000.0, 16, 32
    pos = zeldovich_particles(n, L, seed=4)
    own = owner_np(pos, L, N, world)

    def body(comm):
        torch.cuda.set_device(0)
        eng = Engine(0)
        ps = DistPowerSpectrum(eng, comm, N, L)
        p = eng.to_device(pos[own == comm.rank])
        edges, counts, sums = ps(p, pos.shape[0])
        torch.cuda.synchronize()
        return counts.cpu().numpy(), sums.cpu().numpy()

    res = run_threaded(world, body)
    kr, pr, cr = restate.power_spectrum_from_positions(pos, L, N, return_counts=True)
    for counts, sums in res:
        assert np.array_equal(counts, cr)                      # bin counts bit-exact on every rank
        valid = counts > 0
        assert np.abs(sums[valid] / counts[valid] / pr - 1).max() <= TOL


def test_deposit_slab_reports_foreign_particles():
    from nbody_b200.engine import Engine
    eng = Engine(0)
    L, N = 100.0, 16
    pos = eng.to_device(np.array([[1.0, 2.0, 3.0], [60.0, 2.0, 3.0]]))
    grid = torch.empty((4 + 2, N, N), dtype=torch.float32, device=eng.device)
    nout = torch.zeros(1, dtype=torch.int64, device=eng.device)
    eng.cic_deposit_slab(pos, L / N, N, 0, 4, 2, grid, n_outside=nout)
    assert int(nout.item()) == 1 and abs(float(grid.sum().item()) - 1.0) < 1e-6


def test_wrongly_sharded_particles_are_reported():
    """A particle handed to the wrong rank gets zero mass and zero mesh force from the slab kernels; both
    accel() and step() must raise instead of returning wrong physics."""
    from nbody_b200._lib import NbodyError
    L, n, N, world = 1000.0, 16, 32, 2
    pos = zeldovich_particles(n, L, seed=11)
    cfg = cfg_for("pm", n, N, L, interlace=(True, True), deconvolve=(True, True), G=1.0, mass=2.0)
    own = owner_np(pos, L, N, world)
    bad = int(np.nonzero(own == 1)[0][0])                # give one of rank 1's particles to rank 0 as well

    def fn(st, p, i):
        if st.rank == 0:
            p = torch.cat([p, st.eng.to_device(pos[bad:bad + 1])])
        out = []
        for call in (lambda: st.accel(p), lambda: st.step(p.clone(), torch.zeros_like(p), 1e-3, 1e-3)):
            try:
                call()
                out.append("ok")
            except NbodyError:
                out.append("raised")
        return out

    res, _ = run_ranks(world, cfg, pos, fn)
    assert res[0][0] == "raised"                          # accel(): the rank that holds the stray particle
    assert res[0][1] == "raised" and res[1][1] == "raised"   # step(): the count is all-reduced in migrate()


@pytest.mark.parametrize("n_in", [0, 30, 100, 170])
def test_migrate_merge_kernels(n_in):
    """nb_migrate_merge on one device: arrivals fill holes / are appended, and with more leavers than arrivals
    the tail is compacted into the remaining holes — the surviving set is exactly (kept + arrived)."""
    from nbody_b200.engine import default_engine
    eng = default_engine()
    rs = np.random.RandomState(n_in)
    n, n_holes, cap = 1000, 100, 256
    pos, mom = rs.random_sample((n + cap, 3)), rs.normal(size=(n + cap, 3))
    holes = rs.choice(n, n_holes, replace=False).astype(np.int32)
    holes[:7] = np.arange(n - 7, n)[::-1]                      # some holes inside the tail region
    holes = np.unique(holes)[rs.permutation(np.unique(holes).size)].astype(np.int32)
    n_holes = holes.size
    arr = rs.normal(size=(n_in, 6)) + 10.0
    na = n_in // 3
    dpos, dmom = eng.to_device(pos.copy()), eng.to_device(mom.copy())
    new_n = eng.migrate_merge(dpos, dmom, n, eng.to_device(holes), n_holes, eng.to_device(arr[:na].copy()) if na else None,
                              na, eng.to_device(arr[na:].copy()) if n_in - na else None, n_in - na)
    assert new_n == n + n_in - n_holes
    keep = np.setdiff1d(np.arange(n), holes)
    want = np.vstack([np.hstack([pos[keep], mom[keep]]), arr])
    got = np.hstack([dpos[:new_n].cpu().numpy(), dmom[:new_n].cpu().numpy()])
    assert got.shape == want.shape
    o1, o2 = np.lexsort(got.T[::-1]), np.lexsort(want.T[::-1])
    assert np.array_equal(got[o1], want[o2])


def test_halo_select_and_nan_rows_in_the_cell_sort():
    """nb_halo_select packs the face particles into fixed-capacity NaN-padded buffers; nb_cell_sort skips the NaN
    rows, so (own + padded halo buffers) sorts exactly like (own + selected rows)."""
    from nbody_b200.engine import default_engine
    eng = default_engine()
    rs = np.random.RandomState(2)
    L, n, cap, ncs = 100.0, 5000, 1024, 8
    pos = rs.random_sample((n, 3)) * L
    lo_edge, hi_edge = 7.5, 92.0
    d = eng.to_device(pos)
    out = torch.empty((2, cap, 3), dtype=torch.float64, device=eng.device)
    cnt = torch.zeros(2, dtype=torch.int32, device=eng.device)
    eng.halo_select(d, lo_edge, hi_edge, out[0], out[1], cnt)
    c = cnt.cpu().numpy()
    sel_lo, sel_hi = pos[pos[:, 0] < lo_edge], pos[pos[:, 0] >= hi_edge]
    assert c[0] == sel_lo.shape[0] and c[1] == sel_hi.shape[0] and c.max() < cap
    for b, want in ((0, sel_lo), (1, sel_hi)):
        got = out[b].cpu().numpy()
        assert np.isnan(got[c[b]:]).all()
        g = got[:c[b]]
        assert np.array_equal(g[np.lexsort(g.T[::-1])], want[np.lexsort(want.T[::-1])])
    both = torch.cat([d, out[0], out[1]])
    s4, perm, cs = eng.cell_sort(both, L, ncs)
    ref4, refperm, refcs = eng.cell_sort(eng.to_device(np.vstack([pos, sel_lo, sel_hi])), L, ncs)
    assert np.array_equal(cs.cpu().numpy(), refcs.cpu().numpy())
    nv = int(cs[-1].item())
    assert nv == n + c.sum()
    # same multiset of cell-relative coordinates per cell
    a, b = s4[:nv].cpu().numpy(), ref4[:nv].cpu().numpy()
    csn = cs.cpu().numpy()
    for cell in rs.randint(0, ncs ** 3, 40):
        x, y = a[csn[cell]:csn[cell + 1]], b[csn[cell]:csn[cell + 1]]
        assert np.array_equal(x[np.lexsort(x.T[::-1])], y[np.lexsort(y.T[::-1])])
    # a buffer that is too small: the count still tells
    small = torch.empty((2, 16, 3), dtype=torch.float64, device=eng.device)
    eng.halo_select(d, lo_edge, hi_edge, small[0], small[1], cnt)
    assert cnt.cpu().numpy()[0] == sel_lo.shape[0] > 16
