"""Multi-GPU (one process per GPU) sharding of the force step and the P(k) estimator
(SURVEY §8e).  The reference has no counterpart: every CONCEPT run it launches is a single
process ("Domain decomposition: 1x1x1", run logs line 17; `-n 1 --local`,
NeuNetværk_test.py:155).

  PM / P3M long range   real mesh in x-slabs, spectrum in y-slabs; one all-to-all per
                        transform (2 forward + 6 inverse with interlacing); ghost planes on
                        the +x side of every slab for the deposit (added to the next slab)
                        and the gather (copied from the next slab)
  P3M short range       halo particles within r_cut of the slab faces come from the two
                        neighbouring slabs (positions only), then the single-GPU cell sort +
                        pair kernel run on own + halo with the halo as sources only
  PP                    i-particles sharded by index, all-gather of the packed float4 j set
  P(k)                  slab FFT, each rank bins its y-slab, all-reduce of (count, sum)

Every numeric step is a kernel of libnbody_b200.so (engine.Engine); torch.distributed (NCCL on
the GPUs, gloo in the CPU tests of the exchange logic) only moves buffers.  `ThreadComm`
emulates the ranks as threads of one process on ONE device, so the whole sharded algorithm is
parity-tested on a single GPU before it is run on several.
"""
from __future__ import annotations

import math
import os
import threading

import numpy as np
import torch

from .params import ForceConfig


# ------------------------------------------------------------------------------- layout
class SlabLayout:
    """x-slabs of the real mesh / y-slabs of the spectrum for rank `rank` of `world`."""

    def __init__(self, N: int, world: int, rank: int, n_ghost: int = 2):
        if N % world:
            raise ValueError(f"mesh size {N} is not divisible by the number of ranks {world}")
        self.N, self.world, self.rank = int(N), int(world), int(rank)
        self.nxl = N // world           # own x-planes (real space)
        self.nyl = N // world           # own ky rows (k-space)
        if self.nxl < n_ghost and world > 1:
            raise ValueError("slab thinner than its ghost layer")
        self.x0 = rank * self.nxl
        self.y0 = rank * self.nyl
        self.nzh = N // 2 + 1
        self.n_ghost = n_ghost
        self.rows = self.nyl * self.nzh   # transforms along x per rank

    def owner_of_plane(self, ix):
        return ix // self.nxl


def shard_particles(pos_all: torch.Tensor, method: str, L: float, rank: int, world: int, N: int = 0):
    """The shard of a full particle set that `rank` starts with: x-slab ownership for pm / p3m /
    pk (same float64 cell index as the deposit), a contiguous index range for pp."""
    n = pos_all.shape[0]
    if method == "pp":
        chunk = (n + world - 1) // world
        return pos_all[rank * chunk:min(n, (rank + 1) * chunk)].contiguous()
    if N <= 0 or N % world:
        raise ValueError("slab sharding needs a mesh size divisible by the number of ranks")
    cell = L / N
    ix = torch.div(pos_all[:, 0], cell).to(torch.int64) % N        # int(p/cell) % N
    own = torch.div(ix, N // world, rounding_mode="floor") == rank
    return pos_all[own].contiguous()


# ------------------------------------------------------------------------------- communicators
class TorchComm:
    """torch.distributed (NCCL over NVLink on the GPUs; gloo in the CPU tests)."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)

    def all_to_all(self, out: torch.Tensor, inp: torch.Tensor, async_op: bool = False):
        """Equal contiguous chunks: out[s-th chunk] = chunk `rank` of rank s's inp.  With
        async_op the exchange runs on the communicator's stream and a handle with .wait() (the
        current stream waits, the host does not) is returned, so transforms of the next mesh
        overlap the transpose of this one."""
        if self.world == 1:
            out.view(-1).copy_(inp.view(-1))
            return _DONE if async_op else out
        o, i = out.view(-1), inp.view(-1)
        if o.is_complex():
            o, i = torch.view_as_real(o).view(-1), torch.view_as_real(i).view(-1)
        w = self.dist.all_to_all_single(o, i, group=self.group, async_op=async_op)
        return w if async_op else out

    def shift(self, send: torch.Tensor, recv: torch.Tensor, delta: int):
        """send -> rank+delta, recv <- rank-delta (ring)."""
        if self.world == 1:
            recv.copy_(send)
            return recv
        dst, src = (self.rank + delta) % self.world, (self.rank - delta) % self.world
        ops = [self.dist.P2POp(self.dist.isend, send, dst, self.group),
               self.dist.P2POp(self.dist.irecv, recv, src, self.group)]
        for w in self.dist.batch_isend_irecv(ops):
            w.wait()
        return recv

    def shift_var(self, send: torch.Tensor, n_send: int, recv: torch.Tensor, n_recv: int, delta: int):
        """First n_send rows of `send` -> rank+delta, n_recv rows <- rank-delta into `recv` (both sides know
        the sizes; zero-size messages are skipped on both ends)."""
        if self.world == 1:
            if n_recv:
                recv[:n_recv].copy_(send[:n_send])
            return recv
        dst, src = (self.rank + delta) % self.world, (self.rank - delta) % self.world
        ops = []
        if n_send:
            ops.append(self.dist.P2POp(self.dist.isend, send[:n_send], dst, self.group))
        if n_recv:
            ops.append(self.dist.P2POp(self.dist.irecv, recv[:n_recv], src, self.group))
        if ops:
            for w in self.dist.batch_isend_irecv(ops):
                w.wait()
        return recv

    def all_gather(self, out: torch.Tensor, inp: torch.Tensor):
        if self.world == 1:
            out.view(-1).copy_(inp.view(-1))
            return out
        self.dist.all_gather_into_tensor(out.view(-1), inp.view(-1), group=self.group)
        return out

    def all_reduce(self, t: torch.Tensor, op="sum"):
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM if op == "sum" else self.dist.ReduceOp.MAX,
                                 group=self.group)
        return t

    def exchange_var(self, send: list):
        """send[d]: (rows_d, C) tensor for rank d (send[rank] is ignored).  Returns the list of
        tensors received from every rank (entry `rank` is an empty tensor)."""
        ref = next((t for t in send if t is not None), None)
        if ref is None:
            return [torch.empty((0, 0)) for _ in range(self.world)]
        C, dev, dt = ref.shape[1], ref.device, ref.dtype
        cnt = torch.tensor([0 if (d == self.rank or send[d] is None) else send[d].shape[0]
                            for d in range(self.world)], dtype=torch.int64, device=dev)
        rcnt = torch.empty_like(cnt)
        self.all_to_all(rcnt, cnt)
        sc, rc = cnt.tolist(), rcnt.tolist()                      # host sync: sizes are data dependent
        parts = [send[d] for d in range(self.world) if sc[d]]
        sbuf = torch.cat(parts) if parts else torch.empty((0, C), dtype=dt, device=dev)
        rbuf = torch.empty((sum(rc), C), dtype=dt, device=dev)
        if self.world > 1:
            self.dist.all_to_all_single(rbuf, sbuf.contiguous(), output_split_sizes=rc, input_split_sizes=sc,
                                        group=self.group)
        return list(torch.split(rbuf, rc))

    def barrier(self):
        if self.world > 1:
            self.dist.barrier(group=self.group)


class _Done:
    def wait(self):
        return True


_DONE = _Done()


class ThreadWorld:
    """Shared state of `world` emulated ranks running as threads of one process."""

    def __init__(self, world: int):
        self.world = world
        self.slots = [None] * world
        self.bar = threading.Barrier(world)


class ThreadComm:
    """Same interface as TorchComm for ranks emulated as threads on one device (tests).  All
    threads issue their kernels on the device's default stream, so buffer hand-over needs only
    the host barrier."""

    def __init__(self, tw: ThreadWorld, rank: int):
        self.tw, self.rank, self.world = tw, rank, tw.world

    def _swap(self, obj):
        self.tw.slots[self.rank] = obj
        self.tw.bar.wait()
        got = list(self.tw.slots)
        return got

    def _done(self):
        self.tw.bar.wait()

    def all_to_all(self, out, inp, async_op=False):
        got = self._swap(inp)
        n = out.numel() // self.world
        o = out.view(-1)
        for s in range(self.world):
            o[s * n:(s + 1) * n].copy_(got[s].view(-1)[self.rank * n:(self.rank + 1) * n])
        self._done()
        return _DONE if async_op else out

    def shift(self, send, recv, delta):
        got = self._swap(send)
        recv.copy_(got[(self.rank - delta) % self.world])
        self._done()
        return recv

    def shift_var(self, send, n_send, recv, n_recv, delta):
        got = self._swap(send[:n_send])
        if n_recv:
            recv[:n_recv].copy_(got[(self.rank - delta) % self.world])
        self._done()
        return recv

    def all_gather(self, out, inp):
        got = self._swap(inp)
        n = inp.numel()
        o = out.view(-1)
        for s in range(self.world):
            o[s * n:(s + 1) * n].copy_(got[s].view(-1))
        self._done()
        return out

    def all_reduce(self, t, op="sum"):
        got = self._swap(t.clone())
        acc = got[0].clone()
        for s in range(1, self.world):
            acc = acc + got[s] if op == "sum" else torch.maximum(acc, got[s])
        self._done()
        t.copy_(acc)
        return t

    def exchange_var(self, send):
        got = self._swap(send)
        ref = next((t for t in send if t is not None), None)
        if ref is None:
            self._done()
            return [torch.empty((0, 0)) for _ in range(self.world)]
        out = []
        for s in range(self.world):
            t = None if s == self.rank else got[s][self.rank]
            out.append(t.clone() if t is not None else ref.new_empty((0, ref.shape[1])))
        self._done()
        return out

    def barrier(self):
        self.tw.bar.wait()


def run_threaded(world: int, fn):
    """Run fn(comm) on `world` emulated ranks; returns the list of results (tests)."""
    tw = ThreadWorld(world)
    res, err = [None] * world, [None] * world

    def work(r):
        try:
            res[r] = fn(ThreadComm(tw, r))
        except BaseException as e:   # noqa: BLE001
            err[r] = e
            tw.bar.abort()

    th = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    for e in err:
        if e is not None and not isinstance(e, threading.BrokenBarrierError):
            raise e
    for e in err:
        if e is not None:
            raise e
    return res


# ------------------------------------------------------------------------------- slab mesh
class SlabMesh:
    """Device buffers and the distributed transforms of one rank's slab."""

    def __init__(self, eng, comm, N: int, L: float, nb_fwd: int, nc_inv: int):
        self.eng, self.comm = eng, comm
        self.lay = lay = SlabLayout(N, comm.world, comm.rank)
        self.N, self.L, self.cell = N, float(L), float(L) / N
        dev = eng.device
        npl = lay.nxl + lay.n_ghost
        self.rho = torch.empty((nb_fwd, npl, N, N), dtype=torch.float32, device=dev)
        self.ghost_rx = torch.empty((nb_fwd, lay.n_ghost, N, N), dtype=torch.float32, device=dev)
        self.s2d = torch.empty((lay.nxl, N, lay.nzh), dtype=torch.complex64, device=dev)
        self.xbufs = [torch.empty((lay.world, lay.nxl, lay.nyl, lay.nzh), dtype=torch.complex64, device=dev)
                      for _ in range(2 if (nb_fwd > 1 or nc_inv > 1) else 1)]
        self.xbuf = self.xbufs[0]
        self.rhok = torch.empty((nb_fwd, N, lay.nyl, lay.nzh), dtype=torch.complex64, device=dev)
        self.n_outside = torch.zeros(max(nb_fwd, 1), dtype=torch.int64, device=dev)   # per deposit pass
        if nc_inv:
            self.g = torch.empty((nc_inv, N, lay.nyl, lay.nzh), dtype=torch.complex64, device=dev)
            self.force = torch.empty((nc_inv, npl, N, N), dtype=torch.float32, device=dev)
            self.fghost = torch.empty((nc_inv, lay.n_ghost, N, N), dtype=torch.float32, device=dev)

    # deposit own particles (+ optional half-cell shifted copy) and fold in the ghost planes
    def deposit(self, pos, nb: int):
        lay, eng = self.lay, self.eng
        for b in range(nb):
            eng.cic_deposit_slab(pos, self.cell, self.N, lay.x0, lay.nxl, lay.n_ghost, self.rho[b],
                                 shift=0.5 * self.cell * b, n_outside=self.n_outside[b:b + 1])
        tx = self.rho[:nb, lay.nxl:].contiguous()
        self.comm.shift(tx, self.ghost_rx[:nb], +1)
        for b in range(nb):
            eng.grid_add(self.rho[b, :lay.n_ghost], self.ghost_rx[b])

    def check_outside(self):
        """Raise if the last deposit was handed particles outside this rank's x-slab: the slab kernels
        give them zero mass and zero mesh force, so a wrong sharding would otherwise go unnoticed
        (one host sync; the step folds the same counter into migrate()'s existing sync)."""
        n = int(self.n_outside.sum().item())
        if n:
            from ._lib import NbodyError
            raise NbodyError(f"{n} particle(s) of rank {self.comm.rank} lie outside its x-slab "
                             "(shard with dist.shard_particles / keep the set returned by step())")

    # real x-slab b -> y-slab spectrum rhok[b]  ([N][nyl][nzh])
    def forward(self, b: int):
        return self.forward_all(1, first=b)[0]

    def forward_all(self, nb: int, first: int = 0):
        """2-D R2C + pack of mesh b+1 overlap the all-to-all of mesh b."""
        lay, eng = self.lay, self.eng
        sub = getattr(self, "submark", None) or (lambda name, fn: fn())   # optional finer phase timing
        works = []
        for k in range(nb):
            b = first + k
            xb = self.xbufs[k % len(self.xbufs)]
            sub("fwd.2d_r2c", lambda: eng.slab_fft2d(self.N, lay.nxl, +1, self.rho[b], self.s2d))
            sub("fwd.pack", lambda: eng.slab_pack(self.s2d, xb, lay.nxl, lay.world, lay.nyl, lay.nzh))
            works.append(self.comm.all_to_all(self.rhok[b], xb, async_op=True))
        for k in range(nb):
            sub("fwd.a2a_wait", lambda: works[k].wait())
            sub("fwd.1d_x", lambda: eng.slab_fft1d_x(self.N, lay.rows, +1, self.rhok[first + k]))
        return [self.rhok[first + k] for k in range(nb)]

    # y-slab spectra g[0:nc] -> real x-slabs force[0:nc] (first nxl planes)
    def inverse_all(self, nc: int):
        """Software pipeline: the x transform of component c+1 and the unpack + 2-D C2R of
        component c-1 run while the all-to-all of component c is in flight."""
        lay, eng = self.lay, self.eng
        works = [None] * nc

        def finish(c):
            works[c].wait()
            eng.slab_unpack(self.xbufs[c % len(self.xbufs)], self.s2d, lay.nxl, lay.world, lay.nyl, lay.nzh)
            eng.slab_fft2d(self.N, lay.nxl, -1, self.force[c], self.s2d)

        for c in range(nc):
            eng.slab_fft1d_x(self.N, lay.rows, -1, self.g[c])
            if c >= len(self.xbufs):
                finish(c - len(self.xbufs))          # frees the exchange buffer this component reuses
            works[c] = self.comm.all_to_all(self.xbufs[c % len(self.xbufs)], self.g[c], async_op=True)
        for c in range(max(0, nc - len(self.xbufs)), nc):
            finish(c)

    def inverse(self, c: int):
        lay, eng = self.lay, self.eng
        eng.slab_fft1d_x(self.N, lay.rows, -1, self.g[c])
        self.comm.all_to_all(self.xbuf, self.g[c])
        eng.slab_unpack(self.xbuf, self.s2d, lay.nxl, lay.world, lay.nyl, lay.nzh)
        eng.slab_fft2d(self.N, lay.nxl, -1, self.force[c], self.s2d)

    # ghost planes of the force meshes: copies of the next slab's first planes
    def fill_force_ghosts(self, nc: int):
        lay = self.lay
        tx = self.force[:nc, :lay.n_ghost].contiguous()
        self.comm.shift(tx, self.fghost[:nc], -1)
        self.force[:nc, lay.nxl:].copy_(self.fghost[:nc])


def bind_host_near_device(device_index: int):
    """Restrict the calling process to the CPU cores next to GPU `device_index` (sysfs `local_cpulist` of its PCI
    function), so that pinned host buffers allocated afterwards are placed (first touch) on the NUMA node the GPU
    hangs off and every rank of a node stages its shard through its own memory controller instead of rank 0's.
    Returns {"cpus": n, "numa_node": k} or None when the topology cannot be read or the cores are not in the
    process's cpuset (nothing is changed then)."""
    try:
        pr = torch.cuda.get_device_properties(device_index)
        bus = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        base = f"/sys/bus/pci/devices/{bus}"
        with open(f"{base}/local_cpulist") as fh:
            txt = fh.read().strip()
        cpus = set()
        for part in txt.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        allowed = os.sched_getaffinity(0)
        use = cpus & allowed
        if not use or use == allowed:
            return None
        os.sched_setaffinity(0, use)
        try:
            with open(f"{base}/numa_node") as fh:
                node = int(fh.read().strip())
        except (OSError, ValueError):
            node = -1
        return {"cpus": len(use), "numa_node": node}
    except Exception:      # no CUDA device, no sysfs entry, cpuset that forbids the change: leave everything as it is
        return None


# ------------------------------------------------------------------------------- force step
class DistForceStepper:
    """One kick-drift step on this rank's particles (pos, mom: contiguous (n,3) float64 on the
    engine's device).  `step` returns the rank's particle tensors AFTER migration — they may be
    new tensors of a different length."""

    def __init__(self, cfg: ForceConfig, engine, rank: int | None = None, world: int | None = None, comm=None):
        self.cfg, self.eng = cfg, engine
        self.comm = comm if comm is not None else TorchComm()
        self.rank, self.world = self.comm.rank, self.comm.world
        if (rank is not None and rank != self.rank) or (world is not None and world != self.world):
            raise ValueError("rank/world do not match the communicator")
        if cfg.method not in ("pp", "pm", "p3m"):
            raise ValueError(f"unknown force method {cfg.method!r}")
        self._acc = None
        self._pbuf = self._mbuf = None          # capacity buffers: rows [0,n) own particles, then halo / arrivals
        self._hbuf = self._hcounts = self._hover = self._n_mean = None
        self._mig = self._owner_buf = None
        # device-side halo selection / migration (nb_halo_select, nb_migrate_*): one host sync per step;
        # False keeps the eager torch exchange (reference implementation of the same bookkeeping, tests)
        self.device_exchange = True
        self.halo_cap_floor = 4096
        self.mesh = None
        self._mark = lambda name, fn: fn()      # bench.py installs a CUDA-event recorder per step
        # pair kernel beside the mesh chain on an SM partition (force.py / DESIGN.md 8.2): the confined pair kernel
        # runs on a side stream, the mesh chain and its transposes stay on the caller's stream.  Half of the SMs by
        # default: per rank the mesh chain (with its all-to-all) is the longer half, and whichever side finishes
        # first hands its SMs to the other (stop flag + unconfined second launch / retiring CTAs).
        ov = os.environ.get("NB_P3M_OVERLAP", "auto")
        self.overlap = cfg.method == "p3m" and (True if ov == "1" else "auto" if ov == "auto" else False)
        self.overlap_min_particles = 1 << 23          # own particles of the rank
        self.p3m_sms = int(os.environ.get("NB_P3M_SMS", "0"))
        self.p3m_delay_us = int(os.environ.get("NB_P3M_DELAY_US", "40"))
        self._sr_stream = None
        self.p3m_near_fix = os.environ.get("NB_P3M_NEAR", "0") == "1"
        if cfg.method in ("pm", "p3m"):
            self.nb = 2 if cfg.interlace[0] else 1
            self.nc = 6 if cfg.interlace[1] else 3
            self.mesh = SlabMesh(engine, self.comm, cfg.gridsize, cfg.boxsize, self.nb, self.nc)
        if cfg.method == "p3m":
            self.ncs = int(math.floor(cfg.boxsize / cfg.r_cut))
            if self.ncs < 3:
                raise ValueError("P3M short range needs boxsize >= 3 r_cut")
            if self.world > 1 and self.mesh.lay.nxl * cfg.cell < 2 * cfg.r_cut:
                raise ValueError("slab thinner than 2 r_cut: use fewer ranks")

    def _acc_buf(self, n):
        if self._acc is None or self._acc.shape[0] < n:
            self._acc = torch.empty((max(n, 1), 3), dtype=torch.float64, device=self.eng.device)
        return self._acc[:n]

    def _halo_cap(self, n):
        """Rows of one halo buffer: twice the uniform-density expectation (N_total / ranks) r_cut / slab width.
        The SAME number on every rank (the neighbours exchange the buffers at this fixed size), hence from the
        global particle count, not from this rank's n."""
        c, lay = self.cfg, self.mesh.lay
        if self._n_mean is None:
            if c.npart:
                self._n_mean = c.npart / self.world
            else:       # unknown total: agree on the largest shard once
                t = torch.tensor([n], dtype=torch.int64, device=self.eng.device)
                self.comm.all_reduce(t, "max")
                self._n_mean = float(t.item())
        return max(self.halo_cap_floor, int(2.0 * self._n_mean * c.r_cut / (lay.nxl * c.cell)) + self.halo_cap_floor // 4)

    def _mig_cap(self, n):
        return max(4096, n // 8)

    def _mig_rows(self, n):
        """Rows reserved behind the own particles for arrivals (sticky like the buffers themselves)."""
        return self._mig["cap"] if self._mig is not None and self._mig["cap"] >= self._mig_cap(n) else \
            self._mig_cap(n) + self._mig_cap(n) // 4

    def _extra_rows(self, n):
        """Rows behind the own particles: both received halo buffers (P3M) or the arrivals of a migration."""
        if self.world == 1 or self.mesh is None:
            return 0
        return max(2 * self._halo_cap(n) if self.cfg.method == "p3m" else 0, 2 * self._mig_rows(n))

    def _own(self, pos, mom=None):
        """pos / mom as views (rows [0,n)) of this stepper's capacity buffers.  Halo copies and
        migrating particles are appended behind row n in place, so a step allocates nothing of
        the particle set's size (a growing torch.cat per step cost cudaMalloc stalls of 10+ ms
        on 4 GPUs).  Tensors that are already such views are returned as they are."""
        n = pos.shape[0]
        need = n + self._extra_rows(n)
        if self._pbuf is None or self._pbuf.shape[0] < need:
            cap = need + max(n // 10, 16384)
            old_p, old_m = self._pbuf, self._mbuf
            self._pbuf = torch.empty((cap, 3), dtype=torch.float64, device=self.eng.device)
            self._mbuf = torch.empty((cap, 3), dtype=torch.float64, device=self.eng.device)
            if old_p is not None and pos.data_ptr() == old_p.data_ptr():      # grown in flight: keep the contents
                self._pbuf[:n].copy_(old_p[:n])
                pos = self._pbuf[:n]
                if mom is not None and mom.data_ptr() == old_m.data_ptr():
                    self._mbuf[:n].copy_(old_m[:n])
                    mom = self._mbuf[:n]
            del old_p, old_m
        if pos.data_ptr() != self._pbuf.data_ptr():
            self._pbuf[:n].copy_(pos)
            pos = self._pbuf[:n]
        if mom is not None and mom.data_ptr() != self._mbuf.data_ptr():
            self._mbuf[:n].copy_(mom)
            mom = self._mbuf[:n]
        return pos, mom

    # -- long range ---------------------------------------------------------------------------
    def pm_mesh(self, pos):
        c, m, mk = self.cfg, self.mesh, self._mark
        N = c.gridsize
        mk("deposit+ghost_add", lambda: m.deposit(pos, self.nb))
        mk("fft_fwd(2D R2C, pack, all-to-all, 1D)", lambda: m.forward_all(self.nb))
        pref = 4.0 * math.pi * c.G * c.mass / c.cell ** 3 / float(N) ** 3
        mk("kspace", lambda: self.eng.kspace_fused_slab(
            m.rhok[0], m.rhok[1] if self.nb == 2 else None, c.boxsize, N, m.lay.y0, m.lay.nyl, c.r_s,
            c.deconv_order, pref, m.g[0:3], m.g[3:6] if self.nc == 6 else None))
        mk("fft_inv(1D, all-to-all, unpack, 2D C2R)", lambda: m.inverse_all(self.nc))
        mk("force_ghosts", lambda: m.fill_force_ghosts(self.nc))
        return m.force[0:3], (m.force[3:6] if self.nc == 6 else None)

    def _gather(self, pos, mom, acc, kick, drift):
        c, lay = self.cfg, self.mesh.lay
        f, fs = self.mesh.force[0:3], (self.mesh.force[3:6] if self.nc == 6 else None)
        self.eng.gather_kick_drift_slab(f, fs, c.gridsize, lay.x0, lay.nxl, lay.n_ghost, c.cell, pos, mom, acc, 1.0,
                                        kick, drift, c.boxsize)

    # -- short range ----------------------------------------------------------------------------
    def halo_exchange(self, pos):
        """Positions of the neighbouring slabs' particles within r_cut of this slab's faces."""
        c, lay = self.cfg, self.mesh.lay
        if self.world == 1:
            return pos.new_empty((0, 3))
        x = pos[:, 0]
        x_lo, x_hi = lay.x0 * c.cell, (lay.x0 + lay.nxl) * c.cell
        to_prev = pos[x < x_lo + c.r_cut]
        to_next = pos[x >= x_hi - c.r_cut]
        send = [None] * self.world
        prev, nxt = (self.rank - 1) % self.world, (self.rank + 1) % self.world
        if prev == nxt:
            send[prev] = torch.cat([to_prev, to_next])
        else:
            send[prev], send[nxt] = to_prev, to_next
        for d in range(self.world):
            if send[d] is None:
                send[d] = pos.new_empty((0, 3))
        got = self.comm.exchange_var(send)
        return torch.cat([g for g in got if g.shape[0]]) if any(g.shape[0] for g in got) else pos.new_empty((0, 3))

    def _halo_device(self, pos):
        """Halo exchange without the host: nb_halo_select fills two fixed-capacity buffers (NaN beyond the
        selected rows), the neighbours swap them, and the received buffers land right behind the own rows of
        the capacity buffer — nb_cell_sort skips the NaN rows.  Returns the (n + 2 cap, 3) view."""
        c, lay, eng = self.cfg, self.mesh.lay, self.eng
        n = pos.shape[0]
        pos, _ = self._own(pos)
        cap = self._halo_cap(n)
        if self._hbuf is None or self._hbuf.shape[1] != cap:
            self._hbuf = torch.empty((2, cap, 3), dtype=torch.float64, device=eng.device)
            self._hcounts = torch.zeros(2, dtype=torch.int32, device=eng.device)
        x_lo, x_hi = lay.x0 * c.cell, (lay.x0 + lay.nxl) * c.cell
        eng.halo_select(pos, x_lo + c.r_cut, x_hi - c.r_cut, self._hbuf[0], self._hbuf[1], self._hcounts)
        over = (self._hcounts > cap).any().to(torch.int64)        # checked at the step's one host sync
        self._hover = over if self._hover is None else torch.maximum(self._hover, over)
        dst = self._pbuf[n:n + 2 * cap]
        self.comm.shift(self._hbuf[0], dst[:cap], -1)       # my low face -> previous slab; I receive the next slab's
        self.comm.shift(self._hbuf[1], dst[cap:], +1)       # my high face -> next slab; I receive the previous slab's
        return self._pbuf[:n + 2 * cap]

    def _short_prepare(self, pos):
        """Halo exchange + cell sort of own and halo particles; returns what the pair kernel needs."""
        c, mk = self.cfg, self._mark
        n = pos.shape[0]
        if self.world > 1 and self.device_exchange:
            allp = mk("halo_exchange", lambda: self._halo_device(pos))
            h = allp.shape[0] - n
        else:
            halo = mk("halo_exchange", lambda: self.halo_exchange(pos))
            h = int(halo.shape[0])
            if h and self._pbuf is not None and pos.data_ptr() == self._pbuf.data_ptr() and n + h <= self._pbuf.shape[0]:
                self._pbuf[n:n + h].copy_(halo)          # halo copies sit right behind the own particles
                allp = self._pbuf[:n + h]
            else:
                allp = torch.cat([pos, halo]) if h else pos
        self.last_halo = h
        sorted4, perm, cell_start = mk("cell_sort", lambda: self.eng.cell_sort(allp, c.boxsize, self.ncs, mass=1.0))
        # only the sort cells that overlap this rank's x-slab can hold own (i-) particles: cells are
        # x-major, so that is one contiguous cell range (one cell of margin on both sides)
        lay = self.mesh.lay
        cs = c.boxsize / self.ncs
        cx_lo = max(0, int(math.floor(lay.x0 * c.cell / cs)) - 1)
        cx_hi = min(self.ncs, int(math.floor((lay.x0 + lay.nxl) * c.cell / cs)) + 2)
        plane = self.ncs * self.ncs
        self.last_sorted = (sorted4, cell_start)
        return sorted4, perm, cell_start, cx_lo * plane, cx_hi * plane, n, allp

    def _short_pairs(self, prep, acc, accumulate):
        c = self.cfg
        sorted4, perm, cell_start, cell_lo, cell_hi, n, allp = prep
        near = {}
        if self.p3m_near_fix and c.eps > 0 and 20.0 * c.eps < c.r_cut:      # as ForceStepper._near
            near = {"pos": allp, "r_near": 20.0 * c.eps}
        return self.eng.p3m_short(sorted4, perm, cell_start, c.boxsize, self.ncs, c.r_s, c.r_cut, c.eps,
                                  scale=c.G * c.mass, out=acc, accumulate=accumulate, cell_lo=cell_lo, cell_hi=cell_hi,
                                  i_limit=n, **near)

    def short_range(self, pos, acc, accumulate=True):
        prep = self._short_prepare(pos)
        self._mark("p3m_short", lambda: self._short_pairs(prep, acc, accumulate))
        return acc

    def _short_and_mesh_side_by_side(self, pos, acc_s):
        """P3M step, overlapped form: halo exchange and cell sort alone, then the pair kernel confined to the first
        p3m_sms SMs on a side stream while the mesh chain (deposit, slab FFTs with their transposes, k-space
        kernel) runs on the caller's stream on the other SMs.  Returns the force meshes."""
        eng, mk = self.eng, self._mark
        dev = eng.device
        main = torch.cuda.current_stream(dev)
        if self._sr_stream is None:
            self._sr_stream = torch.cuda.Stream(device=dev)
            if self.p3m_sms <= 0:
                self.p3m_sms = max(1, torch.cuda.get_device_properties(dev).multi_processor_count // 2)
        sr = self._sr_stream
        prep = self._short_prepare(pos)
        sr.wait_stream(main)
        eng.p3m_set_sm_limit(self.p3m_sms)
        try:
            with torch.cuda.stream(sr):
                mk("p3m_short(side stream, confined)", lambda: self._short_pairs(prep, acc_s, False))
        finally:
            eng.p3m_set_sm_limit(0)
        for t in prep[:3] + (prep[6],):
            t.record_stream(sr)
        eng.delay_us(self.p3m_delay_us)      # the confined launch gets its CTAs resident before the deposit asks for SMs
        f, fs = self.pm_mesh(pos)
        eng.p3m_mesh_done()
        mk("p3m_short(join)", lambda: main.wait_stream(sr))
        return f, fs

    def pp(self, pos, acc, n_total: int, periodic=False):
        c = self.cfg
        half = 0.5 * c.boxsize
        chunk = (n_total + self.world - 1) // self.world
        mk = self._mark
        mine = torch.zeros((chunk, 4), dtype=torch.float32, device=self.eng.device)   # padding: zero mass
        mk("pack", lambda: self.eng.pp_pack(pos, origin=(half, half, half), mass=1.0, out=mine[:pos.shape[0]]))
        allj = torch.empty((self.world * chunk, 4), dtype=torch.float32, device=self.eng.device)
        mk("all_gather_j", lambda: self.comm.all_gather(allj, mine))
        mk("pp_accel", lambda: self.eng.pp_accel(mine[:pos.shape[0]], allj, c.eps,
                                                 box=c.boxsize if periodic else 0.0, scale=c.G * c.mass, out=acc))
        if c.eps > 0 and int(c.boxsize / (20.0 * c.eps)) >= 3:
            # float64 near-field fix (force.ForceStepper.pp): needs the float64 positions of all j — the ranks
            # hold contiguous index ranges, so the gathered array is the full set with padding at its very end
            def fix():
                minep = torch.zeros((chunk, 3), dtype=torch.float64, device=self.eng.device)
                minep[:pos.shape[0]].copy_(pos)
                allp = torch.empty((self.world * chunk, 3), dtype=torch.float64, device=self.eng.device)
                self.comm.all_gather(allp, minep)
                self.eng.pp_near_fix(allp[:n_total], c.boxsize, c.eps, 20.0 * c.eps, (half, half, half), acc,
                                     periodic=bool(periodic), scale=c.G * c.mass, mass=1.0, i0=self.rank * chunk,
                                     ni=pos.shape[0])
            mk("pp_near_fix", fix)
        return acc

    # -- migration ----------------------------------------------------------------------------------
    def _flags(self, far=None):
        """(far movers, particles outside the slab at the deposit, halo-buffer overflow) as a device int64[3]."""
        z = torch.zeros(1, dtype=torch.int64, device=self.eng.device)
        hov = self._hover.reshape(1) if self._hover is not None else z
        self._hover = None
        return torch.cat([z if far is None else far.reshape(1).to(torch.int64), self.mesh.n_outside.sum().reshape(1), hov])

    def _raise_flags(self, n_out_slab, h_over):
        from ._lib import NbodyError
        if n_out_slab:
            raise NbodyError(f"{n_out_slab} particle(s) were outside their rank's x-slab at the deposit of this step "
                             "(shard with dist.shard_particles / keep the set returned by step())")
        if h_over:
            raise NbodyError("halo buffer overflow: more particles within r_cut of a slab face than twice the "
                             "uniform-density expectation (raise DistForceStepper._halo_cap)")

    def migrate(self, pos, mom):
        """Hand particles that drifted out of this rank's x-slab to their new owner.  Only the
        leavers and the arrivals are moved: arrivals fill the leavers' slots, a surplus is
        appended, a deficit is closed by moving particles from the tail into the open slots.
        Device path (default): nb_slab_owner + nb_migrate_split pick and pack the leavers, the counts go to the
        neighbours as two 4-byte messages, ONE host synchronisation reads them together with the step's error
        flags, then exact-size neighbour messages and nb_migrate_merge.  Particles that crossed more than one
        slab in a step send every rank through the general (eager) exchange below."""
        c, lay = self.cfg, self.mesh.lay
        if self.world == 1:
            return pos, mom
        n = pos.shape[0]
        if self.device_exchange:
            eng, dev = self.eng, self.eng.device
            pos, mom = self._own(pos, mom)
            cap = self._mig_cap(n)
            if self._mig is not None and self._mig["cap"] >= cap:
                cap = self._mig["cap"]          # sticky: n changes every step, a re-allocation costs a cudaMalloc stall
            else:
                cap += cap // 4
            if self._mig is None or self._mig["cap"] != cap:
                self._mig = {"cap": cap, "send": torch.empty((2, cap, 6), dtype=torch.float64, device=dev),
                             "recv": torch.empty((2, cap, 6), dtype=torch.float64, device=dev),
                             "holes": torch.empty(2 * cap, dtype=torch.int32, device=dev),
                             "counts": torch.zeros(4, dtype=torch.int32, device=dev),
                             "rc": torch.zeros(2, dtype=torch.int32, device=dev)}
            mg = self._mig
            prev, nxt = (self.rank - 1) % self.world, (self.rank + 1) % self.world
            if self._owner_buf is None or self._owner_buf.shape[0] < self._pbuf.shape[0]:
                self._owner_buf = torch.empty(self._pbuf.shape[0], dtype=torch.int32, device=dev)
            owner, _ = eng.slab_owner(pos, c.cell, c.gridsize, lay.nxl, self.world, want_counts=False,
                                      out=self._owner_buf)
            eng.migrate_split(pos, mom, owner, self.rank, prev, nxt, mg["send"][0], mg["send"][1], mg["holes"],
                              mg["counts"])
            # my "to previous" count -> previous rank, I receive the next rank's (its particles coming down to me)
            self.comm.shift(mg["counts"][0:1], mg["rc"][0:1], -1)
            self.comm.shift(mg["counts"][1:2], mg["rc"][1:2], +1)
            # general path wanted by anyone: a particle crossed more than one slab, or a buffer is too small
            general = mg["counts"][3].to(torch.int64) + (mg["counts"][:2] > cap).any() + (mg["rc"] > cap).any()
            flags = self._flags(far=general)
            self.comm.all_reduce(flags, "max")
            host = torch.cat([mg["counts"].to(torch.int64), mg["rc"].to(torch.int64), flags]).tolist()   # THE host sync
            n_prev, n_next, n_holes, _far, r_next, r_prev, general_any, n_out_slab, h_over = (int(v) for v in host)
            self._raise_flags(n_out_slab, h_over)
            if general_any == 0:
                self.comm.shift_var(mg["send"][0], n_prev, mg["recv"][0], r_next, -1)
                self.comm.shift_var(mg["send"][1], n_next, mg["recv"][1], r_prev, +1)
                new_n = eng.migrate_merge(self._pbuf, self._mbuf, n, mg["holes"], n_holes, mg["recv"][0], r_next,
                                          mg["recv"][1], r_prev)
                return self._pbuf[:new_n], self._mbuf[:new_n]
        owner, counts = self.eng.slab_owner(pos, c.cell, c.gridsize, lay.nxl, self.world)
        # leavers of this step and the particles the deposit found outside the slab (wrong sharding on entry)
        moved = torch.stack([n - counts[self.rank], self.mesh.n_outside.sum()])
        self.comm.all_reduce(moved, "max")
        n_moved, n_out_slab = (int(v) for v in moved.tolist())      # host sync, once per step
        if n_out_slab:
            from ._lib import NbodyError
            raise NbodyError(f"{n_out_slab} particle(s) were outside their rank's x-slab at the deposit of this step "
                             "(shard with dist.shard_particles / keep the set returned by step())")
        if n_moved == 0:
            return pos, mom
        out_idx = (owner != self.rank).nonzero().squeeze(1)        # ascending
        dest = owner[out_idx]
        leaving = torch.cat([pos[out_idx], mom[out_idx]], dim=1)
        send = [leaving[dest == d] if d != self.rank else None for d in range(self.world)]
        got = [g for g in self.comm.exchange_var(send) if g.shape[0]]
        arrivals = torch.cat(got) if got else leaving.new_empty((0, 6))
        n_out, n_in = int(out_idx.shape[0]), int(arrivals.shape[0])
        k = min(n_out, n_in)
        if k:
            pos[out_idx[:k]] = arrivals[:k, :3]
            mom[out_idx[:k]] = arrivals[:k, 3:]
        if n_in > n_out:
            extra = n_in - n_out
            if (self._pbuf is not None and pos.data_ptr() == self._pbuf.data_ptr()
                    and mom.data_ptr() == self._mbuf.data_ptr() and n + extra <= self._pbuf.shape[0]):
                self._pbuf[n:n + extra].copy_(arrivals[k:, :3])
                self._mbuf[n:n + extra].copy_(arrivals[k:, 3:])
                pos, mom = self._pbuf[:n + extra], self._mbuf[:n + extra]
            else:
                pos = torch.cat([pos, arrivals[k:, :3]])
                mom = torch.cat([mom, arrivals[k:, 3:]])
        elif n_out > n_in:
            holes = out_idx[k:]
            n_new = n - holes.shape[0]
            tail = torch.arange(n_new, n, device=pos.device)
            tail_keep = tail[~torch.isin(tail, holes)]
            low = holes[holes < n_new]
            pos[low] = pos[tail_keep]
            mom[low] = mom[tail_keep]
            pos, mom = pos[:n_new], mom[:n_new]
        return pos, mom

    # -- public ---------------------------------------------------------------------------------------
    def accel(self, pos, n_total: int = 0, periodic_pp=False):
        c = self.cfg
        acc = self._acc_buf(pos.shape[0])
        if c.method == "pp":
            return self.pp(pos, acc, n_total or c.npart, periodic_pp)
        if self.world > 1 and c.method == "p3m":
            pos, _ = self._own(pos)
        self.pm_mesh(pos)
        if self.world > 1:
            self.mesh.check_outside()
        self._gather(pos, None, acc, 0.0, 0.0)
        if c.method == "p3m":
            self.short_range(pos, acc, accumulate=True)
            if self._hover is not None:
                h_over, self._hover = int(self._hover.item()), None
                self._raise_flags(0, h_over)
        return acc

    def step(self, pos, mom, dt_kick, dt_drift, mark=None, n_total: int = 0, periodic_pp=False, tail_waits=None):
        """tail_waits: a stream the momenta are still being written on (step_host's upload); the step waits for it
        right before the kick, the first use of `mom`."""
        c = self.cfg
        mark = mark or (lambda name, fn: fn())
        self._mark = mark

        def mom_ready():
            if tail_waits is not None:
                torch.cuda.current_stream(self.eng.device).wait_stream(tail_waits)

        try:
            kick = c.mass * dt_kick
            drift = dt_drift / (c.mass * c.a * c.a)
            if c.method == "pp":
                acc = self._acc_buf(pos.shape[0])
                self.pp(pos, acc, n_total or c.npart, periodic_pp)
                mom_ready()
                mark("kick_drift", lambda: self.eng.kick_drift(pos, mom, acc, kick, drift, c.boxsize))
                return pos, mom
            if self.world > 1:
                pos, mom = self._own(pos, mom)
            if c.method == "pm":
                self.pm_mesh(pos)
                mom_ready()
                mark("gather_kick_drift", lambda: self._gather(pos, mom, None, kick, drift))
            else:   # short range into its own buffer, mesh chain, one fused gather + add + kick + drift
                acc_s = self._acc_buf(pos.shape[0])
                if self.overlap is True or (self.overlap == "auto" and pos.shape[0] >= self.overlap_min_particles):
                    f, fs = self._short_and_mesh_side_by_side(pos, acc_s)
                else:
                    self.short_range(pos, acc_s, False)
                    f, fs = self.pm_mesh(pos)
                lay = self.mesh.lay
                mom_ready()
                mark("gather_kick_drift", lambda: self.eng.gather_kick_drift_add(
                    f, fs, c.gridsize, c.cell, pos, mom, acc_s, None, 1.0, kick, drift, c.boxsize, x0=lay.x0,
                    nx_own=lay.nxl, n_ghost=lay.n_ghost))
            return mark("migrate", lambda: self.migrate(pos, mom))
        finally:
            self._mark = lambda name, fn: fn()

    def step_host(self, hpos, hmom, out_pos, out_mom, dt_kick, dt_drift, mark=None, n_total: int = 0,
                  periodic_pp=False):
        """One step of this rank's shard on HOST buffers (pinned (n,3) float64 CPU tensors; out_pos / out_mom with
        room for the particles that migrate in).  Like ForceStepper.step_host: positions are uploaded first, the
        momenta - first needed by the kick - on a side stream under the force computation, the updated shard is
        copied back behind the migration.  Asynchronous on the current stream; returns the new particle count
        of the rank (rows of out_pos / out_mom that are valid once the stream is synchronised)."""
        dev = self.eng.device
        n = hpos.shape[0]
        if self.cfg.method == "pp" or self.world == 1:
            if self._pbuf is None or self._pbuf.shape[0] < n:
                self._pbuf = torch.empty((n, 3), dtype=torch.float64, device=dev)
                self._mbuf = torch.empty((n, 3), dtype=torch.float64, device=dev)
        else:
            self._own(self._pbuf[:n] if self._pbuf is not None and self._pbuf.shape[0] >= n
                      else torch.empty((n, 3), dtype=torch.float64, device=dev))       # reserves the capacity
        if getattr(self, "_side", None) is None:
            self._side = torch.cuda.Stream(device=dev)
        main = torch.cuda.current_stream(dev)
        pos, mom = self._pbuf[:n], self._mbuf[:n]
        pos.copy_(hpos, non_blocking=True)
        self._side.wait_stream(main)                      # behind the position upload (and earlier readers of mom)
        with torch.cuda.stream(self._side):
            mom.copy_(hmom, non_blocking=True)
        pos2, mom2 = self.step(pos, mom, dt_kick, dt_drift, mark=mark, n_total=n_total, periodic_pp=periodic_pp,
                               tail_waits=self._side)
        k = pos2.shape[0]
        if k > out_pos.shape[0] or k > out_mom.shape[0]:
            from ._lib import NbodyError
            raise NbodyError(f"step_host: {k} particles after migration, output buffers hold {out_pos.shape[0]}")
        out_pos[:k].copy_(pos2, non_blocking=True)
        out_mom[:k].copy_(mom2, non_blocking=True)
        return k


# ------------------------------------------------------------------------------- P(k)
class DistPowerSpectrum:
    """Snapshot P(k) (Power_compare.py:135-200, use_fac=False, no deconvolution) on slabs: each
    rank deposits its particles, the slab FFT transposes, each rank bins its ky-slab and the
    299 (count, sum) pairs are all-reduced.  Returns (edges, counts int64, sums float64)."""

    def __init__(self, engine, comm, N: int, L: float, n_edges: int = 300):
        self.eng, self.comm, self.N, self.L, self.n_edges = engine, comm, int(N), float(L), n_edges
        self.mesh = SlabMesh(engine, comm, N, L, 1, 0)
        self.counts = torch.zeros(n_edges - 1, dtype=torch.int64, device=engine.device)
        self.sums = torch.zeros(n_edges - 1, dtype=torch.float64, device=engine.device)

    def __call__(self, pos, n_total: int, mark=None, check=True):
        mark = mark or (lambda name, fn: fn())
        m, N, L = self.mesh, self.N, self.L
        mark("deposit + ghost add", lambda: m.deposit(pos, 1))
        if check and self.comm.world > 1:
            m.check_outside()
        own = m.rho[0, :m.lay.nxl]
        mark("overdensity", lambda: self.eng.grid_affine(own, a=(L / N) ** 3, b=n_total / L ** 3, c=1.0))
        mark("slab fft + all-to-all", lambda: m.forward(0))
        self.counts.zero_()
        self.sums.zero_()
        edges = mark("shell_bin", lambda: self.eng.pk_bin_slab(m.rhok[0], L, N, m.lay.y0, m.lay.nyl, self.counts,
                                                              self.sums, self.n_edges, norm=L ** 3 / N ** 6))
        mark("all-reduce", lambda: (self.comm.all_reduce(self.counts), self.comm.all_reduce(self.sums)))
        return edges, self.counts, self.sums
