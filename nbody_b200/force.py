"""PP / PM / P3M force step driven by the reference's `.param` files (SURVEY §8a', §8b
"+ new: force_step").  The reference has no source for this path (it launches CONCEPT as a
subprocess, NeuNetværk_test.py:155-171); the step here is the in-process replacement:

    PM   deposit -> R2C -> fused k-space kernel -> 3 C2R -> fused gather+kick+drift
    P3M  PM with the Gaussian long-range filter  +  cell sort -> short-range pair kernel
    PP   float4 pack -> shared-memory j-tile direct sum

Single-GPU orchestration lives here; dist.py shards it over ranks.  Every numeric step is
a kernel of libnbody_b200.so (via engine.Engine); there is no CPU path.
"""
from __future__ import annotations

import math
import os

import numpy as np
import torch

from .engine import Engine, default_engine
from .params import ForceConfig, force_config


class ForceStepper:
    """Holds the device buffers of one force configuration and runs steps on resident
    particle tensors (pos, mom: contiguous (Np,3) float64 on the engine's device)."""

    def __init__(self, cfg: ForceConfig, engine: Engine | None = None, pp_variant: int = 0):
        self.cfg = cfg
        self.eng = engine or default_engine()
        self.pp_variant = pp_variant
        self.pp_near_fix = True
        # P3M short range with the pairs inside 20 softening lengths in float64 (nb_p3m_short_exact): 99.9 % of the
        # particles within 1e-6 of the float64 sum instead of 3e-5 and 1e-5 in the max norm, for +21 % pair-kernel time
        #  - off by
        # default, NB_P3M_NEAR=1 or this attribute turns it on
        self.p3m_near_fix = os.environ.get("NB_P3M_NEAR", "0") == "1"
        self.p3m_near_eps = float(os.environ.get("NB_P3M_RNEAR", "20"))     # r_near in softening lengths
        self._force = None
        self._force_s = None
        self._acc = None
        if cfg.method not in ("pp", "pm", "p3m"):
            raise ValueError(f"unknown force method {cfg.method!r}")
        if cfg.method in ("pm", "p3m") and cfg.gridsize < 2:
            raise ValueError("PM/P3M need a potential grid size >= 2")
        if cfg.method == "p3m":
            self.ncs = int(math.floor(cfg.boxsize / cfg.r_cut))
            if self.ncs < 3:
                raise ValueError("P3M short range needs boxsize >= 3 r_cut")
        # P3M: the short-range part goes into its own buffer and ONE fused tail kernel does
        # gather(mesh) + short -> kick -> drift.  At large N the pair kernel and the mesh chain run side by
        # side on an SM partition: the pair kernel, a persistent grid confined to the first
        # `p3m_sms` SMs, against the memory-bound mesh chain on the rest.  Measured on B200 at config 5:
        # 140.4 ms per step at 88 of 148 SMs against 162.6 ms one after the other (80-104 SMs: 140-143 ms).
        # NB_P3M_OVERLAP=0/1 forces it off/on; the default turns it on from 2^26 particles (the size it was
        # measured at is 2^27; small steps are launch-bound and gain nothing from it).
        self.fused_tail = cfg.method == "p3m"
        ov = os.environ.get("NB_P3M_OVERLAP", "auto")
        self.overlap = cfg.method == "p3m" and (True if ov == "1" else "auto" if ov == "auto" else False)
        self.overlap_min_particles = 1 << 26
        self.p3m_sms = int(os.environ.get("NB_P3M_SMS", "0"))      # 0: 0.595 x the device's SM count (88 of 148)
        self.p3m_delay_us = int(os.environ.get("NB_P3M_DELAY_US", "40"))
        self.p3m_sms_predeposited = int(os.environ.get("NB_P3M_SMS_HOST", "0"))   # 0: 1.07 x p3m_sms
        self.host_predeposit = os.environ.get("NB_HOST_PREDEPOSIT", "1") == "1"
        self.host_predeposit_min = 1 << 22        # particles; below, one upload and the plain mesh solve
        self._acc_s = None
        self._streams = None

    # -- buffers --------------------------------------------------------------------------
    def _acc_buf(self, n):
        if self._acc is None or self._acc.shape[0] != n:
            self._acc = torch.empty((n, 3), dtype=torch.float64, device=self.eng.device)
        return self._acc

    def _force_bufs(self):
        N = self.cfg.gridsize
        if self._force is None:
            self._force = torch.empty((3, N, N, N), dtype=torch.float32, device=self.eng.device)
            if self.cfg.interlace[1]:
                self._force_s = torch.empty((3, N, N, N), dtype=torch.float32, device=self.eng.device)
        return self._force, self._force_s

    # -- pieces ---------------------------------------------------------------------------
    def pm_mesh(self, pos, rho=None):
        """Long-range force meshes (acceleration per unit... in the caller's units)."""
        c = self.cfg
        N = c.gridsize
        f, fs = self._force_bufs()
        pref = 4.0 * math.pi * c.G * c.mass / c.cell ** 3 / float(N) ** 3
        if rho is not None:           # meshes deposited by the caller (step_host: chunk by chunk under the upload)
            self.eng.pm_force_mesh_from_rho(rho[0], rho[1], c.boxsize, N, c.r_s, c.deconv_order, bool(c.interlace[1]),
                                            pref, f, fs)
            return f, fs
        self.eng.pm_force_mesh(pos, c.boxsize, N, r_s=c.r_s, deconv_order=c.deconv_order,
                               interlace_up=bool(c.interlace[0]), interlace_down=bool(c.interlace[1]),
                               pref=pref, out=f, out_shifted=fs)
        return f, fs

    def _near(self, pos):
        """Keyword arguments of engine.p3m_short for the float64 near-pair pass (pairs inside 20 softening lengths,
        as for PP), or nothing when it is switched off / there is no softening scale."""
        c = self.cfg
        r_near = self.p3m_near_eps * c.eps
        if not self.p3m_near_fix or c.eps <= 0 or r_near >= c.r_cut:
            return {}
        return {"pos": pos, "r_near": r_near}

    def short_range(self, pos, acc, accumulate=True):
        c = self.cfg
        sorted4, perm, cell_start = self.eng.cell_sort(pos, c.boxsize, self.ncs, mass=1.0)
        self.eng.p3m_short(sorted4, perm, cell_start, c.boxsize, self.ncs, c.r_s, c.r_cut, c.eps,
                           scale=c.G * c.mass, out=acc, accumulate=accumulate, **self._near(pos))
        return acc

    def pp(self, pos, acc, periodic=False):
        """periodic: False (open), True (nearest image only) or "ewald" (nearest image + Ewald
        correction table, CONCEPT's periodic PP)."""
        c = self.cfg
        half = 0.5 * c.boxsize
        p4 = self.eng.pp_pack(pos, origin=(half, half, half), mass=1.0)
        if periodic == "ewald":
            g = int(getattr(self, "ewald_gridsize", 64))
            self.eng.pp_accel_ewald(p4, p4, c.eps, c.boxsize, self.eng.ewald_table(c.boxsize, g), g,
                                    scale=c.G * c.mass, out=acc)
            return acc
        self.eng.pp_accel(p4, p4, c.eps, box=c.boxsize if periodic else 0.0, scale=c.G * c.mass,
                          variant=self.pp_variant, out=acc)
        if self.pp_near_fix and c.eps > 0:
            # pairs closer than 20 softening lengths once more in float64 (the packed coordinates keep 24 bits)
            self.eng.pp_near_fix(pos, c.boxsize, c.eps, 20.0 * c.eps, (half, half, half), acc, periodic=bool(periodic),
                                 scale=c.G * c.mass, mass=1.0)
        return acc

    def _tail_to_host(self, tail, pos, mom, out_pos, out_mom, chunks=8, join=True):
        """Run the per-particle tail kernel `tail(a, b)` chunk by chunk and copy each finished chunk of
        pos/mom to the host buffers on a side stream, so that the device->host copies (the longest part
        of step_host at the C5 size) start as soon as the first chunk is done."""
        dev = self.eng.device
        n = pos.shape[0]
        main = torch.cuda.current_stream(dev)
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
        cs = self._copy_stream
        step = max(1, -(-n // chunks))
        for a in range(0, n, step):
            b = min(n, a + step)
            tail(a, b)
            ev = torch.cuda.Event()
            ev.record(main)
            cs.wait_event(ev)
            with torch.cuda.stream(cs):
                out_pos[a:b].copy_(pos[a:b], non_blocking=True)
                out_mom[a:b].copy_(mom[a:b], non_blocking=True)
        if join:
            main.wait_stream(cs)

    def _p3m_fused(self, pos, mom, acc_out, kick, drift, tail_waits=None, host_out=None, join=True, rho=None):
        """Cell sort + pair kernel into a separate buffer, mesh chain, then ONE fused kernel:
        g = gather(mesh) + short; acc_out <- g and/or kick + drift.  With self.overlap the two
        parts run on two streams and two disjoint sets of SMs."""
        c, eng = self.cfg, self.eng
        dev = eng.device
        n = pos.shape[0]
        if self._acc_s is None or self._acc_s.shape[0] != n:
            self._acc_s = torch.empty((n, 3), dtype=torch.float64, device=dev)
        main = torch.cuda.current_stream(dev)
        if self.overlap is True or (self.overlap == "auto" and n >= self.overlap_min_particles):
            if self._streams is None:
                self._streams = (torch.cuda.Stream(device=dev, priority=-1), torch.cuda.Stream(device=dev))
                if self.p3m_sms <= 0:
                    self.p3m_sms = max(1, round(0.595 * torch.cuda.get_device_properties(dev).multi_processor_count))
            pm_stream, sr_stream = self._streams
            sr_stream.wait_stream(main)
            # the pair kernel takes the first `p3m_sms` SMs for itself (persistent grid, CTAs elsewhere exit),
            # the memory-bound mesh chain runs on the others: FP32-bound and HBM-bound halves of the step side by
            # side.  The pair kernel must find the GPU idle when it starts (a CTA that exits on a foreign SM is not
            # replaced): the cell sort runs alone, and the mesh stream starts with a short delay kernel.
            # with the deposits done already (rho) the mesh side is shorter: the pair kernel gets more of the SMs
            eng.p3m_set_sm_limit(self.p3m_sms if rho is None else
                                 (self.p3m_sms_predeposited or round(self.p3m_sms * 1.07)))
            try:
                with torch.cuda.stream(sr_stream):
                    sorted4, perm, cell_start = eng.cell_sort(pos, c.boxsize, self.ncs, mass=1.0)
                    sorted_ev = torch.cuda.Event()
                    sorted_ev.record(sr_stream)
                    eng.p3m_short(sorted4, perm, cell_start, c.boxsize, self.ncs, c.r_s, c.r_cut, c.eps,
                                  scale=c.G * c.mass, out=self._acc_s, accumulate=False, **self._near(pos))
                pm_stream.wait_event(sorted_ev)
                with torch.cuda.stream(pm_stream):
                    eng.delay_us(self.p3m_delay_us)
                    f, fs = self.pm_mesh(pos, rho=rho)
                    eng.p3m_mesh_done()
            finally:
                eng.p3m_set_sm_limit(0)
            for t in (sorted4, perm, cell_start):          # freed on this stream, used on sr_stream
                t.record_stream(sr_stream)
            main.wait_stream(pm_stream)
            main.wait_stream(sr_stream)
        else:
            self.short_range(pos, self._acc_s, accumulate=False)
            f, fs = self.pm_mesh(pos, rho=rho)
        if tail_waits is not None:            # e.g. the momentum upload of step_host
            main.wait_stream(tail_waits)
        if host_out is not None and mom is not None and acc_out is None:
            self._tail_to_host(lambda a, b: eng.gather_kick_drift_add(f, fs, c.gridsize, c.cell, pos[a:b], mom[a:b],
                                                                      self._acc_s[a:b], None, 1.0, kick, drift,
                                                                      c.boxsize),
                               pos, mom, host_out[0], host_out[1], join=join)
            return
        eng.gather_kick_drift_add(f, fs, c.gridsize, c.cell, pos, mom, self._acc_s, acc_out, 1.0, kick, drift,
                                  c.boxsize)

    # -- public ---------------------------------------------------------------------------
    def accel(self, pos: torch.Tensor, periodic_pp: bool = False) -> torch.Tensor:
        """Acceleration at the particles, (Np,3) float64 (no kick/drift)."""
        c = self.cfg
        acc = self._acc_buf(pos.shape[0])
        if c.method == "pp":
            return self.pp(pos, acc, periodic_pp)
        if self.fused_tail:
            self._p3m_fused(pos, None, acc, 0.0, 0.0)
            return acc
        f, fs = self.pm_mesh(pos)
        self.eng.gather_kick_drift(f, fs, c.gridsize, c.cell, pos, None, acc, 1.0, 0.0, 0.0, c.boxsize)
        if c.method == "p3m":
            self.short_range(pos, acc, accumulate=True)
        return acc

    # -- split interface of the sub-cycled time loop (timeloop.evolve_subcycled) ----------------
    def accel_long(self, pos: torch.Tensor):
        """Mesh (long-range) part of the acceleration, or None for a PP configuration."""
        c = self.cfg
        if c.method == "pp":
            return None
        f, fs = self.pm_mesh(pos)
        acc = self._acc_buf(pos.shape[0])
        self.eng.gather_kick_drift(f, fs, c.gridsize, c.cell, pos, None, acc, 1.0, 0.0, 0.0, c.boxsize)
        return acc

    def accel_short(self, pos: torch.Tensor):
        """Pair (short-range) part: the P³M short range, the whole force for PP, None for PM."""
        c = self.cfg
        if c.method == "pm":
            return None
        n = pos.shape[0]
        if self._acc_s is None or self._acc_s.shape[0] != n:
            self._acc_s = torch.empty((n, 3), dtype=torch.float64, device=self.eng.device)
        if c.method == "pp":
            return self.pp(pos, self._acc_s, False)
        return self.short_range(pos, self._acc_s, accumulate=False)

    def kick(self, pos, mom, acc, k):
        self.eng.kick_drift(pos, mom, acc, float(k), 0.0, self.cfg.boxsize)

    def kick_drift(self, pos, mom, acc, k, d):
        if acc is None:                      # drift only (PM between two mesh kicks)
            z = getattr(self, "_zero_acc", None)
            if z is None or z.shape[0] != pos.shape[0]:
                z = self._zero_acc = torch.zeros((pos.shape[0], 3), dtype=torch.float64, device=self.eng.device)
            acc, k = z, 0.0
        self.eng.kick_drift(pos, mom, acc, float(k), float(d), self.cfg.boxsize)

    def max_norm2(self, acc) -> float:
        return float(self.eng.max_norm2(acc).item())

    def step(self, pos: torch.Tensor, mom: torch.Tensor, dt_kick: float, dt_drift: float,
             periodic_pp: bool = False):
        """One kick-drift step in place: mom += m·g·Δt_kick; pos += mom/(m a²)·Δt_drift; pos %= L."""
        c = self.cfg
        kick = c.mass * dt_kick
        drift = dt_drift / (c.mass * c.a * c.a)
        if c.method == "pm":
            f, fs = self.pm_mesh(pos)
            self.eng.gather_kick_drift(f, fs, c.gridsize, c.cell, pos, mom, None, 1.0, kick, drift, c.boxsize)
            return pos, mom
        if self.fused_tail:
            self._p3m_fused(pos, mom, None, kick, drift)
            return pos, mom
        acc = self.accel(pos, periodic_pp)
        self.eng.kick_drift(pos, mom, acc, kick, drift, c.boxsize)
        return pos, mom


    def _upload_and_deposit(self, hpos, dpos, chunks=8):
        """Position upload in chunks on the current stream; on a second stream each chunk is deposited into the
        (two, with upstream interlacing) density meshes as soon as it has arrived, so the deposits - 11 ms of the mesh
        chain at config 5 - are done when the last chunk lands instead of starting then.  Returns the meshes."""
        c, eng = self.cfg, self.eng
        dev, N = eng.device, c.gridsize
        if getattr(self, "_rho", None) is None:
            self._rho = [torch.empty((N, N, N), dtype=torch.float32, device=dev),
                         torch.empty((N, N, N), dtype=torch.float32, device=dev) if c.interlace[0] else None]
            self._dep_stream = torch.cuda.Stream(device=dev)
        main, dep = torch.cuda.current_stream(dev), self._dep_stream
        dep.wait_stream(main)                       # the previous step's transforms are done reading the meshes
        n = hpos.shape[0]
        step = max(1, -(-n // chunks))
        for a in range(0, n, step):
            b = min(n, a + step)
            dpos[a:b].copy_(hpos[a:b], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(main)
            dep.wait_event(ev)
            with torch.cuda.stream(dep):
                eng.cic_deposit(dpos[a:b], c.cell, N, out=self._rho[0], accumulate=a > 0)
                if self._rho[1] is not None:
                    eng.cic_deposit(dpos[a:b], c.cell, N, out=self._rho[1], shift=0.5 * c.cell, accumulate=a > 0)
        main.wait_stream(dep)
        return self._rho

    def step_host_pipelined(self, hpos: torch.Tensor, hmom: torch.Tensor, out_pos: torch.Tensor, out_mom: torch.Tensor,
                            dt_kick: float, dt_drift: float):
        """step_host for a stream of INDEPENDENT particle sets (PM / P3M): two sets of device buffers, so the upload of
        call k+1 runs while call k computes and downloads (PCIe is full duplex and the copy engines run beside the
        kernels).  Every call still uploads its own positions and momenta and downloads its own result; calls return
        at once, and `finish_host()` makes the current stream wait for the downloads in flight - call it (and
        synchronise) before reading out_pos / out_mom or reusing hpos / hmom.  The steps of ONE time loop depend on
        each other and keep their particles on the device: this is for batches (many snapshots, many runs)."""
        c, eng = self.cfg, self.eng
        if c.method not in ("pm", "p3m"):
            raise ValueError("step_host_pipelined: PM / P3M only")
        n = hpos.shape[0]
        dev = eng.device
        if getattr(self, "_pipe", None) is None or self._pipe["n"] != n:
            self._pipe = {"n": n, "k": 0, "up": torch.cuda.Stream(device=dev),
                          "sets": [(torch.empty((n, 3), dtype=torch.float64, device=dev),
                                    torch.empty((n, 3), dtype=torch.float64, device=dev)) for _ in range(2)],
                          "free": [None, None]}
            if getattr(self, "_copy_stream", None) is None:
                self._copy_stream = torch.cuda.Stream(device=dev)
        pp = self._pipe
        b = pp["k"] & 1
        pp["k"] += 1
        dpos, dmom = pp["sets"][b]
        up, main = pp["up"], torch.cuda.current_stream(dev)
        if pp["free"][b] is not None:
            up.wait_event(pp["free"][b])                  # the download of the call that used this set is done
        else:
            up.wait_stream(main)
        with torch.cuda.stream(up):
            dpos.copy_(hpos, non_blocking=True)
            pos_ev = torch.cuda.Event()
            pos_ev.record(up)
            dmom.copy_(hmom, non_blocking=True)
        main.wait_event(pos_ev)
        kick = c.mass * dt_kick
        drift = dt_drift / (c.mass * c.a * c.a)
        if c.method == "pm":
            f, fs = self.pm_mesh(dpos)
            main.wait_stream(up)
            self._tail_to_host(lambda a, e: eng.gather_kick_drift(f, fs, c.gridsize, c.cell, dpos[a:e], dmom[a:e], None,
                                                                  1.0, kick, drift, c.boxsize),
                               dpos, dmom, out_pos, out_mom, join=False)
        else:
            self._p3m_fused(dpos, dmom, None, kick, drift, tail_waits=up, host_out=(out_pos, out_mom), join=False)
        ev = torch.cuda.Event()
        ev.record(self._copy_stream)
        pp["free"][b] = ev
        return out_pos, out_mom

    def finish_host(self):
        """Current stream waits for the downloads of step_host_pipelined calls in flight."""
        if getattr(self, "_copy_stream", None) is not None:
            torch.cuda.current_stream(self.eng.device).wait_stream(self._copy_stream)

    def step_host(self, hpos: torch.Tensor, hmom: torch.Tensor, out_pos: torch.Tensor, out_mom: torch.Tensor,
                  dt_kick: float, dt_drift: float, periodic_pp: bool = False):
        """One kick-drift step on HOST buffers: hpos/hmom/out_pos/out_mom are (Np,3) float64 CPU tensors
        (pinned for asynchronous copies).  Positions are uploaded first; the momenta — first needed
        by the kick at the end of the step — are uploaded on a side stream while the force is
        computed; the updated pos/mom are then copied back.  Asynchronous on the current stream:
        the caller synchronises before reading out_pos/out_mom."""
        n = hpos.shape[0]
        dev = self.eng.device
        if getattr(self, "_dpos", None) is None or self._dpos.shape[0] != n:
            self._dpos = torch.empty((n, 3), dtype=torch.float64, device=dev)
            self._dmom = torch.empty((n, 3), dtype=torch.float64, device=dev)
            self._side = torch.cuda.Stream(device=dev)
        main = torch.cuda.current_stream(dev)
        rho = None
        if self.host_predeposit and self.cfg.method in ("pm", "p3m") and n >= self.host_predeposit_min:
            rho = self._upload_and_deposit(hpos, self._dpos)
        else:
            self._dpos.copy_(hpos, non_blocking=True)
        # the momenta go up behind the positions, not beside them: two uploads released at the same moment (calls
        # enqueued back to back) share the link and the force computation starts late (measured: +20 ms per call)
        self._side.wait_stream(main)          # ... and the previous step is done with _dmom
        with torch.cuda.stream(self._side):
            self._dmom.copy_(hmom, non_blocking=True)
        c = self.cfg
        kick = c.mass * dt_kick
        drift = dt_drift / (c.mass * c.a * c.a)
        if c.method == "pm":
            f, fs = self.pm_mesh(self._dpos, rho=rho)
            main.wait_stream(self._side)
            self._tail_to_host(lambda a, b: self.eng.gather_kick_drift(f, fs, c.gridsize, c.cell, self._dpos[a:b],
                                                                       self._dmom[a:b], None, 1.0, kick, drift,
                                                                       c.boxsize),
                               self._dpos, self._dmom, out_pos, out_mom)
            return out_pos, out_mom
        elif self.fused_tail:
            self._p3m_fused(self._dpos, self._dmom, None, kick, drift, tail_waits=self._side,
                            host_out=(out_pos, out_mom), rho=rho)
            return out_pos, out_mom
        else:
            acc = self.accel(self._dpos, periodic_pp)
            main.wait_stream(self._side)
            self.eng.kick_drift(self._dpos, self._dmom, acc, kick, drift, c.boxsize)
        out_pos.copy_(self._dpos, non_blocking=True)
        out_mom.copy_(self._dmom, non_blocking=True)
        return out_pos, out_mom


def _cfg_from(param):
    if isinstance(param, ForceConfig):
        return param
    return force_config(param)


def accelerations(param, pos, **kw):
    """numpy in / numpy out: accelerations of all particles under `param` (a ForceConfig, a
    `.param` path or its text)."""
    cfg = _cfg_from(param)
    eng = default_engine()
    p = np.ascontiguousarray(pos, dtype=np.float64)
    if p.ndim != 2 or p.shape[1] != 3:
        raise ValueError("pos must have shape (Np,3)")
    if not cfg.npart:
        cfg.npart = p.shape[0]
    st = ForceStepper(cfg, eng)
    return st.accel(eng.to_device(p), **kw).cpu().numpy()


def force_step(param, pos, mom, dt_kick, dt_drift, **kw):
    """numpy in / numpy out: one kick-drift step; returns new (pos, mom) arrays, the inputs
    are not modified (the reference's functions return fresh arrays)."""
    cfg = _cfg_from(param)
    eng = default_engine()
    p = np.ascontiguousarray(pos, dtype=np.float64)
    m = np.ascontiguousarray(mom, dtype=np.float64)
    if p.ndim != 2 or p.shape[1] != 3 or m.shape != p.shape:
        raise ValueError("pos and mom must both have shape (Np,3)")
    if not cfg.npart:
        cfg.npart = p.shape[0]
    st = ForceStepper(cfg, eng)
    pt, mt = eng.to_device(p), eng.to_device(m)
    st.step(pt, mt, dt_kick, dt_drift, **kw)
    return pt.cpu().numpy(), mt.cpu().numpy()
