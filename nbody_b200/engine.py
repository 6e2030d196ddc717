"""Device-level operators: torch CUDA tensors in, torch CUDA tensors out, every operator a
call into libnbody_b200.so on torch's current stream.  PyTorch is used for device memory,
streams and (in dist.py) torch.distributed only.

The numpy-in / numpy-out functions with the reference's names live in api.py and are thin
wrappers over this class.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib, tables

F32, F64 = 0, 1


def _prec_of(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.float64:
        return F64
    raise TypeError(f"grid dtype must be float32 or float64, got {t.dtype}")


def _ptr(t):
    return C.c_void_p(0) if t is None else C.c_void_p(t.data_ptr())


class Engine:
    """One context per CUDA device (cuFFT plans + scratch live in it)."""

    def __init__(self, device: int | None = None):
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise RuntimeError("nbody_b200 needs a CUDA device (there is no CPU fallback)")
        if device is None:
            device = torch.cuda.current_device()
        self.device = torch.device("cuda", int(device))
        h = C.c_void_p()
        _lib.check(self.lib.nb_init(int(device), C.byref(h)), "nb_init")
        self.ctx = h
        self._tab_cache: dict = {}

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.nb_destroy(self.ctx)
            self.ctx = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    # -- plumbing ---------------------------------------------------------------------
    @property
    def stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def launch_count(self) -> int:
        return int(self.lib.nb_launch_count(self.ctx))

    PHASES = ("deposit", "fft_fwd", "kspace", "fft_inv", "pk_bin", "sort")

    def profile_enable(self, on: bool = True):
        _lib.check(self.lib.nb_profile_enable(self.ctx, int(bool(on))), "nb_profile_enable")

    def profile_read(self) -> dict:
        """{phase: (total ms, intervals)} recorded since the last read (device-side CUDA events)."""
        n = len(self.PHASES)
        ms = (C.c_double * n)()
        cnt = (C.c_int64 * n)()
        _lib.check(self.lib.nb_profile_read(self.ctx, ms, cnt, n), "nb_profile_read")
        return {self.PHASES[i]: (float(ms[i]), int(cnt[i])) for i in range(n) if cnt[i]}

    def fp32_peak(self, packed: bool = True, iters: int = 4096, reps: int = 5) -> float:
        """Measured FP32 FMA throughput of this device in TFLOP/s (FFMA2 if packed, scalar FFMA else)."""
        out = C.c_double(0.0)
        _lib.check(self.lib.nb_fp32_peak(self.ctx, int(bool(packed)), int(iters), int(reps), C.byref(out),
                                         self.stream), "nb_fp32_peak")
        return float(out.value)

    def release_scratch(self):
        _lib.check(self.lib.nb_release_scratch(self.ctx), "nb_release_scratch")

    def to_device(self, arr, dtype=None) -> torch.Tensor:
        """numpy (pageable or pinned) -> device tensor on the current stream."""
        if isinstance(arr, torch.Tensor):
            t = arr
        else:
            a = np.ascontiguousarray(arr, dtype=dtype)
            t = torch.from_numpy(a)
        return t.to(self.device, non_blocking=True)

    def table(self, key, fn) -> torch.Tensor:
        """Cached device copy of a small host table."""
        t = self._tab_cache.get(key)
        if t is None:
            t = torch.from_numpy(np.ascontiguousarray(fn(), dtype=np.float64)).to(self.device)
            self._tab_cache[key] = t
        return t

    def _check_pos(self, pos: torch.Tensor):
        if pos.dtype != torch.float64 or pos.dim() != 2 or pos.shape[1] != 3 or not pos.is_contiguous():
            raise ValueError("positions must be a contiguous (Np,3) float64 tensor")
        if pos.device != self.device:
            raise ValueError(f"positions live on {pos.device}, engine on {self.device}")

    # -- K1 ---------------------------------------------------------------------------
    def cic_deposit(self, pos: torch.Tensor, cell: float, N: int, out: torch.Tensor | None = None,
                    dtype=torch.float32, shift: float = 0.0, accumulate: bool = False,
                    dbg_index: torch.Tensor | None = None) -> torch.Tensor:
        self._check_pos(pos)
        if out is None:
            out = torch.empty((N, N, N), dtype=dtype, device=self.device)
            accumulate = False
        if out.shape != (N, N, N) or not out.is_contiguous():
            raise ValueError("grid must be a contiguous (N,N,N) tensor")
        if dbg_index is not None and (dbg_index.dtype != torch.int32 or dbg_index.shape != pos.shape):
            raise ValueError("dbg_index must be (Np,3) int32")
        _lib.check(self.lib.nb_cic_deposit(self.ctx, _ptr(pos), pos.shape[0], float(cell), int(N), _ptr(out),
                                           _prec_of(out), float(shift), int(bool(accumulate)),
                                           _ptr(dbg_index), self.stream), "nb_cic_deposit")
        return out

    # -- K8 ---------------------------------------------------------------------------
    def grid_sum(self, grid: torch.Tensor) -> torch.Tensor:
        out = torch.empty(1, dtype=torch.float64, device=self.device)
        _lib.check(self.lib.nb_grid_sum(self.ctx, _ptr(grid), grid.numel(), _prec_of(grid), _ptr(out),
                                        self.stream), "nb_grid_sum")
        return out

    def grid_affine(self, grid: torch.Tensor, a: float = 1.0, b: float = 1.0, c: float = 0.0,
                    a_dev: torch.Tensor | None = None, a_div: float = 1.0) -> torch.Tensor:
        """grid <- (grid / a) / b - c in place (a may come from a device scalar / a_div)."""
        _lib.check(self.lib.nb_grid_affine(self.ctx, _ptr(grid), grid.numel(), _prec_of(grid), float(a),
                                           _ptr(a_dev), float(a_div), float(b), float(c), self.stream),
                   "nb_grid_affine")
        return grid

    def max_norm2(self, vec: torch.Tensor) -> torch.Tensor:
        """max_i |v_i|² of an (n,3) float64 tensor as a device scalar (nb_vec3_max_norm2)."""
        self._check_pos(vec)
        out = torch.empty(1, dtype=torch.float64, device=self.device)
        _lib.check(self.lib.nb_vec3_max_norm2(self.ctx, _ptr(vec), vec.shape[0], _ptr(out), self.stream),
                   "nb_vec3_max_norm2")
        return out

    def compute_delta(self, grid: torch.Tensor) -> torch.Tensor:
        """density/density.mean() - 1 in place (FraP_kTilDelta.py:126-127)."""
        s = self.grid_sum(grid)
        return self.grid_affine(grid, a_dev=s, a_div=float(grid.numel()), b=1.0, c=1.0)

    # -- a3 ---------------------------------------------------------------------------
    def deconvolve_cic(self, grid: torch.Tensor, L: float, N: int, guard: bool = True,
                       kq_form: str = "inv_d") -> torch.Tensor:
        s = self.sinc_tab(N, L, kq_form)
        _lib.check(self.lib.nb_deconvolve_cic(self.ctx, _ptr(grid), int(N), _prec_of(grid), _ptr(s),
                                              int(bool(guard)), self.stream), "nb_deconvolve_cic")
        return grid

    # -- §8f row 4: batched snapshot -> delta ----------------------------------------------
    def snapshot_delta_batch(self, pos_cat: torch.Tensor, offsets, boxsizes, N: int, subtract_mean=None,
                             guard: bool = True, dtype=torch.float32) -> torch.Tensor:
        """B snapshots -> (B,N,N,N) delta in one call (Feeder_pipe.py:163-195).  pos_cat: concatenated
        (total,3) float64 device tensor; offsets: B+1 ints; boxsizes: B floats; subtract_mean: None or
        B flags (delta -= delta.mean() for the training/validation set, Feeder_pipe.py:193)."""
        offsets = np.asarray(offsets, dtype=np.int64)
        B = offsets.size - 1
        boxsizes = np.asarray(boxsizes, dtype=np.float64).reshape(B)
        if B < 1 or offsets[0] != 0 or np.any(np.diff(offsets) < 0) or offsets[-1] != pos_cat.shape[0]:
            raise ValueError("offsets must start at 0, be non-decreasing and end at the particle count")
        if pos_cat.dtype != torch.float64 or not pos_cat.is_contiguous():
            raise ValueError("pos_cat must be a contiguous float64 (total,3) tensor")
        N = int(N)
        d_off = torch.from_numpy(offsets).to(self.device)
        d_cell = torch.from_numpy(boxsizes / N).to(self.device)
        tabs = np.stack([tables.sinc_table(N, float(L)) for L in boxsizes])
        d_tabs = torch.from_numpy(np.ascontiguousarray(tabs)).to(self.device)
        d_flags = None
        if subtract_mean is not None:
            d_flags = torch.from_numpy(np.asarray(subtract_mean, dtype=np.int32).reshape(B)).to(self.device)
        grids = torch.empty((B, N, N, N), dtype=dtype, device=self.device)
        _lib.check(self.lib.nb_snapshot_delta_batch(self.ctx, _ptr(pos_cat), _ptr(d_off), int(offsets[-1]), int(B),
                                                    _ptr(d_cell), N, _ptr(d_tabs), int(bool(guard)),
                                                    _ptr(d_flags) if d_flags is not None else None, _ptr(grids),
                                                    _prec_of(grids), self.stream), "nb_snapshot_delta_batch")
        return grids

    def snapshot_pk_batch(self, pos_cat: torch.Tensor, offsets, boxsizes, N: int, n_edges: int = 300,
                          window_power: int = 0, clip_mode: int = 0, clip_val: float = 0.0, closed_last: bool = True,
                          kq_form: str = "inv_d", norms=None):
        """B snapshots sharing the mesh size N -> (edges list, counts (B, n_edges-1) int64, sums float64) in one
        library call (nb_snapshot_pk_batch; Plotter.py:205-238).  norms: per-snapshot factor on |δ_k|²
        (default V / N⁶)."""
        offsets = np.asarray(offsets, dtype=np.int64)
        B = offsets.size - 1
        boxsizes = np.asarray(boxsizes, dtype=np.float64).reshape(B)
        if B < 1 or offsets[0] != 0 or np.any(np.diff(offsets) <= 0) or offsets[-1] != pos_cat.shape[0]:
            raise ValueError("offsets must start at 0, increase and end at the particle count")
        if pos_cat.dtype != torch.float64 or not pos_cat.is_contiguous():
            raise ValueError("pos_cat must be a contiguous float64 (total,3) tensor")
        N = int(N)
        fn = {"inv_d": tables.sinc_table, "piN_L": tables.sinc_table_kq}[kq_form]
        edges, thr, t0, t1, thi, k2t, st = [], [], [], [], [], [], []
        for L in boxsizes:
            e = tables.pk_edges(N, float(L), n_edges)
            th, th_hi = tables.sqrt_thresholds(e)
            edges.append(e); thr.append(th); t0.append(th[0]); t1.append(th[-1]); thi.append(th_hi)
            k2t.append(tables.k_vals(N, float(L)) ** 2); st.append(fn(N, float(L)))
        dev = self.device
        up = lambda a, dt=np.float64: torch.from_numpy(np.ascontiguousarray(a, dtype=dt)).to(dev)   # noqa: E731
        d_off, d_cell, d_np = up(offsets, np.int64), up(boxsizes / N), up(np.diff(offsets).astype(np.float64))
        d_k2, d_s, d_thr = up(np.stack(k2t)), up(np.stack(st)), up(np.stack(thr))
        h = lambda a: np.ascontiguousarray(a, dtype=np.float64)                                      # noqa: E731
        h0, h1, hhi = h(t0), h(t1), h(thi)
        hn = h(boxsizes ** 3 / float(N) ** 6 if norms is None else norms)
        counts = torch.empty((B, n_edges - 1), dtype=torch.int64, device=dev)
        sums = torch.empty((B, n_edges - 1), dtype=torch.float64, device=dev)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)                                                  # noqa: E731
        _lib.check(self.lib.nb_snapshot_pk_batch(self.ctx, _ptr(pos_cat), _ptr(d_off), int(offsets[-1]), int(B),
                                                 _ptr(d_cell), _ptr(d_np), N, _ptr(d_k2), _ptr(d_s), _ptr(d_thr),
                                                 vp(h0), vp(h1), vp(hhi), int(n_edges), int(window_power),
                                                 int(clip_mode), float(clip_val), vp(hn), int(bool(closed_last)),
                                                 _ptr(counts), _ptr(sums), self.stream), "nb_snapshot_pk_batch")
        return edges, counts, sums

    def sinc_tab(self, N, L, kq_form="inv_d"):
        fn = {"inv_d": tables.sinc_table, "piN_L": tables.sinc_table_kq}[kq_form]
        return self.table(("sinc", kq_form, N, float(L)), lambda: fn(N, L))

    # -- K6 ---------------------------------------------------------------------------
    def power_spectrum_bins(self, delta: torch.Tensor, L: float, N: int, n_edges: int = 300,
                            window_power: int = 0, clip_mode: int = 0, clip_val: float = 0.0,
                            norm: float = 1.0, closed_last: bool = True, kq_form: str = "inv_d"):
        """R2C FFT of delta, |δ_k|²·norm/W^p binned over the reference's log-spaced shells.
        Returns (edges numpy, counts int64 tensor, sums float64 tensor)."""
        if delta.shape != (N, N, N) or not delta.is_contiguous():
            raise ValueError("delta must be a contiguous (N,N,N) tensor")
        key = ("edges", N, float(L), n_edges)
        if key not in self._tab_cache:
            e = tables.pk_edges(N, L, n_edges)
            thr, thr_hi = tables.sqrt_thresholds(e)
            self._tab_cache[key] = (e, torch.from_numpy(thr).to(self.device), float(thr[0]), float(thr[-1]), thr_hi)
        edges, thr_d, t0, t1, thr_hi = self._tab_cache[key]
        k2 = self.table(("k2", N, float(L)), lambda: tables.k_vals(N, L) ** 2)
        s = self.sinc_tab(N, L, kq_form)
        counts = torch.empty(n_edges - 1, dtype=torch.int64, device=self.device)
        sums = torch.empty(n_edges - 1, dtype=torch.float64, device=self.device)
        _lib.check(self.lib.nb_power_spectrum(self.ctx, _ptr(delta), int(N), _prec_of(delta), _ptr(k2), _ptr(s),
                                              _ptr(thr_d), t0, t1, thr_hi, int(n_edges), int(window_power),
                                              int(clip_mode), float(clip_val), float(norm),
                                              int(bool(closed_last)), _ptr(counts), _ptr(sums), self.stream),
                   "nb_power_spectrum")
        return edges, counts, sums

    def power_spectrum_custom(self, delta: torch.Tensor, delta_shifted: torch.Tensor | None, L: float, N: int,
                              thresholds, window_power: int = 0, norm: float = 1.0, kq_form: str = "inv_d"):
        """|δ_k|²·norm/W^p summed over caller-defined shells: `thresholds` are increasing float64
        values in k² (bin b holds thresholds[b] <= k² < thresholds[b+1]).  With `delta_shifted` (the
        deposit on the half-cell-shifted mesh) the two spectra are interlaced first
        (nb_power_spectrum_interlaced).  Returns (counts int64, sums float64) device tensors."""
        for g in (delta, delta_shifted):
            if g is not None and (g.shape != (N, N, N) or not g.is_contiguous() or g.dtype != delta.dtype):
                raise ValueError("grids must be contiguous (N,N,N) tensors of one dtype")
        thr = np.ascontiguousarray(thresholds, dtype=np.float64)
        if thr.ndim != 1 or thr.size < 2 or not np.all(np.diff(thr) > 0) or thr[0] <= 0:
            raise ValueError("thresholds must be positive and strictly increasing")
        n_edges = int(thr.size)
        thr_d = torch.from_numpy(thr).to(self.device)
        k2 = self.table(("k2", N, float(L)), lambda: tables.k_vals(N, L) ** 2)
        s = self.sinc_tab(N, L, kq_form)
        counts = torch.empty(n_edges - 1, dtype=torch.int64, device=self.device)
        sums = torch.empty(n_edges - 1, dtype=torch.float64, device=self.device)
        args = (_ptr(k2), _ptr(s), _ptr(thr_d), float(thr[0]), float(thr[-1]), float(thr[-1]), n_edges,
                int(window_power), 0, 0.0, float(norm), 0, _ptr(counts), _ptr(sums), self.stream)
        if delta_shifted is None:
            _lib.check(self.lib.nb_power_spectrum(self.ctx, _ptr(delta), int(N), _prec_of(delta), *args),
                       "nb_power_spectrum")
        else:
            _lib.check(self.lib.nb_power_spectrum_interlaced(self.ctx, _ptr(delta), _ptr(delta_shifted), int(N),
                                                             _prec_of(delta), *args),
                       "nb_power_spectrum_interlaced")
        return counts, sums

    # -- K7 ---------------------------------------------------------------------------
    def hermitian_R(self, c1: torch.Tensor, c2: torch.Tensor, guard: bool) -> torch.Tensor:
        N = c1.shape[0]
        out = torch.empty((N, N, N), dtype=torch.complex128, device=self.device)
        _lib.check(self.lib.nb_hermitian_R(self.ctx, _ptr(c1), _ptr(c2), int(N), int(bool(guard)), _ptr(out),
                                           self.stream), "nb_hermitian_R")
        return out

    def grf_synthesize(self, pk_grid: torch.Tensor, c1: torch.Tensor, c2: torch.Tensor, L: float, N: int,
                       guard: bool, use_fac: bool = True, dtype=torch.float32, want_delta=True, want_psi=False):
        for t in (pk_grid, c1, c2):
            if t.dtype != torch.float64 or t.shape != (N, N, N) or not t.is_contiguous():
                raise ValueError("pk_grid, c1, c2 must be contiguous (N,N,N) float64 tensors")
        kv = self.table(("kv", N, float(L)), lambda: tables.k_vals(N, L))
        fac = float(np.sqrt(2 * np.pi / L ** 3) * N ** 3) if use_fac else 1.0
        delta = torch.empty((N, N, N), dtype=dtype, device=self.device) if want_delta else None
        psi = torch.empty((3, N, N, N), dtype=dtype, device=self.device) if want_psi else None
        _lib.check(self.lib.nb_grf_synthesize(self.ctx, _ptr(pk_grid), _ptr(c1), _ptr(c2), int(N), _ptr(kv), fac,
                                              int(bool(guard)), F64 if dtype == torch.float64 else F32,
                                              _ptr(delta), _ptr(psi), self.stream), "nb_grf_synthesize")
        return delta, psi

    def lpt2_displacement(self, delta: torch.Tensor, L: float) -> torch.Tensor:
        """(3,N,N,N) float64 second-order displacement field of the linear delta (nb_lpt2_displacement)."""
        N = delta.shape[0]
        if delta.dtype != torch.float64 or delta.shape != (N, N, N) or not delta.is_contiguous():
            raise ValueError("delta must be a contiguous (N,N,N) float64 tensor")
        kv = self.table(("kv", N, float(L)), lambda: tables.k_vals(N, L))
        work = torch.empty((6, N, N, N), dtype=torch.float64, device=self.device)
        out = torch.empty((3, N, N, N), dtype=torch.float64, device=self.device)
        _lib.check(self.lib.nb_lpt2_displacement(self.ctx, _ptr(delta), int(N), _ptr(kv), _ptr(work), _ptr(out),
                                                 self.stream), "nb_lpt2_displacement")
        return out

    # -- K3 ---------------------------------------------------------------------------
    def cic_gather(self, field: torch.Tensor, pos: torch.Tensor, cell: float, N: int, shift: float = 0.0,
                   scale: float = 1.0, out: torch.Tensor | None = None, accumulate: bool = False) -> torch.Tensor:
        self._check_pos(pos)
        if field.dim() == 3:
            field = field.unsqueeze(0)
        ncomp = field.shape[0]
        if field.shape[1:] != (N, N, N) or not field.is_contiguous():
            raise ValueError("field must be contiguous (ncomp,N,N,N)")
        if out is None:
            out = torch.empty((pos.shape[0], ncomp), dtype=torch.float64, device=self.device)
            accumulate = False
        _lib.check(self.lib.nb_cic_gather(self.ctx, _ptr(field), int(ncomp), int(N), _prec_of(field), _ptr(pos),
                                          pos.shape[0], float(cell), float(shift), float(scale),
                                          int(bool(accumulate)), _ptr(out), self.stream), "nb_cic_gather")
        return out

    def displace_wrap(self, pos: torch.Tensor, disp: torch.Tensor, L: float) -> torch.Tensor:
        self._check_pos(pos)
        self._check_pos(disp)
        _lib.check(self.lib.nb_displace_wrap(self.ctx, _ptr(pos), _ptr(disp), pos.shape[0], float(L), self.stream),
                   "nb_displace_wrap")
        return pos

    def kick_drift(self, pos, mom, acc, kick: float, drift: float, L: float):
        for t in (pos, mom, acc):
            self._check_pos(t)
        _lib.check(self.lib.nb_kick_drift(self.ctx, _ptr(pos), _ptr(mom), _ptr(acc), pos.shape[0], float(kick),
                                          float(drift), float(L), self.stream), "nb_kick_drift")
        return pos, mom

    def gather_kick_drift(self, force3, force3_shifted, N, cell, pos, mom, acc_out, gscale, kick, drift, L):
        self._check_pos(pos)
        _lib.check(self.lib.nb_gather_kick_drift(self.ctx, _ptr(force3), _ptr(force3_shifted), int(N), float(cell),
                                                 _ptr(pos), _ptr(mom), _ptr(acc_out), pos.shape[0], float(gscale),
                                                 float(kick), float(drift), float(L), self.stream),
                   "nb_gather_kick_drift")

    # -- K2 ---------------------------------------------------------------------------
    def gather_kick_drift_add(self, force3, force3_shifted, N, cell, pos, mom, acc_add, acc_out, gscale, kick, drift,
                              L, x0=0, nx_own=None, n_ghost=0):
        """g = gscale·gather + acc_add; acc_out (nullable) <- g; kick/drift if mom is given."""
        self._check_pos(pos)
        _lib.check(self.lib.nb_gather_kick_drift_add(
            self.ctx, _ptr(force3), _ptr(force3_shifted), int(N), int(x0), int(N if nx_own is None else nx_own),
            int(n_ghost), float(cell), _ptr(pos), _ptr(mom), _ptr(acc_add), _ptr(acc_out), pos.shape[0],
            float(gscale), float(kick), float(drift), float(L), self.stream), "nb_gather_kick_drift_add")

    def p3m_set_sm_limit(self, n: int):
        _lib.check(self.lib.nb_p3m_set_sm_limit(self.ctx, int(n)), "nb_p3m_set_sm_limit")

    def p3m_mesh_done(self):
        _lib.check(self.lib.nb_p3m_mesh_done(self.ctx, self.stream), "nb_p3m_mesh_done")

    def delay_us(self, us: int):
        _lib.check(self.lib.nb_delay_us(self.ctx, int(us), self.stream), "nb_delay_us")

    def p3m_set_ctas_per_sm(self, n: int):
        _lib.check(self.lib.nb_p3m_set_ctas_per_sm(self.ctx, int(n)), "nb_p3m_set_ctas_per_sm")

    def pm_force_mesh(self, pos: torch.Tensor, L: float, N: int, r_s: float = 0.0, deconv_order: int = 2,
                      interlace_up: bool = False, interlace_down: bool = False, pref: float = 1.0,
                      out: torch.Tensor | None = None, out_shifted: torch.Tensor | None = None):
        """3 (+3) float32 force meshes: deposit -> R2C -> K2 -> C2R."""
        self._check_pos(pos)
        kv = self.table(("kv", N, float(L)), lambda: tables.k_vals(N, L))
        s = self.sinc_tab(N, L)
        if out is None:
            out = torch.empty((3, N, N, N), dtype=torch.float32, device=self.device)
        if interlace_down and out_shifted is None:
            out_shifted = torch.empty((3, N, N, N), dtype=torch.float32, device=self.device)
        _lib.check(self.lib.nb_pm_force_mesh(self.ctx, _ptr(pos), pos.shape[0], L / N, int(N), _ptr(kv), _ptr(s),
                                             float(r_s), int(deconv_order), int(bool(interlace_up)),
                                             int(bool(interlace_down)), float(pref), _ptr(out),
                                             _ptr(out_shifted) if interlace_down else C.c_void_p(0), self.stream),
                   "nb_pm_force_mesh")
        return out, (out_shifted if interlace_down else None)

    def pm_force_mesh_from_rho(self, rho: torch.Tensor, rho_shifted, L: float, N: int, r_s: float, deconv_order: int,
                               interlace_down: bool, pref: float, out: torch.Tensor, out_shifted=None):
        """pm_force_mesh from float32 count meshes the caller has deposited (cic_deposit with shift 0 / cell/2)."""
        kv = self.table(("kv", N, float(L)), lambda: tables.k_vals(N, L))
        s = self.sinc_tab(N, L)
        for g in (rho, rho_shifted):
            if g is not None and (g.dtype != torch.float32 or g.shape != (N, N, N) or not g.is_contiguous()):
                raise ValueError("rho meshes must be contiguous (N,N,N) float32")
        _lib.check(self.lib.nb_pm_force_mesh_rho(self.ctx, _ptr(rho), _ptr(rho_shifted), int(N), _ptr(kv), _ptr(s),
                                                 float(r_s), int(deconv_order), int(bool(interlace_down)), float(pref),
                                                 _ptr(out), _ptr(out_shifted) if interlace_down else C.c_void_p(0),
                                                 self.stream), "nb_pm_force_mesh_rho")
        return out, (out_shifted if interlace_down else None)

    def kspace_fused(self, rho1, rho2, L, N, r_s, deconv_order, pref, want_shifted=False):
        """K2 alone on caller-provided half spectra (complex64 [N][N][N/2+1])."""
        kv = self.table(("kv", N, float(L)), lambda: tables.k_vals(N, L))
        s = self.sinc_tab(N, L)
        g = torch.empty((3, N, N, N // 2 + 1), dtype=torch.complex64, device=self.device)
        h = torch.empty_like(g) if want_shifted else None
        _lib.check(self.lib.nb_kspace_fused(self.ctx, _ptr(rho1), _ptr(rho2), int(N), _ptr(kv), _ptr(s), float(r_s),
                                            int(deconv_order), float(pref), _ptr(g), _ptr(h), self.stream),
                   "nb_kspace_fused")
        return g, h

    # -- K4 ---------------------------------------------------------------------------
    def pp_pack(self, pos: torch.Tensor, origin=(0.0, 0.0, 0.0), mass: float = 1.0,
                masses: torch.Tensor | None = None, out: torch.Tensor | None = None) -> torch.Tensor:
        self._check_pos(pos)
        if out is None:
            out = torch.empty((pos.shape[0], 4), dtype=torch.float32, device=self.device)
        _lib.check(self.lib.nb_pp_pack(self.ctx, _ptr(pos), pos.shape[0], float(origin[0]), float(origin[1]),
                                       float(origin[2]), float(mass), _ptr(masses), _ptr(out), self.stream),
                   "nb_pp_pack")
        return out

    def pp_accel(self, ipos4: torch.Tensor, jpos4: torch.Tensor, eps: float, box: float = 0.0,
                 scale: float = 1.0, variant: int = 0, out: torch.Tensor | None = None,
                 accumulate: bool = False) -> torch.Tensor:
        for t in (ipos4, jpos4):
            if t.dtype != torch.float32 or t.dim() != 2 or t.shape[1] != 4 or not t.is_contiguous():
                raise ValueError("packed positions must be contiguous (n,4) float32")
        if out is None:
            out = torch.empty((ipos4.shape[0], 3), dtype=torch.float64, device=self.device)
            accumulate = False
        _lib.check(self.lib.nb_pp_accel(self.ctx, _ptr(ipos4), ipos4.shape[0], _ptr(jpos4), jpos4.shape[0],
                                        float(eps), float(box), float(scale), int(variant),
                                        int(bool(accumulate)), _ptr(out), self.stream), "nb_pp_accel")
        return out

    def pp_near_fix(self, pos: torch.Tensor, L: float, eps: float, r_c: float, origin, acc: torch.Tensor,
                    periodic: bool = False, scale: float = 1.0, mass: float = 1.0, max_cells: int = 256,
                    i0: int = 0, ni: int | None = None):
        """Adds the float64 - float32 difference of every pair closer than r_c to `acc` (nb_pp_near_fix; the
        all-pairs kernel keeps 24-bit coordinates).  No-op when the box holds fewer than 3 cells of side r_c."""
        self._check_pos(pos)
        ncs = min(int(L / r_c), int(max_cells))
        if ncs < 3 or pos.shape[0] == 0:
            return acc
        _s4, perm, cell_start = self.cell_sort(pos, L, ncs, mass=mass)
        ni = pos.shape[0] - int(i0) if ni is None else int(ni)
        _lib.check(self.lib.nb_pp_near_fix(self.ctx, _ptr(pos), pos.shape[0], int(i0), ni, _ptr(perm), _ptr(cell_start), ncs,
                                           float(L), float(origin[0]), float(origin[1]), float(origin[2]), float(mass),
                                           float(eps), float(r_c), int(bool(periodic)), float(scale), _ptr(acc),
                                           self.stream), "nb_pp_near_fix")
        return acc

    # -- K4e: periodic PP with Ewald summation ------------------------------------------
    def ewald_table(self, box: float, g: int = 64, alpha_box: float = 2.0, nreal: int = 4, nfour: int = 5,
                    want_f64: bool = False):
        """Device Ewald-correction table for a box (cached per (box, g, alpha)): float4 (g+1)^3 nodes
        for nb_pp_accel_ewald; with want_f64 also the (g+1,g+1,g+1,3) float64 values."""
        key = ("ewald", float(box), int(g), float(alpha_box), int(nreal), int(nfour))
        if not hasattr(self, "_ewald_tables"):
            self._ewald_tables = {}
        if key not in self._ewald_tables or (want_f64 and self._ewald_tables[key][1] is None):
            g1 = int(g) + 1
            t4 = torch.empty((g1 ** 3, 4), dtype=torch.float32, device=self.device)
            t8 = torch.empty((g1, g1, g1, 3), dtype=torch.float64, device=self.device) if want_f64 else None
            _lib.check(self.lib.nb_ewald_table(self.ctx, float(box), int(g), float(alpha_box) / float(box), int(nreal),
                                               int(nfour), _ptr(t8), _ptr(t4), self.stream), "nb_ewald_table")
            self._ewald_tables[key] = (t4, t8)
        t4, t8 = self._ewald_tables[key]
        return (t4, t8) if want_f64 else t4

    def pp_accel_ewald(self, ipos4: torch.Tensor, jpos4: torch.Tensor, eps: float, box: float, table4: torch.Tensor,
                       g: int, scale: float = 1.0, out: torch.Tensor | None = None,
                       accumulate: bool = False) -> torch.Tensor:
        for t in (ipos4, jpos4):
            if t.dtype != torch.float32 or t.dim() != 2 or t.shape[1] != 4 or not t.is_contiguous():
                raise ValueError("packed positions must be contiguous (n,4) float32")
        if table4.dtype != torch.float32 or table4.shape != ((int(g) + 1) ** 3, 4) or not table4.is_contiguous():
            raise ValueError("table must be the contiguous ((g+1)^3, 4) float32 tensor of ewald_table")
        if out is None:
            out = torch.empty((ipos4.shape[0], 3), dtype=torch.float64, device=self.device)
            accumulate = False
        _lib.check(self.lib.nb_pp_accel_ewald(self.ctx, _ptr(ipos4), ipos4.shape[0], _ptr(jpos4), jpos4.shape[0],
                                              float(eps), float(box), float(scale), _ptr(table4), int(g),
                                              int(bool(accumulate)), _ptr(out), self.stream), "nb_pp_accel_ewald")
        return out

    # -- K5 ---------------------------------------------------------------------------
    def cell_sort(self, pos: torch.Tensor, L: float, ncs: int, mass: float = 1.0):
        self._check_pos(pos)
        n = pos.shape[0]
        sorted4 = torch.empty((n, 4), dtype=torch.float32, device=self.device)
        perm = torch.empty(n, dtype=torch.int32, device=self.device)
        cell_start = torch.empty(ncs ** 3 + 1, dtype=torch.int32, device=self.device)
        _lib.check(self.lib.nb_cell_sort(self.ctx, _ptr(pos), n, float(L), int(ncs), float(mass), _ptr(sorted4),
                                         _ptr(perm), _ptr(cell_start), self.stream), "nb_cell_sort")
        return sorted4, perm, cell_start

    def cell_sort_static(self, pos: torch.Tensor, L: float, ncs: int, sorted4: torch.Tensor, perm: torch.Tensor,
                         cell_start: torch.Tensor, mass: float = 1.0):
        """nb_cell_sort_static: the cell sort into caller-provided outputs with the library's static device scratch
        (no allocation anywhere in the call; at most 2^21 particles and 64^3 cells)."""
        self._check_pos(pos)
        n = pos.shape[0]
        if (sorted4.shape != (n, 4) or sorted4.dtype != torch.float32 or perm.shape != (n,) or perm.dtype != torch.int32
                or cell_start.shape != (ncs ** 3 + 1,) or cell_start.dtype != torch.int32):
            raise ValueError("cell_sort_static: sorted4 (n,4) float32, perm (n,) int32, cell_start (ncs^3+1,) int32")
        _lib.check(self.lib.nb_cell_sort_static(self.ctx, _ptr(pos), n, float(L), int(ncs), float(mass), _ptr(sorted4),
                                                _ptr(perm), _ptr(cell_start), self.stream), "nb_cell_sort_static")
        return sorted4, perm, cell_start

    def p3m_short(self, sorted4, perm, cell_start, L: float, ncs: int, r_s: float, r_cut: float, eps: float,
                  scale: float, out: torch.Tensor, accumulate: bool = False, cell_lo: int = 0,
                  cell_hi: int | None = None, i_limit: int = 0, pos: torch.Tensor | None = None,
                  r_near: float = 0.0, near_count: torch.Tensor | None = None) -> torch.Tensor:
        """pos (the (n,3) float64 array the sort was made from) and r_near > 0: nb_p3m_short_exact — pairs closer than
        r_near are evaluated in float64 from pos; near_count (int32 device scalar) receives their number."""
        u_max = r_cut / (2 * r_s) * 1.0001
        key = ("split", round(u_max, 9))
        if key not in self._tab_cache:
            self._tab_cache[key] = np.ascontiguousarray(tables.split_poly(u_max, 10), dtype=np.float64)
        coef = self._tab_cache[key]
        if cell_hi is None:
            cell_hi = ncs ** 3
        if int(cell_lo) == int(cell_hi) and not accumulate:
            out.zero_()          # nb_p3m_short returns before touching `out` when the cell range is empty
        if pos is not None and r_near > 0.0:
            self._check_pos(pos)
            _lib.check(self.lib.nb_p3m_short_exact(self.ctx, _ptr(sorted4), _ptr(cell_start), _ptr(perm), int(ncs),
                                                   int(cell_lo), int(cell_hi), L / ncs, float(r_cut), float(r_s),
                                                   float(eps), coef.ctypes.data_as(C.c_void_p), int(coef.size),
                                                   float(u_max), float(scale), int(bool(accumulate)), int(i_limit),
                                                   _ptr(out), _ptr(pos), pos.shape[0], float(L), float(r_near),
                                                   _ptr(near_count) if near_count is not None else C.c_void_p(0),
                                                   self.stream), "nb_p3m_short_exact")
            return out
        _lib.check(self.lib.nb_p3m_short(self.ctx, _ptr(sorted4), _ptr(cell_start), _ptr(perm), int(ncs),
                                         int(cell_lo), int(cell_hi), L / ncs, float(r_cut), float(r_s), float(eps),
                                         coef.ctypes.data_as(C.c_void_p), int(coef.size), float(u_max),
                                         float(scale), int(bool(accumulate)), int(i_limit), _ptr(out), self.stream),
                   "nb_p3m_short")
        return out

    def p3m_count_pairs(self, sorted4, cell_start, L: float, ncs: int, r_cut: float, cell_lo=0, cell_hi=None,
                        perm=None, i_limit: int = 0):
        out = torch.empty(2, dtype=torch.int64, device=self.device)
        _lib.check(self.lib.nb_p3m_count_pairs(self.ctx, _ptr(sorted4), _ptr(cell_start), _ptr(perm),
                                               int(i_limit if perm is not None else 0), int(ncs), int(cell_lo),
                                               int(ncs ** 3 if cell_hi is None else cell_hi), L / ncs,
                                               float(r_cut), _ptr(out), self.stream), "nb_p3m_count_pairs")
        c = out.cpu().numpy()
        return int(c[0]), int(c[1])

    # -- slab decomposition (dist.py) ----------------------------------------------------------
    def cic_deposit_slab(self, pos, cell, N, x0, nx_own, n_ghost, out, shift=0.0, accumulate=False,
                         n_outside: torch.Tensor | None = None):
        self._check_pos(pos)
        if out.dtype != torch.float32 or out.numel() != (nx_own + n_ghost) * N * N or not out.is_contiguous():
            raise ValueError("slab grid must be contiguous float32 [nx_own+n_ghost][N][N]")
        _lib.check(self.lib.nb_cic_deposit_slab(self.ctx, _ptr(pos), pos.shape[0], float(cell), int(N), int(x0),
                                                int(nx_own), int(n_ghost), _ptr(out), float(shift),
                                                int(bool(accumulate)), _ptr(n_outside), self.stream),
                   "nb_cic_deposit_slab")
        return out

    def gather_kick_drift_slab(self, force3, force3_shifted, N, x0, nx_own, n_ghost, cell, pos, mom, acc_out,
                               gscale, kick, drift, L):
        self._check_pos(pos)
        _lib.check(self.lib.nb_gather_kick_drift_slab(self.ctx, _ptr(force3), _ptr(force3_shifted), int(N), int(x0),
                                                      int(nx_own), int(n_ghost), float(cell), _ptr(pos), _ptr(mom),
                                                      _ptr(acc_out), pos.shape[0], float(gscale), float(kick),
                                                      float(drift), float(L), self.stream),
                   "nb_gather_kick_drift_slab")

    def kspace_fused_slab(self, rho1, rho2, L, N, y0, nyl, r_s, deconv_order, pref, g3, h3=None):
        kv = self.table(("kv", N, float(L)), lambda: tables.k_vals(N, L))
        s = self.sinc_tab(N, L)
        _lib.check(self.lib.nb_kspace_fused_slab(self.ctx, _ptr(rho1), _ptr(rho2), int(N), int(y0), int(nyl), _ptr(kv),
                                                 _ptr(s), float(r_s), int(deconv_order), float(pref), _ptr(g3),
                                                 _ptr(h3), self.stream), "nb_kspace_fused_slab")
        return g3, h3

    def pk_tables(self, N, L, n_edges, kq_form="inv_d"):
        key = ("edges", N, float(L), n_edges)
        if key not in self._tab_cache:
            e = tables.pk_edges(N, L, n_edges)
            thr, thr_hi = tables.sqrt_thresholds(e)
            self._tab_cache[key] = (e, torch.from_numpy(thr).to(self.device), float(thr[0]), float(thr[-1]), thr_hi)
        k2 = self.table(("k2", N, float(L)), lambda: tables.k_vals(N, L) ** 2)
        return self._tab_cache[key] + (k2, self.sinc_tab(N, L, kq_form))

    def pk_bin_slab(self, spec, L, N, y0, nyl, counts, sums, n_edges=300, window_power=0, clip_mode=0,
                    clip_val=0.0, norm=1.0, closed_last=True, kq_form="inv_d"):
        """Adds the shell sums of the y-slab spectrum [N][nyl][N/2+1] (complex64) into counts/sums."""
        edges, thr_d, t0, t1, thr_hi, k2, s = self.pk_tables(N, L, n_edges, kq_form)
        _lib.check(self.lib.nb_pk_bin_slab(self.ctx, _ptr(spec), int(N), int(y0), int(nyl), _ptr(k2), _ptr(s),
                                           _ptr(thr_d), t0, t1, thr_hi, int(n_edges), int(window_power),
                                           int(clip_mode), float(clip_val), float(norm), int(bool(closed_last)),
                                           _ptr(counts), _ptr(sums), self.stream), "nb_pk_bin_slab")
        return edges

    def slab_fft2d(self, N, nplanes, direction, real, spec):
        _lib.check(self.lib.nb_slab_fft2d(self.ctx, int(N), int(nplanes), int(direction), _ptr(real), _ptr(spec),
                                          self.stream), "nb_slab_fft2d")

    def slab_fft1d_x(self, N, rows, direction, spec):
        _lib.check(self.lib.nb_slab_fft1d_x(self.ctx, int(N), int(rows), int(direction), _ptr(spec), self.stream),
                   "nb_slab_fft1d_x")

    def slab_pack(self, src, dst, nxl, G, nyl, nzh):
        _lib.check(self.lib.nb_slab_pack(self.ctx, _ptr(src), _ptr(dst), int(nxl), int(G), int(nyl), int(nzh),
                                         self.stream), "nb_slab_pack")

    def slab_unpack(self, src, dst, nxl, G, nyl, nzh):
        _lib.check(self.lib.nb_slab_unpack(self.ctx, _ptr(src), _ptr(dst), int(nxl), int(G), int(nyl), int(nzh),
                                           self.stream), "nb_slab_unpack")

    def slab_owner(self, pos, cell, N, planes_per_rank, G, want_counts=True, out=None):
        self._check_pos(pos)
        owner = out[:pos.shape[0]] if out is not None else torch.empty(pos.shape[0], dtype=torch.int32,
                                                                      device=self.device)
        counts = torch.empty(G, dtype=torch.int64, device=self.device) if want_counts else None
        _lib.check(self.lib.nb_slab_owner(self.ctx, _ptr(pos), pos.shape[0], float(cell), int(N),
                                          int(planes_per_rank), _ptr(owner), _ptr(counts), int(G), self.stream),
                   "nb_slab_owner")
        return owner, counts

    def halo_select(self, pos, lo_edge: float, hi_edge: float, out_lo, out_hi, counts2):
        """Rows with x < lo_edge -> out_lo, x >= hi_edge -> out_hi ((cap,3) float64, unused rows NaN);
        counts2: device int32[2] (true counts, may exceed cap)."""
        self._check_pos(pos)
        cap = out_lo.shape[0]
        _lib.check(self.lib.nb_halo_select(self.ctx, _ptr(pos), pos.shape[0], float(lo_edge), float(hi_edge),
                                           _ptr(out_lo), _ptr(out_hi), int(cap), _ptr(counts2), self.stream),
                   "nb_halo_select")

    def migrate_split(self, pos, mom, owner, rank: int, prev: int, nxt: int, send_prev, send_next, holes, counts4):
        self._check_pos(pos)
        _lib.check(self.lib.nb_migrate_split(self.ctx, _ptr(pos), _ptr(mom), pos.shape[0], _ptr(owner), int(rank),
                                             int(prev), int(nxt), _ptr(send_prev), _ptr(send_next),
                                             int(send_prev.shape[0]), _ptr(holes), _ptr(counts4), self.stream),
                   "nb_migrate_split")

    def migrate_merge(self, pos_buf, mom_buf, n: int, holes, n_holes: int, recv_a, na: int, recv_b, nb: int) -> int:
        """Arrivals into the holes / behind row n of the capacity buffers; returns the new row count."""
        n_rem = max(0, int(n_holes) - int(na) - int(nb))
        work = torch.empty(3 * n_rem + 2, dtype=torch.int32, device=self.device) if n_rem else None
        out = C.c_int64(0)
        _lib.check(self.lib.nb_migrate_merge(self.ctx, _ptr(pos_buf), _ptr(mom_buf), int(n), _ptr(holes), int(n_holes),
                                             _ptr(recv_a), int(na), _ptr(recv_b), int(nb), _ptr(work), C.byref(out),
                                             self.stream), "nb_migrate_merge")
        return int(out.value)

    def grid_add(self, dst, src):
        if dst.dtype != torch.float32 or src.dtype != torch.float32 or dst.numel() != src.numel():
            raise ValueError("grid_add needs float32 tensors of equal size")
        _lib.check(self.lib.nb_grid_add(self.ctx, _ptr(dst), _ptr(src), dst.numel(), self.stream), "nb_grid_add")
        return dst


_default: dict[int, Engine] = {}


def default_engine(device: int | None = None) -> Engine:
    if not torch.cuda.is_available():
        _lib.load()   # raises if the library is missing; otherwise report the missing GPU
        raise RuntimeError("nbody_b200 needs a CUDA device (there is no CPU fallback)")
    d = torch.cuda.current_device() if device is None else int(device)
    if d not in _default:
        _default[d] = Engine(d)
    return _default[d]
