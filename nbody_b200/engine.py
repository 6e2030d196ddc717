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


_default: dict[int, Engine] = {}


def default_engine(device: int | None = None) -> Engine:
    if not torch.cuda.is_available():
        _lib.load()   # raises if the library is missing; otherwise report the missing GPU
        raise RuntimeError("nbody_b200 needs a CUDA device (there is no CPU fallback)")
    d = torch.cuda.current_device() if device is None else int(device)
    if d not in _default:
        _default[d] = Engine(d)
    return _default[d]
