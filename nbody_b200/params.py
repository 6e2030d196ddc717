"""Loader for the reference's CONCEPT `.param` files:
Different_Sizes/Giant.param, TheFeedingPipeline/NN.param, SnapVsPK/Uncertain.param.

A `.param` file is Python source that CONCEPT evaluates in a namespace with units and
helpers.  This loader does the same with a small namespace of its own (unit system
Mpc / Gyr / 10^10 M_sun like the snapshots), executing the top-level statements in a
fixpoint loop so that forward references resolve (NN.param:4 uses `primordial_spectrum`
defined at :92; `h` derives from `H0`).  Statements that never resolve (e.g. `_tag` in
Uncertain.param, which the reference's make_paramfile prepends at run time,
NeuNetværk_test.py:79-120) are reported in `unresolved`, not fatal.
"""
from __future__ import annotations

import ast
import math
from dataclasses import dataclass, field

# --- unit system (CONCEPT's): length Mpc, time Gyr, mass 10^10 M_sun ---------------------
_M = 1.0 / 3.0856775814913673e22          # metre in Mpc
_S = 1.0 / (1e9 * 365.25 * 86400.0)       # second in Gyr
UNITS = {
    "Mpc": 1.0, "kpc": 1e-3, "Gpc": 1e3, "pc": 1e-6, "m": _M, "km": 1e3 * _M, "cm": 1e-2 * _M,
    "Gyr": 1.0, "Myr": 1e-3, "kyr": 1e-6, "yr": 1e-9, "s": _S, "hr": 3600.0 * _S, "minutes": 60.0 * _S,
    "day": 86400.0 * _S, "m_sun": 1e-10, "km_s": 1e3 * _M / _S,
}
#: G·(10^10 M_sun) from GM_sun = 1.32712440018e20 m³/s², in Mpc³ Gyr⁻² (10^10 M_sun)⁻¹
G_NEWTON = 1.32712440018e20 * 1e10 * _M ** 3 / _S ** 2


class _Path:
    """Stand-in for CONCEPT's `path` object (only used inside f-strings)."""

    def __getattr__(self, name):
        return f"<{name}>"


def _namespace(overrides):
    ns = dict(UNITS)
    ns.update(pi=math.pi, π=math.pi, sqrt=math.sqrt, cbrt=lambda x: round(x ** (1.0 / 3.0), 12),
              exp=math.exp, log=math.log, inf=math.inf, path=_Path(), param=_Path())
    ns.update(overrides or {})
    return ns


def load_param(source, overrides=None, is_text=False):
    """Evaluate a `.param` file.  Returns (values, unresolved) where values maps every
    assigned top-level name to its value and unresolved lists the source of statements that
    could not be evaluated."""
    text = source if is_text else open(source, "r", encoding="utf-8").read()
    tree = ast.parse(text)
    ns = _namespace(overrides)
    base_keys = set(ns)
    pending = list(tree.body)
    user_h = "h" in (overrides or {})
    for _ in range(len(pending) + 2):
        if not pending:
            break
        still = []
        for node in pending:
            code = compile(ast.Module(body=[node], type_ignores=[]), "<param>", "exec")
            try:
                exec(code, ns)
            except NameError:
                still.append(node)
                continue
            if "H0" in ns and not user_h:
                ns["h"] = ns["H0"] / (100.0 * UNITS["km"] / (UNITS["s"] * UNITS["Mpc"]))
        if len(still) == len(pending):
            break
        pending = still
    values = {k: v for k, v in ns.items() if k not in base_keys and not k.startswith("__")}
    if "h" in ns:
        values["h"] = ns["h"]
    unresolved = [ast.get_source_segment(text, n) or ast.dump(n) for n in pending]
    return values, unresolved


@dataclass
class ForceConfig:
    """What a force step needs from a `.param` file (SURVEY §8a')."""
    method: str = "pm"                 # 'pp' | 'pm' | 'p3m'   select_forces['matter']['gravity']
    npart: int = 0                     # initial_conditions['N']
    boxsize: float = 1000.0            # boxsize (Mpc)
    gridsize: int = 0                  # potential mesh (cbrt(N) in all three reference files)
    interlace: tuple = (False, False)  # potential_options['interlace']['gravity'][method]
    deconvolve: tuple = (True, True)   # potential_options['deconvolve']['gravity'][method]
    differentiation: str = "Fourier"
    r_s_cells: float = 1.25            # P3M split scale in mesh cells (CONCEPT default)
    r_cut_over_rs: float = 4.5         # P3M range in units of r_s (CONCEPT default)
    softening_frac: float = 0.025      # Plummer ε = frac · L / cbrt(Np)
    G: float = 1.0
    mass: float = 1.0
    a: float = 1.0
    H0: float | None = None
    Omega_b: float | None = None
    Omega_cdm: float | None = None
    a_begin: float | None = None
    extra: dict = field(default_factory=dict)

    @property
    def cell(self):
        return self.boxsize / self.gridsize

    @property
    def r_s(self):
        return self.r_s_cells * self.cell if self.method == "p3m" else 0.0

    @property
    def r_cut(self):
        return self.r_cut_over_rs * self.r_s

    @property
    def eps(self):
        return self.softening_frac * self.boxsize / round(self.npart ** (1.0 / 3.0))

    @property
    def deconv_order(self):
        return int(bool(self.deconvolve[0])) + int(bool(self.deconvolve[1]))


def _dig(d, *keys, default=None):
    for k in keys:
        if isinstance(d, dict) and k in d:
            d = d[k]
        else:
            return default
    return d


def _eval_gridsize(expr, npart, global_size):
    if isinstance(expr, (int, float)):
        return int(expr)
    if isinstance(expr, (set, frozenset, list, tuple)) and len(expr):
        return _eval_gridsize(next(iter(expr)), npart, global_size)
    if isinstance(expr, str):
        return int(round(eval(expr, {"__builtins__": {}},
                              {"N": npart, "Ñ": npart, "cbrt": lambda x: x ** (1.0 / 3.0),
                               "gridsize": global_size})))
    raise ValueError(f"cannot interpret gridsize {expr!r}")


def force_config(source, overrides=None, mesh=None, method=None, physical_units=False, **kw) -> ForceConfig:
    """ForceConfig from a `.param` file path / text / already-loaded dict.  `mesh` overrides
    the potential grid size (BASELINE configs 4/5 use a mesh 2x finer than cbrt(N))."""
    if isinstance(source, dict):
        vals = source
    else:
        is_text = isinstance(source, str) and "\n" in source
        vals, _ = load_param(source, overrides, is_text=is_text)
    cfg = ForceConfig()
    cfg.method = method or _dig(vals, "select_forces", "matter", "gravity", default="pm")
    if isinstance(cfg.method, dict):
        cfg.method = next(iter(cfg.method))
    cfg.npart = int(_dig(vals, "initial_conditions", "N", default=0))
    cfg.boxsize = float(vals.get("boxsize", cfg.boxsize))
    po = vals.get("potential_options", {})
    gs = po.get("gridsize", round(cfg.npart ** (1 / 3)) if cfg.npart else 0)
    global_size = None
    if isinstance(gs, dict):
        global_size = _eval_gridsize(gs.get("global", 0), cfg.npart, None) if "global" in gs else None
        g2 = _dig(gs, "particles", "gravity", cfg.method)
        gs = g2 if g2 is not None else (global_size or 0)
    cfg.gridsize = int(mesh) if mesh else _eval_gridsize(gs, cfg.npart, global_size)
    cfg.interlace = tuple(_dig(po, "interlace", "gravity", cfg.method, default=cfg.interlace))
    cfg.deconvolve = tuple(_dig(po, "deconvolve", "gravity", cfg.method, default=cfg.deconvolve))
    cfg.differentiation = _dig(po, "differentiation", "all", "gravity", cfg.method, default="Fourier")
    for key, attr in (("H0", "H0"), ("Ωb", "Omega_b"), ("Ωcdm", "Omega_cdm"), ("a_begin", "a_begin")):
        if key in vals:
            setattr(cfg, attr, float(vals[key]))
    if physical_units:
        cfg.G = G_NEWTON
        if cfg.H0 and cfg.Omega_b is not None and cfg.Omega_cdm is not None and cfg.npart:
            rho_crit = 3.0 * cfg.H0 ** 2 / (8.0 * math.pi * G_NEWTON)
            cfg.mass = rho_crit * (cfg.Omega_b + cfg.Omega_cdm) * cfg.boxsize ** 3 / cfg.npart
    for k, v in kw.items():
        if not hasattr(cfg, k):
            raise TypeError(f"unknown ForceConfig field {k!r}")
        setattr(cfg, k, v)
    return cfg
