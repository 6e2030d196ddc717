"""TEST INFRASTRUCTURE ONLY — loads the reference's own functions without copying them.

The reference scripts cannot be imported (they import classy / h5py / matplotlib /
tensorflow at the top and run work at import time), so this module `ast.parse`s a
reference file where it lies under /root/reference, keeps only the wanted
`FunctionDef` nodes, and `exec`s them in a namespace holding numpy (and numba's
decorators, or pass-through stand-ins).  No reference source text enters this repo.

Only `oracle/gen_golden.py` and the `-m "not gpu"` tests use it, and only in the
build container: /root/reference does not exist on the GPU box (`available()` is
False there and callers skip).  Nothing in the product package imports `oracle/`.
"""
from __future__ import annotations

import ast
import os

import numpy as np

REF_ROOT = os.environ.get("NBODY_REFERENCE_ROOT", "/root/reference")

FILES = {
    "frap": "TheFeedingPipeline/FraP_kTilDelta.py",
    "power_compare": "SnapVsPK/Power_compare.py",
    "plotter": "Different_Sizes/Plotter.py",
    "lpt": "1LPTvs2LPT/1vs2tlptplotter.py",
    "grf": "TheFeedingPipeline/Cosmo data create (new).py",
    "cic3d": "CiC/3D startbetingelser1.py",
}


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, FILES["frap"]))


def _passthrough(*args, **kwargs):
    """Stand-in for numba's @jit / @njit (with or without arguments)."""
    if len(args) == 1 and callable(args[0]) and not kwargs:
        return args[0]
    return lambda f: f


def load(which: str, names, use_numba: bool = True, extra_ns: dict | None = None) -> dict:
    """Return {name: function} for FunctionDefs `names` of reference file `which`.

    use_numba=False replaces @jit/@njit by pass-through decorators (still the
    reference's own code, interpreted by CPython; SURVEY §8c workaround 1).
    """
    path = os.path.join(REF_ROOT, FILES[which])
    with open(path, "r", encoding="utf-8") as fh:
        tree = ast.parse(fh.read(), filename=path)
    wanted = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in set(names)]
    missing = set(names) - {n.name for n in wanted}
    if missing:
        raise KeyError(f"{path}: no top-level function(s) {sorted(missing)}")
    # several reference files define a function twice (e.g. Plotter.py has two
    # compute_power_spectrum variants); the later definition wins, as in Python.
    mod = ast.Module(body=wanted, type_ignores=[])
    ns: dict = {"np": np}
    if use_numba:
        from numba import jit, njit
        ns.update(jit=jit, njit=njit)
    else:
        ns.update(jit=_passthrough, njit=_passthrough)
    try:
        from scipy.interpolate import RegularGridInterpolator
        import scipy.interpolate as scint
        ns.update(RegularGridInterpolator=RegularGridInterpolator, scint=scint)
    except Exception:  # pragma: no cover
        pass
    if extra_ns:
        ns.update(extra_ns)
    exec(compile(mod, path, "exec"), ns)
    return {n: ns[n] for n in names} | {"__ns__": ns}


def default_args_patch(fn_ns: dict, **globals_):
    """Inject module-level globals a reference function reads (e.g. k_mag for find_delta)."""
    fn_ns["__ns__"].update(globals_)
