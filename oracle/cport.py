"""TEST INFRASTRUCTURE ONLY — ctypes loader of oracle/liboracle_c.so (oracle_c.c), the plain-C
float64 restatement used where numpy loops are too slow (bench cpu_baseline, larger parity
cases).  Validated against oracle/restate.py and oracle/force.py in tests/test_oracle_c.py."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "liboracle_c.so")
_lib = None
_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")


def lib():
    global _lib
    if _lib is None:
        if not os.path.isfile(_PATH):
            subprocess.check_call(["make", "-C", _HERE])
        L = C.CDLL(_PATH)
        L.oc_num_threads.restype = C.c_int
        L.oc_set_num_threads.argtypes = [C.c_int]
        L.oc_cic_deposit.argtypes = [_dp, C.c_int64, C.c_double, C.c_int64, C.c_double, _dp]
        L.oc_cic_gather.argtypes = [_dp, C.c_int64, _dp, C.c_int64, C.c_double, C.c_double, _dp]
        L.oc_pp_accel.argtypes = [_dp, C.c_int64, C.c_int64, C.c_int64, C.c_double, C.c_double, C.c_double,
                                  C.c_double, _dp]
        L.oc_p3m_short.argtypes = [_dp, C.c_int64, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double,
                                   C.c_double, _dp, C.POINTER(C.c_int64)]
        L.oc_p3m_short.restype = C.c_int
        L.oc_shell_bin.argtypes = [_dp, C.c_int64, _dp, _dp, C.c_int64, C.c_int, _ip, _dp]
        _lib = L
    return _lib


def num_threads():
    return int(lib().oc_num_threads())


def set_num_threads(n):
    """OpenMP threads of the C loops (overrides an inherited OMP_NUM_THREADS)."""
    lib().oc_set_num_threads(int(n))
    return num_threads()


def cic_density(pos, boxsize, N, shift=0.0):
    pos = np.ascontiguousarray(pos, dtype=np.float64)
    grid = np.zeros((N, N, N))
    lib().oc_cic_deposit(pos, pos.shape[0], boxsize / N, N, float(shift), grid)
    return grid


def cic_gather(field, pos, boxsize, N, shift=0.0):
    pos = np.ascontiguousarray(pos, dtype=np.float64)
    out = np.empty(pos.shape[0])
    lib().oc_cic_gather(np.ascontiguousarray(field, dtype=np.float64), N, pos, pos.shape[0], boxsize / N,
                        float(shift), out)
    return out


def pp_accel(pos, G=1.0, mass=1.0, eps=0.0, box=0.0, i0=0, i1=None):
    pos = np.ascontiguousarray(pos, dtype=np.float64)
    i1 = pos.shape[0] if i1 is None else i1
    acc = np.empty((i1 - i0, 3))
    lib().oc_pp_accel(pos, pos.shape[0], i0, i1, G, mass, eps, box, acc)
    return acc


def p3m_short_accel(pos, L, r_s, r_cut, G=1.0, mass=1.0, eps=0.0, return_pairs=False):
    pos = np.ascontiguousarray(pos, dtype=np.float64)
    acc = np.empty_like(pos)
    n = C.c_int64(0)
    rc = lib().oc_p3m_short(pos, pos.shape[0], L, r_s, r_cut, G, mass, eps, acc, C.byref(n))
    if rc:
        raise ValueError("oc_p3m_short: boxsize must be >= 3 r_cut")
    return (acc, n.value) if return_pairs else acc


def shell_bin(P, N, kvals, edges, closed_last=True):
    counts = np.zeros(len(edges) - 1, dtype=np.int64)
    sums = np.zeros(len(edges) - 1)
    lib().oc_shell_bin(np.ascontiguousarray(P, dtype=np.float64), N, np.ascontiguousarray(kvals),
                       np.ascontiguousarray(edges), len(edges), int(closed_last), counts, sums)
    return counts, sums
