"""ctypes binding of libnbody_b200.so (C-ABI: include/nbody_b200.h).

There is no CPU fallback: if the shared library is missing or a call fails, an exception
is raised (north_star: "no CPU fallback").  The library is built in-tree by
`__graft_entry__.build()` / `make -C nbody_b200/csrc`.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnbody_b200.so")

c_i32, c_i64, c_f64, c_f32, c_vp = C.c_int, C.c_int64, C.c_double, C.c_float, C.c_void_p

# name -> (restype, argtypes); kept in the order of include/nbody_b200.h.  tests/test_abi.py
# checks that every prototype declared in the header is listed here and exported.
PROTOTYPES = {
    "nb_abi_version": (c_i32, []),
    "nb_init": (c_i32, [c_i32, C.POINTER(c_vp)]),
    "nb_destroy": (None, [c_vp]),
    "nb_last_error": (C.c_char_p, []),
    "nb_release_scratch": (c_i32, [c_vp]),
    "nb_launch_count": (c_i64, [c_vp]),
    "nb_profile_enable": (c_i32, [c_vp, c_i32]),
    "nb_profile_read": (c_i32, [c_vp, c_vp, c_vp, c_i32]),
    "nb_fp32_peak": (c_i32, [c_vp, c_i32, c_i32, c_i32, C.POINTER(c_f64), c_vp]),
    "nb_cic_deposit": (c_i32, [c_vp, c_vp, c_i64, c_f64, c_i32, c_vp, c_i32, c_f64, c_i32, c_vp, c_vp]),
    "nb_grid_sum": (c_i32, [c_vp, c_vp, c_i64, c_i32, c_vp, c_vp]),
    "nb_grid_affine": (c_i32, [c_vp, c_vp, c_i64, c_i32, c_f64, c_vp, c_f64, c_f64, c_f64, c_vp]),
    "nb_vec3_max_norm2": (c_i32, [c_vp, c_vp, c_i64, c_vp, c_vp]),
    "nb_deconvolve_cic": (c_i32, [c_vp, c_vp, c_i32, c_i32, c_vp, c_i32, c_vp]),
    "nb_power_spectrum": (c_i32, [c_vp, c_vp, c_i32, c_i32, c_vp, c_vp, c_vp, c_f64, c_f64, c_f64, c_i32,
                                  c_i32, c_i32, c_f64, c_f64, c_i32, c_vp, c_vp, c_vp]),
    "nb_power_spectrum_interlaced": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_i32, c_vp, c_vp, c_vp, c_f64, c_f64, c_f64,
                                             c_i32, c_i32, c_i32, c_f64, c_f64, c_i32, c_vp, c_vp, c_vp]),
    "nb_hermitian_R": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_i32, c_vp, c_vp]),
    "nb_grf_synthesize": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i32, c_vp, c_f64, c_i32, c_i32, c_vp, c_vp, c_vp]),
    "nb_lpt2_displacement": (c_i32, [c_vp, c_vp, c_i32, c_vp, c_vp, c_vp, c_vp]),
    "nb_cic_gather": (c_i32, [c_vp, c_vp, c_i32, c_i32, c_i32, c_vp, c_i64, c_f64, c_f64, c_f64, c_i32, c_vp, c_vp]),
    "nb_displace_wrap": (c_i32, [c_vp, c_vp, c_vp, c_i64, c_f64, c_vp]),
    "nb_kick_drift": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i64, c_f64, c_f64, c_f64, c_vp]),
    "nb_gather_kick_drift": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_f64, c_vp, c_vp, c_vp, c_i64, c_f64, c_f64,
                                     c_f64, c_f64, c_vp]),
    "nb_gather_kick_drift_add": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_i32, c_i32, c_i32, c_f64, c_vp, c_vp, c_vp, c_vp,
                                         c_i64, c_f64, c_f64, c_f64, c_f64, c_vp]),
    "nb_kspace_fused": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_f64, c_i32, c_f64, c_vp, c_vp, c_vp]),
    "nb_pm_force_mesh": (c_i32, [c_vp, c_vp, c_i64, c_f64, c_i32, c_vp, c_vp, c_f64, c_i32, c_i32, c_i32, c_f64,
                                 c_vp, c_vp, c_vp]),
    "nb_pm_force_mesh_rho": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_f64, c_i32, c_i32, c_f64, c_vp, c_vp,
                                     c_vp]),
    "nb_pp_pack": (c_i32, [c_vp, c_vp, c_i64, c_f64, c_f64, c_f64, c_f64, c_vp, c_vp, c_vp]),
    "nb_pp_accel": (c_i32, [c_vp, c_vp, c_i64, c_vp, c_i64, c_f64, c_f64, c_f64, c_i32, c_i32, c_vp, c_vp]),
    "nb_pp_near_fix": (c_i32, [c_vp, c_vp, c_i64, c_i64, c_i64, c_vp, c_vp, c_i32, c_f64, c_f64, c_f64, c_f64, c_f64, c_f64, c_f64,
                               c_i32, c_f64, c_vp, c_vp]),
    "nb_cell_sort": (c_i32, [c_vp, c_vp, c_i64, c_f64, c_i32, c_f64, c_vp, c_vp, c_vp, c_vp]),
    "nb_cell_sort_static": (c_i32, [c_vp, c_vp, c_i64, c_f64, c_i32, c_f64, c_vp, c_vp, c_vp, c_vp]),
    "nb_p3m_short": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i32, c_i64, c_i64, c_f64, c_f64, c_f64, c_f64, c_vp, c_i32,
                             c_f64, c_f64, c_i32, c_i64, c_vp, c_vp]),
    "nb_p3m_short_exact": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i32, c_i64, c_i64, c_f64, c_f64, c_f64, c_f64, c_vp, c_i32,
                                   c_f64, c_f64, c_i32, c_i64, c_vp, c_vp, c_i64, c_f64, c_f64, c_vp, c_vp]),
    "nb_p3m_set_ctas_per_sm": (c_i32, [c_vp, c_i32]),
    "nb_p3m_set_sm_limit": (c_i32, [c_vp, c_i32]),
    "nb_delay_us": (c_i32, [c_vp, c_i32, c_vp]),
    "nb_p3m_mesh_done": (c_i32, [c_vp, c_vp]),
    "nb_p3m_count_pairs": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i64, c_i32, c_i64, c_i64, c_f64, c_f64, c_vp, c_vp]),
    "nb_cic_deposit_slab": (c_i32, [c_vp, c_vp, c_i64, c_f64, c_i32, c_i32, c_i32, c_i32, c_vp, c_f64, c_i32, c_vp,
                                    c_vp]),
    "nb_gather_kick_drift_slab": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_i32, c_i32, c_i32, c_f64, c_vp, c_vp, c_vp,
                                          c_i64, c_f64, c_f64, c_f64, c_f64, c_vp]),
    "nb_kspace_fused_slab": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_i32, c_i32, c_vp, c_vp, c_f64, c_i32, c_f64, c_vp,
                                     c_vp, c_vp]),
    "nb_pk_bin_slab": (c_i32, [c_vp, c_vp, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp, c_f64, c_f64, c_f64, c_i32, c_i32,
                               c_i32, c_f64, c_f64, c_i32, c_vp, c_vp, c_vp]),
    "nb_slab_fft2d": (c_i32, [c_vp, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp]),
    "nb_slab_fft1d_x": (c_i32, [c_vp, c_i32, c_i64, c_i32, c_vp, c_vp]),
    "nb_slab_pack": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_i32, c_i32, c_i32, c_vp]),
    "nb_slab_unpack": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_i32, c_i32, c_i32, c_vp]),
    "nb_slab_owner": (c_i32, [c_vp, c_vp, c_i64, c_f64, c_i32, c_i32, c_vp, c_vp, c_i32, c_vp]),
    "nb_halo_select": (c_i32, [c_vp, c_vp, c_i64, c_f64, c_f64, c_vp, c_vp, c_i32, c_vp, c_vp]),
    "nb_migrate_split": (c_i32, [c_vp, c_vp, c_vp, c_i64, c_vp, c_i32, c_i32, c_i32, c_vp, c_vp, c_i32, c_vp, c_vp, c_vp]),
    "nb_migrate_merge": (c_i32, [c_vp, c_vp, c_vp, c_i64, c_vp, c_i32, c_vp, c_i32, c_vp, c_i32, c_vp,
                                 C.POINTER(c_i64), c_vp]),
    "nb_grid_add": (c_i32, [c_vp, c_vp, c_vp, c_i64, c_vp]),
    "nb_snapshot_delta_batch": (c_i32, [c_vp, c_vp, c_vp, c_i64, c_i32, c_vp, c_i32, c_vp, c_i32, c_vp, c_vp, c_i32,
                                        c_vp]),
    "nb_snapshot_pk_batch": (c_i32, [c_vp, c_vp, c_vp, c_i64, c_i32, c_vp, c_vp, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp,
                                     c_i32, c_i32, c_i32, c_f64, c_vp, c_i32, c_vp, c_vp, c_vp]),
    "nb_ewald_table": (c_i32, [c_vp, c_f64, c_i32, c_f64, c_i32, c_i32, c_vp, c_vp, c_vp]),
    "nb_pp_accel_ewald": (c_i32, [c_vp, c_vp, c_i64, c_vp, c_i64, c_f64, c_f64, c_f64, c_vp, c_i32, c_i32, c_vp, c_vp]),
}


class NbodyError(RuntimeError):
    pass


_lib = None


def load():
    """Load the shared library once; raise loudly if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C nbody_b200/csrc` (there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    got = lib.nb_abi_version()
    if got != ABI_VERSION:
        raise ImportError(f"libnbody_b200.so ABI {got} != binding ABI {ABI_VERSION}")
    _lib = lib
    return lib


ABI_VERSION = 6


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = load().nb_last_error()
        raise NbodyError(f"{what or 'nbody_b200 call'} failed (code {rc}): {msg.decode() if msg else ''}")
