"""nb_cell_sort_static (the allocation-free cell sort: scratch in a static device array of the library) against
nb_cell_sort, both called through the C-ABI, on two calls in a row.  Index work: bit-exact.  cell_start must be
identical; inside a cell both sorts place the particles with atomics, so slots are compared per cell after ordering
them by the original particle index.  Collected last (file name) so that a failure here hides no other test."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _canonical(sorted4, perm, cell_start):
    """Slots ordered by (cell, original index)."""
    cell = np.repeat(np.arange(cell_start.size - 1), np.diff(cell_start))
    n_valid = int(cell_start[-1])
    order = np.lexsort((perm[:n_valid], cell))
    return perm[:n_valid][order], sorted4[:n_valid][order]


def _sort_through_abi(entry, eng, dpos, L, ncs, mass, out):
    from nbody_b200 import _lib
    sorted4, perm, cell_start = out
    rc = entry(eng.ctx, C.c_void_p(dpos.data_ptr()), dpos.shape[0], float(L), int(ncs), float(mass),
               C.c_void_p(sorted4.data_ptr()), C.c_void_p(perm.data_ptr()), C.c_void_p(cell_start.data_ptr()), eng.stream)
    _lib.check(rc, "cell sort")


@pytest.mark.parametrize("nan_rows", [False, True])
def test_static_cell_sort_equals_cell_sort_on_two_calls_in_a_row(nan_rows):
    import torch
    from nbody_b200.engine import default_engine
    eng = default_engine()
    lib = eng.lib
    dev = eng.device
    rs = np.random.RandomState(31)
    # two different particle sets and cell counts: the second call finds the first call's contents in the scratch
    cases = [(50_000, 250.0, 11, 1.0), (20_011, 1000.0, 7, 2.5)]
    for n, L, ncs, mass in cases:
        pos = rs.uniform(0, L, size=(n, 3))
        pos[:100] = rs.normal(0.3 * L, 0.01 * L, size=(100, 3)) % L          # a crowded cell
        pos[100] = [0.0, L * (1 - 1e-16), -1e-9]                              # edges and a coordinate below zero
        if nan_rows:
            pos[rs.choice(n, 500, replace=False)] = np.nan                    # unused halo-buffer slots are skipped
        dpos = eng.to_device(pos)
        outs = []
        for entry in (lib.nb_cell_sort, lib.nb_cell_sort_static):
            out = (torch.zeros((n, 4), dtype=torch.float32, device=dev), torch.full((n,), -7, dtype=torch.int32, device=dev),
                   torch.full((ncs ** 3 + 1,), -7, dtype=torch.int32, device=dev))
            _sort_through_abi(entry, eng, dpos, L, ncs, mass, out)
            torch.cuda.synchronize()
            outs.append(tuple(t.cpu().numpy() for t in out))
        (s_ref, p_ref, c_ref), (s_new, p_new, c_new) = outs
        assert np.array_equal(c_new, c_ref)
        assert c_ref[0] == 0 and c_ref[-1] == n - (500 if nan_rows else 0)
        pr, sr = _canonical(s_ref, p_ref, c_ref)
        pn, sn = _canonical(s_new, p_new, c_new)
        assert np.array_equal(pn, pr)
        assert np.array_equal(sn, sr)


def test_static_cell_sort_rejects_sizes_beyond_its_scratch():
    import torch
    from nbody_b200 import _lib
    from nbody_b200.engine import default_engine
    eng = default_engine()
    dev = eng.device
    small = torch.zeros((8, 3), dtype=torch.float64, device=dev)
    out = (torch.zeros((8, 4), dtype=torch.float32, device=dev), torch.zeros(8, dtype=torch.int32, device=dev),
           torch.zeros(65 ** 3 + 1, dtype=torch.int32, device=dev))
    with pytest.raises(_lib.NbodyError, match="exceed the static scratch"):
        _sort_through_abi(eng.lib.nb_cell_sort_static, eng, small, 100.0, 65, 1.0, out)          # 65^3 > 2^18 cells
    n_big = (1 << 21) + 1
    with pytest.raises(_lib.NbodyError, match="exceed the static scratch"):
        rc = eng.lib.nb_cell_sort_static(eng.ctx, C.c_void_p(small.data_ptr()), n_big, 100.0, 4, 1.0,
                                         C.c_void_p(out[0].data_ptr()), C.c_void_p(out[1].data_ptr()),
                                         C.c_void_p(out[2].data_ptr()), eng.stream)              # rejected before any access
        _lib.check(rc, "nb_cell_sort_static")
    # the engine-level wrapper goes through the same entry point
    n, L, ncs = 4096, 64.0, 5
    pos = eng.to_device(np.random.RandomState(5).uniform(0, L, size=(n, 3)))
    ref = eng.cell_sort(pos, L, ncs)
    got = eng.cell_sort_static(pos, L, ncs, torch.empty((n, 4), dtype=torch.float32, device=dev),
                               torch.empty(n, dtype=torch.int32, device=dev),
                               torch.empty(ncs ** 3 + 1, dtype=torch.int32, device=dev))
    torch.cuda.synchronize()
    assert torch.equal(ref[2], got[2])
