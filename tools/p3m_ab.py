"""Developer A/B check of the short-range kernels: run with NB_P3M_KERNEL=<v> and a tag; compares
against the other tag's saved output when it exists."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nbody_b200.engine import Engine  # noqa: E402

tag = sys.argv[1]
other = sys.argv[2] if len(sys.argv) > 2 else None
eng = Engine(0)
rs = np.random.RandomState(3)
out = {}
for name, L, ncs, npart, ilim_frac, half in (("u20", 100.0, 6, 6 ** 3 * 20, 1.0, False), ("odd7", 100.0, 7, 7 ** 3 * 25, 1.0, False), ("clump", 100.0, 9, 9 ** 3 * 12, 0.9, True), ("halo", 100.0, 5, 5 ** 3 * 14, 0.8, True),
                                       ("dense", 100.0, 5, 5 ** 3 * 40, 1.0, False), ("big", 100.0, 40, 40 ** 3 * 22, 1.0, False)):
    pos = rs.uniform(0, L, size=(npart, 3))
    if half:
        pos[:, 0] *= 0.6
    cs = L / ncs
    r_cut = cs * 0.999
    r_s = r_cut / 4.5
    p = eng.to_device(pos)
    s4, perm, cstart = eng.cell_sort(p, L, ncs, mass=1.0)
    acc = torch.zeros_like(p)
    eng.p3m_short(s4, perm, cstart, L, ncs, r_s, r_cut, 0.01 * cs, scale=1.0, out=acc, accumulate=False,
                  i_limit=int(npart * ilim_frac))
    torch.cuda.synchronize()
    out[name] = acc.cpu().numpy()
    out[name + "_cell"] = (np.floor(pos / cs).astype(int) % ncs)
np.savez(f"gpurun_out/p3m_ab_{tag}.npz", **out)
if other:
    o = np.load(f"gpurun_out/p3m_ab_{other}.npz")
    for name in ("u20", "odd7", "clump", "halo", "dense", "big"):
        a, b = out[name], o[name]
        d = np.abs(a - b).max(axis=1)
        rms = np.sqrt((b ** 2).mean())
        bad = np.nonzero(d > 1e-4 * rms)[0]
        print(name, "max diff / rms", d.max() / rms, "bad", bad.size, "of", a.shape[0])
        if bad.size:
            cells = out[name + "_cell"][bad]
            print("  bad cells z histogram", np.bincount(cells[:, 2], minlength=8)[:40])
            print("  first bad", bad[:5], a[bad[:3]], b[bad[:3]])
