"""Throughput of the Ewald PP kernel (pairs/s) beside the plain nearest-image kernel."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from nbody_b200.engine import default_engine

eng = default_engine()
n, L, g = int(sys.argv[1]) if len(sys.argv) > 1 else 65536, 1000.0, 64
pos = torch.rand((n, 3), dtype=torch.float64, device="cuda") * L
p4 = eng.pp_pack(pos, origin=(L / 2, L / 2, L / 2), mass=1.0)
t4 = eng.ewald_table(L, g)
out = torch.empty((n, 3), dtype=torch.float64, device="cuda")
for name, fn in (("ewald", lambda: eng.pp_accel_ewald(p4, p4, 0.1, L, t4, g, out=out)),
                 ("nearest image", lambda: eng.pp_accel(p4, p4, 0.1, box=L, out=out))):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print(f"{name}: n = {n}, {ms:.2f} ms, {n * n / ms / 1e6:.1f} G pairs/s")
