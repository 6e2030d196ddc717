import sys; sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np, torch
from conftest import zeldovich_particles
from nbody_b200 import force as F
from nbody_b200.engine import default_engine
from nbody_b200.params import ForceConfig
eng = default_engine()
L, N, n = 100.0, 64, 40
pos = zeldovich_particles(n, L, seed=11)
cfg = ForceConfig(method="p3m", npart=n ** 3, boxsize=L, gridsize=N)
st = F.ForceStepper(cfg, eng)
dpos = eng.to_device(pos)
def run():
    a = st.short_range(dpos, eng.to_device(np.zeros_like(pos)), accumulate=False).cpu().numpy()
    torch.cuda.synchronize(); return a
def err(a, r): return np.abs(a - r).max() / np.sqrt((r ** 2).mean())
p1, p2 = run(), run()
print("plain vs plain", err(p2, p1), "ncs", st.ncs)
for lim in (24, 100, 147):
    eng.p3m_set_sm_limit(lim)
    c1 = run(); c2 = run()
    eng.p3m_mesh_done(); c3 = run()
    eng.p3m_set_sm_limit(0)
    d = np.abs(c1 - p1).max(axis=1); w = np.argsort(d)[-3:]
    print("limit", lim, "conf vs plain", err(c1, p1), "conf vs conf", err(c2, c1), "after mesh_done", err(c3, p1), "worst rows", w, d[w], p1[w[-1]], c1[w[-1]])
    print("   rows differing > 1e-6:", int((d > 1e-6 * np.sqrt((p1**2).mean())).sum()), "of", len(d))
