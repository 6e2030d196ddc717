"""Developer micro-benchmark of the force steps (resident data, CUDA events)."""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nbody_b200.engine import Engine  # noqa: E402
from nbody_b200.force import ForceStepper  # noqa: E402
from nbody_b200.params import ForceConfig  # noqa: E402
from tools.kbench import make_pos, timeit  # noqa: E402


def ev():
    return torch.cuda.Event(enable_timing=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pm", default="256:512")
    ap.add_argument("--p3m", default="256:512")
    ap.add_argument("--pp", default="64")
    ap.add_argument("--interlace", type=int, default=1)
    a = ap.parse_args()
    eng = Engine(0)
    L = 1000.0
    il = (bool(a.interlace), bool(a.interlace))
    for method, spec in (("pm", a.pm), ("p3m", a.p3m)):
        for sz in [s for s in spec.split(",") if s]:
            n, N = (int(x) for x in sz.split(":"))
            pos = make_pos(n, L, "glass")
            mom = torch.zeros_like(pos)
            cfg = ForceConfig(method=method, npart=n ** 3, boxsize=L, gridsize=N, interlace=il)
            st = ForceStepper(cfg, eng)
            res = {"method": method, "n": n, "N": N, "interlace": il[0]}
            # phase timings
            marks = []
            def mark(name):
                e = ev(); e.record(); marks.append((name, e))
            for rep in range(2):
                marks.clear()
                mark("start")
                f, fs = st.pm_mesh(pos); mark("pm_mesh")
                acc = st._acc_buf(pos.shape[0])
                eng.gather_kick_drift(f, fs, N, cfg.cell, pos, None, acc, 1.0, 0.0, 0.0, L); mark("gather")
                if method == "p3m":
                    s4, perm, cs = eng.cell_sort(pos, L, st.ncs); mark("sort")
                    eng.p3m_short(s4, perm, cs, L, st.ncs, cfg.r_s, cfg.r_cut, cfg.eps, 1.0, acc, True); mark("short")
                eng.kick_drift(pos, mom, acc, 0.0, 0.0, L); mark("kickdrift")
                torch.cuda.synchronize()
            for (n0, e0), (n1, e1) in zip(marks[:-1], marks[1:]):
                res[n1 + "_ms"] = round(e0.elapsed_time(e1), 3)
            tmin, tavg = timeit(lambda: st.step(pos, mom, 0.0, 0.0), iters=3, warm=1)
            res["step_ms"] = round(tmin, 3)
            res["particle_steps_per_s"] = round(n ** 3 / tmin * 1e3, 1)
            print(json.dumps(res), flush=True)
            del st, pos, mom
            eng.release_scratch()
            torch.cuda.empty_cache()
    for sz in [s for s in a.pp.split(",") if s]:
        n = int(sz)
        pos = make_pos(n, 500.0, "glass")
        p4 = eng.pp_pack(pos, origin=(250.0,) * 3)
        out = torch.empty((n ** 3, 3), dtype=torch.float64, device="cuda")
        for variant in (1, 2, 3):
            tmin, _ = timeit(lambda: eng.pp_accel(p4, p4, 0.1, variant=variant, out=out), iters=2, warm=1)
            inter = float(n ** 3) ** 2
            print(json.dumps({"method": "pp", "n": n, "variant": variant, "ms": round(tmin, 3),
                              "Ginter_per_s": round(inter / tmin / 1e6, 1),
                              "fp32_frac_of_74.4TF": round(inter * 20 / tmin / 1e9 / 74.4e3, 4)}), flush=True)


if __name__ == "__main__":
    main()
