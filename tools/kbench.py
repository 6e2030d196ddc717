"""Developer micro-benchmark: times individual operators with CUDA events on resident data.
Not the driver's bench (that is bench.py)."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nbody_b200.engine import Engine  # noqa: E402

HBM = 6548.5


def timeit(fn, iters=5, warm=2, flush=None):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts), sum(ts) / len(ts)


def make_pos(n, L, kind, seed=7):
    g = torch.Generator(device="cuda").manual_seed(seed)
    lat = torch.arange(n, device="cuda", dtype=torch.float64) * (L / n)
    pos = torch.stack(torch.meshgrid(lat, lat, lat, indexing="ij"), -1).reshape(-1, 3)
    pos = (pos + (torch.rand(pos.shape, generator=g, device="cuda", dtype=torch.float64) - 0.5) * (L / n)) % L
    if kind == "shuffled":
        pos = pos[torch.randperm(pos.shape[0], device="cuda", generator=g)]
    return pos.contiguous()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="128:256,256:512")
    ap.add_argument("--ops", default="deposit,pk,deconv")
    a = ap.parse_args()
    eng = Engine(0)
    L = 1000.0
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    for sz in a.sizes.split(","):
        n, N = (int(x) for x in sz.split(":"))
        res = {"n": n, "N": N}
        if "deposit" in a.ops:
            for kind in ("glass", "shuffled"):
                pos = make_pos(n, L, kind)
                grid = torch.empty((N, N, N), dtype=torch.float32, device="cuda")
                tmin, tavg = timeit(lambda: eng.cic_deposit(pos, L / N, N, out=grid), flush=flush)
                alg = 24.0 * n ** 3 + 8.0 * N ** 3
                res[f"deposit_{kind}_ms"] = round(tmin, 4)
                res[f"deposit_{kind}_GBs"] = round(alg / tmin / 1e6, 1)
                res[f"deposit_{kind}_frac"] = round(alg / tmin / 1e6 / HBM, 3)
                del pos
        if "pk" in a.ops or "deconv" in a.ops:
            grid = torch.randn((N, N, N), dtype=torch.float32, device="cuda")
            if "pk" in a.ops:
                eng.power_spectrum_bins(grid, L, N)
                tmin, _ = timeit(lambda: eng.power_spectrum_bins(grid, L, N), flush=flush)
                res["pk_total_ms"] = round(tmin, 4)
            if "deconv" in a.ops:
                tmin, _ = timeit(lambda: eng.deconvolve_cic(grid, L, N), flush=flush)
                res["deconv_total_ms"] = round(tmin, 4)
        print(json.dumps(res), flush=True)
    if "batch" in a.ops:
        # §8f row 4: the 300-snapshot corpus (32^3 particles on a 32^3 mesh each), one call vs 300
        import numpy as np
        B, n, N, Lb = 300, 32, 32, 500.0
        pos = torch.cat([make_pos(n, Lb, "glass", seed=s) for s in range(B)])
        off = np.arange(B + 1, dtype=np.int64) * n ** 3
        box = np.full(B, Lb)
        tb, _ = timeit(lambda: eng.snapshot_delta_batch(pos, off, box, N), iters=5, warm=2)

        def loop():
            for b in range(B):
                g = eng.cic_deposit(pos[b * n ** 3:(b + 1) * n ** 3], Lb / N, N)
                eng.compute_delta(g)
                eng.deconvolve_cic(g, Lb, N, guard=True)
        tl, _ = timeit(loop, iters=3, warm=1)
        print(json.dumps({"op": "snapshot_delta_batch", "B": B, "n": n, "N": N, "batched_ms": round(tb, 3),
                          "per_snapshot_loop_ms": round(tl, 3), "snapshots_per_s_batched": round(B / tb * 1e3, 1)}))


if __name__ == "__main__":
    main()
