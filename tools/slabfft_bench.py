"""Developer micro-benchmark: the per-rank pieces of the slab FFT (no communication) at the
shapes rank 0 would have for a mesh N split over G ranks, against the single-GPU 3-D cuFFT."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nbody_b200.engine import Engine  # noqa: E402


def timeit(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--N", type=int, default=1024)
    ap.add_argument("--G", default="1,2,4,8")
    a = ap.parse_args()
    eng = Engine(0)
    N = a.N
    nzh = N // 2 + 1
    for G in (int(x) for x in a.G.split(",")):
        nxl = nyl = N // G
        real = torch.randn((nxl, N, N), dtype=torch.float32, device="cuda")
        s2d = torch.empty((nxl, N, nzh), dtype=torch.complex64, device="cuda")
        xbuf = torch.empty((G, nxl, nyl, nzh), dtype=torch.complex64, device="cuda")
        spec = torch.randn((N, nyl, nzh), dtype=torch.complex64, device="cuda")
        res = {"N": N, "G": G, "spec_GB": round(spec.numel() * 8 / 1e9, 3)}
        res["fft2d_r2c_ms"] = round(timeit(lambda: eng.slab_fft2d(N, nxl, +1, real, s2d)), 3)
        res["pack_ms"] = round(timeit(lambda: eng.slab_pack(s2d, xbuf, nxl, G, nyl, nzh)), 3)
        res["fft1d_x_ms"] = round(timeit(lambda: eng.slab_fft1d_x(N, nyl * nzh, +1, spec)), 3)
        res["unpack_ms"] = round(timeit(lambda: eng.slab_unpack(xbuf, s2d, nxl, G, nyl, nzh)), 3)
        res["fft2d_c2r_ms"] = round(timeit(lambda: eng.slab_fft2d(N, nxl, -1, real, s2d)), 3)
        res["a2a_ms_at_770GBs"] = round(spec.numel() * 8 * (G - 1) / G / 770e9 * 1e3, 3)
        print(json.dumps(res), flush=True)
        del real, s2d, xbuf, spec
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
