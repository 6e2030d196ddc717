"""Worker of tests/test_gpu_nccl.py: launched by torch.distributed.run with one rank per GPU, runs the
sharded steps of nbody_b200.dist under REAL NCCL (TorchComm: all_to_all_single with uneven splits,
neighbour shifts, all-gather, all-reduce) and compares with the single-GPU ForceStepper on the same
inputs (which tests/test_gpu_force.py pins to the oracle).  Rank 0 prints one JSON line."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def gather_rows(t, world):
    """All ranks' (n_r, k) float64 tensors stacked on every rank (uneven n_r)."""
    n = torch.tensor([t.shape[0]], device=t.device, dtype=torch.int64)
    ns = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(ns, n)
    cap = int(max(int(x.item()) for x in ns))
    pad = torch.zeros((cap, t.shape[1]), dtype=t.dtype, device=t.device)
    pad[:t.shape[0]] = t
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad)
    return torch.cat([b[:int(k.item())] for b, k in zip(bufs, ns)]).cpu().numpy()


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from conftest import zeldovich_particles
    from nbody_b200.dist import DistForceStepper, DistPowerSpectrum, TorchComm, shard_particles
    from nbody_b200.engine import Engine
    from nbody_b200.force import ForceStepper
    from nbody_b200.params import ForceConfig
    eng = Engine(local)
    out = {"world": world}
    L = 500.0
    for method, n, N, steps in (("pm", 32, 64, 3), ("p3m", 32, 64, 3), ("pp", 16, 0, 2)):
        pos0 = zeldovich_particles(n, L, seed=21)
        rs = np.random.RandomState(1)
        mom0 = rs.normal(scale=30.0, size=pos0.shape)          # large drifts: particles cross slab faces
        cfg = ForceConfig(method=method, npart=n ** 3, boxsize=L, gridsize=N, interlace=(True, True),
                          deconvolve=(True, True), G=1.0, mass=1.5, a=0.5)
        # single-GPU result on this device
        p1, m1 = eng.to_device(pos0.copy()), eng.to_device(mom0.copy())
        ref = ForceStepper(cfg, eng)
        for _ in range(steps):
            ref.step(p1, m1, 0.3, 0.7)
        p1, m1 = p1.cpu().numpy(), m1.cpu().numpy()
        # the same steps sharded over the NCCL ranks
        full = eng.to_device(pos0)
        shard = shard_particles(full, method, L, rank, world, N)
        if method == "pp":
            chunk = (pos0.shape[0] + world - 1) // world
            msh = eng.to_device(mom0[rank * chunk:min(pos0.shape[0], (rank + 1) * chunk)])
        else:
            ix = (pos0[:, 0] / (L / N)).astype(np.int64) % N
            msh = eng.to_device(mom0[ix // (N // world) == rank])
        st = DistForceStepper(cfg, eng, rank, world)
        p, m = shard, msh
        moved = 0
        for _ in range(steps):
            n_before = p.shape[0]
            p, m = st.step(p, m, 0.3, 0.7, n_total=n ** 3)
            moved += abs(p.shape[0] - n_before)
        allp, allm = gather_rows(p.contiguous(), world), gather_rows(m.contiguous(), world)
        o1 = np.lexsort((allm[:, 2], allm[:, 1], allm[:, 0]))
        o2 = np.lexsort((m1[:, 2], m1[:, 1], m1[:, 0]))
        d = np.abs(allp[o1] - p1[o2])
        d = np.minimum(d, L - d)
        out[method] = {"n_total": int(allp.shape[0]), "n_expected": int(n ** 3),
                       "mom_err": float(np.abs(allm[o1] - m1[o2]).max() / np.abs(m1).max()),
                       "pos_err_cells": float(d.max() / (L / max(N, n))), "count_change_this_rank": int(moved)}
    # slab P(k) against the single-GPU estimator
    n, N, L = 32, 64, 1000.0
    pos0 = zeldovich_particles(n, L, seed=4)
    from nbody_b200 import api
    k1, pk1, c1 = api.power_spectrum_from_positions(pos0, L, N, return_counts=True)
    ps = DistPowerSpectrum(eng, TorchComm(), N, L)
    _e, counts, sums = ps(shard_particles(eng.to_device(pos0), "pk", L, rank, world, N), pos0.shape[0])
    counts, sums = counts.cpu().numpy(), sums.cpu().numpy()
    v = counts > 0
    out["pk"] = {"counts_equal": bool(np.array_equal(counts, c1)), "p_err": float(np.abs(sums[v] / counts[v] / pk1 - 1).max())}
    if rank == 0:
        print("NCCL_RESULT " + json.dumps(out), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
