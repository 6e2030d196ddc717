"""Host-side logic of the multi-GPU path on CPU: world_size-2 (and 4) `gloo` process groups
exercising nbody_b200.dist.TorchComm exactly as the NCCL run does — the slab-FFT transpose
layouts (pack -> all-to-all -> x transform, and back), the ghost-plane ring shift, the
variable-size halo / migration exchange, all-gather and all-reduce — plus the thread-emulated
communicator the single-GPU parity tests use.  No CUDA kernels are involved here; numpy stands
in for the per-rank transforms so that the exchange logic is checked against np.fft.rfftn."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _slab_fft_roundtrip(comm, N, seed):
    """numpy model of SlabMesh.forward / inverse with the real exchanges."""
    from nbody_b200.dist import SlabLayout
    lay = SlabLayout(N, comm.world, comm.rank)
    rs = np.random.RandomState(seed)
    full = rs.normal(size=(N, N, N))                      # every rank builds the same field
    mine = full[lay.x0:lay.x0 + lay.nxl]
    s2d = np.fft.rfft2(mine, axes=(1, 2))                 # [nxl][N][nzh]
    # pack: [nxl][G][nyl][nzh] -> [G][nxl][nyl][nzh]
    send = np.ascontiguousarray(s2d.reshape(lay.nxl, lay.world, lay.nyl, lay.nzh).transpose(1, 0, 2, 3))
    recv = torch.empty(send.shape, dtype=torch.complex128)
    comm.all_to_all(recv, torch.from_numpy(send))
    spec = np.fft.fft(recv.numpy().reshape(N, lay.nyl, lay.nzh), axis=0)       # [N kx][nyl][nzh]
    ref = np.fft.rfftn(full)[:, lay.y0:lay.y0 + lay.nyl, :]
    err_fwd = np.abs(spec - ref).max() / np.abs(ref).max()
    # inverse: x transform, the result is the send layout [G][nxl][nyl][nzh]
    back = np.ascontiguousarray(np.fft.ifft(spec, axis=0))
    got = torch.empty(back.shape, dtype=torch.complex128)
    comm.all_to_all(got.view(lay.world, lay.nxl, lay.nyl, lay.nzh), torch.from_numpy(back))
    # unpack: [G][nxl][nyl][nzh] -> [nxl][N][nzh]
    un = got.numpy().reshape(lay.world, lay.nxl, lay.nyl, lay.nzh).transpose(1, 0, 2, 3).reshape(lay.nxl, N, lay.nzh)
    real = np.fft.irfft2(un, s=(N, N), axes=(1, 2))
    err_inv = np.abs(real - mine).max()
    return err_fwd, err_inv


def _exercise(comm):
    from nbody_b200.dist import shard_particles
    r, w = comm.rank, comm.world
    out = {}
    out["fft"] = _slab_fft_roundtrip(comm, 8 if w <= 4 else 16, seed=3)
    # ring shifts (ghost planes): +1 for the deposit, -1 for the gather
    recv = torch.empty(4)
    comm.shift(torch.full((4,), float(r)), recv, +1)
    out["shift+"] = recv[0].item() == float((r - 1) % w)
    comm.shift(torch.full((4,), float(r)), recv, -1)
    out["shift-"] = recv[0].item() == float((r + 1) % w)
    # variable-size exchange: rank r sends (d + 1) rows tagged (r, d) to every other rank d
    send = [torch.tensor([[float(r), float(d)]] * (d + 1), dtype=torch.float64) if d != r else None for d in range(w)]
    got = comm.exchange_var(send)
    ok = True
    for s in range(w):
        if s == r:
            ok &= got[s].shape[0] == 0
        else:
            ok &= got[s].shape == (r + 1, 2) and bool((got[s][:, 0] == s).all()) and bool((got[s][:, 1] == r).all())
    out["exchange"] = ok
    # all-gather / all-reduce
    ag = torch.empty(w * 3)
    comm.all_gather(ag, torch.full((3,), float(r)))
    out["gather"] = ag.view(w, 3)[:, 0].tolist() == [float(s) for s in range(w)]
    t = torch.tensor([r + 1], dtype=torch.int64)
    comm.all_reduce(t)
    out["reduce"] = int(t.item()) == w * (w + 1) // 2
    m = torch.tensor([r], dtype=torch.int64)
    comm.all_reduce(m, "max")
    out["max"] = int(m.item()) == w - 1
    # slab ownership partitions the particle set, pp sharding is a contiguous split
    rs = np.random.RandomState(0)
    pos = torch.from_numpy(rs.uniform(0, 100.0, size=(1000, 3)))
    mine = shard_particles(pos, "pm", 100.0, r, w, N=16)
    cnt = torch.tensor([mine.shape[0]], dtype=torch.int64)
    comm.all_reduce(cnt)
    lo, hi = r * 100.0 / w, (r + 1) * 100.0 / w
    out["shard"] = int(cnt.item()) == 1000 and bool(((mine[:, 0] >= lo) & (mine[:, 0] < hi)).all())
    ppm = shard_particles(pos, "pp", 100.0, r, w)
    c2 = torch.tensor([ppm.shape[0]], dtype=torch.int64)
    comm.all_reduce(c2)
    out["shard_pp"] = int(c2.item()) == 1000
    return out


def _check(out):
    ef, ei = out.pop("fft")
    assert ef < 1e-12 and ei < 1e-12, (ef, ei)
    bad = [k for k, v in out.items() if not v]
    assert not bad, bad


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from nbody_b200.dist import TorchComm
        q.put((rank, _exercise(TorchComm())))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_torchcomm_gloo(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for _rank, out in res:
        _check(out)


@pytest.mark.parametrize("world", [1, 2, 3])
def test_threadcomm_matches(world):
    from nbody_b200.dist import run_threaded
    if world == 3:
        # 8 is not divisible by 3: only the collectives, not the slab layout
        def fn(comm):
            t = torch.tensor([comm.rank + 1], dtype=torch.int64)
            comm.all_reduce(t)
            return int(t.item())
        assert run_threaded(3, fn) == [6, 6, 6]
        return
    for out in run_threaded(world, _exercise):
        _check(out)


def test_layout_errors():
    from nbody_b200.dist import SlabLayout
    with pytest.raises(ValueError):
        SlabLayout(10, 4, 0)
    with pytest.raises(ValueError):
        SlabLayout(4, 4, 0)          # 1 plane per rank < 2 ghost planes
    lay = SlabLayout(16, 4, 3)
    assert (lay.x0, lay.nxl, lay.y0, lay.nyl, lay.nzh, lay.rows) == (12, 4, 12, 4, 9, 36)


def test_bind_host_near_device_leaves_the_process_alone_when_it_cannot_read_the_topology():
    """No CUDA device here: the helper returns None and the CPU affinity of the process is unchanged."""
    import os
    from nbody_b200.dist import bind_host_near_device
    before = os.sched_getaffinity(0)
    assert bind_host_near_device(0) is None
    assert os.sched_getaffinity(0) == before
