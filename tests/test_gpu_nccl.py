"""The multi-GPU path under REAL NCCL ("the real NCCL path has no correctness check on
hardware"): needs >= 2 CUDA devices, otherwise skipped (the single-GPU box runs the same per-rank code
with thread-emulated ranks in test_gpu_dist.py).  Launches tests/nccl_worker.py with one rank per GPU
(2, and 4 when present) and checks its report: PM / P³M / PP kick-drift steps with particle migration
and the slab P(k) equal the single-GPU result."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _devices():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_steps_under_nccl_match_single_gpu(world):
    if _devices() < world:
        pytest.skip(f"needs {world} CUDA devices")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29511 + world), os.path.join(HERE, "nccl_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("NCCL_RESULT ")]
    assert r.returncode == 0 and lines, r.stdout[-2000:] + r.stderr[-4000:]
    res = json.loads(lines[-1][len("NCCL_RESULT "):])
    assert res["world"] == world
    for method in ("pm", "p3m", "pp"):
        m = res[method]
        assert m["n_total"] == m["n_expected"], method                  # no particle lost or duplicated in migration
        assert m["mom_err"] <= 1e-5 and m["pos_err_cells"] <= 1e-5, (method, m)
    assert res["pm"]["count_change_this_rank"] + res["p3m"]["count_change_this_rank"] > 0   # migration happened
    assert res["pk"]["counts_equal"] and res["pk"]["p_err"] <= 1e-5
