# scratch driver for gpurun calls (development only)
set -x
python -m pytest tests/test_gpu_dist.py tests/test_gpu_nccl.py tests/test_gpu_timeloop.py tests/test_gpu_density.py tests/test_gpu_force.py tests/test_gpu_force_large.py -q 2>&1 | tail -40 > gpurun_out/r2j_gputests.log; cat gpurun_out/r2j_gputests.log | tail -25
python bench.py --steps 5 --warmup 3 --no-secondary > gpurun_out/r2j_bench_p3m.json 2> gpurun_out/r2j_bench_p3m.err; tail -c 300 gpurun_out/r2j_bench_p3m.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 --no-secondary > gpurun_out/r2j_bench_p3m_2gpu.json 2> gpurun_out/r2j_bench_p3m_2gpu.err; tail -c 600 gpurun_out/r2j_bench_p3m_2gpu.err
python - <<'PY'
import json
for f in ("gpurun_out/r2j_bench_p3m.json","gpurun_out/r2j_bench_p3m_2gpu.json"):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["ms_per_step"], d["phase_ms"], d.get("phase_ms_by_rank",{}).get("halo_exchange"), d.get("phase_ms_by_rank",{}).get("migrate"), d["e2e"]["ms_per_step"])
    except Exception as e: print(f, "ERR", e)
PY
