# scratch driver for gpurun calls (development only)
set -x
timeout 420 python -m pytest tests/test_gpu_nccl.py -q -x 2>&1 | tail -30 > gpurun_out/r2k_nccl.log; tail -12 gpurun_out/r2k_nccl.log
for n in 4 2; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/r2k_bench_p3m_${n}gpu.json 2> gpurun_out/r2k_bench_p3m_${n}gpu.err; tail -c 400 gpurun_out/r2k_bench_p3m_${n}gpu.err
done
python - <<'PY'
import json
for n in (4,2):
    f=f"gpurun_out/r2k_bench_p3m_{n}gpu.json"
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(n, d["ms_per_step"], d["phase_ms"], d["e2e"]["ms_per_step"], d["invariants"], {k:(v.get("ms_per_step"),v.get("error")) for k,v in d.get("secondary",{}).items()})
    except Exception as e: print(f, "ERR", e)
PY
