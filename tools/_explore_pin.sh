#!/bin/bash
# scratch driver for one gpurun call: full GPU suite, smoke, default bench lines
mkdir -p gpurun_out
python -m pytest tests -q -m gpu -x 2>&1 | tail -8
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
python bench.py > gpurun_out/r2q_bench_p3m.json 2> gpurun_out/r2q_bench_p3m.err; tail -c 600 gpurun_out/r2q_bench_p3m.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r2q_bench_p3m.json").read().strip().splitlines()[-1])
print("p3m", d["ms_per_step"], d["value"], d["e2e"], d.get("overlap"), d["clocks"], d["gpu_launches"])
print({k: d.get(k) for k in ("cpu_baseline",)})
PY
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2q_bench_ref.json 2> gpurun_out/r2q_bench_ref.err; tail -c 300 gpurun_out/r2q_bench_ref.json
