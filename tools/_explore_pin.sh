#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_force.py -q -m gpu -x 2>&1 | tail -3
python bench.py --steps 5 --warmup 3 --no-secondary --no-cpu > gpurun_out/r2y_bench.json 2> gpurun_out/r2y.err || tail -5 gpurun_out/r2y.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r2y_bench.json").read().strip().splitlines()[-1])
print(round(d["ms_per_step"],2), round(d["e2e"]["ms_per_step"],1), d["e2e"].get("pipelined"))
PY
