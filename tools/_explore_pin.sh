#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_baseline_sizes.py -q -m gpu -x -s -k "exact" 2>&1 | grep -i "P3M 128\|near pairs\|passed\|failed\|rror" | head -4
NB_P3M_NEAR=1 python -m pytest tests/test_gpu_force.py tests/test_gpu_dist.py tests/test_gpu_force_large.py -q -m gpu -x 2>&1 | tail -2
NB_P3M_NEAR=1 python bench.py --steps 5 --warmup 3 --no-secondary --no-cpu > gpurun_out/r3a_bench_near1.json 2> gpurun_out/r3a.err || tail -3 gpurun_out/r3a.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r3a_bench_near1.json").read().strip().splitlines()[-1])
print("near 1:", round(d["ms_per_step"],2), round(d["phase_ms"]["p3m_short"],2), d["roofline"].get("exact_mode"))
PY
