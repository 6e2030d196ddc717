#!/bin/bash
# scratch driver for one gpurun call (development only)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_force.py tests/test_gpu_dist.py -q -m gpu -x 2>&1 | tail -2
python bench.py > gpurun_out/r2final2_bench_p3m.json 2> gpurun_out/r2final2.err; tail -c 300 gpurun_out/r2final2.err
python -c "
import json
d=json.loads(open('gpurun_out/r2final2_bench_p3m.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['e2e']['value'], d['e2e']['pipelined']['ms_per_step'], d['roofline'].get('exact_mode',{}).get('ms'), d['clocks'])"
