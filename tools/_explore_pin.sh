#!/bin/bash
# scratch driver for one gpurun call (development only): full GPU suite, smoke, the default bench line
mkdir -p gpurun_out
python -m pytest tests -q -m gpu -x 2>&1 | tail -4
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py --no-cpu > gpurun_out/last_bench.json 2> gpurun_out/last_bench.err; tail -c 300 gpurun_out/last_bench.err
python -c "
import json
d=json.loads(open('gpurun_out/last_bench.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['pipelined']['ms_per_step'], d['roofline'].get('exact_mode',{}).get('ms'))"
