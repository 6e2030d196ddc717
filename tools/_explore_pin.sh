#!/bin/bash
# scratch driver for one gpurun call: ncu launch list + one full capture of the confined pair kernel
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-cpu --no-secondary > gpurun_out/r2t_plain.json 2> gpurun_out/r2t_plain.err || { tail -5 gpurun_out/r2t_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2t_launches_p3m_c5.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-secondary > gpurun_out/r2t_ncu.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:p3m_short_v6 --launch-skip 2 -c 2 -o gpurun_out/r2t_k5_confined python bench.py --steps 2 --warmup 1 --no-cpu --no-secondary > gpurun_out/r2t_ncu2.log 2>&1
echo "full rc=$?"; ls -la gpurun_out/*.ncu-rep
