# scratch driver for gpurun calls (development only)
set -x
python -m pytest tests -m gpu -q 2>&1 | tail -15 > gpurun_out/r2i_gputests.log; cat gpurun_out/r2i_gputests.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r2i_bench_p3m.json 2> gpurun_out/r2i_bench_p3m.err
for w in pm pp pk; do python bench.py --workload $w --steps 5 --warmup 3 > gpurun_out/r2i_bench_$w.json 2> gpurun_out/r2i_bench_$w.err; done
B="python bench.py --steps 1 --warmup 1 --no-cpu --no-secondary"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2i_launches_p3m_c5.csv $B > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"kspace_fused|sort_count|sort_scatter|cic_deposit|gather_kick|p3m_short_v6|grf_kernel" -c 14 -o gpurun_out/r2i_prof_step $B > gpurun_out/ncu2.log 2>&1
P="python bench.py --workload pp --steps 1 --warmup 1 --no-cpu"
ncu --set full --clock-control none --import-source on -k regex:"pp_accel|pp_near" -s 2 -c 2 -o gpurun_out/r2i_prof_pp $P > gpurun_out/ncu3.log 2>&1
K="python bench.py --workload pk --steps 1 --warmup 1 --no-cpu"
ncu --set full --clock-control none --import-source on -k regex:"pk_shell|grid_affine" -s 2 -c 2 -o gpurun_out/r2i_prof_pk $K > gpurun_out/ncu4.log 2>&1
ls -la gpurun_out/*.ncu-rep
tail -3 gpurun_out/ncu2.log
