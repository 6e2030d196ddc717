set -x
for off in 0.0 0.5; do
  python tools/pin_concept.py --backend gpu --seeds 4 --methods pm,p3m,snapvspk --lattice-offset $off --estimator-interlace 0 --out gpurun_out/r2b_pin_off${off}_il0.json > gpurun_out/r2b_pin_off${off}_il0.log 2>&1
done
python tools/pin_concept.py --backend gpu --seeds 4 --methods pm,p3m --lattice-offset 0.5 --estimator-interlace 0 --dlna 0.0125 --out gpurun_out/r2b_pin_off0.5_il0_dlna0125.json > gpurun_out/r2b_pin_off0.5_il0_dlna0125.log 2>&1
python tools/pin_concept.py --backend gpu --seeds 4 --methods pm,p3m --lattice-offset 0.5 --estimator-interlace 0 --dlna 0.00625 --out gpurun_out/r2b_pin_off0.5_il0_dlna00625.json > gpurun_out/r2b_pin_off0.5_il0_dlna00625.log 2>&1
python tools/pin_concept.py --backend gpu --seeds 4 --methods p3m --lattice-offset 0.5 --estimator-interlace 0 --softening 0.05 --out gpurun_out/r2b_pin_soft05.json > gpurun_out/r2b_pin_soft05.log 2>&1
python tools/pin_concept.py --backend gpu --seeds 4 --methods pm,p3m --lattice-offset 0.5 --estimator-interlace 0 --lpt 2 --out gpurun_out/r2b_pin_lpt2.json > gpurun_out/r2b_pin_lpt2.log 2>&1
python tools/pin_concept.py --backend gpu --seeds 2 --methods pp --lattice-offset 0.5 --estimator-interlace 0 --dlna 0.05 --out gpurun_out/r2b_pin_pp.json > gpurun_out/r2b_pin_pp.log 2>&1
tail -n 12 gpurun_out/r2b_pin_*.log
