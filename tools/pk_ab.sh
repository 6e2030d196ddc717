#!/bin/bash
# Developer A/B of the shell-binning kernel's occupancy knobs: rebuild pkbin.o with
# -DPKF_MIN_CTAS / -DPKF_CHUNK on the GPU box, time the kernel alone with ncu's launch list.
cd "$(dirname "$0")/../nbody_b200/csrc" || exit 1
for v in "$@"; do
  c=${v%%:*}; k=${v##*:}
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC \
      -DPKF_MIN_CTAS=$c -DPKF_CHUNK=$k $PK_EXTRA -c pkbin.cu -o pkbin.o 2>/dev/null || exit 1
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libnbody_b200.so *.o \
      -L/usr/local/cuda/lib64 -lcufft -Xlinker -rpath -Xlinker /usr/local/cuda/lib64 || exit 1
  (cd ../.. && timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:pk_shell_bin_rows \
      --csv python tools/kbench.py --sizes 512:1024 --ops pk 2>/dev/null | grep pk_shell_bin_rows | \
      awk -F'","' -v tag="ctas=$c chunk=$k" '{gsub(/"/,"",$NF); s+=$NF; n++} END {printf "%s: %d launches, mean %.4f (unit of the csv)\n", tag, n, s/n}')
  (cd ../.. && timeout 200 python -m pytest tests/test_gpu_density.py -x -q -m gpu -k "pk" 2>&1 | tail -1)
done
