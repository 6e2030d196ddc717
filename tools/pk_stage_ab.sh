#!/bin/bash
# Developer A/B: 2 vs 3 cp.async stages of the shell-binning kernel (NB_PK_STAGES), kernel time from ncu's launch list.
cd "$(dirname "$0")/.." || exit 1
for st in 2 3; do
  NB_PK_STAGES=$st timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:pk_shell_bin_rows \
      --csv python tools/kbench.py --sizes 512:1024 --ops pk 2>/dev/null | grep pk_shell_bin_rows | \
      awk -F'","' -v tag="stages=$st" '{gsub(/"/,"",$NF); s+=$NF; n++} END {printf "%s: %d launches, mean %.1f ns\n", tag, n, s/n}'
done
