#!/bin/bash
# r02 session AA (1 GPU): full ncu captures of the FINAL sweep kernel on the shard one rank of 2 / 4 / 8 holds (per-N DRAM traffic, residual limiter at N = 8)
mkdir -p gpurun_out
G="timeout 300 python scripts/sweep_grid.py --reps 1"
for cfg in "2 4x8" "4 4x4" "8 4x2"; do
  set -- $cfg
  $G --worlds $1 --variants flags --tiles $2 --out gpurun_out/aa_plain.txt > /dev/null 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -s 2 -c 1 -o gpurun_out/aa_prof_w$1 -f \
      $G --worlds $1 --variants flags --tiles $2 --out gpurun_out/aa_ncu_runs.txt > gpurun_out/aa_ncu_w$1.log 2>&1
  tail -1 gpurun_out/aa_ncu_w$1.log | cut -c1-200
  python scripts/ncu_summary.py gpurun_out/aa_prof_w$1.ncu-rep gpurun_out/aa_prof_w$1.json
  rm -f gpurun_out/aa_prof_w$1.ncu-rep
done
ls -la gpurun_out | grep " aa_"
