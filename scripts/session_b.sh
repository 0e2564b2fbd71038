#!/bin/bash
# r02 session B: why is the LL kernel slower per warp?  ncu captures of both protocols, poll back-off variants
mkdir -p gpurun_out
G="timeout 600 python scripts/sweep_grid.py --reps 1"
for cfg in "1 flags 4x8" "1 ll 4x8" "8 flags 4x2" "8 ll 4x2"; do
  set -- $cfg
  $G --worlds $1 --variants $2 --tiles $3 --out gpurun_out/b_plain.txt > /dev/null 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -s 2 -c 1 -o gpurun_out/b_prof_w$1_$2 -f \
      $G --worlds $1 --variants $2 --tiles $3 --out gpurun_out/b_ncu_runs.txt > gpurun_out/b_ncu_w$1_$2.log 2>&1
  tail -1 gpurun_out/b_ncu_w$1_$2.log
done
timeout 600 python scripts/sweep_grid.py --worlds 8,4,1 --variants ll --tiles 4x1,4x2,4x4,4x8 --out gpurun_out/b_grid_canary.txt 2>&1 | grep "^n "
timeout 300 python scripts/task_latency.py ll | tee gpurun_out/b_task_latency_ll.txt
for lib in sleep256 sleep1024; do
  DOMA_B200_LIB=$PWD/doma_b200/exp_$lib.so timeout 600 python scripts/sweep_grid.py --worlds 8,1 --variants ll --tiles 4x2,4x8 --out gpurun_out/b_grid_$lib.txt 2>&1 | grep "^n "
done
