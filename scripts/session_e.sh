#!/bin/bash
# r02 session E: small-case latency regime - who is slowing the chain?  canary polling, CTA caps; inline sync variants
mkdir -p gpurun_out
for lib in default canary canary32; do
  if [ "$lib" = default ]; then unset DOMA_B200_LIB; else export DOMA_B200_LIB=$PWD/doma_b200/exp_$lib.so; fi
  for mb in 0 37 74 148 296; do
    export DOMA_B200_MAX_BLOCKS=$mb
    echo "== lib $lib max_blocks $mb"
    timeout 300 python scripts/small_case_tiles.py tr2d asm --quick 2>&1 | grep -v "^$" | tail -12
  done
done 2>&1 | tee gpurun_out/e_small.txt
unset DOMA_B200_MAX_BLOCKS
for lib in sync0i sync1i sync2i sync2; do
  export DOMA_B200_LIB=$PWD/doma_b200/exp_$lib.so
  echo "== lib $lib"
  timeout 300 python scripts/task_latency.py flags 2>&1 | tail -1
  for cfg in "1 4x8" "8 4x2"; do set -- $cfg
    timeout 600 python scripts/sweep_grid.py --worlds $1 --variants flags --tiles $2 --out gpurun_out/e_grid_$lib.txt 2>&1 | grep "^n "
  done
done 2>&1 | tee gpurun_out/e_sync.txt
