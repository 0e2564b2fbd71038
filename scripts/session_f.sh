#!/bin/bash
# r02 session F: two rows per step (R2) - parity under the forced variant, then timing against the one-row kernel
mkdir -p gpurun_out
( DOMA_B200_ROWS2=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -8 ) | tee gpurun_out/f_pytest_r2.txt
for cfg in "default 0" "default 1" "r2mb4 1" "r2mb2 1"; do set -- $cfg
  if [ "$1" = default ]; then unset DOMA_B200_LIB; else export DOMA_B200_LIB=$PWD/doma_b200/exp_$1.so; fi
  export DOMA_B200_ROWS2=$2
  echo "== lib $1 rows2 $2"
  timeout 300 python scripts/task_latency.py flags 2>&1 | tail -1
  timeout 300 python scripts/small_case_tiles.py tr2d asm k1 --quick 2>&1 | grep "sweep us"
  for g in "1 4x8" "8 4x2" "8 4x4" "4 4x4"; do set -- $g
    timeout 600 python scripts/sweep_grid.py --worlds $1 --variants flags --tiles $2 --out gpurun_out/f_grid.txt 2>&1 | grep "^n "
  done
done 2>&1 | tee gpurun_out/f_r2.txt
