#!/bin/bash
# r02 session N (1 GPU): default library = bulk z plane for tz >= 2, next-cell prefetch, |mu|/dx table, 32-bit cell offsets; baseline exp_nobulk.so (session M's)
mkdir -p gpurun_out
( DOMA_B200_CHECK=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -4 ) | tee gpurun_out/n_pytest.txt
for lib in default nobulk default; do
  if [ "$lib" = default ]; then unset DOMA_B200_LIB; else export DOMA_B200_LIB=$PWD/doma_b200/exp_$lib.so; fi
  echo "== lib $lib"
  timeout 300 python scripts/task_latency.py flags 2>&1 | tail -1
  timeout 300 python scripts/small_case_tiles.py asm k1 tr2d --quick 2>&1 | grep "sweep us" | grep auto
  for g in "1 4x8" "8 4x2" "4 4x4" "2 4x8"; do set -- $g
    timeout 600 python scripts/sweep_grid.py --worlds $1 --variants flags --tiles $2 --out gpurun_out/n_grid.txt 2>&1 | grep "^n "
  done
done 2>&1 | tee gpurun_out/n_ab.txt
