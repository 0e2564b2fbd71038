#!/bin/bash
# r02 session V (1 GPU): look-ahead x-face load through a walking pointer (exp_xptr.so) against the default library; 4-cell rows on the 8-rank shard
mkdir -p gpurun_out
for lib in default xptr default xptr; do
  if [ "$lib" = default ]; then unset DOMA_B200_LIB; else export DOMA_B200_LIB=$PWD/doma_b200/exp_$lib.so; fi
  echo "== lib $lib"
  for g in "1 4x8" "8 4x2" "4 4x4"; do set -- $g
    timeout 600 python scripts/sweep_grid.py --worlds $1 --variants flags --tiles $2 --out gpurun_out/v_grid.txt 2>&1 | grep "^n "
  done
  timeout 300 python scripts/small_case_tiles.py k1 --quick 2>&1 | grep "sweep us" | grep auto
done 2>&1 | tee gpurun_out/v_ab.txt
( DOMA_B200_LIB=$PWD/doma_b200/exp_xptr.so DOMA_B200_CHECK=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3 ) | tee gpurun_out/v_pytest.txt
