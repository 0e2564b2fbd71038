#!/bin/bash
# r02 session M (1 GPU): z-face plane of a tile by bulk asynchronous copies (cp.async.bulk + mbarrier; default library) against the
# per-lane ld/st path (exp_nobulk.so): parity under the dependency checker, task latency, 256^3 and its 4- / 8-rank shards, small cases
mkdir -p gpurun_out
( DOMA_B200_CHECK=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -4 ) | tee gpurun_out/m_pytest_bulk.txt
for lib in default nobulk default nobulk; do
  if [ "$lib" = default ]; then unset DOMA_B200_LIB; else export DOMA_B200_LIB=$PWD/doma_b200/exp_$lib.so; fi
  echo "== lib $lib"
  timeout 300 python scripts/task_latency.py flags 2>&1 | tail -1
  timeout 300 python scripts/small_case_tiles.py asm k1 --quick 2>&1 | grep "sweep us"
  for g in "1 4x8" "8 4x2" "4 4x4"; do set -- $g
    timeout 600 python scripts/sweep_grid.py --worlds $1 --variants flags --tiles $2 --out gpurun_out/m_grid.txt 2>&1 | grep "^n "
  done
done 2>&1 | tee gpurun_out/m_ab.txt
