#!/bin/bash
# r02 session C: full GPU test-suite on the current tree; FLAGS sync variants (out-of-line wait/publish)
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 ) | tee gpurun_out/c_pytest.txt
for lib in default sync1 sync2; do
  if [ "$lib" = default ]; then unset DOMA_B200_LIB; else export DOMA_B200_LIB=$PWD/doma_b200/exp_$lib.so; fi
  echo "== lib $lib"
  timeout 300 python scripts/task_latency.py flags 2>&1 | tail -1
  for cfg in "1 4x8" "2 4x8" "4 4x4" "8 4x2"; do set -- $cfg
    timeout 600 python scripts/sweep_grid.py --worlds $1 --variants flags --tiles $2 --out gpurun_out/c_grid_$lib.txt 2>&1 | grep "^n "
  done
done
unset DOMA_B200_LIB
timeout 600 python scripts/shipped_cases.py --out gpurun_out/c_shipped_cases.json 2>&1 | tail -20
