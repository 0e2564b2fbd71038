#!/bin/bash
# r02 session O (1 GPU): which of {next-cell prefetch, |mu|/dx table} pays on top of bulk z + 32-bit offsets
mkdir -p gpurun_out
for lib in a b c default nobulk a b c default; do
  if [ "$lib" = default ]; then unset DOMA_B200_LIB; else export DOMA_B200_LIB=$PWD/doma_b200/exp_$lib.so; fi
  echo "== lib $lib"
  for g in "1 4x8" "8 4x2"; do set -- $g
    timeout 600 python scripts/sweep_grid.py --worlds $1 --variants flags --tiles $2 --out gpurun_out/o_grid.txt 2>&1 | grep "^n "
  done
done 2>&1 | tee gpurun_out/o_ab.txt
