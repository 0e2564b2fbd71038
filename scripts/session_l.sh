#!/bin/bash
# r02 session L (1 GPU): (1) the issue ceiling of the cell solve's own instruction stream (scripts/cell_ceiling.cu);
# (2) A/B of sweep.cu builds: next-cell operand prefetch (pf), 20 warps per SM at 96 registers (mb5), 18 warps at 112 (w3mb6)
mkdir -p gpurun_out
doma_b200/bin/cell_ceiling | tee gpurun_out/l_cell_ceiling.txt
for lib in default pf mb5 pfmb5 w3mb6 pfw3mb6 default; do
  if [ "$lib" = default ]; then unset DOMA_B200_LIB; else export DOMA_B200_LIB=$PWD/doma_b200/exp_$lib.so; fi
  echo "== lib $lib"
  for g in "1 4x8" "8 4x2"; do set -- $g
    timeout 600 python scripts/sweep_grid.py --worlds $1 --variants flags --tiles $2 --out gpurun_out/l_grid.txt 2>&1 | grep "^n "
  done
done 2>&1 | tee gpurun_out/l_ab.txt
