#!/bin/bash
# r02 session R (1 GPU): rows of 4 cells (template parameter TX): parity incl. the checker, then the small shipped cases over tiles
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -6 ) | tee gpurun_out/r_pytest.txt
timeout 900 python scripts/small_case_tiles.py tr2d asm k1 2>&1 | grep "sweep us" | tee gpurun_out/r_small_tiles.txt
for g in "1 4x8" "8 4x2"; do set -- $g
  timeout 600 python scripts/sweep_grid.py --worlds $1 --variants flags --tiles $2 --out gpurun_out/r_grid.txt 2>&1 | grep "^n "
done
