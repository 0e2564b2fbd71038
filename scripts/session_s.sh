#!/bin/bash
# r02 session S (1 GPU): full GPU suite with the TX-templated kernel and the refitted tile chooser; automatic choices on every case
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 ) | tee gpurun_out/s_gputests.txt
timeout 600 python scripts/small_case_tiles.py tr2d asm k1 --quick 2>&1 | grep "sweep us" | grep auto | tee gpurun_out/s_small_auto.txt
timeout 600 python scripts/sweep_grid.py --worlds 1,2,4,8 --variants auto --tiles auto --out gpurun_out/s_grid.txt 2>&1 | grep "^n "
timeout 1500 python scripts/shipped_cases.py --full --out gpurun_out/s_shipped_cases.json 2>&1 | cut -c1-330 | grep -v '"inner_loop": "host"'
