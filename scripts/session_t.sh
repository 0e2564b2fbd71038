#!/bin/bash
# r02 session T (1 GPU): the new T3 digests through both front ends; automatic tile on the small cases after the R1 correction
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests/test_gpu_full_cases.py tests/test_gpu_driver.py -m gpu -x -q 2>&1 | tail -6 ) | tee gpurun_out/t_pytest.txt
timeout 600 python scripts/small_case_tiles.py tr2d asm k1 --quick 2>&1 | grep "sweep us" | grep auto | tee gpurun_out/t_small_auto.txt
