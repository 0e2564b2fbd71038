#!/bin/bash
# r02 session A: validate the refactor + both protocols, baseline numbers, protocol/tile grid
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv
nproc
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 ) | tee gpurun_out/a_pytest.txt
timeout 600 python scripts/task_latency.py flags 2>&1 | tee gpurun_out/a_task_latency_flags.txt
timeout 600 python scripts/task_latency.py ll 2>&1 | tee gpurun_out/a_task_latency_ll.txt
timeout 600 python scripts/sweep_grid.py --out gpurun_out/a_sweep_grid.txt 2>&1 | tail -60
timeout 600 python scripts/small_case_tiles.py 2>&1 | tee gpurun_out/a_small_case_tiles.txt
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err; tail -c 2500 gpurun_out/a_bench.json; tail -3 gpurun_out/a_bench.err
