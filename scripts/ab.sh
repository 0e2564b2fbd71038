#!/bin/bash
# A/B of library variants (LIBS) at one GPU and in the 8-rank angle-shard emulation
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_cases.py -m gpu -x -q 2>&1 | tail -4
LIBS="${LIBS:-default old}" TILES="4x8" bash scripts/sweep_exp.sh 2>&1 | tee gpurun_out/ab.txt
LIBS="${LIBS:-default old}" WORLDS="8" TILES="${ETILES:-4x4 4x2 2x2}" bash scripts/emul.sh
