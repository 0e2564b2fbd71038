#!/bin/bash
# r02 session Q (1 GPU): full solves of the shipped cases (bounded budgets of tests/golden/full + the unmodified files to convergence)
mkdir -p gpurun_out
timeout 1500 python scripts/shipped_cases.py --full --out gpurun_out/q_shipped_cases.json 2>&1 | tail -30
