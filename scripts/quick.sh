#!/bin/bash
# quick GPU iteration: parity subset + short bench (no CPU baseline / e2e). Args: extra bench flags.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -15
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e "$@" 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('V', d['sweep_variant'], 'tile', d['tile'], 'value %.4g' % d['value'], 'kernel %.4g' % d['roofline']['kernel_updates_per_s'], 'launch_ms %.2f' % d['roofline']['launch_ms'], d['clocks'])
"
