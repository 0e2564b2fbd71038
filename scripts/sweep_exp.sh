#!/bin/bash
# run the short bench over (lib, tile) combinations: LIBS="mb5 mb6" TILES="4x8 2x8" scripts/sweep_exp.sh [bench flags]
for lib in ${LIBS:-default}; do
  for tile in ${TILES:-4x8}; do
    if [ "$lib" = default ]; then unset DOMA_B200_LIB; else export DOMA_B200_LIB=$PWD/doma_b200/exp_$lib.so; fi
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --tile $tile "$@" 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('lib $lib', 'V', d['sweep_variant'], 'tile', d['tile'], 'value %.4g' % d['value'], 'kernel %.4g' % d['roofline']['kernel_updates_per_s'], 'launch_ms %.2f' % d['roofline']['launch_ms'], 'W', d['clocks']['power_w_max'], d['clocks']['sm_mhz'], d['clocks']['reasons'])
"
  done
done
