#!/bin/bash
# per-rank sweep time of a W-rank angle-sharded run, emulated on one GPU: W in $WORLDS, tiles in $TILES
mkdir -p gpurun_out
for lib in ${LIBS:-default}; do if [ "$lib" = default ]; then unset DOMA_B200_LIB; else export DOMA_B200_LIB=$PWD/doma_b200/exp_$lib.so; fi
for W in ${WORLDS:-8}; do for variant in ${VARIANTS:-auto}; do for tile in ${TILES:-4x8}; do
if [ "$tile" = auto ]; then TF=""; else TF="--tile $tile"; fi
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e $TF --variant $variant --emulate-world $W "$@" 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('lib $lib W $W', d['sweep_variant'], 'tile', d['tile'], 'units', d['units_per_rank'], 'sweep_ms %.2f' % d['roofline']['launch_ms'], 'ms_per_step %.1f' % d['ms_per_step'], 'other_ms_per_group %.2f' % ((d['ms_per_step'] - 8 * d['roofline']['launch_ms']) / 8), 'W', d['clocks']['power_w_max'])
"
done; done; done; done 2>&1 | tee gpurun_out/emul.txt
