#!/bin/bash
# One full ncu capture of the sweep kernel (after a plain run of the same command exits 0).
# usage: scripts/ncu_sweep.sh <tag> [bench flags...]   -> gpurun_out/prof_<tag>.ncu-rep
tag=$1; shift
mkdir -p gpurun_out
P="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e $*"
$P > gpurun_out/plain_$tag.log 2>&1 || { tail -5 gpurun_out/plain_$tag.log; exit 1; }
# warm-up = 3 steps x 8 groups = 24 sweeps; capture the 27th sweep launch
ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -s 26 -c 1 -o gpurun_out/prof_$tag -f $P > gpurun_out/ncu_$tag.log 2>&1
tail -2 gpurun_out/ncu_$tag.log
