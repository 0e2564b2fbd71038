#!/bin/bash
# r02 session H: full GPU suite on the tree with R2 (automatic) + single-copy row body; co-issue microbenchmark;
# arithmetic attribution; flat-builder start-up; shipped cases; bench
mkdir -p gpurun_out
( timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 ) | tee gpurun_out/h_pytest.txt
doma_b200/bin/fp64_coissue 2>&1 | tee gpurun_out/h_fp64_coissue.txt
for l in default ieee nofma ieee_nofma; do
  if [ "$l" = default ]; then unset DOMA_B200_LIB; else export DOMA_B200_LIB=$PWD/doma_b200/exp_$l.so; fi
  timeout 300 python tests/attribution.py $l 2>&1 | grep "rel err"
done | tee gpurun_out/h_attribution.txt
unset DOMA_B200_LIB
for f in --flat --no-flat; do
  echo "== doma_b200_driver -i full_syn256_b.xml $f"
  ( /usr/bin/time -v timeout 900 doma_b200/bin/doma_b200_driver -i tests/golden/cases/full_syn256_b.xml $f 2>&1 | grep -E "^DRIVER|Maximum resident|Elapsed" )
done 2>&1 | tee gpurun_out/h_driver_flat.txt
timeout 900 python scripts/shipped_cases.py --out gpurun_out/h_shipped_cases.json 2>&1 | tail -30 | cut -c1-400
timeout 600 python scripts/small_case_tiles.py tr2d asm k1 2>&1 | grep "sweep us" | tee gpurun_out/h_small_case_tiles.txt
timeout 300 python scripts/task_latency.py flags 2>&1 | tail -3 | tee gpurun_out/h_task_latency.txt
timeout 900 python bench.py > gpurun_out/h_bench.json 2> gpurun_out/h_bench.err; tail -c 3000 gpurun_out/h_bench.json
