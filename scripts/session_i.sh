#!/bin/bash
# r02 session I (1 GPU): ncu evidence for the shipped kernels - full captures of the sweep at W = 1, 2, 4, 8 (per-N DRAM traffic),
# of the R2 kernel on the assembly case, the launch list of the bench command; cost of the IEEE / no-FMA debug builds
mkdir -p gpurun_out
G="timeout 600 python scripts/sweep_grid.py --reps 1"
for cfg in "1 4x8" "2 4x8" "4 4x4" "8 4x2"; do
  set -- $cfg
  $G --worlds $1 --variants flags --tiles $2 --out gpurun_out/i_plain.txt > /dev/null 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -s 2 -c 1 -o gpurun_out/i_prof_w$1 -f \
      $G --worlds $1 --variants flags --tiles $2 --out gpurun_out/i_ncu_runs.txt > gpurun_out/i_ncu_w$1.log 2>&1
  tail -1 gpurun_out/i_ncu_w$1.log
  python scripts/ncu_summary.py gpurun_out/i_prof_w$1.ncu-rep gpurun_out/i_prof_w$1.json
  if [ $1 != 1 ]; then rm -f gpurun_out/i_prof_w$1.ncu-rep; fi
done
timeout 300 python scripts/small_case_tiles.py asm --quick > /dev/null 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -s 2 -c 1 -o gpurun_out/i_prof_asm_r2 -f \
    python scripts/small_case_tiles.py asm --quick > gpurun_out/i_ncu_asm.log 2>&1
python scripts/ncu_summary.py gpurun_out/i_prof_asm_r2.ncu-rep gpurun_out/i_prof_asm_r2.json; rm -f gpurun_out/i_prof_asm_r2.ncu-rep
timeout 300 python scripts/small_case_tiles.py tr2d --quick > /dev/null 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -s 2 -c 1 -o gpurun_out/i_prof_tr2d -f \
    python scripts/small_case_tiles.py tr2d --quick > gpurun_out/i_ncu_tr2d.log 2>&1
python scripts/ncu_summary.py gpurun_out/i_prof_tr2d.ncu-rep gpurun_out/i_prof_tr2d.json; rm -f gpurun_out/i_prof_tr2d.ncu-rep
# launch list of the bench command (after the plain run exits 0)
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-parity"
timeout 600 $B > gpurun_out/i_bench_plain.json 2> gpurun_out/i_bench_plain.err &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/i_launches_n256.csv $B > gpurun_out/i_ncu_list.log 2>&1
tail -1 gpurun_out/i_bench_plain.json | cut -c1-300
for l in ieee nofma ieee_nofma; do
  echo "== lib $l"
  DOMA_B200_LIB=$PWD/doma_b200/exp_$l.so timeout 600 python scripts/sweep_grid.py --worlds 1 --variants flags --tiles 4x8 --out gpurun_out/i_grid_$l.txt 2>&1 | grep "^n "
done | tee gpurun_out/i_debug_builds.txt
ls -la gpurun_out/ | grep " i_"
