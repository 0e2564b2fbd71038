#!/bin/bash
# r02 session AC (1 GPU): evidence for the final r02 library (bulk z plane, next-cell prefetch, 32-bit offsets) - GPU tests, both bench arms, the ncu launch list of the bench command and
# one full capture of the sweep kernel at 256^3 (after the plain command exited 0)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv,noheader; nproc
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/ac_gputests.txt
python bench.py --impl reference > gpurun_out/ac_bench_ref.json 2> gpurun_out/ac_bench_ref.err; cut -c1-400 gpurun_out/ac_bench_ref.json
python bench.py > gpurun_out/ac_bench_n256.json 2> gpurun_out/ac_bench.err; cut -c1-1500 gpurun_out/ac_bench_n256.json; tail -3 gpurun_out/ac_bench.err
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-parity"
timeout 600 $B > gpurun_out/ac_bench_plain.json 2> gpurun_out/ac_bench_plain.err &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/ac_launches_n256.csv $B > gpurun_out/ac_ncu_list.log 2>&1
timeout 600 $B > /dev/null 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -s 26 -c 1 -o gpurun_out/ac_prof_sweep -f $B > gpurun_out/ac_ncu_full.log 2>&1
tail -2 gpurun_out/ac_ncu_full.log
python scripts/ncu_summary.py gpurun_out/ac_prof_sweep.ncu-rep gpurun_out/ac_prof_sweep.json
ls -la gpurun_out | grep " ac_"
