#!/bin/bash
# scripts/lib_eval_all.sh: evaluate every doma_b200/exp_*.so and the default library
mkdir -p gpurun_out; : > gpurun_out/lib_eval.txt
unset DOMA_B200_LIB; python scripts/lib_eval.py | tee -a gpurun_out/lib_eval.txt
for f in doma_b200/exp_*.so; do DOMA_B200_LIB=$PWD/$f python scripts/lib_eval.py | tee -a gpurun_out/lib_eval.txt; done
unset DOMA_B200_LIB; python scripts/lib_eval.py | tee -a gpurun_out/lib_eval.txt
sort -k2 -n gpurun_out/lib_eval.txt | head -8
