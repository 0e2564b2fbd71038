#!/bin/bash
# build an experimental libdoma_b200 variant: scripts/build_exp.sh <name> <extra nvcc flags...> -> doma_b200/exp_<name>.so
name=$1; shift
cd "$(dirname "$0")/../doma_b200/csrc"
F="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -I ../../include -I ."
mkdir -p /tmp/exp_$name
nvcc $F "$@" -c sweep.cu -o /tmp/exp_$name/sweep.o && nvcc $F -c api.cu -o /tmp/exp_$name/api.o && nvcc $F -fmad=false -c elementwise.cu -o /tmp/exp_$name/elementwise.o && nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../exp_$name.so /tmp/exp_$name/*.o -ldl && echo built exp_$name.so
