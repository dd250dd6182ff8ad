#!/bin/bash
# Build a variant of the library that differs in ONE translation unit: tools/mkvariant.sh <name> <unit.cu> [nvcc -D flags...]
# -> .ab/lib_<name>.so (select it with SCG_LIB=$PWD/.ab/lib_<name>.so; tools/ab_bench.sh alternates bench runs)
NAME=$1; UNIT=$2; shift 2
mkdir -p .ab
B=$(basename $UNIT .cu)
nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC "$@" -c -o .ab/$NAME.$B.o safe_control_gym_b200/csrc/$UNIT 2> .ab/$NAME.log && \
nvcc -shared -o .ab/lib_$NAME.so .ab/$NAME.$B.o $(ls safe_control_gym_b200/build/*.o | grep -v "/$B.o")
