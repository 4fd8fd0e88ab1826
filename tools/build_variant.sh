#!/bin/bash
# usage: tools/build_variant.sh NAME "EXTRA NVCC FLAGS"  -> multigrid_b200/variants/libmgb200_NAME.so (degree 4 only swapped)
set -e
cd "$(dirname "$0")/../multigrid_b200/csrc"
mkdir -p ../variants _build/var_$1
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 --extended-lambda -Xcompiler -fPIC -DMGB_P=4 $2 -c degree_tu.cu -o _build/var_$1/degree_4.o
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/libmgb200_$1.so _build/solver.o _build/dg.o _build/adaptive.o _build/degree_1.o _build/degree_2.o _build/degree_3.o _build/var_$1/degree_4.o _build/degree_5.o _build/degree_6.o _build/degree_7.o _build/degree_8.o
