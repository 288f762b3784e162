#!/bin/bash
# Build a variant of libb200lk.so with extra -D flags for ONE translation unit (A/B experiments; load it with B200LK_LIB).
#   tools/build_variant.sh <name> <file.cu> "<-D flags>"   ->  iqtree2_b200/libb200lk_<name>.so
set -e
cd "$(dirname "$0")/.."
NAME=$1; SRC=$2; DEFS=$3
python iqtree2_b200/build.py > /dev/null
mkdir -p iqtree2_b200/csrc/build/var_$NAME
OBJ=iqtree2_b200/csrc/build/var_$NAME/$(basename ${SRC%.cu}).o
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fopenmp $DEFS -c -o $OBJ iqtree2_b200/csrc/$SRC
OTHERS=$(ls iqtree2_b200/csrc/build/*.o | grep -v "/$(basename ${SRC%.cu}).o")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -Xcompiler -fPIC,-fopenmp -o iqtree2_b200/libb200lk_$NAME.so $OBJ $OTHERS
cuobjdump -res-usage $OBJ 2>/dev/null | grep -A1 "dense20_rp_kernelILb0" | grep -o "REG:[0-9]*\|STACK:[0-9]*" | paste - - | head -2
