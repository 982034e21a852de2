#!/bin/bash
# Experiment helper: build/libsvb_<tag>.so = libsvb with sgbm.cu recompiled under extra -D flags.
#   tools/build_variant.sh r16 -DSW_ROWS4=16        then   SVB_LIBRARY=build/libsvb_r16.so python bench.py ...
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
TAG=$1; shift
SRC=$ROOT/stereovision_b200/csrc
mkdir -p $ROOT/build
ARCH="-gencode arch=compute_100a,code=sm_100a"
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo $ARCH -Xcompiler -fPIC "$@" -c $SRC/sgbm.cu -o $ROOT/build/sgbm_$TAG.o
/usr/local/cuda/bin/nvcc -shared $ARCH -o $ROOT/build/libsvb_$TAG.so $SRC/bm.o $SRC/core.o $SRC/display.o $SRC/fuse.o $SRC/imgproc.o $ROOT/build/sgbm_$TAG.o
echo built $ROOT/build/libsvb_$TAG.so
