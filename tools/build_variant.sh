#!/bin/bash
# A/B builds of the CUDA library: tools/build_variant.sh NAME [-DFLAG ...] -> build/variants/NAME.so
# (run with MVOLPS_LIB=build/variants/NAME.so python profiles/prof_solve.py)
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
name=$1; shift
mkdir -p $ROOT/build/variants
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC -I$ROOT/include "$@" \
  -c $ROOT/mvolps_b200/csrc/engine_cuda.cu -o $ROOT/build/variants/$name.o
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $ROOT/build/variants/$name.so $ROOT/build/variants/$name.o \
  $ROOT/build/glpk_shim.o $ROOT/build/lp_reader.o $ROOT/build/bnb_driver.o -Xlinker -Bsymbolic -ldl
