#!/bin/bash
# registers / shared memory / spills of every kernel in one CUDA source:  tools/ptxas_info.sh dyn_tend
cd "$(dirname "$0")/../mpas-model_b200/csrc/cuda"
nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -fmad=false -prec-div=true -prec-sqrt=true -Xptxas -v -c $1.cu -o /dev/null 2>&1 | \
  grep -E "Compiling entry|registers|spill" | sed -E 's/.*Compiling entry function .(_ZN[0-9]+_GLOBAL__N__[0-9a-z_]+)/\1/' | paste - - - | \
  sed -E "s/' for 'sm_100a'//; s/ptxas info\s+: //g" | awk '{print}' | cut -c1-260
