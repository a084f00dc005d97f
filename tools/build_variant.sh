#!/bin/bash
# Builds an experimental variant of the CUDA library with extra -D flags on kernels.cu:
#   tools/build_variant.sh <name> "<nvcc -D flags>"   ->  heatray_b200/_variants/libheatray_cuda_<name>.so
# Select it at run time with HEATRAY_CUDA_LIB=<path> (heatray_b200/capi.py). Experiments only; the product is the default build.
set -e
cd "$(dirname "$0")/.."
N=$1; shift
mkdir -p heatray_b200/_variants
python -m heatray_b200.build > /dev/null
nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -Xcompiler -fPIC,-fvisibility=hidden,-ffp-contract=off,-fwrapv,-pthread --expt-relaxed-constexpr ${HR_FMAD:--fmad=false} \
     "$@" -c heatray_b200/csrc/kernels.cu -o heatray_b200/_variants/kernels_$N.o
OBJS=$(ls heatray_b200/_build/*.o | grep -v kernels.cu.o)
nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o heatray_b200/_variants/libheatray_cuda_$N.so heatray_b200/_variants/kernels_$N.o $OBJS -Xcompiler -pthread
echo heatray_b200/_variants/libheatray_cuda_$N.so
