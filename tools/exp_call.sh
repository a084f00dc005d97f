#!/bin/bash
# usage: tools/exp_call.sh <out-name> <variant> ...  — parity tests under each variant, then bench default + variants on configs #2/#3
O=gpurun_out; mkdir -p $O; N=$1; shift
for v in "$@"; do
  echo "[$v] parity:"; HEATRAY_CUDA_LIB=heatray_b200/_variants/libheatray_cuda_$v.so python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -1
done > $O/$N.txt
bash tools/exp_variants.sh "$@" >> $O/$N.txt 2>&1
cat $O/$N.txt
