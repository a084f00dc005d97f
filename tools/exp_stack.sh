#!/bin/bash
# Experiments: (1) the first N traversal-stack entries in shared memory (HR_SMEM_STACK = 4 / 8 / 12) vs all in local memory;
# (2) k_shade pulls the next trip's ray / hit records into L2 (HR_SHADE_PREFETCH=1).
# Variants are built with tools/build_variant.sh <name> "-D..." before the call.
O=gpurun_out; mkdir -p $O
for v in ss8 pf; do
  HEATRAY_CUDA_LIB=heatray_b200/_variants/libheatray_cuda_$v.so python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
done > $O/exp_stack.txt
bash tools/exp_variants.sh ss4 ss8 ss12 pf >> $O/exp_stack.txt 2>&1
cat $O/exp_stack.txt
