#!/bin/bash
# Experiment: the first N traversal-stack entries in shared memory (HR_SMEM_STACK = 4 / 8 / 12) vs all in local memory.
# Variants are built with tools/build_variant.sh ssN "-DHR_SMEM_STACK=N" before the call.
O=gpurun_out; mkdir -p $O
HEATRAY_CUDA_LIB=heatray_b200/_variants/libheatray_cuda_ss8.so python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2 > $O/exp_stack.txt
bash tools/exp_variants.sh ss4 ss8 ss12 >> $O/exp_stack.txt 2>&1
cat $O/exp_stack.txt
