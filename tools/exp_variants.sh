#!/bin/bash
# bench each experimental library variant on configs #2 and #3 (no CPU baseline), print value + shares
for v in default "$@"; do
  for cfg in sphere_grid material_grid; do
    if [ "$v" = default ]; then unset HEATRAY_CUDA_LIB; else export HEATRAY_CUDA_LIB=heatray_b200/_variants/libheatray_cuda_$v.so; fi
    python bench.py --config $cfg --no-cpu-baseline --steps 16 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
print('$v $cfg value %.1f ms %.3f shares %s closest %.0f any %.0f' % (d['value'], d['ms_per_step'], {k:round(x,3) for k,x in r['kernel_share_of_step'].items()}, r['per_ray']['closest']['Mrays_per_s'], r['per_ray']['any']['Mrays_per_s']))"
  done
done
