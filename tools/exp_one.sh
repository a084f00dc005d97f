#!/bin/bash
# usage: tools/exp_one.sh <label> [ENV=val ...]  — bench configs #2/#3 once under the given environment
L=$1; shift
for cfg in sphere_grid material_grid; do
  env "$@" python bench.py --config $cfg --no-cpu-baseline --steps 16 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; c=r['per_ray']['closest']; a=r['per_ray']['any']
print('[$L] $cfg value %.1f ms %.3f closest %.0f any %.0f shares %s' % (d['value'], d['ms_per_step'], c['Mrays_per_s'], a['Mrays_per_s'], {k:round(x,3) for k,x in r['kernel_share_of_step'].items()}))"
done
