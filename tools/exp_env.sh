#!/bin/bash
# usage: tools/exp_env.sh "VAR=a VAR2=b" "VAR=c" ...   — bench configs #2/#3 under each environment setting
for envs in "" "$@"; do
  for cfg in sphere_grid material_grid; do
    env $envs python bench.py --config $cfg --no-cpu-baseline --steps 16 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; c=r['per_ray']['closest']; a=r['per_ray']['any']
print('[$envs] $cfg value %.1f ms %.3f nodes/tris %.2f/%.2f any %.2f/%.2f closest %.0f any %.0f bvh_nodes %d' % (d['value'], d['ms_per_step'], c['nodes'], c['tris'], a['nodes'], a['tris'], c['Mrays_per_s'], a['Mrays_per_s'], d['config']['bvh_nodes']))"
  done
done
