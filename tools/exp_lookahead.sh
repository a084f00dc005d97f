#!/bin/bash
# e2e with a look-ahead of 1 / 2 / 3 deferred passes (bench.py HR_BENCH_LOOKAHEAD), configs #2 and #1; GPU tests first
O=gpurun_out; mkdir -p $O
: > $O/exp_lookahead.txt
for cfg in sphere_grid material_grid cornell; do for la in 1 2 3; do
  HR_BENCH_LOOKAHEAD=$la python bench.py --config $cfg --no-cpu-baseline --steps 32 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']
print('$cfg lookahead $la value %.1f ms %.3f e2e %.1f ms %.3f serial %.1f' % (d['value'], d['ms_per_step'], e['value'], e['ms_per_step'], e['serial']['value']))"
done; done >> $O/exp_lookahead.txt 2>&1
cat $O/exp_lookahead.txt
