#!/bin/bash
# Tail threshold sweep (HRC_TUNE_TAIL_RAYS): whole-pass value, serial caller loop and speculative e2e per config.
T=${1:-r02_exp_tail}
out=gpurun_out/$T.txt; : > $out
for cfg in sphere_grid cornell material_grid; do
  for tail in 0 16384 49152 131072 1048576; do
    HRC_TUNE_TAIL_RAYS=$tail python bench.py --config $cfg --steps 32 --warmup 3 > gpurun_out/tmp_tail.json 2> gpurun_out/tmp_tail.err || { tail -3 gpurun_out/tmp_tail.err >> $out; continue; }
    python - "$cfg" "$tail" >> $out <<PY
import json, sys
d = json.loads(open("gpurun_out/tmp_tail.json").read().strip().splitlines()[-1])
e = d["e2e"]
print("%-14s tail %8s value %8.0f ms/step %.3f e2e %8.0f serial %8.0f launches %d" % (sys.argv[1], sys.argv[2], d["value"], d["ms_per_step"], e["value"], e["serial"]["value"], d["gpu_launches"]))
PY
  done
done
cat $out
