#!/bin/bash
# Deferred light shading (HRC_TUNE_DEFER_LIGHTS): whole-pass value / e2e and the occlusion kernel's rate per config.
out=gpurun_out/r02_exp_defer_lights.txt; : > $out
for cfg in sphere_grid material_grid terrain cornell; do
  for v in 0 1; do
    steps=32; [ $cfg = terrain ] && steps=16; [ $cfg = cornell ] && steps=64
    HRC_TUNE_DEFER_LIGHTS=$v python bench.py --config $cfg --steps $steps --warmup 3 --no-cpu-baseline > gpurun_out/tmp_dl.json 2> gpurun_out/tmp_dl.err || { tail -3 gpurun_out/tmp_dl.err >> $out; continue; }
    python - "$cfg" "$v" >> $out <<PY
import json, sys
d = json.loads(open("gpurun_out/tmp_dl.json").read().strip().splitlines()[-1])
pr = d["roofline"]["per_ray"]; sh = d["roofline"]["kernel_share_of_step"]
print("%-14s defer %s value %7.0f e2e %7.0f serial %7.0f | any %6.0f Mrays/s (%.1f ms) closest %6.0f | instrumented ms/step %.3f" % (sys.argv[1], sys.argv[2], d["value"], d["e2e"]["value"], d["e2e"]["serial"]["value"], pr["any"]["Mrays_per_s"], pr["any"]["ms"], pr["closest"]["Mrays_per_s"], d["roofline"]["instrumented_ms_per_step"]))
PY
  done
done
cat $out
