#!/bin/bash
# Pixel tile of a warp's primary rays (HRC_TUNE_RAY_TILE_W): value, primary / secondary closest-hit rates per config.
out=gpurun_out/r02_exp_raytile.txt; : > $out
for cfg in sphere_grid material_grid terrain; do
  for w in 0 16 8 4; do
    steps=32; [ $cfg = terrain ] && steps=16
    HRC_TUNE_RAY_TILE_W=$w python bench.py --config $cfg --steps $steps --warmup 3 --no-cpu-baseline > gpurun_out/tmp_rt.json 2> gpurun_out/tmp_rt.err || { tail -3 gpurun_out/tmp_rt.err >> $out; continue; }
    python - "$cfg" "$w" >> $out <<PY
import json, sys
d = json.loads(open("gpurun_out/tmp_rt.json").read().strip().splitlines()[-1])
pr = d["roofline"]["per_ray"]; c, s = pr["closest"], pr["secondary"]
prim_rays, prim_ms = c["rays"] - s["rays"], c["ms"] - s["ms"]
print("%-14s tile_w %2s value %7.0f e2e %7.0f | primary %6.0f Mrays/s  secondary %6.0f  any %6.0f" % (sys.argv[1], sys.argv[2], d["value"], d["e2e"]["value"], prim_rays / prim_ms / 1e3, s["Mrays_per_s"], pr["any"]["Mrays_per_s"]))
PY
  done
done
cat $out
