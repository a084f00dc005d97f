#!/bin/bash
# Final bench lines of every config (no ncu), tag $1.
T=${1:-r02_v4}; O=gpurun_out
python bench.py > $O/${T}_bench_sphere_grid.json 2> $O/${T}_bench_sphere_grid.err || tail -5 $O/${T}_bench_sphere_grid.err
python bench.py --config sphere_grid_1m --no-cpu-baseline > $O/${T}_bench_sphere_grid_1m.json 2>/dev/null
python bench.py --config material_grid --no-cpu-baseline > $O/${T}_bench_material_grid.json 2>/dev/null
python bench.py --config terrain --steps 16 --no-cpu-baseline > $O/${T}_bench_terrain.json 2>/dev/null
python bench.py --config cornell --steps 64 --no-cpu-baseline > $O/${T}_bench_cornell.json 2>/dev/null
python - <<PY
import json
for c in ("sphere_grid","sphere_grid_1m","material_grid","terrain","cornell"):
    try: d=json.loads(open(f"gpurun_out/${T}_bench_{c}.json").read().strip().splitlines()[-1])
    except Exception as e: print(c,"ERR",e); continue
    e=d["e2e"]; print(c, "value %.0f ms %.3f e2e %.0f serial %.0f frac %.3f" % (d["value"], d["ms_per_step"], e["value"], e["serial"]["value"], d["roofline"]["frac"]))
PY
