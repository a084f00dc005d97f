#!/bin/bash
# Round-end style GPU check: parity tests, bench lines for the configs, ncu launch list + one full capture of k_trace.
# usage: tools/gpu_check.sh <tag>     (outputs under gpurun_out/<tag>_*)
T=${1:-chk}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > $O/${T}_pytest.txt; cat $O/${T}_pytest.txt
python bench.py > $O/${T}_bench_sphere_grid.json 2> $O/${T}_bench_sphere_grid.err || { tail -20 $O/${T}_bench_sphere_grid.err; exit 1; }
python bench.py --config sphere_grid_1m --no-cpu-baseline > $O/${T}_bench_sphere_grid_1m.json 2>/dev/null
python bench.py --config material_grid --no-cpu-baseline > $O/${T}_bench_material_grid.json 2>/dev/null
python bench.py --config terrain --steps 8 --no-cpu-baseline > $O/${T}_bench_terrain.json 2>/dev/null
python bench.py --config cornell --steps 64 --no-cpu-baseline > $O/${T}_bench_cornell.json 2>/dev/null
python bench.py --impl reference --steps 16 --warmup 1 > $O/${T}_bench_reference.json 2>/dev/null
# ncu serialises kernels; one lane keeps the launch list in plain pass order (3 warm-up passes, 2 counter passes with the STATS variants, 2 timed, 2 instrumented, 2 e2e)
export HRC_TUNE_PASS_LANES=1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${T}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/${T}_ncu_list.log 2>&1
# one whole pass of traversal launches (9 closest + 9 any) in steady state: skip six passes' worth (warm-up + counter passes) = 108 k_trace launches
ncu --set full --clock-control none --import-source on -k regex:k_trace -s 108 -c 18 -o $O/${T}_trace_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/${T}_ncu_full.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 108 -c 6 -o $O/${T}_shade_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/${T}_ncu_full2.log 2>&1
unset HRC_TUNE_PASS_LANES
ls -la $O | tail -20
