#!/bin/bash
# DRAM and L2 bytes per launch of the traversal and shade kernels for the three single-GPU configs (profiles/traffic.json; SURVEY 8d asks for both).
# One lane, no graphs, so the launch list is in plain pass order; 3 warm-up passes are skipped, one whole pass of each kernel is captured.
# usage: tools/ncu_traffic.sh <tag>   -> gpurun_out/<tag>_traffic_<config>.csv
T=${1:-r02}
export HRC_TUNE_PASS_LANES=1 HRC_TUNE_GRAPHS=0
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active
for cfg in sphere_grid material_grid terrain; do
  ncu --metrics $M --clock-control none -k regex:"k_trace|k_shade" -s 81 -c 48 --csv --log-file gpurun_out/${T}_traffic_$cfg.csv python bench.py --config $cfg --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
  echo "$cfg: $(grep -c k_trace gpurun_out/${T}_traffic_$cfg.csv) metric rows"
done
