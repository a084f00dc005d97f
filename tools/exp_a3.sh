#!/bin/bash
# Row A3 experiments: top BVH levels in shared memory (top9 / top73) and TMA bulk refill (tma) against the default build:
# parity subset, bench lines, and the counters that explain them for the closest-hit launches of one pass under ncu.
O=gpurun_out
for v in default top9 top73 tma; do
  if [ "$v" = default ]; then unset HEATRAY_CUDA_LIB; else export HEATRAY_CUDA_LIB=heatray_b200/_variants/libheatray_cuda_$v.so; fi
  echo "== $v: parity"; python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fixed_ray_set or image_parity_equal" 2>&1 | tail -1
  HRC_TUNE_PASS_LANES=1 ncu --metrics gpu__time_duration.sum,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum \
      --clock-control none -k regex:k_trace -s 108 -c 18 --csv --log-file $O/r02_a3_$v.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
done
unset HEATRAY_CUDA_LIB
tools/exp_variants.sh top9 top73 tma
python - <<'PY'
import csv, glob, collections
for f in sorted(glob.glob("gpurun_out/r02_a3_*.csv")):
    rows = [r for r in csv.reader(open(f)) if len(r) > 10]
    hdr = rows[0]; ki, mi, vi = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"); idi = hdr.index("ID")
    per = collections.defaultdict(dict)
    for r in rows[1:]:
        per[(r[idi], r[ki])][r[mi]] = float(r[vi].replace(",", ""))
    clo = [m for (i, k), m in per.items() if "Lb0ELb0ELb0" in k or "false, false, false" in k]
    def avg(name, w=None): 
        return sum(m[name] for m in clo) / max(1, len(clo))
    if clo:
        t = sum(m["gpu__time_duration.sum"] for m in clo)
        print(f, "closest launches %d total %.1f us | L1 hit %.1f L2 hit %.1f long_sb %.2f issue %.1f lanes %.2f inst %.0f M" % (len(clo), t / 1e3, avg("l1tex__t_sector_hit_rate.pct"), avg("lts__t_sector_hit_rate.pct"),
              avg("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"), avg("smsp__issue_active.avg.pct_of_peak_sustained_active"), avg("smsp__thread_inst_executed_per_inst_executed.ratio"), sum(m["smsp__inst_executed.sum"] for m in clo) / 1e6))
PY
