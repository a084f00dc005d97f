#!/usr/bin/env python
"""profiles/traffic.json from the CSVs of tools/ncu_traffic.sh: per config and kernel, DRAM bytes (read + write) and L2 bytes per launch, averaged over
the captured launches, with the launch time and the hit rates / lane counts captured in the same pass."""
import collections
import csv
import json
import sys
from pathlib import Path

NAMES = {"k_trace<0, 0, 0>": "k_trace<closest>", "k_trace<1, 0, 0>": "k_trace<any>", "k_trace<1, 0, 1>": "k_trace<any>", "k_shade<0>": "k_shade<miss>", "k_shade<1>": "k_shade<pbr>", "k_shade<2>": "k_shade<glass>"}


def main(tag):
    out = {"_source": f"ncu --metrics (dram__bytes_read/write.sum, lts__t_bytes.sum, ...) --clock-control none, bench.py --steps 2 --warmup 3 with one lane and graphs off "
                      f"(tools/ncu_traffic.sh {tag}); 48 launches after the warm-up passes = one whole pass and a half of every kernel; bytes are per launch"}
    for cfg in ("sphere_grid", "material_grid", "terrain"):
        p = Path(f"gpurun_out/{tag}_traffic_{cfg}.csv")
        if not p.exists():
            continue
        rows = [r for r in csv.reader(open(p)) if len(r) > 10]
        hdr = rows[0]
        ki, mi, vi, idi = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("ID")
        per = collections.defaultdict(dict)
        for r in rows[1:]:
            per[(r[idi], r[ki])][r[mi]] = float(r[vi].replace(",", ""))
        agg = collections.defaultdict(list)
        for (_, k), m in per.items():
            for pat, name in NAMES.items():
                if pat in k:
                    agg[name].append(m)
        out[cfg] = {}
        for name, ms in sorted(agg.items()):
            n = len(ms)
            avg = lambda key: sum(m[key] for m in ms) / n
            t = avg("gpu__time_duration.sum") / 1e9
            dram = avg("dram__bytes_read.sum") + avg("dram__bytes_write.sum")
            out[cfg][name] = {"launches_captured": n, "dram_bytes_per_launch": dram, "l2_bytes_per_launch": avg("lts__t_bytes.sum"), "avg_launch_us": t * 1e6,
                              "dram_GBps": dram / t / 1e9, "l2_GBps": avg("lts__t_bytes.sum") / t / 1e9, "l2_hit_pct": avg("lts__t_sector_hit_rate.pct"),
                              "l1_hit_pct": avg("l1tex__t_sector_hit_rate.pct"), "lanes_per_instruction": avg("smsp__thread_inst_executed_per_inst_executed.ratio"),
                              "issue_active_pct": avg("smsp__issue_active.avg.pct_of_peak_sustained_active"), "warps_active_pct": avg("sm__warps_active.avg.pct_of_peak_sustained_active")}
    Path("profiles/traffic.json").write_text(json.dumps(out, indent=1))
    for cfg in out:
        if cfg.startswith("_"):
            continue
        for k, v in out[cfg].items():
            print(f"{cfg:14s} {k:18s} n={v['launches_captured']:2d} {v['avg_launch_us']:8.1f} us  DRAM {v['dram_bytes_per_launch'] / 1e6:7.1f} MB ({v['dram_GBps']:6.0f} GB/s)  L2 {v['l2_bytes_per_launch'] / 1e6:8.1f} MB ({v['l2_GBps']:6.0f} GB/s)  "
                  f"L2 hit {v['l2_hit_pct']:4.1f} L1 hit {v['l1_hit_pct']:4.1f} lanes {v['lanes_per_instruction']:4.1f} issue {v['issue_active_pct']:4.1f} warps {v['warps_active_pct']:4.1f}")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "r02")
