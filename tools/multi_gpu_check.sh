#!/bin/bash
# N-GPU bench lines: default (sample-index sharding, NCCL reduce, + the strong-scaling object), the own peer-memory exchange kernel, terrain by 8-row bands (NCCL reduce / peer kernel).
N=${1:-8}; T=${2:-r02}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:3}" > gpurun_out/${T}_$2_n$N.json 2> gpurun_out/${T}_$2_n$N.err || tail -5 gpurun_out/${T}_$2_n$N.err; }
run 29601 bench --steps 32
HR_BENCH_COLLECTIVE=peer run 29602 bench_peer --steps 32 --no-strong
HR_BENCH_BAND_ROWS=8 run 29603 bench_terrain_rows8 --config terrain --steps 32
HR_BENCH_BAND_ROWS=8 HR_BENCH_COLLECTIVE=peer run 29604 bench_terrain_rows8_peer --config terrain --steps 32
python - <<PY
import json
for name in ("bench", "bench_peer", "bench_terrain_rows8", "bench_terrain_rows8_peer"):
    try:
        d = json.loads(open(f"gpurun_out/${T}_{name}_n$N.json").read().strip().splitlines()[-1])
    except Exception as e:
        print(name, "no line", e); continue
    print(name, "value %.0f ms %.3f reduce_ms %.3f (%s) e2e %.0f serial %.0f spec3 %.0f" % (d["value"], d["ms_per_step"], d.get("reduce_ms", 0), d.get("exchange"), d["e2e"]["value"], d["e2e"]["serial"]["value"], d["e2e"]["speculative3"]["value"]), "overlap-only %.0f" % d["e2e"].get("overlapped_exchange", {}).get("value", 0))
    print("   reduce_check", d.get("reduce_check"))
    if d.get("strong"): print("   strong", {k: v for k, v in d["strong"].items() if k != "config"})
PY
