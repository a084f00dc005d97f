#!/bin/bash
# N-GPU default bench line only (no strong-scaling object), with and without the high-priority NCCL stream.
N=${1:-2}; T=${2:-r02c}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:3}" > gpurun_out/${T}_$2_n$N.json 2> gpurun_out/${T}_$2_n$N.err || tail -5 gpurun_out/${T}_$2_n$N.err; }
run 29611 prio --steps 32 --no-strong
HR_BENCH_NCCL_PRIORITY=0 run 29612 noprio --steps 32 --no-strong
python - <<PY
import json
for name in ("prio", "noprio"):
    try:
        d = json.loads(open(f"gpurun_out/${T}_{name}_n$N.json").read().strip().splitlines()[-1])
    except Exception as e:
        print(name, "no line", e); continue
    e = d["e2e"]
    print(name, "value %.0f e2e %.0f serial %.0f spec3 %.0f overlap %.0f" % (d["value"], e["value"], e["serial"]["value"], e["speculative3"]["value"], e["overlapped_exchange"]["value"]))
PY
