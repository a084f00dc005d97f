#!/bin/bash
N=${1:-2}
for load in 0 1; do for mult in 1 2 4; do
  echo "== load $load grid mult $mult"
  HRC_EXCHANGE_LOAD=$load HRC_EXCHANGE_GRID_MULT=$mult python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29655 tools/exp_exchange.py 2>&1 | grep "GPUs" | grep "peer"
done; done
