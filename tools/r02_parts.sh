#!/bin/bash
# Parts of the exchange alone, and whole steps with the SH sum next to the per-Gaussian backward (0) or after it (1).
set -u
export GGS_BENCH_WATCHDOG_S=60
mkdir -p gpurun_out
N=${N:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for W in c3 c4; do for LATE in 0 1; do
  GGS_PEER_SH_LATE=$LATE PEER_WORKLOAD=$W timeout 200 $TR tools/peer_parts.py > gpurun_out/peer_parts_${W}_late$LATE.log 2>&1; echo "$W late=$LATE rc=$?"
  grep "^{" gpurun_out/peer_parts_${W}_late$LATE.log || tail -n 5 gpurun_out/peer_parts_${W}_late$LATE.log
done; done
