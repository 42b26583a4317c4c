#!/bin/bash
# Round 2 multi-GPU pass (gpurun --gpus N): peer exchange against NCCL (push and switch-reduced variants), the bench line
# with exchange_check (+ the c4 sub-object), the C5 point-range sharding.  Every run has a 60 s watchdog.
set -u
N=${N:-2}
export GGS_BENCH_WATCHDOG_S=${GGS_BENCH_WATCHDOG_S:-60}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
show() { python - "$1" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], "n", d.get("n_gpus"), "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"], 1),
          "| check", (d.get("exchange_check") or {}).get("max_rel_vs_nccl"), "| c4", {k: round(v, 3) for k, v in (d.get("c4") or {}).items() if isinstance(v, float)},
          "|", d["config"].get("parallelism", "")[-75:])
except Exception as ex:
    print(sys.argv[1], "unreadable:", ex)
PY
}
if [ "${CHECK:-1}" = "1" ]; then
  PEER_WORKLOAD=${CHECK_WORKLOAD:-c3} timeout 300 $TR tools/peer_check.py > gpurun_out/peer_check_n${N}.log 2>&1; echo "peer_check rc=$?"
  grep "rank 0\|nccl\|MISMATCH\|Error\|error" gpurun_out/peer_check_n${N}.log | grep -v Warning | tail -n 12
fi
for PULL in ${PULLS:-0 1}; do
  GGS_PEER_PULL=$PULL GGS_BENCH_C4=${C4:-1} timeout 600 $TR bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_n${N}_pull${PULL}.json 2> gpurun_out/bench_n${N}_pull${PULL}.err; echo "bench n=$N pull=$PULL rc=$?"
  show gpurun_out/bench_n${N}_pull${PULL}.json; tail -n 2 gpurun_out/bench_n${N}_pull${PULL}.err
done
if [ "${NCCL:-1}" = "1" ]; then
  GGS_PEER=0 GGS_BENCH_C4=0 timeout 400 $TR bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_n${N}_nccl.json 2> gpurun_out/bench_n${N}_nccl.err; echo "bench nccl rc=$?"
  show gpurun_out/bench_n${N}_nccl.json
fi
if [ "${N1:-1}" = "1" ]; then
  timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench n1 rc=$?"
  show gpurun_out/bench_n1.json
fi
if [ "${C5:-1}" = "1" ]; then
  timeout 600 $TR bench.py --gpus $N --workload c5 --steps 3 --warmup 3 ${C5_ARGS:-} > gpurun_out/bench_c5_n${N}.json 2> gpurun_out/bench_c5_n${N}.err; echo "c5 n=$N rc=$?"
  show gpurun_out/bench_c5_n${N}.json; tail -n 2 gpurun_out/bench_c5_n${N}.err
fi
