#!/bin/bash
# Fused colour-gradient push (inside the per-Gaussian backward) against the separate push kernel: correctness, then timing.
set -u
N=${N:-2}
export GGS_BENCH_WATCHDOG_S=${GGS_BENCH_WATCHDOG_S:-60}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
show() { python - "$1" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    st = {k: round(v["ms"], 3) for k, v in d.get("stages", {}).items() if k in ("render_backward", "gaussian_backward")}
    print(sys.argv[1], "n", d.get("n_gpus"), "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"], 1),
          "| check", (d.get("exchange_check") or {}).get("max_rel_vs_nccl"), "| c4", {k: round(v, 3) for k, v in (d.get("c4") or {}).items() if isinstance(v, float)}, st)
except Exception as ex:
    print(sys.argv[1], "unreadable:", ex)
PY
}
for W in ${CHECKS:-c1}; do
  PEER_WORKLOAD=$W timeout 200 $TR tools/peer_check.py > gpurun_out/peer_check_${W}_n${N}.log 2>&1; echo "peer_check $W rc=$?"
  grep "rank 0\|nccl\|MISMATCH\|Error\|error" gpurun_out/peer_check_${W}_n${N}.log | grep -v Warning | tail -n 8
done
for F in ${FUSE:-1 0}; do
  GGS_PEER_FUSE_RGB=$F GGS_BENCH_C4=${C4:-1} timeout 600 $TR bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_n${N}_fuse${F}.json 2> gpurun_out/bench_n${N}_fuse${F}.err; echo "bench n=$N fuse=$F rc=$?"
  show gpurun_out/bench_n${N}_fuse${F}.json; grep -v "OMP_NUM\|\*\*\*" gpurun_out/bench_n${N}_fuse${F}.err | tail -n 2
done
if [ "${N1:-0}" = "1" ]; then
  timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench n1 rc=$?"
  show gpurun_out/bench_n1.json
fi
