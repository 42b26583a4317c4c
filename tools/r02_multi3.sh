#!/bin/bash
# N = 8: chunked (pipelined) sum against the unchunked one, C4 sub-object, C5 at 15 M and at the 50 M points of BASELINE
# config 5 (N = 8 and one GPU).
set -u
N=${N:-8}
export GGS_BENCH_WATCHDOG_S=${GGS_BENCH_WATCHDOG_S:-60}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
show() { python - "$1" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], "n", d.get("n_gpus"), "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"], 1),
          "| check", (d.get("exchange_check") or {}).get("max_rel_vs_nccl"), "| c4", {k: round(v, 3) for k, v in (d.get("c4") or {}).items() if isinstance(v, float)},
          "| later", round((d.get("later_rounds") or {}).get("value", 0), 1))
except Exception as ex:
    print(sys.argv[1], "unreadable:", ex)
PY
}
for CH in ${CHUNKS:-4 1 2}; do
  GGS_PEER_CHUNKS=$CH GGS_BENCH_C4=1 timeout 400 $TR bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_n${N}_chunks${CH}.json 2> gpurun_out/bench_n${N}_chunks${CH}.err; echo "bench n=$N chunks=$CH rc=$?"
  show gpurun_out/bench_n${N}_chunks${CH}.json; grep -v "OMP_NUM\|\*\*\*" gpurun_out/bench_n${N}_chunks${CH}.err | tail -n 2
done
timeout 300 $TR bench.py --gpus $N --workload c5 --steps 3 --warmup 3 > gpurun_out/bench_c5_n${N}.json 2> gpurun_out/bench_c5_n${N}.err; echo "c5 n=$N rc=$?"
show gpurun_out/bench_c5_n${N}.json
timeout 400 $TR bench.py --gpus $N --workload c5 --c5-gaussians 3333334 --steps 3 --warmup 3 > gpurun_out/bench_c5_50m_n${N}.json 2> gpurun_out/bench_c5_50m_n${N}.err; echo "c5 50M n=$N rc=$?"
show gpurun_out/bench_c5_50m_n${N}.json; grep -v "OMP_NUM\|\*\*\*" gpurun_out/bench_c5_50m_n${N}.err | tail -n 2
timeout 400 python bench.py --workload c5 --c5-gaussians 3333334 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5_50m_n1.json 2> gpurun_out/bench_c5_50m_n1.err; echo "c5 50M n=1 rc=$?"
show gpurun_out/bench_c5_50m_n1.json; tail -n 2 gpurun_out/bench_c5_50m_n1.err
