#!/bin/bash
# Both arms on the other scene shapes (one GPU): the speed-up as a curve, not a point.
set -u
mkdir -p gpurun_out
for WS in ${WORKLOADS:-c2:1 c4:1 c3_3m:1 c3:2 c3:0.5 c2:3}; do
  W=${WS%%:*}; S=${WS##*:}
  timeout 300 python bench.py --workload $W --scale-mult $S --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/curve_${W}_s${S}_ours.json 2> gpurun_out/curve_${W}_s${S}_ours.err; ro=$?
  timeout 300 python bench.py --workload $W --scale-mult $S --impl reference --steps 20 --warmup 5 > gpurun_out/curve_${W}_s${S}_ref.json 2> gpurun_out/curve_${W}_s${S}_ref.err; rr=$?
  python - ${W}_s${S} $ro $rr <<'PY'
import json, sys
w = sys.argv[1]
try:
    o = json.loads(open(f"gpurun_out/curve_{w}_ours.json").read().strip().splitlines()[-1])
    r = json.loads(open(f"gpurun_out/curve_{w}_ref.json").read().strip().splitlines()[-1])
    print(w, o["config"]["workload"][:60], "| D", o["scene"]["num_rendered"], "| ours", round(o["ms_per_step"], 3), "ms ref", round(r["ms_per_step"], 3), "ms | ratio", round(o["value"] / r["value"], 2),
          "e2e ratio", round(o["e2e"]["value"] / r["e2e"]["value"], 2))
except Exception as ex:
    print(w, "failed", sys.argv[2:], ex)
PY
done
