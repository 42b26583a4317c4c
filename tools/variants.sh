#!/bin/bash
# A/B timing of library variants on one GPU:  tools/variants.sh [--workload W] base NAME...
# ("base" = the in-tree libggs_rasterizer.so; NAME = variants/libggs_rasterizer_NAME.so, see csrc/Makefile `variant`).
# One bench.py run per variant (30 steps), stage times side by side.  Output: gpurun_out/variant_NAME.json
set -u
mkdir -p gpurun_out
V=$PWD/geometry-grounded-gaussian-splatting_b200/diff_gaussian_rasterization/variants
EXTRA=""
if [ "${1:-}" = "--workload" ]; then EXTRA="--workload $2"; shift 2; fi
for n in "$@"; do
  if [ "$n" = "base" ]; then unset GGS_RASTERIZER_LIB; else export GGS_RASTERIZER_LIB=$V/libggs_rasterizer_$n.so; fi
  timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline $EXTRA > gpurun_out/variant_$n.json 2> gpurun_out/variant_$n.err || echo "variant $n failed: $(tail -n 2 gpurun_out/variant_$n.err)"
done
unset GGS_RASTERIZER_LIB
python - "$@" <<'PY'
import json, sys
for n in sys.argv[1:]:
    try:
        d = json.loads(open(f"gpurun_out/variant_{n}.json").read().strip().splitlines()[-1])
        st = {k: round(v["ms"], 4) for k, v in d.get("stages", {}).items()}
        print(f"{n:10s} ms/step {d['ms_per_step']:.4f}  e2e {d['e2e']['value']:.1f}  {st}")
    except Exception as ex:
        print(n, "unreadable:", ex)
PY
