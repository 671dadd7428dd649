#!/bin/bash
# short device-only bench; prints Mpix/s and per-stage ms.  Usage: scripts/quick_bench.sh <tag>
python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-aux > gpurun_out/bench_$1.log 2>&1
python - <<PY
import json
l=[x for x in open("gpurun_out/bench_$1.log") if x.startswith("{")][-1]
d=json.loads(l); print(round(d["value"],2), {k.split("::",1)[1] if k.startswith("render::") else k: round(v,3) for k,v in d["stage_ms"].items()})
PY
