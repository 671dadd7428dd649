#!/bin/bash
# parity of the k-buffer kernels + the K=16 auxiliary figure of bench.py.  Usage: scripts/k16_bench.sh <tag>
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "kbuffer or origin" 2>&1 | tail -2
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench_k16_$1.log 2>&1
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/bench_k16_$1.log") if l.startswith("{")][-1]); k=d["sorted_k16"]
print(round(d["value"],1), round(k["mpix_per_s"],1), {a.split("::",1)[1] if a.startswith("render::") else a: round(b,3) for a,b in k["stage_ms"].items()})
PY
