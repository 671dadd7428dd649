#!/usr/bin/env python
"""Prints (and writes to gpurun_out/parity_ref.json) the three-way comparison statistics of tests/test_gpu_reference.py for
every case: reference + shim (oracle/_ref) vs C oracle vs libgut_b200.  GPU box only."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from tests import test_gpu_reference as T  # noqa: E402

if __name__ == "__main__":
    names = sys.argv[1:] or sorted(T.CASES)
    out = {}
    runs = [(n, 0) for n in names] + ([] if sys.argv[1:] else list(T.KB_CASES))
    for name, k in runs:
        key = name if k == 0 else f"{name}@k{k}"
        try:
            out[key] = T.three_way(name, k)
        except Exception as e:  # keep going: the report is diagnostic
            import traceback
            out[key] = {"error": repr(e), "trace": traceback.format_exc()}
        print(key, json.dumps(out[key], indent=1), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "parity_ref.json"), "w") as f:
        json.dump(out, f, indent=1)
