#!/bin/bash
# GPU box: parity tests, then a short bench; prints the headline and the per-stage table.
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python bench.py --steps ${1:-10} --no-cpu-baseline > gpurun_out/bench_quick.log 2>&1; echo bench exit=$?
python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/bench_quick.log") if l.startswith("{")][-1])
    print("value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "ref_cuda", d.get("ref_cuda_baseline",{}).get("value"))
    for k,v in d["stages"].items(): print(f"  {k:16s} {v['ms']:.4f} ms  share {v['share']:.3f}")
except Exception as e:
    print("no bench line:", e); print(open("gpurun_out/bench_quick.log").read()[-2000:])
PY
