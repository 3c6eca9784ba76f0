#!/bin/bash
# times bench.py for each library variant in lib/var_*.so (kernel experiments); extra arguments go to bench.py
for so in 3d-gaussian-splat-attack_b200/lib/var_*.so; do
  GSR_B200_LIB=$PWD/$so timeout 200 python bench.py --steps 10 --no-cpu-baseline --no-e2e --no-drop-in "$@" > gpurun_out/var.log 2>&1
  python - "$so" <<PY
import json,sys
try:
    d=json.loads([l for l in open("gpurun_out/var.log") if l.startswith("{")][-1])
    print(sys.argv[1].split("/")[-1], "value", round(d["value"],1), {k: round(v["ms"],4) for k,v in d["stages"].items() if k in ("preprocess_fwd","preprocess_bwd","duplicate_keys","sort","blend_fwd","blend_bwd")})
except Exception as e:
    print(sys.argv[1], "FAILED", e, open("gpurun_out/var.log").read()[-500:])
PY
done
