#!/bin/bash
# GPU box: duration and DRAM bytes of the multi-view per-Gaussian backward kernel (ncu, C2, 8 views per launch),
# then the fused-backward parity tests and a short bench.  Extra arguments go to bench.py.
timeout 400 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
  -k regex:preprocess_bwd_views --launch-skip 2 -c 2 --csv --log-file gpurun_out/views_k.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-drop-in --no-c3 "$@" > /dev/null 2>&1
python - <<'PY'
import csv
rows = [r for r in csv.reader(l for l in open("gpurun_out/views_k.csv") if not l.startswith("==")) if len(r) > 5]
for r in rows[1:]:
    print(r[4][:60], r[-3], r[-1])
PY
timeout 300 python -m pytest tests/test_gpu_views_fused.py -x -q 2>&1 | tail -2
timeout 250 python bench.py --steps 20 --no-cpu-baseline --no-c3 --no-drop-in "$@" > gpurun_out/fused_bench.json 2> gpurun_out/fused_bench.err; echo bench=$?
python -c "
import json; d=json.load(open('gpurun_out/fused_bench.json')); print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'ms/step', round(d['ms_per_step'],3))"
