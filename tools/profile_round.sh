#!/bin/bash
# GPU box: the round's measurement set -- parity tests, the default bench line, the ncu launch list of the same
# command and one `--set full` capture of every product kernel of one view (summarised here with tools/ncu_*.py).
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/bench_full.log 2> gpurun_out/bench_full.err; echo bench=$?
tail -1 gpurun_out/bench_full.log > gpurun_out/bench_line.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1; echo list=$?
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^(blend_|preprocess_|onesweep_|duplicate_tiles|tile_ranges|scan_tiles|sort_|publish_)" --launch-skip 45 -c 18 -f -o gpurun_out/full_round \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1; echo full=$?
