#!/bin/bash
# GPU box: the round's measurement set -- parity tests, the default bench line, the ncu launch list of the same
# command and one `--set full` capture of every product kernel of one view (summarised here with tools/ncu_*.py).
# usage: tools/profile_round.sh [tag]   -> gpurun_out/<tag>_*
T=${1:-round}
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/${T}_bench_full.log 2> gpurun_out/${T}_bench_full.err; echo bench=$?
grep '^{' gpurun_out/${T}_bench_full.log | tail -1 > gpurun_out/${T}_bench_line.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${T}_launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-c3 > gpurun_out/${T}_ncu_list.log 2>&1; echo list=$?
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^(blend_|preprocess_|onesweep_|duplicate_tiles|tile_ranges|scan_tiles|sort_)" --launch-skip 45 -c 18 -f -o gpurun_out/${T}_full \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-drop-in --no-c3 > gpurun_out/${T}_ncu_full.log 2>&1; echo full=$?
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:preprocess_(bwd|fwd)_views" --launch-skip 2 -c 4 -f -o gpurun_out/${T}_full_views \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-drop-in --no-c3 > gpurun_out/${T}_ncu_full_views.log 2>&1; echo full_views=$?
