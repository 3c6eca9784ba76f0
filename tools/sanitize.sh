#!/bin/bash
# GPU box: compute-sanitizer (memcheck, racecheck, initcheck, synccheck) over the small GPU tests -- every product
# kernel, both blend variants, the sort (also through its stage-level entry point), the three-stream ViewPipeline and
# the host-buffer pipeline.  Writes one log per tool and a summary to gpurun_out/; copy the summary to profiles/.
SEL='tiny or small or sort or golden or odd_gaussian or misaligned or overflow or two_streams or reference_style'
mkdir -p gpurun_out
: > gpurun_out/sanitize_summary.txt
for tool in memcheck racecheck initcheck synccheck; do
  extra=""
  [ $tool = memcheck ] && extra="--leak-check no"
  [ $tool = racecheck ] && extra="--racecheck-report all"
  timeout 1500 compute-sanitizer --tool $tool $extra --target-processes all --print-limit 20 \
    python -m pytest tests -m gpu -x -q -k "$SEL" -p no:cacheprovider > gpurun_out/sanitize_$tool.log 2>&1
  rc=$?
  {
    echo "== compute-sanitizer --tool $tool (exit $rc)"
    grep -E "passed|failed|error" gpurun_out/sanitize_$tool.log | tail -2
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY" gpurun_out/sanitize_$tool.log | sort | uniq -c
    grep -E "^========= (Invalid|Uninitialized|Race|Barrier|Error:|Warning:)" gpurun_out/sanitize_$tool.log | sort | uniq -c | head -20
  } >> gpurun_out/sanitize_summary.txt
done
cat gpurun_out/sanitize_summary.txt
