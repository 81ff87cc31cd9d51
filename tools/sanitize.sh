#!/bin/bash
# compute-sanitizer over every kernel family (SURVEY.md section 5 "race detection").
#   tools/sanitize.sh [outdir]        run on a B200 box: gpurun -- 'bash tools/sanitize.sh gpurun_out'
# Writes <outdir>/sanitizer_<tool>.{log,out}; summaries are copied to profiles/ by hand after reading them.
# racecheck: shared-memory hazards (the exchange slots, reduction cells, descriptor / staging arrays);
# synccheck: barrier misuse (named barriers of the multi-warp pipelines, __syncwarp masks);
# memcheck:  out-of-bounds / misaligned global + shared accesses (ragged and oversized counts, odd alignment).
set -u
OUT=${1:-gpurun_out}
mkdir -p "$OUT"
cd "$(dirname "$0")/.."
status=0
for tool in memcheck synccheck racecheck; do
    extra=""
    [ "$tool" = racecheck ] && extra="--racecheck-report all"
    timeout ${SANITIZE_TIMEOUT:-1500} compute-sanitizer --tool $tool $extra --print-limit 40 --log-file "$OUT/sanitizer_$tool.log" \
        python tools/sanitize_driver.py > "$OUT/sanitizer_$tool.out" 2>&1
    rc=$?
    echo "== $tool: exit $rc" | tee -a "$OUT/sanitizer_summary.txt"
    tail -n 3 "$OUT/sanitizer_$tool.out" | tee -a "$OUT/sanitizer_summary.txt"
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Error:" "$OUT/sanitizer_$tool.log" | sort | uniq -c | sort -rn | head -20 | tee -a "$OUT/sanitizer_summary.txt"
    [ $rc -ne 0 ] && status=1
done
exit $status
