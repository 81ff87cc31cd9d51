#!/bin/bash
# Round-2 (second half) ncu captures of the lean route: --set full, one launch of demod_lean_kernel and of the certified
# kernel's flagged-packet pass after warm-up.  Every command first runs plain (must exit 0), then under ncu.
set -u
OUT=${1:-gpurun_out}
WHICH=${2:-"7 10 12"}              # which captures (three reports are ~60 MB: gpurun brings back at most 64 MiB per call)
cap() {   # name, kernel regex, skip, count, command...
    local name=$1 k=$2 skip=$3 cnt=$4; shift 4
    "$@" > $OUT/ncu_${name}_plain.log 2>&1 || { echo "$name: plain run failed"; return; }
    ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c $cnt -f -o $OUT/r02_${name} "$@" > $OUT/ncu_${name}.log 2>&1
    tail -1 $OUT/ncu_${name}.log
}
[[ " $WHICH " == *" 7 "* ]] && cap sf7_lean 'demod_lean|demod_cert' 6 2 python bench.py --workload sf7 --steps 2 --warmup 3 --no-cpu --no-also --packets 524288 --e2e-packets 64
[[ " $WHICH " == *" 10 "* ]] && cap sf10_lean 'demod_lean|demod_cert' 6 2 python bench.py --workload sf10 --steps 2 --warmup 3 --no-cpu --no-also --packets 32768 --e2e-packets 64
[[ " $WHICH " == *" 12 "* ]] && cap sf12_lean 'demod_lean|demod_cert' 6 2 python bench.py --workload sf12 --steps 2 --warmup 3 --no-cpu --no-also --packets 4096 --e2e-packets 16
