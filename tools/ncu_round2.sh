#!/bin/bash
# Round-2 ncu captures (run on the B200 box through gpurun; summaries are made here with tools/ncu_summary.py).
# Every command first runs plain (must exit 0), then under ncu: --set full, one launch of the named kernel after warm-up.
set -u
OUT=${1:-gpurun_out}
cap() {   # name, kernel regex, skip, command...
    local name=$1 k=$2 skip=$3; shift 3
    "$@" > $OUT/ncu_${name}_plain.log 2>&1 || { echo "$name: plain run failed"; return; }
    ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -f -o $OUT/r02_${name} "$@" > $OUT/ncu_${name}.log 2>&1
    tail -1 $OUT/ncu_${name}.log
}
cap sf7_cert demod_cert 3 python bench.py --workload sf7 --steps 2 --warmup 3 --no-cpu --no-also --packets 524288 --e2e-packets 64
cap sf10_cert demod_cert 3 python bench.py --workload sf10 --steps 2 --warmup 3 --no-cpu --no-also --packets 32768 --e2e-packets 64
cap sf12_cert demod_cert 3 python bench.py --workload sf12 --steps 2 --warmup 3 --no-cpu --no-also --packets 4096 --e2e-packets 16
cap stream_sf9 stream_sync 1 python tools/bench_stream.py --sf 9 --bytes 2e9 --samples-per-stream 131072 --steps 2
