#!/bin/bash
# Runs bench.py (device-resident leg only) on the product library and on each ablated variant; one line each.
cd "$(dirname "$0")/.."
pick='import sys,json; d=json.loads(sys.stdin.read()); print(sys.argv[1], d["ms_per_step"], "ms", round(d["roofline"]["frac"],4), d["clocks"]["sm_mhz"], d["clocks"]["reasons"])'
for w in ${WORKLOADS:-sf7 sf9 sf12}; do
  python bench.py --workload $w --steps 5 --warmup 3 --no-cpu 2>/dev/null | python -c "$pick" "$w exact"
  for v in nosincos nofft; do
    LORA_PHY_LIB=$PWD/tools/ubench/ablate/liblora_$v.so python bench.py --workload $w --steps 5 --warmup 3 --no-cpu 2>/dev/null | python -c "$pick" "$w $v"
  done
done
