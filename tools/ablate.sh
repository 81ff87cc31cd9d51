#!/bin/bash
# Measurement-only build variants of the demodulator (never the product; outputs are git-ignored .so files
# under tools/ubench/ablate/).  Each variant swaps ONE part of the per-sample chain for a cheap, inexact
# stand-in so that `LORA_PHY_LIB=<variant> python bench.py --no-cpu` times the chain without it.
# (round 1's nosincos / nofft / neither variants of demod_packet.cu went with that kernel: profiles/r01_ablation.txt)
# Usage: tools/ablate.sh   (after `make -C <pkg>/csrc`, which provides the other objects)
set -euo pipefail
ROOT=$(cd "$(dirname "$0")/.." && pwd)
C=$ROOT/lora-sdr-lightweight-standalone-library-clean_b200/csrc
OUT=$ROOT/tools/ubench/ablate
mkdir -p "$OUT"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -Xcompiler -fPIC"
# certified demodulator (round 2): the walk without its transform (keys taken from the rotated samples, every window
# counted as certified) -- the streaming skeleton's own cost at the product's occupancy
for v in noxform; do
  $NVCC $FLAGS -DCERT_ABLATE=1 -c "$C/demod_cert.cu" -o "$OUT/demod_cert_$v.o"
  $NVCC -shared -cudart static -gencode arch=compute_100a,code=sm_100a -o "$OUT/liblora_cert_$v.so" \
      "$OUT/demod_cert_$v.o" $(ls "$C"/build/*.o | grep -v '/demod_cert.o$')
  echo "built $OUT/liblora_cert_$v.so"
done
