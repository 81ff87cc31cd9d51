#!/bin/bash
# Measurement-only build variants of the demodulator (never the product; outputs are git-ignored .so files
# under tools/ubench/ablate/).  Each variant swaps ONE part of the per-sample chain for a cheap, inexact
# stand-in so that `LORA_PHY_LIB=<variant> python bench.py --no-cpu` times the chain without it.
#   nosincos : the glibc-exact double-precision sincosf of the CFO rotation -> 3 float instructions
#   neither  : both of the above (what is left: bulk copies, scale, complex multiply, |X|^2 first-max, bookkeeping)
#   nofft    : the exact-order FFT of the data windows skipped (pipeline kernel, SF7-11), argmax on the rotated samples
# Usage: tools/ablate.sh   (after `make -C <pkg>/csrc`, which provides the other objects)
set -euo pipefail
ROOT=$(cd "$(dirname "$0")/.." && pwd)
C=$ROOT/lora-sdr-lightweight-standalone-library-clean_b200/csrc
OUT=$ROOT/tools/ubench/ablate
mkdir -p "$OUT"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -Xcompiler -fPIC"
$NVCC $FLAGS -DLORA_ABLATE_SINCOS -c "$C/demod_packet.cu" -o "$OUT/demod_packet_nosincos.o"
$NVCC -shared -cudart static -gencode arch=compute_100a,code=sm_100a -o "$OUT/liblora_nosincos.so" \
    "$OUT/demod_packet_nosincos.o" $(ls "$C"/build/*.o | grep -v '/demod_packet.o$')
echo "built $OUT/liblora_nosincos.so"
$NVCC $FLAGS -DLORA_ABLATE_FFT -c "$C/demod_packet.cu" -o "$OUT/demod_packet_nofft.o"
$NVCC -shared -cudart static -gencode arch=compute_100a,code=sm_100a -o "$OUT/liblora_nofft.so" \
    "$OUT/demod_packet_nofft.o" $(ls "$C"/build/*.o | grep -v '/demod_packet.o$')
echo "built $OUT/liblora_nofft.so"
$NVCC $FLAGS -DLORA_ABLATE_FFT -DLORA_ABLATE_SINCOS -c "$C/demod_packet.cu" -o "$OUT/demod_packet_neither.o"
$NVCC -shared -cudart static -gencode arch=compute_100a,code=sm_100a -o "$OUT/liblora_neither.so" \
    "$OUT/demod_packet_neither.o" $(ls "$C"/build/*.o | grep -v '/demod_packet.o$')
echo "built $OUT/liblora_neither.so"
# certified demodulator (round 2): the walk without its transform (keys taken from the rotated samples, every window
# counted as certified) -- the streaming skeleton's own cost at the product's occupancy
for v in noxform; do
  $NVCC $FLAGS -DCERT_ABLATE=1 -c "$C/demod_cert.cu" -o "$OUT/demod_cert_$v.o"
  $NVCC -shared -cudart static -gencode arch=compute_100a,code=sm_100a -o "$OUT/liblora_cert_$v.so" \
      "$OUT/demod_cert_$v.o" $(ls "$C"/build/*.o | grep -v '/demod_cert.o$')
  echo "built $OUT/liblora_cert_$v.so"
done
