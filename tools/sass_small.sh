#!/bin/bash
# Compiles ONE instantiation of the render kernel (default: the small-scene kernel the bench times) in a few seconds
# and prints its ptxas line and opcode histogram: tools/sass_small.sh [out.sass]
cd "$(dirname "$0")/../openglraytracer_b200/csrc"
T=${TU:-/tmp/small_tu.cu}
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -I. -Xptxas -v -c $T -o /tmp/small_tu.o 2>&1 | grep -E "registers|spill|error"
cuobjdump -sass /tmp/small_tu.o > ${1:-/tmp/small_new.sass}
grep -cE "^\s+/\*[0-9a-f]{4}\*/" ${1:-/tmp/small_new.sass}
