#!/bin/bash
# Build the C restatement of the D = 1 oracle (test infrastructure / timed CPU baseline) -> oracle/_build/
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
mkdir -p "$HERE/_build"
gcc -O2 -fopenmp -shared -fPIC -o "$HERE/_build/libcvi_d1.so" "$HERE/cvi_d1.c" -lm
echo "built $HERE/_build/libcvi_d1.so"
