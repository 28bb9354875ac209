#!/bin/bash
# Build libvdp_b200.so in-tree for sm_100a (cross-compiles without a GPU).
set -e
HERE="$(cd "$(dirname "$0")/.." && pwd)"
SRC="$HERE/vi_diffusion_processes_b200/csrc"
OUT="${VDP_OUT:-$HERE/vi_diffusion_processes_b200/libvdp_b200.so}"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC \
     -diag-suppress 177 ${VDP_NVCC_EXTRA} -o "$OUT" "$SRC/vdp_abi.cu"
echo "built $OUT"
