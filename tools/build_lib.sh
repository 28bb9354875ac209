#!/bin/bash
# Build libvdp_b200.so in-tree for sm_100a (cross-compiles without a GPU).
# The library is five translation units (csrc/tu_*.cu, each selects a group of entry points of csrc/vdp_all.inl);
# they are compiled in parallel and linked into one shared object.  VDP_TUS="cvi core" rebuilds only those units
# (the objects of the others are reused from $VDP_OBJ).
set -e
HERE="$(cd "$(dirname "$0")/.." && pwd)"
SRC="$HERE/vi_diffusion_processes_b200/csrc"
OUT="${VDP_OUT:-$HERE/vi_diffusion_processes_b200/libvdp_b200.so}"
OBJ="${VDP_OBJ:-/tmp/vdp_b200_obj}"
ALL="core cvi prior vgp dist"
TUS="${VDP_TUS:-$ALL}"
mkdir -p "$OBJ"
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -diag-suppress 177 ${VDP_NVCC_EXTRA}"
pids=()
for tu in $TUS; do
    nvcc $FLAGS -c -o "$OBJ/tu_$tu.o" "$SRC/tu_$tu.cu" > "$OBJ/tu_$tu.log" 2>&1 &
    pids+=($!)
done
fail=0
for p in "${pids[@]}"; do wait "$p" || fail=1; done
for tu in $TUS; do cat "$OBJ/tu_$tu.log"; done
[ $fail -eq 0 ] || { echo "compile failed"; exit 1; }
objs=""
for tu in $ALL; do objs="$objs $OBJ/tu_$tu.o"; done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o "$OUT.tmp" $objs -ldl
mv "$OUT.tmp" "$OUT"
echo "built $OUT"
