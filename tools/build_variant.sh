#!/bin/bash
# Tuning helper: build variants/libvdp_<name>.so with extra nvcc flags for the "cvi" translation unit only (the unit
# bench.py's site update runs from); the other units are taken from the last full build.
#   tools/build_variant.sh l9 "-DVDP_LC1=9 -DVDP_LN1=5 -DVDP_LM1=5"
set -e
HERE="$(cd "$(dirname "$0")/.." && pwd)"
NAME="$1"; FLAGS="$2"
OBJ=/tmp/vdp_b200_obj
VOBJ=/tmp/vdp_b200_var_$NAME
mkdir -p "$VOBJ" "$HERE/variants"
cp $OBJ/tu_core.o $OBJ/tu_prior.o $OBJ/tu_vgp.o $OBJ/tu_dist.o $OBJ/tu_cvi.o "$VOBJ/"
# VDP_VARIANT_TUS="cvi dist": the units rebuilt with the extra flags (default: cvi)
for tu in ${VDP_VARIANT_TUS:-cvi}; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -diag-suppress 177 $FLAGS \
       -c -o "$VOBJ/tu_$tu.o" "$HERE/vi_diffusion_processes_b200/csrc/tu_$tu.cu" &
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o "$HERE/variants/libvdp_$NAME.so" "$VOBJ"/tu_*.o
echo "built variants/libvdp_$NAME.so"
