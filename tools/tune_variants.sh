#!/bin/bash
# Tuning helper (GPU box): bench.py --quick for the shipped library and every variants/libvdp_*.so; one line each.
cd "$(dirname "$0")/.."
run() {
  VDP_B200_LIB="$2" python bench.py --quick --no-c5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1', round(d['ms_per_step'],4), {k:round(v['ms_per_launch'],4) for k,v in d['phases'].items() if v['ms_per_launch']>0})"
}
run base ""
for f in variants/libvdp_*.so; do n=$(basename $f .so); run ${n#libvdp_} "$PWD/$f"; done
