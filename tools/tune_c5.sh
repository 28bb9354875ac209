#!/bin/bash
# Tuning helper (GPU box): one-GPU C5 update time for the shipped library and every variants/libvdp_*.so
cd "$(dirname "$0")/.."
run() {
  VDP_B200_LIB="$2" timeout 120 python tools/c5_time_sharded.py --steps 3 --warmup 2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['ms_per_step'],3), d['elbo'])"
}
run base ""
for f in variants/libvdp_*.so; do n=$(basename $f .so); run ${n#libvdp_} "$PWD/$f"; done
