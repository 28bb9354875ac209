#!/bin/bash
# Tuning helper (GPU box): tools/quad_profile.py for the shipped library and every variants/libvdp_*.so
cd "$(dirname "$0")/.."
run() {
  VDP_B200_LIB="$2" timeout 120 python tools/quad_profile.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1', {k:(round(v['ms_per_update'],3), round(v['phases_ms']['kl_grad'],3), round(v['phases_ms']['elbo'],3)) for k,v in d.items() if isinstance(v,dict)})"
}
run base ""
for f in variants/libvdp_*.so; do n=$(basename $f .so); run ${n#libvdp_} "$PWD/$f"; done
