#!/bin/bash
# run the bench once per tile-shape variant library under variants/
for f in variants/lib_*.so; do
  VDP_B200_LIB=$PWD/$f python bench.py --quick --steps 5 --warmup 3 2>/dev/null | tail -1 | python -c "
import json,sys
l=json.loads(sys.stdin.read())
print('$f', round(l['ms_per_step'],3), ' '.join('%s=%.3f'%(k,v['ms_per_launch']) for k,v in l['phases'].items()))"
done
