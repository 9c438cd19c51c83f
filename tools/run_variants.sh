#!/bin/bash
# Kernel / pipeline experiments: runs the bench with the shipped library and with every orb-ygz-slam_b200/variants/lib_*.so
# (built by hand with different -D flags; git-ignored) and prints value, e2e and the stage times.  Extra bench flags: "$@".
for lib in "" /root/repo/orb-ygz-slam_b200/variants/lib_*.so; do
  [ -z "$lib" ] || [ -f "$lib" ] || continue
  echo -n "variant $(basename ${lib:-shipped}): "
  YGZORB_LIB=$lib python bench.py --steps 10 --warmup 3 --no-cpu --no-sharded --no-mixed "$@" 2>/dev/null | python -c "
import json,sys; j=json.loads(sys.stdin.read()); s=j['stage_ms_per_step']; print(round(j['value']), round(j['e2e']['value']), ' '.join('%s=%.3f'%(k[:4],s[k]) for k in ['pyramid','fast','quadtree','blur','orient_desc','search_by_projection']))"
done
