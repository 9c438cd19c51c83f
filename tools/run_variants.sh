for v in "" a b c d e f g h; do
  if [ -z "$v" ]; then lib=""; else lib=/root/repo/orb-ygz-slam_b200/variants/lib_$v.so; [ -f $lib ] || continue; fi
  echo -n "variant ${v:-base}: "
  YGZORB_LIB=$lib python bench.py --steps 10 --warmup 3 --no-cpu --no-sharded --no-mixed 2>/dev/null | python -c "
import json,sys; j=json.loads(sys.stdin.read()); s=j['stage_ms_per_step']; print(round(j['value']), ' '.join('%s=%.3f'%(k[:4],s[k]) for k in ['pyramid','fast','quadtree','blur','orient_desc','search_by_projection']))"
done
