#!/bin/bash
# k_fast time vs resident CTAs per SM (dynamic shared memory padding)
for pad in 0 4000 8000 12000 16000 24000; do
  YGZ_FAST_SMEM_PAD=$pad python bench.py --no-cpu --steps 5 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('pad', $pad, 'device', round(d['value']), 'fast ms', round(d['stage_ms_per_step']['fast'],3))"
done
