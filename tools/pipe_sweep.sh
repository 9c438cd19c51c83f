#!/bin/bash
# e2e throughput of bench.py for several upload-pipeline chunk sizes
for p in 32 64 128 256 0; do
  python bench.py --no-cpu --pipe $p 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('pipe', $p, 'device', round(d['value']), 'e2e', round(d['e2e']['value']))"
done
