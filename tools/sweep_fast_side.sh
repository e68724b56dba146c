#!/bin/bash
# FAST's later launches (the tall-cell class) on the side stream vs in line (tuning aid; ORBFE_FAST_SIDE)
ORBFE_FAST_SIDE=2 timeout 600 python -m pytest tests/test_extract_gpu.py tests/test_pipeline_gpu.py -x -q 2>&1 | tail -1
for s in 1 2 1 2; do
  ORBFE_FAST_SIDE=$s timeout 300 python bench.py --no-cpu --no-match --no-legs --steps 60 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('fast_side', $s, 'device', round(d['value']), 'e2e', round(d['e2e']['value']), 'latency', d.get('shapes',{}).get('latency'))"
done
