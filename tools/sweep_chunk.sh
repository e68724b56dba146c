#!/bin/bash
# e2e throughput of the pipelined host path for several chunk sizes (tuning aid)
for c in "$@"; do
  ORBFE_PIPE_CHUNK=$c timeout 300 python bench.py --no-cpu --no-match --no-legs --steps 30 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('chunk', $c, 'device', round(d['value']), 'e2e', round(d['e2e']['value']))"
done
