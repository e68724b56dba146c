#!/bin/bash
# e2e throughput of orbfe_extract_batch for compute-stream counts x chunk sizes (environment knobs, no rebuild)
for s in 2 3 4; do for c in 48 64 96 128; do
  ORBFE_PIPE_STREAMS=$s ORBFE_PIPE_CHUNK=$c timeout 300 python bench.py --no-cpu --no-match --no-legs --steps 30 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('streams', $s, 'chunk', $c, 'device', round(d['value']), 'e2e', round(d['e2e']['value']))"
done; done
