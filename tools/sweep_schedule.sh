#!/bin/bash
# e2e throughput of orbfe_extract_batch (512 frames) for explicit chunk schedules and compute-stream counts
run() { ORBFE_PIPE_STREAMS=$1 ORBFE_PIPE_SCHEDULE=$2 timeout 300 python bench.py --no-cpu --no-match --no-legs --steps 30 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('streams $1 schedule $2: e2e', round(d['e2e']['value']), 'ms', round(d['e2e']['ms_per_step'],3))"; }
for s in 2 3 4; do
  run $s "16,62,62,62,62,62,62,62,62"
  run $s "16,64,64,64,64,64,48,40,32,24,16,8,8"
  run $s "16,64,64,64,64,56,48,40,32,24,16,12,8,4"
  run $s "8,24,48,64,64,64,64,48,40,32,24,16,8,8"
  run $s "16,48,48,48,48,48,48,48,40,32,24,16,8,8,8,8,8,8"
  run $s "32,96,96,96,64,48,32,24,16,8"
done
