#!/bin/bash
# Blur-under-FAST sweep (environment knobs, no rebuild): ORBFE_BLUR_OVERLAP x ORBFE_BLUR_PERSIST (resident blur CTAs per SM).
#   gpurun -- 'bash tools/sweep_blur.sh > gpurun_out/blur_sweep.txt 2>&1'
for cfg in "0 0" "1 0" "1 1" "1 2" "1 3" "1 4" "1 6"; do
  set -- $cfg
  export ORBFE_BLUR_OVERLAP=$1 ORBFE_BLUR_PERSIST=$2
  echo "== overlap $1 resident CTAs per SM $2"
  python -m pytest tests/test_extract_gpu.py tests/test_pipeline_gpu.py -x -q 2>&1 | tail -1
  python bench.py --no-cpu --no-match --no-legs --steps 30 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().split('\n')[-1]); print(round(d['value']), round(d['e2e']['value']), d['ms_per_step'])"
done
