#!/bin/bash
# tensor blur: parity, then ALU vs tensor timing, then (optional $1=ncu) one ncu capture of k_blur_tc
timeout 600 python -m pytest tests/test_blur_tensor_gpu.py -x -q 2>&1 | tail -3
for b in alu tensor; do ORBFE_BLUR=$b python bench.py --no-cpu --no-match --no-legs --steps 30 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().split(chr(10))[-1]); print('$b', round(d['value']), round(d['e2e']['value']), {k: round(v,3) for k,v in d['roofline']['stage_ms_per_step'].items()})"; done
if [ "$1" = ncu ]; then
  ORBFE_BLUR=tensor timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_blur_tc --launch-skip 3 --launch-count 1 -f -o gpurun_out/blur_tc python bench.py --steps 1 --warmup 3 --no-cpu --no-match --no-legs > gpurun_out/blur_tc_ncu.log 2>&1
  tail -2 gpurun_out/blur_tc_ncu.log
fi
