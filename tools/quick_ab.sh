#!/bin/bash
# Shortest possible A/B of variant libraries (tools/_variants/*.so): bench only - its own 4-frame byte comparison with the
# reference extractor is the parity check.  usage: tools/quick_ab.sh trig main
C=orb-slam3-noted_b200/csrc
for v in "$@"; do
  cp tools/_variants/$v.so $C/liborbfe.so
  timeout 40 python bench.py --no-cpu --no-match --no-legs --steps 30 2>gpurun_out/quick_$v.err | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$v', round(d['value']), round(d['e2e']['value']), {k: round(x,3) for k,x in d['roofline']['stage_ms_per_step'].items()}, d['config_detail'].get('parity_checked_frames') if 'config_detail' in d else d['config'].get('parity_checked_frames'))" || tail -3 gpurun_out/quick_$v.err
done
