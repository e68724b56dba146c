#!/bin/bash
# Build-time tunable sweep: tools/variant_sweep.sh build "NAME:-DFOO=1 -DBAR=2" ...   (here, needs nvcc)
#                           tools/variant_sweep.sh run                                  (on the GPU box)
# Variants are whole product libraries under gpurun_out/variants/ (scratch, not tracked); `run` swaps each in
# as csrc/liborbfe.so, runs the extractor parity tests and a short bench, and restores the original.
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
CSRC=$ROOT/orb-slam3-noted_b200/csrc
VDIR=$ROOT/tools/_variants
if [ "$1" = build ]; then
  shift; rm -rf $VDIR; mkdir -p $VDIR
  for spec in "$@"; do
    name=${spec%%:*}; flags=${spec#*:}
    ( /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared $flags \
        -o $VDIR/$name.so $CSRC/*.cu && echo "built $name ($flags)" ) &
  done
  wait
else
  cp $CSRC/liborbfe.so /tmp/liborbfe.orig.so
  for so in $VDIR/*.so; do
    cp $so $CSRC/liborbfe.so; touch $CSRC/liborbfe.so
    echo "== $(basename $so .so)"
    python -m pytest $ROOT/tests/test_extract_gpu.py -x -q 2>&1 | tail -1
    python $ROOT/bench.py --no-cpu --no-match --no-legs --steps 60 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['e2e']['value']), {k: round(v,3) for k,v in d['roofline']['stage_ms_per_step'].items()})"
  done
  cp /tmp/liborbfe.orig.so $CSRC/liborbfe.so
fi
