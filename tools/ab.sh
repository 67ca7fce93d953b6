#!/bin/bash
# usage (on the GPU box): tools/ab.sh <reps> -- A/B of build_ab/old.so vs build_ab/new.so on the same box,
# alternating, device-leg fps of bench.py
L=mlt_b200/lib/libmltcuda_rt.so
for i in $(seq ${1:-3}); do
  for v in old new; do
    cp build_ab/$v.so $L
    python bench.py --steps ${STEPS:-20} --warmup 3 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['value'], d['e2e']['value'])"
  done
done
cp build_ab/new.so $L
