#!/bin/bash
# usage (on the GPU box): tools/ab.sh <reps> <variant>...   -- same-box comparison of runtime builds
# build_ab/<variant>.so, alternating; prints bench.py's device-leg and e2e frames/s.  BENCH_ARGS adds flags.
L=mlt_b200/lib/libmltcuda_rt.so
cp $L build_ab/.saved.so
R=${1:-3}; shift
for i in $(seq $R); do
  for v in "$@"; do
    cp build_ab/$v.so $L
    python bench.py --steps ${STEPS:-20} --warmup 3 $BENCH_ARGS 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['value'], d['e2e']['value'])"
  done
done
cp build_ab/.saved.so $L
