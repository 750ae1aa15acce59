#!/bin/bash
# env-knob sweep on rank 0's 1/8 share of C4 (emulated) and on the whole frame
cd "$(dirname "$0")/.."
run() { env "$@" python bench.py --no-cpu --no-dense --steps 10 --warmup 3 2>/dev/null | python -c "import json,sys; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('  ms', round(j['ms_per_step'],4), {k: round(v,3) for k,v in j['roofline']['exact_culled']['kernel_ms'].items()})"; }
for w in 8 1; do
  echo "== world $w"
  for knobs in "X=1" "B200RT_NR0=2" "B200RT_NR=2" "B200RT_NR0=8" "B200RT_NO_CONCURRENCY=1" "B200RT_SORT_LAST=1" "B200RT_SS_GRID=16" "B200RT_SS_GRID=32" "B200RT_GRAPH_ALL=1"; do
    echo " $knobs"; run B200RT_EMULATE_WORLD=$w $knobs
  done
done
