#!/bin/bash
cd "$(dirname "$0")/.."
run() { env "$@" python bench.py --no-cpu --no-dense --steps 10 --warmup 3 2>/dev/null | python -c "import json,sys; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('  ms', round(j['ms_per_step'],4), 'e2e', round(j['e2e']['ms_per_step'],4), {k: round(v,3) for k,v in j['roofline']['exact_culled']['kernel_ms'].items()})"; }
for w in 1; do
  echo "== world $w"
  for knobs in "X=1" "B200RT_CGRID_MAX_PATHS=1000000000 B200RT_CGRID_FROM=3" "B200RT_CGRID_MAX_PATHS=1000000000 B200RT_CGRID_FROM=2" "B200RT_CGRID_MAX_PATHS=1000000000 B200RT_CGRID_FROM=1"; do echo " $knobs"; run B200RT_EMULATE_WORLD=$w $knobs; done
done
