#!/bin/bash
# beam-culling statistics (diagnostic kernel build) of the whole C4 frame and of rank 0's 1/8 share
cd "$(dirname "$0")/.."
for cfg in "1 16" "8 16" "8 960"; do
  set -- $cfg
  echo "== world $1 strip $2"
  B200RT_DEBUG_CULL=1 B200RT_EMULATE_WORLD=$1 B200RT_STRIP=$2 python bench.py --no-cpu --no-dense --steps 1 --warmup 1 2>&1 >/dev/null | grep "cull\]" | head -8
done
